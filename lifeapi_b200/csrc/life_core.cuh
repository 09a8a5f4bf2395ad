// life_core.cuh -- register-resident 64x64 torus tile and the B3/S23 generation step (sm_100a).
//
// Layout is the reference's (LifeAPI.hpp:39-40,131): a pattern is uint64_t state[64]; word index
// = column x, bit index = row y.  A pattern is owned by LPP consecutive lanes of one warp
// ("lanes per pattern"); each lane keeps W = 64/LPP consecutive words as (lo,hi) 32-bit register
// pairs.  Bit neighbours (y+-1) come from 64-bit rotates built from SHF funnel shifts; word
// neighbours (x+-1) are the adjacent registers, and at the lane boundary the 2-bit column counts
// are exchanged with __shfl_sync whose width=LPP argument gives the torus wrap.
//
// The rule is LifeState::StepAlt (LifeAPI.hpp:1218-1254) restated as 3-input LUTs:
//   per 32-cell half-word and generation: 2 SHF + 7.5 LOP3 (SURVEY.md 8d counts 2 + 10 for the literal form).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace lifeb200 {

constexpr int kN = 64;  // LifeAPI.hpp:12

template <int LUT>
__device__ __forceinline__ uint32_t lop3(uint32_t a, uint32_t b, uint32_t c) {
  uint32_t d;
  asm("lop3.b32 %0, %1, %2, %3, %4;" : "=r"(d) : "r"(a), "r"(b), "r"(c), "n"(LUT));
  return d;
}
// truth tables with a=0xF0, b=0xCC, c=0xAA
constexpr int kXor3 = 0x96;       // a^b^c
constexpr int kMaj = 0xE8;        // (a&b)|(c&(a^b))
constexpr int kXorAnd = 0x78;     // a^(b&c)
constexpr int kXorThenAnd = 0x28; // (a^b)&c
constexpr int kAndOr = 0xE0;      // a&(b|c)
constexpr int kOr3 = 0xFE;        // a|b|c
constexpr int kAndXor = 0x6A;     // (a&b)^c
constexpr int kOrAnd = 0xEA;      // (a&b)|c

// 2-bit count of the (y-1, y, y+1) triple inside one column word: LifeAPI.hpp:897-907 CountRows.
struct Cnt {
  uint32_t b0lo, b0hi, b1lo, b1hi;
};

__device__ __forceinline__ Cnt count_rows(uint32_t lo, uint32_t hi) {
  // rotl64(a,1) / rotr64(a,1) on the (lo,hi) pair: 4 SHF
  uint32_t l_lo = __funnelshift_l(hi, lo, 1), l_hi = __funnelshift_l(lo, hi, 1);
  uint32_t r_lo = __funnelshift_r(lo, hi, 1), r_hi = __funnelshift_r(hi, lo, 1);
  Cnt c;
  c.b0lo = lop3<kXor3>(l_lo, r_lo, lo);
  c.b0hi = lop3<kXor3>(l_hi, r_hi, hi);
  c.b1lo = lop3<kMaj>(l_lo, r_lo, lo);
  c.b1hi = lop3<kMaj>(l_hi, r_hi, hi);
  return c;
}

// ---- row-interleaved register layout ("IL") -----------------------------------------------------
// In the reference layout the low register of a column holds rows 0..31 and the high one rows
// 32..63, so both row neighbours of a cell sit in the SAME register one bit away: two funnel
// shifts per register and generation.  Interleaved, the low register holds the EVEN rows (bit k =
// row 2k) and the high one the ODD rows (bit k = row 2k+1).  Then the row below an even row is the
// same bit of the odd register and the row above it the odd register rotated by one; for an odd
// row it is the even register as is, and the even register rotated the other way.  One 32-bit
// rotate per register and generation instead of two; the torus wrap 63 -> 0 is the rotate's own
// wrap.  Everything after the column count is bit-position-wise and does not change.  Kernels that
// run many generations convert once after the load and once before the store (about 35 integer
// instructions per word each way, against 17 per word and generation).
__device__ __forceinline__ Cnt count_rows_il(uint32_t ev, uint32_t od) {
  const uint32_t up = __funnelshift_l(od, od, 1);  // row 2k-1 at bit k
  const uint32_t dn = __funnelshift_r(ev, ev, 1);  // row 2k+2 at bit k
  Cnt c;
  c.b0lo = lop3<kXor3>(up, ev, od);
  c.b1lo = lop3<kMaj>(up, ev, od);
  c.b0hi = lop3<kXor3>(ev, od, dn);
  c.b1hi = lop3<kMaj>(ev, od, dn);
  return c;
}
// even bits of x to the low half, odd bits to the high half (outer perfect unshuffle) and back
__device__ __forceinline__ uint32_t unshuffle32(uint32_t x) {
  uint32_t t;
  t = (x ^ (x >> 1)) & 0x22222222u; x ^= t ^ (t << 1);
  t = (x ^ (x >> 2)) & 0x0C0C0C0Cu; x ^= t ^ (t << 2);
  t = (x ^ (x >> 4)) & 0x00F000F0u; x ^= t ^ (t << 4);
  t = (x ^ (x >> 8)) & 0x0000FF00u; x ^= t ^ (t << 8);
  return x;
}
__device__ __forceinline__ uint32_t shuffle32(uint32_t x) {
  uint32_t t;
  t = (x ^ (x >> 8)) & 0x0000FF00u; x ^= t ^ (t << 8);
  t = (x ^ (x >> 4)) & 0x00F000F0u; x ^= t ^ (t << 4);
  t = (x ^ (x >> 2)) & 0x0C0C0C0Cu; x ^= t ^ (t << 2);
  t = (x ^ (x >> 1)) & 0x22222222u; x ^= t ^ (t << 1);
  return x;
}
// (rows 0..31, rows 32..63) -> (even rows, odd rows) of one column word, and back
__device__ __forceinline__ void to_interleaved(uint32_t& lo, uint32_t& hi) {
  const uint32_t a = unshuffle32(lo), b = unshuffle32(hi);
  lo = __byte_perm(a, b, 0x5410);  // even rows of 0..31 below even rows of 32..63
  hi = __byte_perm(a, b, 0x7632);  // odd rows likewise
}
__device__ __forceinline__ void from_interleaved(uint32_t& lo, uint32_t& hi) {
  const uint32_t a = __byte_perm(lo, hi, 0x5410), b = __byte_perm(lo, hi, 0x7632);
  lo = shuffle32(a);
  hi = shuffle32(b);
}

// B3/S23 select over the three column counts (the job of the StepAlt tail, LifeAPI.hpp:1236-1252).
// X = u0+c0+d0 and Y = u1+c1+d1 are each encoded in two LUTs -- X as (not-all-equal, majority),
// a Gray code of 0..3, Y as (parity, majority) -- and next = [X+2Y == 3] | a & [X+2Y == 4] is then
// three more LUTs found by exhaustive search (tools/lut_search.c) and checked on all 2^7 inputs
// (tests/test_oracle.py::test_rule_luts_exhaustive).  7 LOP3 per half-word instead of the 8 of the
// literal two-full-adder form.
constexpr int kNotAllEq = 0x7E;  // !(a==b==c)
__device__ __forceinline__ uint32_t rule32(uint32_t a, uint32_t u0, uint32_t u1, uint32_t c0, uint32_t c1,
                                           uint32_t d0, uint32_t d1) {
  uint32_t xa = lop3<kNotAllEq>(u0, c0, d0);
  uint32_t xb = lop3<kMaj>(u0, c0, d0);
  uint32_t ya = lop3<kXor3>(u1, c1, d1);
  uint32_t yb = lop3<kMaj>(u1, c1, d1);
  uint32_t g1 = lop3<0x16>(xb, xa, a);
  uint32_t g2 = lop3<0x6D>(g1, yb, a);
  return lop3<0x09>(g2, ya, xa);
}

// Pair-shared form of the same select, found by SAT-based exact synthesis (tools/lut_synth_pair_cnf.py,
// `0 3 4 --dc --raw`; every smaller split of the same family is UNSAT) and checked on all
// inputs in tests/test_oracle.py: three LUTs turn the 2-bit counts (a, b) of two adjacent columns into
// a 3-signal code of a+b that is shared by BOTH columns of the pair; four more per column give
// next = [a+b+c == 3] | cell & [a+b+c == 4] from that code, the count c of the column on the far side,
// the low count bits of the pair and the centre cell.  3 + 2 x 4 = 11 LOP3 per two half-words.
// The select uses the don't-cares of the count: the cell is the middle bit of its own column's
// triple, so cell = 1 implies count >= 1 and count = 3 implies cell = 1; it is exact on every
// (a, b, c, cell) a real column can produce, for the left and for the right column of the pair alike.
struct PairCode {
  uint32_t p0, p1, p2;
};
__device__ __forceinline__ PairCode pair_code(uint32_t a0, uint32_t a1, uint32_t b0, uint32_t b1) {
  PairCode p;
  p.p0 = lop3<0xB2>(a1, b0, b1);
  p.p1 = lop3<0xD4>(a1, b1, p.p0);
  p.p2 = lop3<kMaj>(a0, b0, p.p0);
  return p;
}
__device__ __forceinline__ uint32_t pair_select(const PairCode& p, uint32_t a0, uint32_t b0, uint32_t c0, uint32_t c1,
                                                uint32_t cell) {
  const uint32_t t0 = lop3<kMaj>(c0, p.p0, p.p2);
  const uint32_t t1 = lop3<0x16>(c1, p.p1, t0);
  const uint32_t t2 = lop3<kXor3>(c0, a0, b0);  // bit 0 of the 3x3 sum
  return lop3<0xC8>(cell, t1, t2);
}

template <int LPP>
__device__ __forceinline__ Cnt halo(const Cnt& c, int src) {
  if constexpr (LPP == 1) {
    return c;
  } else {
    Cnt r;
    r.b0lo = __shfl_sync(0xffffffffu, c.b0lo, src, LPP);
    r.b0hi = __shfl_sync(0xffffffffu, c.b0hi, src, LPP);
    r.b1lo = __shfl_sync(0xffffffffu, c.b1lo, src, LPP);
    r.b1hi = __shfl_sync(0xffffffffu, c.b1hi, src, LPP);
    return r;
  }
}

// One pattern slice held by one lane.
template <int LPP>
struct Tile {
  static constexpr int W = kN / LPP;
  static_assert(LPP == 1 || LPP == 2 || LPP == 4 || LPP == 8 || LPP == 16 || LPP == 32, "LPP");
  uint32_t lo[W], hi[W];

  // sub = lane index inside the pattern's lane group (0..LPP-1)
  // plain (coherent) vector loads, no __restrict__: several entry points step IN PLACE (in == out, records
  // updated where they lie), and a non-coherent ld.global.nc of memory the same kernel later writes is undefined
  __device__ __forceinline__ void load(const uint64_t* pat, int sub) {
    const uint4* p = reinterpret_cast<const uint4*>(pat + sub * W);
#pragma unroll
    for (int k = 0; k < W / 2; k++) {
      uint4 v = p[k];
      lo[2 * k] = v.x; hi[2 * k] = v.y; lo[2 * k + 1] = v.z; hi[2 * k + 1] = v.w;
    }
  }
  __device__ __forceinline__ void store(uint64_t* pat, int sub) const {
    uint4* p = reinterpret_cast<uint4*>(pat + sub * W);
#pragma unroll
    for (int k = 0; k < W / 2; k++) p[k] = make_uint4(lo[2 * k], hi[2 * k], lo[2 * k + 1], hi[2 * k + 1]);
  }

  // One generation in place == LifeState::Step() (LifeAPI.hpp:1196-1216).  Words are taken two at a
  // time: the pair code of (cnt[w], cnt[w+1]) is shared by both words, so a generation costs, per
  // 32-cell register, 2 LOP3 (column count) + 1.5 (half a pair code) + 4 (select) = 7.5 LOP3 plus the
  // row shifts: 2 SHF in the reference layout (9.5 integer-pipe instructions), 1 SHF when the tile is
  // row-interleaved (IL = true: 8.5).  Updating in place is safe because every neighbour enters
  // through column counts taken before the word is overwritten.
  template <bool IL = false>
  __device__ __forceinline__ void step(int sub) {
    static_assert(W % 2 == 0, "words are processed in pairs");
    const int up_src = sub + LPP - 1, dn_src = sub + 1;  // taken modulo LPP by shfl width
    auto rows = [](uint32_t l, uint32_t h) { return IL ? count_rows_il(l, h) : count_rows(l, h); };
    Cnt first = rows(lo[0], hi[0]);
    Cnt last = rows(lo[W - 1], hi[W - 1]);
    Cnt prev = halo<LPP>(last, up_src);
    Cnt after = halo<LPP>(first, dn_src);
    Cnt c0 = first;
#pragma unroll
    for (int w = 0; w < W; w += 2) {
      const Cnt c1 = (w + 1 == W - 1) ? last : rows(lo[w + 1], hi[w + 1]);
      const Cnt c2 = (w + 2 == W) ? after : (w + 2 == W - 1) ? last : rows(lo[w + 2], hi[w + 2]);
      const PairCode plo = pair_code(c0.b0lo, c0.b1lo, c1.b0lo, c1.b1lo);
      const PairCode phi = pair_code(c0.b0hi, c0.b1hi, c1.b0hi, c1.b1hi);
      lo[w] = pair_select(plo, c0.b0lo, c1.b0lo, prev.b0lo, prev.b1lo, lo[w]);
      hi[w] = pair_select(phi, c0.b0hi, c1.b0hi, prev.b0hi, prev.b1hi, hi[w]);
      lo[w + 1] = pair_select(plo, c0.b0lo, c1.b0lo, c2.b0lo, c2.b1lo, lo[w + 1]);
      hi[w + 1] = pair_select(phi, c0.b0hi, c1.b0hi, c2.b0hi, c2.b1hi, hi[w + 1]);
      prev = c1;
      c0 = c2;
    }
  }
  // reference layout <-> row-interleaved layout of every word of the tile (see count_rows_il)
  __device__ __forceinline__ void interleave() {
#pragma unroll
    for (int w = 0; w < W; w++) to_interleaved(lo[w], hi[w]);
  }
  __device__ __forceinline__ void deinterleave() {
#pragma unroll
    for (int w = 0; w < W; w++) from_interleaved(lo[w], hi[w]);
  }

  // NeighbourCount(const LifeState&) (NeighbourCount.hpp:40-70 == LifeAPI.hpp:909-952): the four
  // bit planes of the 3x3 sum INCLUDING the centre (0..9).  The planes are the binary digits of a
  // sum, so any exact adder gives the reference's bits: two 3:2 compressors over the column
  // counts, then total = s0 + 2(k0 + s1) + 4 k1.
  __device__ __forceinline__ void neighbour_count(int sub, Tile& b3, Tile& b2, Tile& b1, Tile& b0) const {
    const int up_src = sub + LPP - 1, dn_src = sub + 1;
    Cnt first = count_rows(lo[0], hi[0]);
    Cnt last = (W == 1) ? first : count_rows(lo[W - 1], hi[W - 1]);
    Cnt prev = halo<LPP>(last, up_src);
    Cnt after = halo<LPP>(first, dn_src);
    Cnt cur = first;
#pragma unroll
    for (int w = 0; w < W; w++) {
      Cnt next;
      if (w == W - 1) next = after;
      else if (w + 1 == W - 1) next = last;
      else next = count_rows(lo[w + 1], hi[w + 1]);
      {
        uint32_t s0 = lop3<kXor3>(prev.b0lo, cur.b0lo, next.b0lo), k0 = lop3<kMaj>(prev.b0lo, cur.b0lo, next.b0lo);
        uint32_t s1 = lop3<kXor3>(prev.b1lo, cur.b1lo, next.b1lo), k1 = lop3<kMaj>(prev.b1lo, cur.b1lo, next.b1lo);
        uint32_t cy = k0 & s1;
        b0.lo[w] = s0; b1.lo[w] = k0 ^ s1; b2.lo[w] = k1 ^ cy; b3.lo[w] = k1 & cy;
      }
      {
        uint32_t s0 = lop3<kXor3>(prev.b0hi, cur.b0hi, next.b0hi), k0 = lop3<kMaj>(prev.b0hi, cur.b0hi, next.b0hi);
        uint32_t s1 = lop3<kXor3>(prev.b1hi, cur.b1hi, next.b1hi), k1 = lop3<kMaj>(prev.b1hi, cur.b1hi, next.b1hi);
        uint32_t cy = k0 & s1;
        b0.hi[w] = s0; b1.hi[w] = k0 ^ s1; b2.hi[w] = k1 ^ cy; b3.hi[w] = k1 & cy;
      }
      prev = cur; cur = next;
    }
  }

  // OR over this lane's words of (state ^ wanted) & (wanted | unwanted): LifeTarget.hpp:44-51.
  // wm = wanted & mask, m = wanted | unwanted (precomputed per target).
  __device__ __forceinline__ uint32_t target_diff(const Tile& wm, const Tile& m) const {
    uint32_t d = 0;
#pragma unroll
    for (int w = 0; w < W; w++) {
      uint32_t t0 = lop3<kAndXor>(lo[w], m.lo[w], wm.lo[w]);
      uint32_t t1 = lop3<kAndXor>(hi[w], m.hi[w], wm.hi[w]);
      d = lop3<kOr3>(d, t0, t1);
    }
    return d;
  }
  __device__ __forceinline__ uint32_t xor_diff(const Tile& o) const {  // operator== LifeAPI.hpp:213-220
    uint32_t d = 0;
#pragma unroll
    for (int w = 0; w < W; w++) d |= (lo[w] ^ o.lo[w]) | (hi[w] ^ o.hi[w]);
    return d;
  }
  __device__ __forceinline__ uint32_t any_bits() const {  // IsEmpty LifeAPI.hpp:281-288
    uint32_t d = 0;
#pragma unroll
    for (int w = 0; w < W; w++) d = lop3<kOr3>(d, lo[w], hi[w]);
    return d;
  }
  __device__ __forceinline__ uint32_t pop() const {  // GetPop LifeAPI.hpp:290-298 (this lane's share)
    uint32_t p = 0;
#pragma unroll
    for (int w = 0; w < W; w++) p += __popc(lo[w]) + __popc(hi[w]);
    return p;
  }
};

// group-wide reductions over the LPP lanes owning one pattern
template <int LPP>
__device__ __forceinline__ bool group_any(bool pred, int lane) {
  if constexpr (LPP == 1) return pred;
  unsigned b = __ballot_sync(0xffffffffu, pred);
  if constexpr (LPP == 32) return b != 0;
  unsigned gmask = ((1u << (LPP & 31)) - 1u) << (lane & ~(LPP - 1));
  return (b & gmask) != 0;
}
template <int LPP>
__device__ __forceinline__ uint32_t group_sum(uint32_t v) {
#pragma unroll
  for (int o = LPP / 2; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o, 32);
  return v;
}

// splitmix64 soup source (build-defined; replaces LifeAPI.hpp:63-69).
__device__ __forceinline__ uint64_t splitmix64(uint64_t z) {
  z += 0x9E3779B97F4A7C15ULL;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
  return z ^ (z >> 31);
}
__device__ __forceinline__ uint64_t soup_word(uint64_t seed, uint64_t id, int i, int dist) {
  uint64_t r = splitmix64(seed + id * 64 + (uint64_t)i);
  if (dist == 1) r = (1ULL << 61) + (r % ((1ULL << 61) + 1));  // shape of LifeAPI.hpp:21-22
  return r;
}

}  // namespace lifeb200

namespace lifeb200 {

// flat view of a tile's 2*W registers (k < W: low halves, else high halves); indices are
// compile-time after unrolling, so everything stays in registers
template <int LPP>
__device__ __forceinline__ uint32_t& reg(Tile<LPP>& t, int k) {
  return k < Tile<LPP>::W ? t.lo[k] : t.hi[k - Tile<LPP>::W];
}
template <int LPP>
__device__ __forceinline__ uint32_t reg(const Tile<LPP>& t, int k) {
  return k < Tile<LPP>::W ? t.lo[k] : t.hi[k - Tile<LPP>::W];
}

// 3x3 dilation of a plane on the torus: ZOI (LifeAPI.hpp:521-536) or, with hollow = true,
// ZOIHollow (LifeAPI.hpp:541-562, the centre cell itself is left out).
template <int LPP>
__device__ __forceinline__ Tile<LPP> dilate(const Tile<LPP>& t, int sub, bool hollow) {
  constexpr int W = Tile<LPP>::W;
  Tile<LPP> mid, full, out;
#pragma unroll
  for (int w = 0; w < W; w++) {
    uint32_t lo = t.lo[w], hi = t.hi[w];
    uint32_t l_lo = __funnelshift_l(hi, lo, 1), l_hi = __funnelshift_l(lo, hi, 1);
    uint32_t r_lo = __funnelshift_r(lo, hi, 1), r_hi = __funnelshift_r(hi, lo, 1);
    mid.lo[w] = l_lo | r_lo; mid.hi[w] = l_hi | r_hi;
    full.lo[w] = mid.lo[w] | lo; full.hi[w] = mid.hi[w] | hi;
  }
  uint32_t up_lo = full.lo[W - 1], up_hi = full.hi[W - 1], dn_lo = full.lo[0], dn_hi = full.hi[0];
  if constexpr (LPP > 1) {
    up_lo = __shfl_sync(0xffffffffu, up_lo, sub + LPP - 1, LPP);
    up_hi = __shfl_sync(0xffffffffu, up_hi, sub + LPP - 1, LPP);
    dn_lo = __shfl_sync(0xffffffffu, dn_lo, sub + 1, LPP);
    dn_hi = __shfl_sync(0xffffffffu, dn_hi, sub + 1, LPP);
  }
#pragma unroll
  for (int w = 0; w < W; w++) {
    uint32_t a_lo = w == 0 ? up_lo : full.lo[w - 1], a_hi = w == 0 ? up_hi : full.hi[w - 1];
    uint32_t b_lo = w == W - 1 ? dn_lo : full.lo[w + 1], b_hi = w == W - 1 ? dn_hi : full.hi[w + 1];
    out.lo[w] = lop3<kOr3>(a_lo, hollow ? mid.lo[w] : full.lo[w], b_lo);
    out.hi[w] = lop3<kOr3>(a_hi, hollow ? mid.hi[w] : full.hi[w], b_hi);
  }
  return out;
}

}  // namespace lifeb200
