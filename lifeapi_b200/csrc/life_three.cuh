// life_three.cuh -- the three-valued step (bitslicing/unknown_step.hpp, generator unknown_step.py:3-38) as
// two B3/S23 selects instead of two 4-bit NeighbourCounts and a 9-term cover.
//
// On planes that keep the LifeStable invariant state & unknown == 0 (LifeStable.hpp:220-221,320-335) the
// generator's truth function depends only on the centre and on two 3x3 sums (centre included):
//     lo = count(state)            -- every unknown neighbour OFF
//     hi = count(state | unknown)  -- every unknown neighbour ON
// B3/S23 splits a sum N into three classes by the (known) centre cell:  L: N <= 2,  M: alive,  H: too many,
//     alive(N) = [N == 3] | cell & [N == 4],     A(N) = [N >= 3]   (A & ~alive is class H)
// and with lo <= hi the classes are ordered too, so
//     next_on      = ~u & alive(lo) & alive(hi)                               (both ends of the range alive)
//     next_unknown =  u | A(hi) & ~(A(lo) & ~alive(lo)) & ~(alive(lo) & alive(hi))   (the range straddles a class edge)
// which is the cover's output on every reachable input (tests/test_oracle.py::test_three_valued_fast_path_exhaustive
// checks all of them against the fragment text).  Planes that overlap are outside that domain: a kernel
// must send those through the literal cover (unknown.cu does, per warp).
//
// Cost per 32-cell register and generation: 1 (state | unknown) + 2 x [2 column count + 2 (half a 4-gate
// pair code) + 5 (tail: alive and A)] + 3 (combine) = 22 LOP3 + the row shifts (4 SHF reference layout,
// 2 SHF row-interleaved), against 36 LOP3 + 4 SHF for counts + cover.  The pair code and the tails are
// from SAT exact synthesis (tools/lut_synth_pair3v_cnf.py `0 4 5 --dc --raw --outs alive,A`; (code, tail)
// = (3,5), (5,4), (6,4), (4,5,4), (4,4,5) are UNSAT, tools/synth_logs/pair3v_cnf_results.txt).
#pragma once
#include "life_core.cuh"

namespace lifeb200 {

struct PairCode4 {
  uint32_t p0, p1, p2, p3;
};
// (a0, a1), (b0, b1): binary column counts of the left and the right column of the pair
__device__ __forceinline__ PairCode4 pair_code4(uint32_t a0, uint32_t a1, uint32_t b0, uint32_t b1) {
  PairCode4 p;
  p.p0 = lop3<0x56>(a0, b0, b1);
  p.p1 = lop3<0xBE>(a1, b1, p.p0);
  p.p2 = lop3<0xE4>(a1, b1, p.p0);
  p.p3 = lop3<0x92>(a0, b0, p.p1);
  return p;
}
struct AliveA {
  uint32_t alive, a;
};
// tail of the pair's LEFT column: (c0, c1) = count of the column to its left, cell = its centre
__device__ __forceinline__ AliveA pair_tail_left(const PairCode4& p, uint32_t c0, uint32_t c1, uint32_t cell) {
  const uint32_t l0 = lop3<0xB2>(c1, cell, p.p2);
  const uint32_t l1 = lop3<0x48>(c1, p.p1, p.p2);
  const uint32_t l2 = lop3<kMaj>(c0, p.p3, l0);
  AliveA o;
  o.a = lop3<kOr3>(p.p2, l1, l2);
  o.alive = lop3<0x46>(p.p2, l1, l2);
  return o;
}
// tail of the pair's RIGHT column: (c0, c1) = count of the column to its right
__device__ __forceinline__ AliveA pair_tail_right(const PairCode4& p, uint32_t c0, uint32_t c1, uint32_t cell) {
  const uint32_t r0 = lop3<0x92>(c1, cell, p.p2);
  const uint32_t r1 = lop3<kMaj>(c0, p.p3, r0);
  const uint32_t r2 = lop3<0xE2>(c1, p.p1, p.p2);
  AliveA o;
  o.a = lop3<kOr3>(p.p2, r1, r2);
  o.alive = lop3<0x16>(p.p2, r1, r2);
  return o;
}
// (u, lo signals, hi signals) -> next_on, next_unknown
__device__ __forceinline__ void three_valued_combine(uint32_t u, const AliveA& lo, const AliveA& hi, uint32_t& next_on,
                                                     uint32_t& next_unknown) {
  next_on = lop3<0x08>(u, lo.alive, hi.alive);               // ~u & alive_lo & alive_hi
  const uint32_t x = lop3<0x47>(lo.a, lo.alive, hi.alive);   // ~(A_lo & ~alive_lo) & ~(alive_lo & alive_hi)
  next_unknown = lop3<0xF8>(u, hi.a, x);                     // u | A_hi & x
}

// One three-valued generation of a tile pair in place; requires state & unknown == 0.  IL: both tiles are
// row-interleaved (count_rows_il).  Same walk as Tile::step: words two at a time, every neighbour enters
// through column counts taken before the word is overwritten.
template <int LPP, bool IL>
__device__ __forceinline__ void three_valued_step(Tile<LPP>& s, Tile<LPP>& u, int sub) {
  constexpr int W = Tile<LPP>::W;
  static_assert(W % 2 == 0, "words are processed in pairs");
  const int up_src = sub + LPP - 1, dn_src = sub + 1;
  auto rows = [](uint32_t l, uint32_t h) { return IL ? count_rows_il(l, h) : count_rows(l, h); };
  uint32_t tlo[W], thi[W];
#pragma unroll
  for (int w = 0; w < W; w++) { tlo[w] = s.lo[w] | u.lo[w]; thi[w] = s.hi[w] | u.hi[w]; }
  const Cnt sfirst = rows(s.lo[0], s.hi[0]), slast = rows(s.lo[W - 1], s.hi[W - 1]);
  const Cnt tfirst = rows(tlo[0], thi[0]), tlast = rows(tlo[W - 1], thi[W - 1]);
  Cnt sprev = halo<LPP>(slast, up_src), tprev = halo<LPP>(tlast, up_src);
  const Cnt safter = halo<LPP>(sfirst, dn_src), tafter = halo<LPP>(tfirst, dn_src);
  Cnt s0 = sfirst, t0 = tfirst;
#pragma unroll
  for (int w = 0; w < W; w += 2) {
    const Cnt s1 = (w + 1 == W - 1) ? slast : rows(s.lo[w + 1], s.hi[w + 1]);
    const Cnt t1 = (w + 1 == W - 1) ? tlast : rows(tlo[w + 1], thi[w + 1]);
    const Cnt s2 = (w + 2 == W) ? safter : (w + 2 == W - 1) ? slast : rows(s.lo[w + 2], s.hi[w + 2]);
    const Cnt t2 = (w + 2 == W) ? tafter : (w + 2 == W - 1) ? tlast : rows(tlo[w + 2], thi[w + 2]);
    const PairCode4 pslo = pair_code4(s0.b0lo, s0.b1lo, s1.b0lo, s1.b1lo), pshi = pair_code4(s0.b0hi, s0.b1hi, s1.b0hi, s1.b1hi);
    const PairCode4 ptlo = pair_code4(t0.b0lo, t0.b1lo, t1.b0lo, t1.b1lo), pthi = pair_code4(t0.b0hi, t0.b1hi, t1.b0hi, t1.b1hi);
    uint32_t on, unk;
    three_valued_combine(u.lo[w], pair_tail_left(pslo, sprev.b0lo, sprev.b1lo, s.lo[w]),
                         pair_tail_left(ptlo, tprev.b0lo, tprev.b1lo, tlo[w]), on, unk);
    s.lo[w] = on; u.lo[w] = unk;
    three_valued_combine(u.hi[w], pair_tail_left(pshi, sprev.b0hi, sprev.b1hi, s.hi[w]),
                         pair_tail_left(pthi, tprev.b0hi, tprev.b1hi, thi[w]), on, unk);
    s.hi[w] = on; u.hi[w] = unk;
    three_valued_combine(u.lo[w + 1], pair_tail_right(pslo, s2.b0lo, s2.b1lo, s.lo[w + 1]),
                         pair_tail_right(ptlo, t2.b0lo, t2.b1lo, tlo[w + 1]), on, unk);
    s.lo[w + 1] = on; u.lo[w + 1] = unk;
    three_valued_combine(u.hi[w + 1], pair_tail_right(pshi, s2.b0hi, s2.b1hi, s.hi[w + 1]),
                         pair_tail_right(pthi, t2.b0hi, t2.b1hi, thi[w + 1]), on, unk);
    s.hi[w + 1] = on; u.hi[w + 1] = unk;
    sprev = s1; tprev = t1;
    s0 = s2; t0 = t2;
  }
}

}  // namespace lifeb200
