// next.cu -- the callers and data formats either side of the stepping path:
//   LifeWeld::Step (LifeWeld.hpp:169-186), InteractionCounts / InteractionCountsAndNext
//   (LifeAPI.hpp:956-1040) and LifeState::GetHash = XXH3_64bits of the 512-byte state
//   (LifeAPI.hpp:373), each batched over whole patterns held in registers.
#include <algorithm>

#include "ctx.h"
#include "host_pipeline.cuh"
#include "life_core.cuh"
#include "life_multi.cuh"
#include "xxh3_keys.h"

namespace lifeb200 {

constexpr int kBlockN = 256;
static inline unsigned grid_n(size_t n, int lpp) { return (unsigned)((n * (size_t)lpp + kBlockN - 1) / kBlockN); }
// The tail after the two full adders is a 6-gate network from exact synthesis (tools/lut_synth_generic_cnf.py weld_tail 6;
// the ripple form above is 9 gates; z3 found 8 gates in 76 minutes, 7 in 82, 6 in 27, tools/synth_logs/weld_tail_*.txt).
// g0, g1 are the full adder of the weight-2 bits (k0, s1, f1), g4 the weight-4 sum bit; g2 and g3 fold the weight-1
// bits, the cell and the rule together, using the count's don't-cares -- the cell is the middle bit of its own column's
// triple, so cell = 1 implies a count >= 1 there and a count of 3 implies cell = 1.  Checked on every reachable
// (left, mid, right, cell, frozen) in tests/test_oracle.py::test_weld_rule_luts_exhaustive.
__device__ __forceinline__ uint32_t weld_rule(uint32_t l0, uint32_t l1, uint32_t m0, uint32_t m1, uint32_t r0, uint32_t r1,
                                              uint32_t cell, uint32_t f2, uint32_t f1, uint32_t f0) {
  const uint32_t s0 = lop3<kXor3>(l0, m0, r0), k0 = lop3<kMaj>(l0, m0, r0);
  const uint32_t s1 = lop3<kXor3>(l1, m1, r1), k1 = lop3<kMaj>(l1, m1, r1);
  const uint32_t g0 = lop3<kXor3>(k0, s1, f1);
  const uint32_t g1 = lop3<kMaj>(k0, s1, f1);
  const uint32_t g2 = lop3<0xDA>(s0, cell, f0);
  const uint32_t g3 = lop3<0x98>(cell, g0, g2);
  const uint32_t g4 = lop3<kXor3>(k1, f2, g1);
  return lop3<0x48>(g2, g3, g4);
}

// ---- LifeWeld::Step, lane groups in the reference layout (launches of fewer than 12 generations): the column counts
// of the walk below feed weld_rule directly -- 2 SHF + 2 + 10 LOP3 = 14 per 32 cells (18 with the four count planes and
// the ripple adder of round 1).
template <int LPP>
__global__ void __launch_bounds__(kBlockN, 1) weld_step_kernel(uint64_t* __restrict__ records, size_t n, unsigned gens) {
  constexpr int W = Tile<LPP>::W;
  const size_t p = ((size_t)blockIdx.x * kBlockN + threadIdx.x) / LPP;
  const int sub = threadIdx.x & (LPP - 1);
  const bool valid = p < n;
  Tile<LPP> s, f2, f1, f0;
  if (valid) {
    uint64_t* r = records + p * 4 * kN;  // state, frozen2, frozen1, frozen0 (LifeWeld.hpp:18-22)
    s.load(r, sub); f2.load(r + kN, sub); f1.load(r + 2 * kN, sub); f0.load(r + 3 * kN, sub);
  } else {
#pragma unroll
    for (int w = 0; w < W; w++) s.lo[w] = s.hi[w] = f2.lo[w] = f2.hi[w] = f1.lo[w] = f1.hi[w] = f0.lo[w] = f0.hi[w] = 0;
  }
  const int up_src = sub + LPP - 1, dn_src = sub + 1;
#pragma unroll 1
  for (unsigned g = 0; g < gens; g++) {
    const Cnt first = count_rows(s.lo[0], s.hi[0]);
    const Cnt last = (W == 1) ? first : count_rows(s.lo[W - 1], s.hi[W - 1]);
    Cnt prev = halo<LPP>(last, up_src);
    const Cnt after = halo<LPP>(first, dn_src);
    Cnt cur = first;
#pragma unroll
    for (int w = 0; w < W; w++) {
      const Cnt next = (w == W - 1) ? after : (w + 1 == W - 1) ? last : count_rows(s.lo[w + 1], s.hi[w + 1]);
      s.lo[w] = weld_rule(prev.b0lo, prev.b1lo, cur.b0lo, cur.b1lo, next.b0lo, next.b1lo, s.lo[w], f2.lo[w], f1.lo[w], f0.lo[w]);
      s.hi[w] = weld_rule(prev.b0hi, prev.b1hi, cur.b0hi, cur.b1hi, next.b0hi, next.b1hi, s.hi[w], f2.hi[w], f1.hi[w], f0.hi[w]);
      prev = cur;
      cur = next;
    }
  }
  if (valid) s.store(records + p * 4 * kN, sub);
}

// ---- LifeWeld::Step in the row-class layout (life_multi.cuh): 4 records per group of 8 lanes ----------------
// The state lives in registers exactly as in step_kernel_x<8> (64 registers per lane, 1/4 shift per register and
// generation).  The three frozen planes are constants of the launch and do not fit beside it (64 + 192), so every
// lane packs them once and parks them in shared memory (768 B per lane, [48][128] uint4 per CTA: a warp's LDS.128
// reads 32 consecutive uint4, conflict-free) and reads them back every generation: 48 LDS.128 against ~790 integer
// instructions.  Per 32 cells and generation: 2 LOP3 column count + 1/4 SHF + 10 LOP3 (12.25).  What is computed:
//   s0, k0 = full adder of the three low count bits; s1, k1 = of the three high ones        (sum = s0 + 2 (k0 + s1) + 4 k1)
//   T0 = s0 ^ f0, c0 = s0 & f0; x = k0 ^ s1; T1 = x ^ f1 ^ c0, c1 = maj(x, f1, c0); S2 = k1 ^ (k0 & s1); T2 = S2 ^ f2 ^ c1
//   next = (T0 ^ T2) & (cell | T0) & (T1 ^ T2)          -- bit3 of the count and the last carry dropped, as the reference does
// (9 gates as written; weld_rule below evaluates the same function in 6)
// A warp whose 16 records have no frozen cell at all (LifeWeld(state), LifeWeld.hpp:24-25) runs the plain step.
constexpr int kWeldXSmem = 128 * 48 * 16;

__global__ void __launch_bounds__(128, 2) weld_step_kernel_x(uint64_t* records, size_t n, unsigned gens) {
  constexpr int M = 8, P = 4;
  using TX = TileX<M>;
  const size_t gid = (size_t)blockIdx.x * 128 + threadIdx.x;
  const size_t p0 = (gid >> 3) * P;
  const int sub = threadIdx.x & 7;
  extern __shared__ uint4 frozen[];  // [48][128]: plane pl, column c, half h at row pl * 16 + c * 2 + h
  TX t;
  uint32_t any = 0;
#pragma unroll
  for (int pl = 0; pl < 4; pl++) {
#pragma unroll
    for (int c = 0; c < 8; c += 2) {
      uint64_t wa[P], wb[P];
#pragma unroll
      for (int q = 0; q < P; q++) {
        uint4 v = make_uint4(0, 0, 0, 0);
        if (p0 + q < n) v = *reinterpret_cast<const uint4*>(records + (p0 + q) * 4 * kN + pl * kN + sub * 8 + c);
        wa[q] = ((uint64_t)v.y << 32) | v.x;
        wb[q] = ((uint64_t)v.w << 32) | v.z;
      }
      if (pl == 0) {
        t.pack(c, wa);
        t.pack(c + 1, wb);
      } else {
        TX tmp;  // only columns c, c + 1 are touched: two columns' worth of registers
        tmp.pack(c, wa);
        tmp.pack(c + 1, wb);
#pragma unroll
        for (int cc = c; cc < c + 2; cc++) {
          frozen[((pl - 1) * 16 + cc * 2) * 128 + threadIdx.x] = make_uint4(tmp.r[cc][0], tmp.r[cc][1], tmp.r[cc][2], tmp.r[cc][3]);
          frozen[((pl - 1) * 16 + cc * 2 + 1) * 128 + threadIdx.x] = make_uint4(tmp.r[cc][4], tmp.r[cc][5], tmp.r[cc][6], tmp.r[cc][7]);
#pragma unroll
          for (int k = 0; k < M; k++) any |= tmp.r[cc][k];
        }
      }
    }
  }
  const bool warp_frozen = __any_sync(0xffffffffu, any != 0);  // uniform over the warp: no divergence below
  if (!warp_frozen) {
#pragma unroll 1
    for (unsigned g = 0; g < gens; g++) t.step(sub);
  } else {
#pragma unroll 1
    for (unsigned g = 0; g < gens; g++) {
      asm volatile("" ::: "memory");  // keep the shared loads inside the loop (hoisted they are 192 registers)
      const TX::Col first = t.count(0), last = t.count(7);
      TX::Col prev = TX::halo(last, sub + 7);
      const TX::Col after = TX::halo(first, sub + 1);
      TX::Col cur = first;
#pragma unroll
      for (int c = 0; c < 8; c++) {
        const TX::Col next = (c == 7) ? after : (c + 1 == 7) ? last : t.count(c + 1);
        uint32_t f[3][M];
#pragma unroll
        for (int pl = 0; pl < 3; pl++) {
          const uint4 a = frozen[(pl * 16 + c * 2) * 128 + threadIdx.x], b = frozen[(pl * 16 + c * 2 + 1) * 128 + threadIdx.x];
          f[pl][0] = a.x; f[pl][1] = a.y; f[pl][2] = a.z; f[pl][3] = a.w;
          f[pl][4] = b.x; f[pl][5] = b.y; f[pl][6] = b.z; f[pl][7] = b.w;
        }
#pragma unroll
        for (int k = 0; k < M; k++)
          t.r[c][k] = weld_rule(prev.b0[k], prev.b1[k], cur.b0[k], cur.b1[k], next.b0[k], next.b1[k], t.r[c][k], f[0][k], f[1][k], f[2][k]);
        prev = cur;
        cur = next;
      }
    }
  }
#pragma unroll
  for (int c = 0; c < 8; c += 2) {
    uint64_t wa[P], wb[P];
    t.unpack(c, wa);
    t.unpack(c + 1, wb);
#pragma unroll
    for (int q = 0; q < P; q++)
      if (p0 + q < n)
        *reinterpret_cast<uint4*>(records + (p0 + q) * 4 * kN + sub * 8 + c) =
            make_uint4((uint32_t)wa[q], (uint32_t)(wa[q] >> 32), (uint32_t)wb[q], (uint32_t)(wb[q] >> 32));
  }
}

// ---- InteractionCounts[AndNext]: out1, out2, outMore (+ next), the reference's formulas verbatim ----
template <int LPP>
__device__ __forceinline__ void interaction_planes(const Tile<LPP>& s, int sub, Tile<LPP>& o1, Tile<LPP>& o2, Tile<LPP>& om,
                                                   Tile<LPP>& nx) {
  constexpr int W = Tile<LPP>::W;
  const int up_src = sub + LPP - 1, dn_src = sub + 1;
  Cnt first = count_rows(s.lo[0], s.hi[0]);
  Cnt last = count_rows(s.lo[W - 1], s.hi[W - 1]);
  Cnt prev = halo<LPP>(last, up_src);
  Cnt after = halo<LPP>(first, dn_src);
  Cnt cur = first;
#pragma unroll
  for (int w = 0; w < W; w++) {
    Cnt next = (w == W - 1) ? after : (w + 1 == W - 1) ? last : count_rows(s.lo[w + 1], s.hi[w + 1]);
#pragma unroll
    for (int h = 0; h < 2; h++) {
      const uint32_t a = h ? s.hi[w] : s.lo[w];
      const uint32_t u0 = h ? prev.b0hi : prev.b0lo, u1 = h ? prev.b1hi : prev.b1lo;
      const uint32_t c0 = h ? cur.b0hi : cur.b0lo, c1 = h ? cur.b1hi : cur.b1lo;
      const uint32_t l0 = h ? next.b0hi : next.b0lo, l1 = h ? next.b1hi : next.b1lo;
      const uint32_t fs = lop3<kXor3>(u0, c0, l0), fc = lop3<kMaj>(u0, c0, l0);
      const uint32_t cs = lop3<kXor3>(u1, c1, l1);
      uint32_t cc = lop3<kMaj>(u1, c1, l1);
      const uint32_t r1 = ~a & ~cc & fs & ~cs & ~fc;
      const uint32_t r2 = ~a & ~cc & ~fs & (cs ^ fc);
      const uint32_t rm = ~a & ~r2 & (fc | cs | cc);
      cc = cc ^ (cs & fc);
      const uint32_t nn = (fs ^ cc) & (fc ^ cs ^ cc) & (a | fs);
      if (h) { o1.hi[w] = r1; o2.hi[w] = r2; om.hi[w] = rm; nx.hi[w] = nn; }
      else { o1.lo[w] = r1; o2.lo[w] = r2; om.lo[w] = rm; nx.lo[w] = nn; }
    }
    prev = cur; cur = next;
  }
}

// A lane owns kN / LPP consecutive words.  Default LPP = 32 (a warp per pattern, lane L owns words 2 L, 2 L + 1):
// the load and each of the four plane stores is ONE contiguous 512-byte request per warp.  With 8 lanes per
// pattern a store instruction writes 16-byte pieces 64 bytes apart, four times the sectors per request:
// 4.9 TB/s against 6.7 TB/s on 1 Mi soups (profiles/r01o_interaction.txt; 512 B in + 2 KiB out per pattern).
template <int LPP>
__global__ void __launch_bounds__(kBlockN) interaction_kernel(const uint64_t* __restrict__ in, uint64_t* __restrict__ out,
                                                              size_t n) {
  constexpr int W = Tile<LPP>::W;
  const size_t p = ((size_t)blockIdx.x * kBlockN + threadIdx.x) / LPP;
  const int sub = threadIdx.x & (LPP - 1);
  const bool valid = p < n;  // invalid groups still take part in the full-warp shuffles
  Tile<LPP> s, o1, o2, om, nx;
  if (valid) s.load(in + p * kN, sub);
  else {
#pragma unroll
    for (int w = 0; w < W; w++) s.lo[w] = s.hi[w] = 0;
  }
  interaction_planes<LPP>(s, sub, o1, o2, om, nx);
  if (!valid) return;
  uint64_t* o = out + p * 4 * kN;
  o1.store(o, sub); o2.store(o + kN, sub); om.store(o + 2 * kN, sub); nx.store(o + 3 * kN, sub);
}

// ---- XXH3_64bits(state, 512): 8 stripes of 64 bytes.  A warp hashes 4 patterns (2 KiB) at a time: every
// load is one contiguous 512-byte request, lane L holding bytes [16 L, 16 L + 16) of each pattern = the two
// accumulator lanes i0 = 2 (L & 3), i0 + 1 of stripe L >> 2.  accumulate_512 only ADDS into the accumulators
// and a 512-byte input has no scramble round, so the stripes commute: partial sums are reduced across the
// eight stripe lanes, four patterns at once (each butterfly step keeps half of the patterns and sends the
// other half: 4 + 2 + 2 64-bit shuffles instead of 3 x 8).
__constant__ uint64_t c_stripe_keys[8][8] = {LIFEB200_XXH3_STRIPE_KEYS};
__constant__ uint64_t c_merge_keys[8] = {LIFEB200_XXH3_MERGE_KEYS};
__constant__ uint64_t c_acc_init[8] = {xxh3::kPrime32_3, xxh3::kPrime64_1, xxh3::kPrime64_2, xxh3::kPrime64_3,
                                       xxh3::kPrime64_4, xxh3::kPrime32_2, xxh3::kPrime64_5, xxh3::kPrime32_1};

__device__ __forceinline__ uint64_t fold128(uint64_t a, uint64_t b) { return (a * b) ^ __umul64hi(a, b); }
__device__ __forceinline__ uint64_t shfl_xor64(uint64_t v, int o) { return __shfl_xor_sync(0xffffffffu, v, o); }

__global__ void __launch_bounds__(kBlockN) hash_kernel(const uint64_t* __restrict__ in, uint64_t* __restrict__ hashes,
                                                       size_t n) {
  const size_t warps = ((size_t)gridDim.x * kBlockN) >> 5, groups = (n + 3) / 4;
  size_t g = ((size_t)blockIdx.x * kBlockN + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31, stripe = lane >> 2, i0 = 2 * (lane & 3);
  if (g >= groups) return;  // warp-uniform
  const uint64_t key0 = c_stripe_keys[stripe][i0], key1 = c_stripe_keys[stripe][i0 + 1];
  const uint64_t init0 = c_acc_init[i0], init1 = c_acc_init[i0 + 1], mk0 = c_merge_keys[i0], mk1 = c_merge_keys[i0 + 1];
  const bool hi4 = lane & 16, hi3 = lane & 8;
  auto fetch = [&](size_t grp, uint4 (&v)[4]) {
#pragma unroll
    for (int k = 0; k < 4; k++) {
      v[k] = make_uint4(0, 0, 0, 0);
      if (grp * 4 + k < n) v[k] = __ldg(reinterpret_cast<const uint4*>(in + (grp * 4 + k) * kN) + lane);
    }
  };
  uint4 v[4], nv[4];
  fetch(g, v);
  // a persistent warp walks groups g, g + warps, ..: the next group's four requests are in flight while this
  // one is reduced (one short-lived CTA per 32 patterns was bound by CTA launch rate, not by HBM)
  for (; g < groups; g += warps) {
    const bool more = g + warps < groups;
    if (more) fetch(g + warps, nv);
    uint64_t a0[4], a1[4];  // this lane's share of acc[i0], acc[i0 + 1] of patterns 4 g .. 4 g + 3
#pragma unroll
    for (int k = 0; k < 4; k++) {
      const uint64_t d0 = ((uint64_t)v[k].y << 32) | v[k].x, d1 = ((uint64_t)v[k].w << 32) | v[k].z;
      const uint64_t k0 = d0 ^ key0, k1 = d1 ^ key1;
      a0[k] = (k0 & 0xffffffffULL) * (k0 >> 32) + d1;  // acc[i] += lo * hi;  acc[i ^ 1] += data
      a1[k] = (k1 & 0xffffffffULL) * (k1 >> 32) + d0;
    }
    // sum over the 8 stripes (lane bits 2..4).  Bit 4 splits patterns {0,1} | {2,3}, bit 3 splits the pair.
    uint64_t b0[2], b1[2];
#pragma unroll
    for (int j = 0; j < 2; j++) {
      const uint64_t keep0 = hi4 ? a0[2 + j] : a0[j], send0 = hi4 ? a0[j] : a0[2 + j];
      const uint64_t keep1 = hi4 ? a1[2 + j] : a1[j], send1 = hi4 ? a1[j] : a1[2 + j];
      b0[j] = keep0 + shfl_xor64(send0, 16);
      b1[j] = keep1 + shfl_xor64(send1, 16);
    }
    uint64_t c0 = (hi3 ? b0[1] : b0[0]) + shfl_xor64(hi3 ? b0[0] : b0[1], 8);
    uint64_t c1 = (hi3 ? b1[1] : b1[0]) + shfl_xor64(hi3 ? b1[0] : b1[1], 8);
    c0 += shfl_xor64(c0, 4);
    c1 += shfl_xor64(c1, 4);
    // lanes with (bit 4, bit 3) = (h, l) now hold the full acc[i0], acc[i0 + 1] of pattern 2 h + l: mergeAccs
    uint64_t r = fold128((c0 + init0) ^ mk0, (c1 + init1) ^ mk1);
    r += shfl_xor64(r, 1);
    r += shfl_xor64(r, 2);
    r += 512ULL * xxh3::kPrime64_1;
    r ^= r >> 37;
    r *= xxh3::kPrimeMx1;
    r ^= r >> 32;
    const size_t p = g * 4 + (hi4 ? 2 : 0) + (hi3 ? 1 : 0);
    if ((lane & 7) == 0 && p < n) hashes[p] = r;
    if (more) {
#pragma unroll
      for (int k = 0; k < 4; k++) v[k] = nv[k];
    }
  }
}

}  // namespace lifeb200

using namespace lifeb200;

// persistent warps for hash_kernel: exactly as many CTAs as are resident at once (one wave, no tail)
static unsigned grid_hash(const lifeb200_ctx* ctx, size_t n) {
  int per_sm = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, hash_kernel, kBlockN, 0) != cudaSuccess || per_sm < 1) per_sm = 2;
  const unsigned resident = (unsigned)(ctx->sm_count > 0 ? ctx->sm_count : 148) * (unsigned)per_sm;
  return std::min(grid_n((n + 3) / 4, 32), resident);
}

extern "C" {

// measured crossover against the lane-group kernel, both with the 6-gate rule and the 11-instruction pack
// (profiles/r02x_thresholds.txt): level at 16 generations, +2.5 % at 20, +7 % at 32, +10 % at 64, +11 % at 128 (the
// lane-group kernel needs no layout conversion)
constexpr unsigned kWeldMultiFrom = 18;
static int launch_weld(lifeb200_ctx* ctx, cudaStream_t st, uint64_t* records, size_t n, unsigned gens) {
  if (ctx->lpp == 0 && (ctx->step_layout == 8 || (ctx->step_layout == 0 && gens >= kWeldMultiFrom))) {
    const size_t groups = (n + 3) / 4;
    LIFEB200_CUDA(cudaFuncSetAttribute(weld_step_kernel_x, cudaFuncAttributeMaxDynamicSharedMemorySize, kWeldXSmem));
    weld_step_kernel_x<<<(unsigned)((groups * 8 + 127) / 128), 128, kWeldXSmem, st>>>(records, n, gens);
    LIFEB200_LAUNCHED(ctx);
    return LIFEB200_OK;
  }
  int lpp = std::max(8, effective_lpp(ctx));  // 8 tiles live: keep W <= 8
  LIFEB200_DISPATCH_LPP_MIN4(lpp, (weld_step_kernel<LPP><<<grid_n(n, LPP), kBlockN, 0, st>>>(records, n, gens)));
  LIFEB200_LAUNCHED(ctx);
  return LIFEB200_OK;
}

int lifeb200_weld_step_batch_dev(lifeb200_ctx* ctx, uint64_t* records, size_t n, unsigned gens) {
  LIFEB200_CHECK_CTX(ctx);
  if (n == 0 || gens == 0) return LIFEB200_OK;
  if (!records || !aligned16(records)) return set_error(LIFEB200_E_ARG, "bad record pointer");
  return launch_weld(ctx, ctx->stream, records, n, gens);
}

int lifeb200_weld_step_batch_host(lifeb200_ctx* ctx, uint64_t* records, size_t n, unsigned gens) {
  LIFEB200_CHECK_CTX(ctx);
  if (n == 0 || gens == 0) return LIFEB200_OK;
  if (!records) return set_error(LIFEB200_E_ARG, "null record pointer");
  return host_pipeline_impl(ctx, records, records, n, 4 * LIFEB200_PATTERN_BYTES, 4 * LIFEB200_PATTERN_BYTES, false, (size_t)1 << 14,
                            [&](cudaStream_t st, const uint64_t* din, uint64_t*, size_t cnt) -> int {
                              return launch_weld(ctx, st, const_cast<uint64_t*>(din), cnt, gens);
                            });
}

static int launch_interaction(lifeb200_ctx* ctx, cudaStream_t st, const uint64_t* in, uint64_t* planes, size_t n) {
  int lpp = ctx->lpp ? std::max(8, ctx->lpp) : 32;
  LIFEB200_DISPATCH_LPP_MIN4(lpp, (interaction_kernel<LPP><<<grid_n(n, LPP), kBlockN, 0, st>>>(in, planes, n)));
  LIFEB200_LAUNCHED(ctx);
  return LIFEB200_OK;
}

int lifeb200_interaction_counts_dev(lifeb200_ctx* ctx, const uint64_t* in, uint64_t* planes, size_t n) {
  LIFEB200_CHECK_CTX(ctx);
  if (n == 0) return LIFEB200_OK;
  if (!in || !planes || !aligned16(in) || !aligned16(planes)) return set_error(LIFEB200_E_ARG, "null or misaligned pointer");
  return launch_interaction(ctx, ctx->stream, in, planes, n);
}

int lifeb200_interaction_counts_host(lifeb200_ctx* ctx, const uint64_t* in, uint64_t* planes, size_t n) {
  LIFEB200_CHECK_CTX(ctx);
  if (n == 0) return LIFEB200_OK;
  if (!in || !planes) return set_error(LIFEB200_E_ARG, "null pointer");
  return host_pipeline_impl(ctx, in, planes, n, LIFEB200_PATTERN_BYTES, 4 * LIFEB200_PATTERN_BYTES, true, (size_t)1 << 15,
                            [&](cudaStream_t st, const uint64_t* din, uint64_t* dout, size_t cnt) -> int {
                              return launch_interaction(ctx, st, din, dout, cnt);
                            });
}

int lifeb200_hash_batch_dev(lifeb200_ctx* ctx, const uint64_t* patterns, uint64_t* hashes, size_t n) {
  LIFEB200_CHECK_CTX(ctx);
  if (n == 0) return LIFEB200_OK;
  if (!patterns || !hashes || !aligned16(patterns)) return set_error(LIFEB200_E_ARG, "null or misaligned pointer");
  hash_kernel<<<grid_hash(ctx, n), kBlockN, 0, ctx->stream>>>(patterns, hashes, n);
  LIFEB200_LAUNCHED(ctx);
  return LIFEB200_OK;
}

int lifeb200_hash_batch_host(lifeb200_ctx* ctx, const uint64_t* patterns, uint64_t* hashes, size_t n) {
  LIFEB200_CHECK_CTX(ctx);
  if (n == 0) return LIFEB200_OK;
  if (!patterns || !hashes) return set_error(LIFEB200_E_ARG, "null pointer");
  const size_t chunk = std::min<size_t>(n, (size_t)1 << 16);
  const size_t nchunks = (n + chunk - 1) / chunk;
  const int used = (int)std::min<size_t>(kSlots, nchunks);
  for (int s = 0; s < used; s++) {
    int rc = ensure_slot(ctx, s, chunk * LIFEB200_PATTERN_BYTES, chunk * sizeof(uint64_t));
    if (rc) return rc;
  }
  for (size_t c = 0; c < nchunks; c++) {
    Slot& sl = ctx->slots[c % kSlots];
    const size_t off = c * chunk, cnt = std::min(chunk, n - off);
    uint64_t* d = static_cast<uint64_t*>(sl.dbuf);
    uint64_t* dh = static_cast<uint64_t*>(sl.dbuf2);
    LIFEB200_CUDA(cudaMemcpyAsync(d, patterns + off * kN, cnt * LIFEB200_PATTERN_BYTES, cudaMemcpyHostToDevice, sl.stream));
    hash_kernel<<<grid_hash(ctx, cnt), kBlockN, 0, sl.stream>>>(d, dh, cnt);
    LIFEB200_LAUNCHED(ctx);
    LIFEB200_CUDA(cudaMemcpyAsync(hashes + off, dh, cnt * sizeof(uint64_t), cudaMemcpyDeviceToHost, sl.stream));
  }
  for (int s = 0; s < used; s++) LIFEB200_CUDA(cudaStreamSynchronize(ctx->slots[s].stream));
  return LIFEB200_OK;
}

}  // extern "C"
