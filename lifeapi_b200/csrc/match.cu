// match.cu -- stepping fused with LifeState::Contains(const LifeTarget&) (LifeTarget.hpp:44-51),
// placement construction (Moved/operator|, LifeAPI.hpp:699-717,247-253), the soup survey with the
// period-2 activity test (operator==, LifeAPI.hpp:213-220) and hit compaction.
#include <algorithm>
#include <vector>

#include "ctx.h"
#include "life_core.cuh"
#include "life_multi.cuh"

namespace lifeb200 {

constexpr int kBlockM = 256;

// Per-pattern recovery state machine, identical in every lane of the pattern's group.
struct Recovery {
  unsigned first_false = 0, run = 0, recovered = 0;
  bool seen_false = false, hit = false, done = false;
  __device__ __forceinline__ void observe(bool mismatch, unsigned g, unsigned stable_interval) {
    if (done) return;
    if (mismatch) {
      if (!seen_false) { seen_false = true; first_false = g; }
      run = 0;
    } else if (seen_false) {
      if (++run == stable_interval) { hit = true; done = true; recovered = g + 1 - stable_interval; }
    }
  }
  // the same transition without branches (selects only), for enumerate_kernel_x: it runs two warps per
  // scheduler, where a divergent branch is an exposed stall and splits the basic block the scheduler
  // interleaves.  A call with mismatch = false before any mismatch was seen changes nothing.
  __device__ __forceinline__ void observe_flat(bool mismatch, unsigned g, unsigned stable_interval) {
    const bool live = !done, mis = live && mismatch;
    first_false = (mis && !seen_false) ? g : first_false;
    seen_false = seen_false || mis;
    const bool ok = live && !mismatch && seen_false;
    run = mis ? 0u : run + (ok ? 1u : 0u);
    const bool fin = ok && run == stable_interval;
    recovered = fin ? g + 1 - stable_interval : recovered;
    hit = hit || fin;
    done = done || fin;
  }
};

// Warp-aggregated append: one atomicAdd per warp, hits beyond `cap` are counted but not written.
__device__ __forceinline__ void append_hit(bool emit, uint64_t id, uint32_t a, uint32_t b, lifeb200_hit* hits,
                                           size_t cap, unsigned long long* n_hits) {
  unsigned m = __ballot_sync(0xffffffffu, emit);
  if (m == 0) return;
  int lane = threadIdx.x & 31;
  int leader = __ffs(m) - 1;
  unsigned long long base = 0;
  if (lane == leader) base = atomicAdd(n_hits, (unsigned long long)__popc(m));
  base = __shfl_sync(0xffffffffu, base, leader);
  if (emit) {
    unsigned long long slot = base + __popc(m & ((1u << lane) - 1u));
    if (slot < cap) {
      lifeb200_hit h;
      h.id = id; h.g_first_false = a; h.g_recovered = b;
      hits[slot] = h;
    }
  }
}

template <int LPP>
__device__ __forceinline__ void run_match(Tile<LPP>& t, Tile<LPP>& wm, Tile<LPP>& m, int sub, int lane,
                                          unsigned gens, unsigned stable_interval, Recovery& r) {
  // even/odd row registers for the whole run (one shift per register and generation, life_core.cuh);
  // Contains is bit-position-wise, so the target planes are converted once and nothing is converted back
  t.interleave();
  wm.interleave();
  m.interleave();
#pragma unroll 1
  for (unsigned g = 1; g <= gens; g++) {
    t.template step<true>(sub);
    uint32_t d = t.target_diff(wm, m);
    bool mismatch = group_any<LPP>(d != 0, lane);
    r.observe(mismatch, g, stable_interval);
    if ((g & 7u) == 0 && __all_sync(0xffffffffu, r.done)) break;
  }
}

template <int LPP>
__global__ void __launch_bounds__(kBlockM, 1) step_match_kernel(const uint64_t* __restrict__ patterns,
                                                             const uint64_t* __restrict__ targets,
                                                             const uint32_t* __restrict__ target_index, size_t n,
                                                             uint64_t first_id, unsigned gens, unsigned stable_interval,
                                                             lifeb200_hit* hits, size_t cap, unsigned long long* n_hits) {
  constexpr int W = Tile<LPP>::W;
  const size_t gid = (size_t)blockIdx.x * kBlockM + threadIdx.x;
  const size_t p = gid / LPP;
  const int lane = threadIdx.x & 31, sub = threadIdx.x & (LPP - 1);
  const bool valid = p < n;
  Tile<LPP> t, wm, m;
  Recovery r;
  if (valid) {
    t.load(patterns + p * kN, sub);
    const uint64_t* tg = targets + (size_t)(target_index ? target_index[p] : 0u) * 2 * kN;
    wm.load(tg, sub);       // wanted
    m.load(tg + kN, sub);   // unwanted
#pragma unroll
    for (int w = 0; w < W; w++) {
      m.lo[w] |= wm.lo[w]; m.hi[w] |= wm.hi[w];  // mask = wanted | unwanted
    }
  } else {
#pragma unroll
    for (int w = 0; w < W; w++) t.lo[w] = t.hi[w] = wm.lo[w] = wm.hi[w] = m.lo[w] = m.hi[w] = 0;
    r.done = true;
  }
  run_match<LPP>(t, wm, m, sub, lane, gens, stable_interval, r);
  append_hit(valid && r.hit && sub == 0, first_id + p, r.first_false, r.recovered, hits, cap, n_hits);
}

__device__ __forceinline__ uint64_t rotl64v(uint64_t v, unsigned s) {
  s &= 63u;
  return s ? (v << s) | (v >> (64u - s)) : v;
}

// tables: [cat | catZOI | proj | projZOI], each entry 64 words.
template <int LPP>
__global__ void __launch_bounds__(kBlockM, 1) enumerate_kernel(const uint64_t* __restrict__ cat,
                                                            const uint64_t* __restrict__ cat_zoi,
                                                            const uint64_t* __restrict__ proj,
                                                            const uint64_t* __restrict__ proj_zoi, unsigned n_proj,
                                                            uint64_t id_begin, uint64_t id_end, unsigned gens,
                                                            unsigned stable_interval, lifeb200_hit* hits, size_t cap,
                                                            unsigned long long* n_hits, unsigned long long* n_skipped) {
  constexpr int W = Tile<LPP>::W;
  const uint64_t gid = (uint64_t)blockIdx.x * kBlockM + threadIdx.x;
  const uint64_t id = id_begin + gid / LPP;
  const int lane = threadIdx.x & 31, sub = threadIdx.x & (LPP - 1);
  const bool valid = id < id_end;
  Tile<LPP> t, wm, m;
  Recovery r;
  uint32_t overlap = 0;
  if (valid) {
    const unsigned dy = (unsigned)(id & 63u), dx = (unsigned)((id >> 6) & 63u);
    const uint64_t cg = id >> 12;
    const unsigned g = (unsigned)(cg % n_proj);
    const uint64_t c = cg / n_proj;
    const uint64_t* cw = cat + c * kN;
    const uint64_t* cz = cat_zoi + c * kN;
    const uint64_t* pw = proj + (size_t)g * kN;
    const uint64_t* pz = proj_zoi + (size_t)g * kN;
#pragma unroll
    for (int w = 0; w < W; w++) {
      const int i = sub * W + w;
      const int src = (i - (int)dx) & 63;  // Moved: result[i] = rotl(state[i - x], y)
      uint64_t cat_w = __ldg(cw + i), zoi_w = __ldg(cz + i);
      uint64_t pm = rotl64v(__ldg(pw + src), dy), pzm = rotl64v(__ldg(pz + src), dy);
      uint64_t s = cat_w | pm;
      t.lo[w] = (uint32_t)s; t.hi[w] = (uint32_t)(s >> 32);
      // LifeTarget(catalyst): wanted = cat, unwanted = ZOI & ~cat  =>  mask = ZOI, wanted&mask = cat
      wm.lo[w] = (uint32_t)cat_w; wm.hi[w] = (uint32_t)(cat_w >> 32);
      m.lo[w] = (uint32_t)zoi_w; m.hi[w] = (uint32_t)(zoi_w >> 32);
      uint64_t ov = zoi_w & pzm;
      overlap |= (uint32_t)ov | (uint32_t)(ov >> 32);
    }
  } else {
#pragma unroll
    for (int w = 0; w < W; w++) t.lo[w] = t.hi[w] = wm.lo[w] = wm.hi[w] = m.lo[w] = m.hi[w] = 0;
  }
  const bool skip = group_any<LPP>(overlap != 0, lane);
  if (!valid || skip) r.done = true;
  {
    unsigned sk = __ballot_sync(0xffffffffu, valid && skip && sub == 0);
    if (lane == 0 && sk) atomicAdd(n_skipped, (unsigned long long)__popc(sk));
  }
  run_match<LPP>(t, wm, m, sub, lane, gens, stable_interval, r);
  append_hit(valid && r.hit && sub == 0, id, r.first_false, r.recovered, hits, cap, n_hits);
}

// The same enumeration with 4 placements per group of 8 lanes in the row-class layout (life_multi.cuh).
// Ids that differ only in their two low bits share catalyst, projectile and dx and have consecutive dy,
// so a group owns ids 4q .. 4q+3: one set of table words, four rotations.  All four share the target
// too; its planes come already packed (every cell replicated into the four pattern bits, built on the
// host per catalyst).  They do not fit in registers next to the state (64 + 128), so every group parks
// them in shared memory (64 KiB per CTA) and its lanes read them back each generation: 32 LDS.128, one
// wavefront each, against ~590 integer instructions.
// Per generation and register: 7.75 (step) + 1.5 (Contains) integer-pipe instructions instead of 8.5 + 1.5.
constexpr int kEnumXSmem = 128 * 32 * 16;  // threads x uint4 per thread x 16 B

// OR over the 8 lanes of a group (xor-shuffles with a full mask: __reduce_or_sync with a partial mask
// compiles to a convergence loop with calls and branches)
__device__ __forceinline__ uint32_t group8_or(uint32_t v) {
  v |= __shfl_xor_sync(0xffffffffu, v, 1);
  v |= __shfl_xor_sync(0xffffffffu, v, 2);
  v |= __shfl_xor_sync(0xffffffffu, v, 4);
  return v;
}

__global__ void __launch_bounds__(128, 1) enumerate_kernel_x(const uint64_t* __restrict__ cat_zoi,
                                                             const uint64_t* __restrict__ proj,
                                                             const uint64_t* __restrict__ proj_zoi,
                                                             const uint32_t* __restrict__ packed_wanted,
                                                             const uint32_t* __restrict__ packed_mask, unsigned n_proj,
                                                             uint64_t id_begin, uint64_t id_end, unsigned gens,
                                                             unsigned stable_interval, lifeb200_hit* hits, size_t cap,
                                                             unsigned long long* n_hits, unsigned long long* n_skipped) {
  constexpr int M = 8, P = 4;
  const uint64_t gid = (uint64_t)blockIdx.x * 128 + threadIdx.x;
  const uint64_t base = (id_begin & ~3ull) + (gid >> 3) * P;
  const int lane = threadIdx.x & 31, sub = threadIdx.x & 7, mine = sub & 3;
  const uint64_t id = base + mine;
  const bool valid = id >= id_begin && id < id_end;
  const bool group_live = base < id_end;
  TileX<M> t;
  const uint4* wanted4 = nullptr;
  const uint4* mask4 = nullptr;
  uint32_t overlap = 0;
  if (group_live) {
    const unsigned dy0 = (unsigned)(base & 63u), dx = (unsigned)((base >> 6) & 63u);
    const uint64_t cg = base >> 12;
    const unsigned g = (unsigned)(cg % n_proj);
    const uint64_t c = cg / n_proj;
    const uint64_t* cz = cat_zoi + c * kN;
    const uint64_t* pw = proj + (size_t)g * kN;
    const uint64_t* pz = proj_zoi + (size_t)g * kN;
    wanted4 = reinterpret_cast<const uint4*>(packed_wanted + (c * kN + sub * 8) * M);
    mask4 = reinterpret_cast<const uint4*>(packed_mask + (c * kN + sub * 8) * M);
#pragma unroll
    for (int col = 0; col < 8; col++) {
      const int i = sub * 8 + col;
      const int src = (i - (int)dx) & 63;  // Moved: result[i] = rotl(state[i - x], y)
      const uint64_t zoi_w = __ldg(cz + i), p_w = __ldg(pw + src), pz_w = __ldg(pz + src);
      uint64_t w[P];
#pragma unroll
      for (int q = 0; q < P; q++) {
        w[q] = rotl64v(p_w, dy0 + q);
        const uint64_t ov = zoi_w & rotl64v(pz_w, dy0 + q);
        overlap |= (ov != 0 ? 1u : 0u) << q;
      }
      t.pack(col, w);
      const uint4 a = __ldg(wanted4 + 2 * col), b = __ldg(wanted4 + 2 * col + 1);
      t.r[col][0] |= a.x; t.r[col][1] |= a.y; t.r[col][2] |= a.z; t.r[col][3] |= a.w;
      t.r[col][4] |= b.x; t.r[col][5] |= b.y; t.r[col][6] |= b.z; t.r[col][7] |= b.w;
    }
  } else {
#pragma unroll
    for (int col = 0; col < 8; col++)
#pragma unroll
      for (int k = 0; k < M; k++) t.r[col][k] = 0;
    wanted4 = reinterpret_cast<const uint4*>(packed_wanted + sub * 8 * M);  // catalyst 0: any valid address
    mask4 = reinterpret_cast<const uint4*>(packed_mask + sub * 8 * M);
  }
  // park the planes: one 4 KiB slice per group, [word][lane of the group] as uint4; groups of a warp that
  // share the catalyst (all four, except across a catalyst edge) then read the FIRST such group's slice, so
  // a warp's request is 8 distinct consecutive uint4 = one conflict-free 128-byte wavefront
  extern __shared__ uint4 parked[];  // [16 groups][32 words][8 lanes]; words 0..15 wanted, 16..31 mask
  const int grp = threadIdx.x >> 3;
  {
    uint4* mine4 = parked + grp * 256 + sub;
#pragma unroll
    for (int q = 0; q < 16; q++) {
      mine4[q * 8] = __ldg(wanted4 + q);
      mine4[(16 + q) * 8] = __ldg(mask4 + q);
    }
  }
  __syncwarp();
  int src_grp = grp;
  {
    const unsigned long long key = group_live ? (unsigned long long)((base >> 12) / n_proj) : ~0ull;
#pragma unroll
    for (int o = 3; o >= 0; o--) {  // lowest group of this warp with my catalyst
      const unsigned long long other = __shfl_sync(0xffffffffu, key, o * 8);
      if (other == key) src_grp = (grp & ~3) + o;
    }
  }
  const uint4* planes = parked + src_grp * 256 + sub;
  const bool skip = (group8_or(overlap) >> mine) & 1u;
  Recovery r;
  if (!valid || skip) r.done = true;
  {
    unsigned sk = __ballot_sync(0xffffffffu, valid && skip && sub < P);
    if (lane == 0 && sk) atomicAdd(n_skipped, (unsigned long long)__popc(sk));
  }
  // The group-wide OR and the recovery state machine of generation g - 1 sit at the top of iteration g, in
  // the same basic block as the step, so their shuffle latency hides under it (two warps per scheduler:
  // nothing else would cover it); the first call sees d_prev = 0 and changes nothing.  The per-register
  // differences are OR-ed as a tree, not a 32-deep chain.
  uint32_t d_prev = 0;
#pragma unroll 1
  for (unsigned g = 1; g <= gens; g++) {
    asm volatile("" ::: "memory");  // keep the shared loads inside the loop (hoisted they are 128 registers)
    const uint32_t all = group8_or(d_prev);  // bit 4j+q set: some cell of placement q differed after g - 1
    r.observe_flat((all & (0x11111111u << mine)) != 0, g - 1, stable_interval);
    t.step(sub);
    uint32_t dc[8];
#pragma unroll
    for (int col = 0; col < 8; col++) {
      const uint4 wa = planes[(2 * col) * 8], wb = planes[(2 * col + 1) * 8];
      const uint4 ma = planes[(16 + 2 * col) * 8], mb = planes[(17 + 2 * col) * 8];
      const uint32_t x0 = lop3<kAndXor>(t.r[col][0], ma.x, wa.x), x1 = lop3<kAndXor>(t.r[col][1], ma.y, wa.y);
      const uint32_t x2 = lop3<kAndXor>(t.r[col][2], ma.z, wa.z), x3 = lop3<kAndXor>(t.r[col][3], ma.w, wa.w);
      const uint32_t x4 = lop3<kAndXor>(t.r[col][4], mb.x, wb.x), x5 = lop3<kAndXor>(t.r[col][5], mb.y, wb.y);
      const uint32_t x6 = lop3<kAndXor>(t.r[col][6], mb.z, wb.z), x7 = lop3<kAndXor>(t.r[col][7], mb.w, wb.w);
      dc[col] = lop3<kOr3>(lop3<kOr3>(x0, x1, x2), lop3<kOr3>(x3, x4, x5), x6 | x7);
    }
    d_prev = lop3<kOr3>(lop3<kOr3>(dc[0], dc[1], dc[2]), lop3<kOr3>(dc[3], dc[4], dc[5]), dc[6] | dc[7]);
    if ((g & 7u) == 0 && __all_sync(0xffffffffu, r.done)) break;  // decided up to g - 1: nothing left to observe
  }
  {
    const uint32_t all = group8_or(d_prev);
    r.observe_flat((all & (0x11111111u << mine)) != 0, gens, stable_interval);
  }
  append_hit(valid && r.hit && sub < P, id, r.first_false, r.recovered, hits, cap, n_hits);
}

// Per-pattern 64-bit fold used as a size-independent checksum (build-defined):
// fold(p) = XOR_i splitmix64(state[i] + i*K);  batch = XOR_p splitmix64(fold(p) + id_p).
constexpr uint64_t kFoldK = 0xD6E8FEB86659FD93ULL;
template <int LPP>
__device__ __forceinline__ uint64_t group_fold(const Tile<LPP>& t, int sub) {
  constexpr int W = Tile<LPP>::W;
  uint64_t f = 0;
#pragma unroll
  for (int w = 0; w < W; w++) {
    uint64_t v = ((uint64_t)t.hi[w] << 32) | t.lo[w];
    f ^= splitmix64(v + (uint64_t)(sub * W + w) * kFoldK);
  }
#pragma unroll
  for (int o = LPP / 2; o > 0; o >>= 1) f ^= __shfl_xor_sync(0xffffffffu, f, o, 32);
  return f;
}

__device__ __forceinline__ void block_accumulate(uint64_t pop, uint64_t h, unsigned long long* pop_sum,
                                                 unsigned long long* hash_xor) {
  // warp reduce, then one atomic pair per warp
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    pop += __shfl_xor_sync(0xffffffffu, pop, o, 32);
    h ^= __shfl_xor_sync(0xffffffffu, h, o, 32);
  }
  if ((threadIdx.x & 31) == 0) {
    atomicAdd(pop_sum, (unsigned long long)pop);
    atomicXor(hash_xor, (unsigned long long)h);
  }
}

template <int LPP>
__global__ void __launch_bounds__(kBlockM, 1) survey_kernel(uint64_t first_id, size_t n, uint64_t seed, int dist,
                                                         unsigned gens, lifeb200_hit* hits, size_t cap,
                                                         unsigned long long* n_hits, unsigned long long* pop_sum,
                                                         unsigned long long* hash_xor) {
  constexpr int W = Tile<LPP>::W;
  const size_t gid = (size_t)blockIdx.x * kBlockM + threadIdx.x;
  const size_t p = gid / LPP;
  const int lane = threadIdx.x & 31, sub = threadIdx.x & (LPP - 1);
  const bool valid = p < n;
  const uint64_t id = first_id + p;
  Tile<LPP> t, before;
#pragma unroll
  for (int w = 0; w < W; w++) {
    uint64_t v = valid ? soup_word(seed, id, sub * W + w, dist) : 0ULL;
    t.lo[w] = (uint32_t)v; t.hi[w] = (uint32_t)(v >> 32);
  }
  t.interleave();  // even/odd row registers; ==, GetPop are layout-blind, the fold needs the words back
#pragma unroll 1
  for (unsigned g = 0; g + 2 < gens; g++) t.template step<true>(sub);
  before = t;
  t.template step<true>(sub);
  t.template step<true>(sub);
  const bool active = group_any<LPP>(t.xor_diff(before) != 0, lane);
  const uint32_t pop = group_sum<LPP>(t.pop());
  t.deinterleave();
  const uint64_t f = group_fold<LPP>(t, sub);
  const bool leader = valid && sub == 0;
  block_accumulate(leader ? pop : 0u, leader ? splitmix64(f + id) : 0ULL, pop_sum, hash_xor);
  append_hit(leader && active, id, pop, 0u, hits, cap, n_hits);
}

// The same survey with P = M/2 soups per group of 8 lanes in the row-class layout (life_multi.cuh).
template <int M>
__global__ void __launch_bounds__(128, 1) survey_kernel_x(uint64_t first_id, size_t n, uint64_t seed, int dist, unsigned gens,
                                                          lifeb200_hit* hits, size_t cap, unsigned long long* n_hits,
                                                          unsigned long long* pop_sum, unsigned long long* hash_xor) {
  constexpr int P = M / 2;
  const size_t gid = (size_t)blockIdx.x * 128 + threadIdx.x;
  const size_t p0 = (gid >> 3) * P;
  const int lane = threadIdx.x & 31, sub = threadIdx.x & 7;
  TileX<M> t, before;
#pragma unroll
  for (int c = 0; c < 8; c++) {
    uint64_t w[P];
#pragma unroll
    for (int s = 0; s < P; s++) w[s] = (p0 + s < n) ? soup_word(seed, first_id + p0 + s, sub * 8 + c, dist) : 0ULL;
    t.pack(c, w);
  }
#pragma unroll 1
  for (unsigned g = 0; g + 2 < gens; g++) t.step(sub);
  before = t;
  t.step(sub);
  t.step(sub);
  // per soup s: its cells are the bits j*P + s of every register
  uint32_t diff = 0, pop[P];
#pragma unroll
  for (int s = 0; s < P; s++) pop[s] = 0;
#pragma unroll
  for (int c = 0; c < 8; c++)
#pragma unroll
    for (int k = 0; k < M; k++) {
      diff |= t.r[c][k] ^ before.r[c][k];
#pragma unroll
      for (int s = 0; s < P; s++) pop[s] += __popc(t.r[c][k] & ((0xFFFFFFFFu / ((1u << P) - 1u)) << s));
    }
  uint64_t f[P];
#pragma unroll
  for (int s = 0; s < P; s++) f[s] = 0;
#pragma unroll
  for (int c = 0; c < 8; c++) {
    uint64_t w[P];
    t.unpack(c, w);
#pragma unroll
    for (int s = 0; s < P; s++) f[s] ^= splitmix64(w[s] + (uint64_t)(sub * 8 + c) * kFoldK);
  }
  // lane `sub` of the group reports soup s = sub (sub < P)
  uint32_t my_pop = 0, my_diff = 0;
  uint64_t my_f = 0;
#pragma unroll
  for (int s = 0; s < P; s++) {
    uint32_t d = diff & ((0xFFFFFFFFu / ((1u << P) - 1u)) << s), q = pop[s];
    uint64_t h = f[s];
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) {
      d |= __shfl_xor_sync(0xffffffffu, d, o, 32);
      q += __shfl_xor_sync(0xffffffffu, q, o, 32);
      h ^= __shfl_xor_sync(0xffffffffu, h, o, 32);
    }
    if (sub == s) { my_pop = q; my_diff = d; my_f = h; }
  }
  (void)lane;
  const bool leader = sub < P && p0 + sub < n;
  const uint64_t id = first_id + p0 + sub;
  block_accumulate(leader ? my_pop : 0u, leader ? splitmix64(my_f + id) : 0ULL, pop_sum, hash_xor);
  append_hit(leader && my_diff != 0, id, my_pop, 0u, hits, cap, n_hits);
}

template <int LPP>
__global__ void __launch_bounds__(kBlockM) fold_kernel(const uint64_t* __restrict__ patterns, size_t n, uint64_t first_id,
                                                       unsigned long long* pop_sum, unsigned long long* hash_xor) {
  constexpr int W = Tile<LPP>::W;
  const size_t gid = (size_t)blockIdx.x * kBlockM + threadIdx.x;
  const size_t p = gid / LPP;
  const int sub = threadIdx.x & (LPP - 1);
  const bool valid = p < n;
  Tile<LPP> t;
  if (valid) t.load(patterns + p * kN, sub);
  else {
#pragma unroll
    for (int w = 0; w < W; w++) t.lo[w] = t.hi[w] = 0;
  }
  const uint32_t pop = group_sum<LPP>(t.pop());
  const uint64_t f = group_fold<LPP>(t, sub);
  const bool leader = valid && sub == 0;
  block_accumulate(leader ? pop : 0u, leader ? splitmix64(f + first_id + p) : 0ULL, pop_sum, hash_xor);
}

static inline unsigned grid_m(uint64_t n, int lpp) { return (unsigned)((n * (uint64_t)lpp + kBlockM - 1) / kBlockM); }

static void host_zoi(const uint64_t* s, uint64_t* out) {  // LifeAPI.hpp:521-536
  uint64_t t[64];
  for (int i = 0; i < 64; i++) t[i] = s[i] | (s[i] << 1 | s[i] >> 63) | (s[i] >> 1 | s[i] << 63);
  for (int i = 0; i < 64; i++) out[i] = t[(i + 63) & 63] | t[i] | t[(i + 1) & 63];
}

// match kernels keep three tiles (state, wanted&mask, mask) in registers: cap W at 16.
// packing 4 placements costs about 6 generations of stepping
// measured crossovers against the lane-group kernels (profiles/r01z_multi_thresholds.txt)
constexpr unsigned kEnumerateMultiFrom = 16, kSurveyMultiFrom = 16;
static inline int match_lpp(const lifeb200_ctx* ctx) { return std::max(4, effective_lpp(ctx)); }
static inline int survey_lpp(const lifeb200_ctx* ctx) { return std::max(2, effective_lpp(ctx)); }

}  // namespace lifeb200

using namespace lifeb200;

extern "C" {

int lifeb200_step_match_dev(lifeb200_ctx* ctx, const uint64_t* patterns, const uint64_t* targets,
                            const uint32_t* target_index, size_t n, uint64_t first_id, unsigned gens,
                            unsigned stable_interval, lifeb200_hit* hits, size_t cap, unsigned long long* n_hits) {
  LIFEB200_CHECK_CTX(ctx);
  if (!n_hits) return set_error(LIFEB200_E_ARG, "null n_hits");
  if (stable_interval == 0) return set_error(LIFEB200_E_ARG, "stable_interval must be >= 1");
  LIFEB200_CUDA(cudaMemsetAsync(n_hits, 0, sizeof(unsigned long long), ctx->stream));
  if (n == 0) return LIFEB200_OK;
  if (!patterns || !targets || (!hits && cap)) return set_error(LIFEB200_E_ARG, "null pointer");
  if (!aligned16(patterns) || !aligned16(targets)) return set_error(LIFEB200_E_ARG, "pointers must be 16-byte aligned");
  int lpp = match_lpp(ctx);
  LIFEB200_DISPATCH_LPP(lpp, (step_match_kernel<LPP><<<grid_m(n, LPP), kBlockM, 0, ctx->stream>>>(
                                 patterns, targets, target_index, n, first_id, gens, stable_interval, hits, cap, n_hits)));
  LIFEB200_LAUNCHED(ctx);
  return LIFEB200_OK;
}

// Hit records leave the device in atomic-reservation order; the caller gets them by id.  Ids are unique and lie
// in [id_lo, id_lo + id_count): when a sizeable share of the range hit (a survey after few generations: most
// soups are still active) one scatter through a slot table orders them in O(range + found) where std::sort
// took 15 ms for 512 Ki records; sparse hit lists (the enumeration: thousands of hits in millions of ids) sort.
static void order_hits(lifeb200_hit* hits, uint64_t found, uint64_t id_lo, uint64_t id_count) {
  auto by_id = [](const lifeb200_hit& a, const lifeb200_hit& b) { return a.id < b.id; };
  if (found < 2 || std::is_sorted(hits, hits + found, by_id)) return;
  if (id_count < 0xFFFFFFFFull && found >= id_count / 16) {
    std::vector<uint32_t> slot(id_count, 0xFFFFFFFFu);
    bool ok = true;
    for (uint64_t i = 0; i < found && ok; i++) {
      const uint64_t k = hits[i].id - id_lo;
      ok = k < id_count && slot[k] == 0xFFFFFFFFu;
      if (ok) slot[k] = (uint32_t)i;
    }
    if (ok) {
      const std::vector<lifeb200_hit> tmp(hits, hits + found);
      uint64_t o = 0;
      for (uint64_t k = 0; k < id_count; k++)
        if (slot[k] != 0xFFFFFFFFu) hits[o++] = tmp[slot[k]];
      return;
    }
  }
  std::sort(hits, hits + found, by_id);
}

static int fetch_hits(lifeb200_ctx* ctx, lifeb200_hit* d_hits, size_t cap, lifeb200_hit* hits, uint64_t* n_hits,
                      uint64_t id_lo, uint64_t id_count) {
  // counters were copied to the pinned mirror and the stream synchronised by the caller
  uint64_t found = ctx->h_counters[0];
  if (n_hits) *n_hits = found;
  if (found > cap) return set_error(LIFEB200_E_OVERFLOW, "hit buffer too small: %llu hits, capacity %llu",
                                    (unsigned long long)found, (unsigned long long)cap);
  if (found) {
    LIFEB200_CUDA(cudaMemcpyAsync(hits, d_hits, found * sizeof(lifeb200_hit), cudaMemcpyDeviceToHost, ctx->stream));
    LIFEB200_CUDA(cudaStreamSynchronize(ctx->stream));
    order_hits(hits, found, id_lo, id_count);
  }
  return LIFEB200_OK;
}

int lifeb200_enumerate_collisions(lifeb200_ctx* ctx, const uint64_t* catalysts, unsigned n_catalysts,
                                  const uint64_t* projectiles, unsigned n_projectiles, uint64_t id_begin,
                                  uint64_t id_end, unsigned gens, unsigned stable_interval, lifeb200_hit* hits,
                                  size_t cap, uint64_t* n_hits, uint64_t* n_skipped) {
  LIFEB200_CHECK_CTX(ctx);
  if (!catalysts || !projectiles || n_catalysts == 0 || n_projectiles == 0)
    return set_error(LIFEB200_E_ARG, "empty catalyst / projectile table");
  if (stable_interval == 0) return set_error(LIFEB200_E_ARG, "stable_interval must be >= 1");
  const uint64_t total = (uint64_t)n_catalysts * n_projectiles * 4096ull;
  if (id_end > total) id_end = total;
  if (n_hits) *n_hits = 0;
  if (n_skipped) *n_skipped = 0;
  if (id_begin >= id_end) return LIFEB200_OK;
  if (!hits && cap) return set_error(LIFEB200_E_ARG, "null hits with non-zero capacity");
  const uint64_t n = id_end - id_begin;

  // tables: cat | catZOI | proj | projZOI
  const size_t words = (size_t)(n_catalysts + n_projectiles) * 2 * 64;
  std::vector<uint64_t> tab(words);
  uint64_t* h_cat = tab.data();
  uint64_t* h_cz = h_cat + (size_t)n_catalysts * 64;
  uint64_t* h_pr = h_cz + (size_t)n_catalysts * 64;
  uint64_t* h_pz = h_pr + (size_t)n_projectiles * 64;
  std::copy(catalysts, catalysts + (size_t)n_catalysts * 64, h_cat);
  std::copy(projectiles, projectiles + (size_t)n_projectiles * 64, h_pr);
  for (unsigned c = 0; c < n_catalysts; c++) host_zoi(h_cat + (size_t)c * 64, h_cz + (size_t)c * 64);
  for (unsigned g = 0; g < n_projectiles; g++) host_zoi(h_pr + (size_t)g * 64, h_pz + (size_t)g * 64);

  // many generations: 4 placements per lane group in the row-class layout (enumerate_kernel_x), unless the
  // caller pinned the lane grouping or another layout.  Its target planes are packed here: every cell of
  // wanted (= catalyst) and mask (= its ZOI) replicated into the four pattern bits of register (column, k).
  const bool multi = ctx->lpp == 0 && (ctx->step_layout == 8 || (ctx->step_layout == 0 && gens >= kEnumerateMultiFrom));
  std::vector<uint32_t> packed;
  if (multi) {
    packed.resize((size_t)n_catalysts * 64 * 8 * 2);
    uint32_t* pk_w = packed.data();
    uint32_t* pk_m = pk_w + (size_t)n_catalysts * 64 * 8;
    for (size_t ci = 0; ci < (size_t)n_catalysts * 64; ci++)
      for (int k = 0; k < 8; k++) {
        uint32_t w = 0, m = 0;
        for (int j = 0; j < 8; j++) {
          if (h_cat[ci] >> (8 * j + k) & 1) w |= 0xFu << (4 * j);
          if (h_cz[ci] >> (8 * j + k) & 1) m |= 0xFu << (4 * j);
        }
        pk_w[ci * 8 + k] = w;
        pk_m[ci * 8 + k] = m;
      }
  }

  uint64_t* d_tab = nullptr;
  uint32_t* d_packed = nullptr;
  lifeb200_hit* d_hits = nullptr;
  // context-owned scratch (ctx.h): no allocation after the first call of a given size
  if (int rc0 = ensure_scratch(ctx->d_tabbuf, ctx->tabbuf_cap, words * 8 + packed.size() * 4)) return rc0;
  d_tab = static_cast<uint64_t*>(ctx->d_tabbuf);
  d_packed = reinterpret_cast<uint32_t*>(d_tab + words);
  if (cap) {
    if (int rc0 = ensure_scratch(ctx->d_hitbuf, ctx->hitbuf_cap, cap * sizeof(lifeb200_hit))) return rc0;
    d_hits = static_cast<lifeb200_hit*>(ctx->d_hitbuf);
  }
  int rc = LIFEB200_OK;
  auto body = [&]() -> int {
    LIFEB200_CUDA(cudaMemcpyAsync(d_tab, tab.data(), words * 8, cudaMemcpyHostToDevice, ctx->stream));
    LIFEB200_CUDA(cudaMemsetAsync(ctx->d_counters, 0, 4 * sizeof(unsigned long long), ctx->stream));
    const uint64_t* d_cat = d_tab;
    const uint64_t* d_cz = d_cat + (size_t)n_catalysts * 64;
    const uint64_t* d_pr = d_cz + (size_t)n_catalysts * 64;
    const uint64_t* d_pz = d_pr + (size_t)n_projectiles * 64;
    if (multi) {
      LIFEB200_CUDA(cudaMemcpyAsync(d_packed, packed.data(), packed.size() * 4, cudaMemcpyHostToDevice, ctx->stream));
      const uint64_t groups = (id_end - (id_begin & ~3ull) + 3) / 4;
      LIFEB200_CUDA(cudaFuncSetAttribute(enumerate_kernel_x, cudaFuncAttributeMaxDynamicSharedMemorySize, kEnumXSmem));
      enumerate_kernel_x<<<(unsigned)((groups * 8 + 127) / 128), 128, kEnumXSmem, ctx->stream>>>(
          d_cz, d_pr, d_pz, d_packed, d_packed + (size_t)n_catalysts * 64 * 8, n_projectiles, id_begin, id_end, gens,
          stable_interval, d_hits, cap, ctx->d_counters + 0, ctx->d_counters + 1);
    } else {
      int lpp = match_lpp(ctx);
      LIFEB200_DISPATCH_LPP(lpp, (enumerate_kernel<LPP><<<grid_m(n, LPP), kBlockM, 0, ctx->stream>>>(
                                     d_cat, d_cz, d_pr, d_pz, n_projectiles, id_begin, id_end, gens, stable_interval, d_hits,
                                     cap, ctx->d_counters + 0, ctx->d_counters + 1)));
    }
    LIFEB200_LAUNCHED(ctx);
    LIFEB200_CUDA(cudaMemcpyAsync(ctx->h_counters, ctx->d_counters, 4 * sizeof(unsigned long long),
                                  cudaMemcpyDeviceToHost, ctx->stream));
    LIFEB200_CUDA(cudaStreamSynchronize(ctx->stream));
    if (n_skipped) *n_skipped = ctx->h_counters[1];
    return fetch_hits(ctx, d_hits, cap, hits, n_hits, id_begin, id_end - id_begin);
  };
  rc = body();
  return rc;
}

int lifeb200_soup_survey(lifeb200_ctx* ctx, uint64_t first_id, size_t n, uint64_t seed, int dist, unsigned gens,
                         lifeb200_hit* hits, size_t cap, uint64_t* n_hits, uint64_t* pop_sum, uint64_t* hash_xor) {
  LIFEB200_CHECK_CTX(ctx);
  if (gens < 2) return set_error(LIFEB200_E_ARG, "survey needs gens >= 2");
  if (dist != 0 && dist != 1) return set_error(LIFEB200_E_ARG, "dist must be 0 or 1");
  if (n_hits) *n_hits = 0;
  if (pop_sum) *pop_sum = 0;
  if (hash_xor) *hash_xor = 0;
  if (n == 0) return LIFEB200_OK;
  if (!hits && cap) return set_error(LIFEB200_E_ARG, "null hits with non-zero capacity");
  lifeb200_hit* d_hits = nullptr;
  if (cap) {
    if (int rc0 = ensure_scratch(ctx->d_hitbuf, ctx->hitbuf_cap, cap * sizeof(lifeb200_hit))) return rc0;
    d_hits = static_cast<lifeb200_hit*>(ctx->d_hitbuf);
  }
  auto body = [&]() -> int {
    LIFEB200_CUDA(cudaMemsetAsync(ctx->d_counters, 0, 4 * sizeof(unsigned long long), ctx->stream));
    // many generations: 4 soups per lane group in the row-class layout (life_multi.cuh), unless the caller
    // pinned the lane grouping or the layout
    const bool multi = ctx->lpp == 0 && (ctx->step_layout == 8 || (ctx->step_layout == 0 && gens >= kSurveyMultiFrom));
    if (multi) {
      const size_t groups = (n + 3) / 4;
      survey_kernel_x<8><<<(unsigned)((groups * 8 + 127) / 128), 128, 0, ctx->stream>>>(
          first_id, n, seed, dist, gens, d_hits, cap, ctx->d_counters + 0, ctx->d_counters + 2, ctx->d_counters + 3);
    } else {
      int lpp = survey_lpp(ctx);
      LIFEB200_DISPATCH_LPP(lpp, (survey_kernel<LPP><<<grid_m(n, LPP), kBlockM, 0, ctx->stream>>>(
                                     first_id, n, seed, dist, gens, d_hits, cap, ctx->d_counters + 0, ctx->d_counters + 2,
                                     ctx->d_counters + 3)));
    }
    LIFEB200_LAUNCHED(ctx);
    LIFEB200_CUDA(cudaMemcpyAsync(ctx->h_counters, ctx->d_counters, 4 * sizeof(unsigned long long),
                                  cudaMemcpyDeviceToHost, ctx->stream));
    LIFEB200_CUDA(cudaStreamSynchronize(ctx->stream));
    if (pop_sum) *pop_sum = ctx->h_counters[2];
    if (hash_xor) *hash_xor = ctx->h_counters[3];
    return fetch_hits(ctx, d_hits, cap, hits, n_hits, first_id, n);
  };
  return body();
}

int lifeb200_fold_dev(lifeb200_ctx* ctx, const uint64_t* patterns, size_t n, uint64_t first_id, uint64_t* pop_sum,
                      uint64_t* hash_xor) {
  LIFEB200_CHECK_CTX(ctx);
  if (pop_sum) *pop_sum = 0;
  if (hash_xor) *hash_xor = 0;
  if (n == 0) return LIFEB200_OK;
  if (!patterns || !aligned16(patterns)) return set_error(LIFEB200_E_ARG, "bad pattern pointer");
  LIFEB200_CUDA(cudaMemsetAsync(ctx->d_counters, 0, 4 * sizeof(unsigned long long), ctx->stream));
  fold_kernel<8><<<grid_m(n, 8), kBlockM, 0, ctx->stream>>>(patterns, n, first_id, ctx->d_counters + 2,
                                                              ctx->d_counters + 3);
  LIFEB200_LAUNCHED(ctx);
  LIFEB200_CUDA(cudaMemcpyAsync(ctx->h_counters, ctx->d_counters, 4 * sizeof(unsigned long long), cudaMemcpyDeviceToHost,
                                ctx->stream));
  LIFEB200_CUDA(cudaStreamSynchronize(ctx->stream));
  if (pop_sum) *pop_sum = ctx->h_counters[2];
  if (hash_xor) *hash_xor = ctx->h_counters[3];
  return LIFEB200_OK;
}

}  // extern "C"
