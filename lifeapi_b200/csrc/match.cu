// match.cu -- stepping fused with LifeState::Contains(const LifeTarget&) (LifeTarget.hpp:44-51),
// placement construction (Moved/operator|, LifeAPI.hpp:699-717,247-253), the soup survey with the
// period-2 activity test (operator==, LifeAPI.hpp:213-220) and hit compaction.
#include <algorithm>
#include <chrono>
#include <cstdlib>
#include <mutex>
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

// Hit records leave the kernels in ID ORDER without a sort: every id writes one 64-bit word of a dense result
// array (bit 63 = hit, bits 32..62 = b, bits 0..31 = a), and three small launches compact it -- per-tile hit
// counts, an exclusive scan of the tile counts, a scatter in which every thread owns 8 consecutive ids.  The
// order is the id order by construction (deterministic, unlike atomic slot reservation), hits beyond `cap` are
// counted but not written, and the list can stay on the device for the gather (8 bytes per id of scratch: 16 Mi
// soups = 128 MiB; the three launches move 2 x 8 bytes per id, about 10 us per 2 Mi ids).
__device__ __forceinline__ uint64_t dense_record(bool hit, uint32_t a, uint32_t b) {
  return hit ? (1ull << 63) | ((uint64_t)(b & 0x7FFFFFFFu) << 32) | a : 0ull;
}

constexpr int kCompactThreads = 256, kCompactPer = 8, kCompactTile = kCompactThreads * kCompactPer;

__device__ __forceinline__ uint32_t compact_load(const uint64_t* __restrict__ dense, size_t n, size_t base, uint64_t (&v)[kCompactPer]) {
  uint32_t c = 0;
#pragma unroll
  for (int k = 0; k < kCompactPer; k += 2) {
    ulonglong2 w = make_ulonglong2(0, 0);
    if (base + k + 1 < n) w = *reinterpret_cast<const ulonglong2*>(dense + base + k);
    else if (base + k < n) w.x = dense[base + k];
    v[k] = w.x; v[k + 1] = w.y;
    c += (uint32_t)(w.x >> 63) + (uint32_t)(w.y >> 63);
  }
  return c;
}

// exclusive prefix of `c` over the CTA's threads in thread order; *total = the CTA's sum
__device__ __forceinline__ uint32_t block_exclusive(uint32_t c, uint32_t* total) {
  __shared__ uint32_t warp_sum[kCompactThreads / 32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint32_t inc = c;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += t;
  }
  if (lane == 31) warp_sum[warp] = inc;
  __syncthreads();
  uint32_t before = 0, all = 0;
#pragma unroll
  for (int w = 0; w < kCompactThreads / 32; w++) {
    const uint32_t t = warp_sum[w];
    if (w < warp) before += t;
    all += t;
  }
  *total = all;
  return before + inc - c;
}

__global__ void __launch_bounds__(kCompactThreads) compact_count_kernel(const uint64_t* __restrict__ dense, size_t n,
                                                                         unsigned long long* __restrict__ tile_base) {
  uint64_t v[kCompactPer];
  const uint32_t c = compact_load(dense, n, (size_t)blockIdx.x * kCompactTile + threadIdx.x * kCompactPer, v);
  uint32_t total;
  block_exclusive(c, &total);
  if (threadIdx.x == 0) tile_base[blockIdx.x] = total;
}

// one CTA: tile counts -> exclusive bases in place, the grand total -> *n_hits
__global__ void __launch_bounds__(1024) compact_scan_kernel(unsigned long long* __restrict__ tile_base, size_t n_tiles,
                                                            unsigned long long* __restrict__ n_hits) {
  __shared__ unsigned long long warp_sum[32];
  __shared__ unsigned long long carry_s;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) carry_s = 0;
  __syncthreads();
  for (size_t off = 0; off < n_tiles; off += 1024) {
    const size_t i = off + threadIdx.x;
    const unsigned long long c = i < n_tiles ? tile_base[i] : 0ull;
    unsigned long long inc = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const unsigned long long t = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += t;
    }
    if (lane == 31) warp_sum[warp] = inc;
    __syncthreads();
    unsigned long long before = carry_s, all = 0;
    for (int w = 0; w < 32; w++) {
      const unsigned long long t = warp_sum[w];
      if (w < warp) before += t;
      all += t;
    }
    if (i < n_tiles) tile_base[i] = before + inc - c;
    __syncthreads();
    if (threadIdx.x == 0) carry_s += all;
    __syncthreads();
  }
  if (threadIdx.x == 0) *n_hits = carry_s;
}

__global__ void __launch_bounds__(kCompactThreads) compact_scatter_kernel(const uint64_t* __restrict__ dense, size_t n,
                                                                           const unsigned long long* __restrict__ tile_base,
                                                                           uint64_t first_id, lifeb200_hit* __restrict__ hits,
                                                                           size_t cap) {
  uint64_t v[kCompactPer];
  const size_t base = (size_t)blockIdx.x * kCompactTile + threadIdx.x * kCompactPer;
  const uint32_t c = compact_load(dense, n, base, v);
  uint32_t total;
  unsigned long long slot = tile_base[blockIdx.x] + block_exclusive(c, &total);
  if (c == 0) return;
#pragma unroll
  for (int k = 0; k < kCompactPer; k++)
    if (v[k] >> 63) {
      if (slot < cap) {
        lifeb200_hit h;
        h.id = first_id + base + k;
        h.g_first_false = (uint32_t)v[k];
        h.g_recovered = (uint32_t)(v[k] >> 32) & 0x7FFFFFFFu;
        hits[slot] = h;
      }
      slot++;
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
                                                             unsigned gens, unsigned stable_interval,
                                                             uint64_t* __restrict__ dense) {
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
  if (valid && sub == 0) dense[p] = dense_record(r.hit, r.first_false, r.recovered);
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
                                                            unsigned stable_interval, uint64_t* __restrict__ dense,
                                                            unsigned long long* n_skipped) {
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
  if (valid && sub == 0) dense[id - id_begin] = dense_record(r.hit, r.first_false, r.recovered);
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

// The generation loop of the layout-8 match kernels: 4 patterns per group of 8 lanes; `planes` points at this lane's
// column of the group's parked target planes ([32 words][8 lanes] uint4: words 0..15 wanted, 16..31 mask), `mine`
// = which of the four patterns this lane keeps the recovery state machine of.
__device__ __forceinline__ void run_match_x(TileX<8>& t, const uint4* planes, int sub, int mine, unsigned gens,
                                            unsigned stable_interval, Recovery& r) {
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
}

__global__ void __launch_bounds__(128, 1) enumerate_kernel_x(const uint64_t* __restrict__ cat_zoi,
                                                             const uint64_t* __restrict__ proj,
                                                             const uint64_t* __restrict__ proj_zoi,
                                                             const uint32_t* __restrict__ packed_wanted,
                                                             const uint32_t* __restrict__ packed_mask, unsigned n_proj,
                                                             uint64_t id_begin, uint64_t id_end, unsigned gens,
                                                             unsigned stable_interval, uint64_t* __restrict__ dense,
                                                             unsigned long long* n_skipped) {
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
  run_match_x(t, planes, sub, mine, gens, stable_interval, r);
  if (valid && sub < P) dense[id - id_begin] = dense_record(r.hit, r.first_false, r.recovered);
}

// The explicit-pattern match (lifeb200_step_match_*) with 4 patterns per group of 8 lanes: patterns p0 .. p0 + 3 are
// packed bit-interleaved (layout 8) and so are THEIR targets -- every pattern has its own, so the group packs the four
// wanted planes and the four wanted | unwanted planes on the device and parks them in its own 4 KiB slice of shared
// memory.  Same loop as the enumeration: 7.75 + 1.5 integer-pipe instructions per register and generation instead
// of the 8.5 + 1.5 of step_match_kernel (layout 2).
__global__ void __launch_bounds__(128, 1) step_match_kernel_x(const uint64_t* __restrict__ patterns,
                                                              const uint64_t* __restrict__ targets,
                                                              const uint32_t* __restrict__ target_index, size_t n,
                                                              unsigned gens, unsigned stable_interval,
                                                              uint64_t* __restrict__ dense) {
  constexpr int M = 8, P = 4;
  const size_t gid = (size_t)blockIdx.x * 128 + threadIdx.x;
  const size_t p0 = (gid >> 3) * P;
  const int sub = threadIdx.x & 7, mine = sub & 3;
  const size_t p = p0 + mine;
  const bool valid = p < n;
  extern __shared__ uint4 parked[];  // [16 groups][32 words][8 lanes]; words 0..15 wanted, 16..31 mask
  uint4* slice = parked + (threadIdx.x >> 3) * 256 + sub;
  const uint64_t* tg[P];
#pragma unroll
  for (int q = 0; q < P; q++)
    tg[q] = (p0 + q < n) ? targets + (size_t)(target_index ? target_index[p0 + q] : 0u) * 2 * kN : nullptr;
  TileX<M> t;
#pragma unroll
  for (int pl = 0; pl < 3; pl++) {  // 0: wanted, 1: wanted | unwanted, 2: the patterns themselves
#pragma unroll
    for (int c = 0; c < 8; c += 2) {
      uint64_t wa[P], wb[P];
#pragma unroll
      for (int q = 0; q < P; q++) {
        uint4 v = make_uint4(0, 0, 0, 0);
        if (p0 + q < n) {
          if (pl == 2) {
            v = __ldg(reinterpret_cast<const uint4*>(patterns + (p0 + q) * kN + sub * 8 + c));
          } else {
            v = __ldg(reinterpret_cast<const uint4*>(tg[q] + sub * 8 + c));
            if (pl == 1) {
              const uint4 u = __ldg(reinterpret_cast<const uint4*>(tg[q] + kN + sub * 8 + c));
              v.x |= u.x; v.y |= u.y; v.z |= u.z; v.w |= u.w;
            }
          }
        }
        wa[q] = ((uint64_t)v.y << 32) | v.x;
        wb[q] = ((uint64_t)v.w << 32) | v.z;
      }
      t.pack(c, wa);
      t.pack(c + 1, wb);
      if (pl < 2) {
#pragma unroll
        for (int cc = c; cc < c + 2; cc++) {
          slice[(pl * 16 + 2 * cc) * 8] = make_uint4(t.r[cc][0], t.r[cc][1], t.r[cc][2], t.r[cc][3]);
          slice[(pl * 16 + 2 * cc + 1) * 8] = make_uint4(t.r[cc][4], t.r[cc][5], t.r[cc][6], t.r[cc][7]);
        }
      }
    }
  }
  __syncwarp();
  Recovery r;
  if (!valid) r.done = true;
  run_match_x(t, slice, sub, mine, gens, stable_interval, r);
  if (valid && sub < P) dense[p] = dense_record(r.hit, r.first_false, r.recovered);
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
                                                         unsigned gens, uint64_t* __restrict__ dense,
                                                         unsigned long long* pop_sum, unsigned long long* hash_xor) {
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
  if (leader) dense[p] = dense_record(active, pop, 0u);
}

// The same survey with P = M/2 soups per group of 8 lanes in the row-class layout (life_multi.cuh).
template <int M>
__global__ void __launch_bounds__(128, 1) survey_kernel_x(uint64_t first_id, size_t n, uint64_t seed, int dist, unsigned gens,
                                                          uint64_t* __restrict__ dense, unsigned long long* pop_sum,
                                                          unsigned long long* hash_xor) {
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
  if (leader) dense[p0 + sub] = dense_record(my_diff != 0, my_pop, 0u);
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
// measured crossovers against the lane-group kernels (profiles/r02x_multi_thresholds.txt; enumeration: level at 8
// generations, 6 % ahead at 16)
constexpr unsigned kEnumerateMultiFrom = 12, kSurveyMultiFrom = 16, kStepMatchMultiFrom = 8;
static inline int match_lpp(const lifeb200_ctx* ctx) { return std::max(4, effective_lpp(ctx)); }
static inline int survey_lpp(const lifeb200_ctx* ctx) { return std::max(4, effective_lpp(ctx)); }

// dense result scratch of a call over `n` ids (context-owned, grow-only)
static int ensure_dense(lifeb200_ctx* ctx, size_t n) {
  if (int rc = ensure_scratch(ctx->d_dense, ctx->dense_cap, n * sizeof(uint64_t))) return rc;
  const size_t tiles = (n + kCompactTile - 1) / kCompactTile;
  return ensure_scratch(ctx->d_tiles, ctx->tiles_cap, tiles * sizeof(unsigned long long));
}

int reserve_ids(lifeb200_ctx* ctx, size_t n) { return ensure_dense(ctx, n); }

// dense[0..n) -> hits[0..min(found, cap)) in id order, found -> *d_n_hits (device); three launches on the stream
static int compact_hits(lifeb200_ctx* ctx, size_t n, uint64_t first_id, lifeb200_hit* d_hits, size_t cap,
                        unsigned long long* d_n_hits) {
  const uint64_t* dense = static_cast<const uint64_t*>(ctx->d_dense);
  unsigned long long* tiles = static_cast<unsigned long long*>(ctx->d_tiles);
  const size_t n_tiles = (n + kCompactTile - 1) / kCompactTile;
  compact_count_kernel<<<(unsigned)n_tiles, kCompactThreads, 0, ctx->stream>>>(dense, n, tiles);
  LIFEB200_LAUNCHED(ctx);
  compact_scan_kernel<<<1, 1024, 0, ctx->stream>>>(tiles, n_tiles, d_n_hits);
  LIFEB200_LAUNCHED(ctx);
  compact_scatter_kernel<<<(unsigned)n_tiles, kCompactThreads, 0, ctx->stream>>>(dense, n, tiles, first_id, d_hits, cap);
  LIFEB200_LAUNCHED(ctx);
  return LIFEB200_OK;
}

static uint64_t table_key(const uint64_t* a, size_t na, const uint64_t* b, size_t nb) {  // FNV-1a style over the words
  uint64_t h = 0xCBF29CE484222325ull;
  auto mix = [&h](const uint64_t* p, size_t n) {
    for (size_t i = 0; i < n; i++) { h ^= p[i]; h *= 0x100000001B3ull; h ^= h >> 29; }
  };
  mix(a, na); mix(b, nb);
  return h ^ ((uint64_t)na << 32) ^ nb;
}

// Enumeration tables on the device: cat | catZOI | proj | projZOI (64 words each) followed by the layout-8 target
// planes (every cell of wanted = catalyst and of mask = its ZOI replicated into the four pattern bits of register
// (column, k)).  Built and uploaded once per distinct table content: a sharded caller walks one table set over
// many id ranges, and the context keeps the last set keyed by a hash of its words.
static int ensure_tables_body(lifeb200_ctx* ctx, const uint64_t* catalysts, unsigned n_catalysts, const uint64_t* projectiles,
                              unsigned n_projectiles);
static int ensure_tables(lifeb200_ctx* ctx, const uint64_t* catalysts, unsigned n_catalysts, const uint64_t* projectiles,
                         unsigned n_projectiles) {
  return guarded([&]() -> int { return ensure_tables_body(ctx, catalysts, n_catalysts, projectiles, n_projectiles); });
}
static int ensure_tables_body(lifeb200_ctx* ctx, const uint64_t* catalysts, unsigned n_catalysts, const uint64_t* projectiles,
                              unsigned n_projectiles) {
  const uint64_t key = table_key(catalysts, (size_t)n_catalysts * 64, projectiles, (size_t)n_projectiles * 64);
  if (ctx->tab_valid && ctx->tab_key == key && ctx->tab_ncat == n_catalysts && ctx->tab_nproj == n_projectiles)
    return LIFEB200_OK;
  ctx->tab_valid = false;
  const size_t words = (size_t)(n_catalysts + n_projectiles) * 2 * 64;
  const size_t packed_words = (size_t)n_catalysts * 64 * 8 * 2;
  std::vector<uint64_t> tab(words);
  uint64_t* h_cat = tab.data();
  uint64_t* h_cz = h_cat + (size_t)n_catalysts * 64;
  uint64_t* h_pr = h_cz + (size_t)n_catalysts * 64;
  uint64_t* h_pz = h_pr + (size_t)n_projectiles * 64;
  std::copy(catalysts, catalysts + (size_t)n_catalysts * 64, h_cat);
  std::copy(projectiles, projectiles + (size_t)n_projectiles * 64, h_pr);
  for (unsigned c = 0; c < n_catalysts; c++) host_zoi(h_cat + (size_t)c * 64, h_cz + (size_t)c * 64);
  for (unsigned g = 0; g < n_projectiles; g++) host_zoi(h_pr + (size_t)g * 64, h_pz + (size_t)g * 64);
  std::vector<uint32_t> packed(packed_words);
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
  if (int rc = ensure_scratch(ctx->d_tabbuf, ctx->tabbuf_cap, words * 8 + packed_words * 4)) return rc;
  uint64_t* d_tab = static_cast<uint64_t*>(ctx->d_tabbuf);
  // pageable sources: the stream is synchronised before `tab` / `packed` go out of scope
  LIFEB200_CUDA(cudaMemcpyAsync(d_tab, tab.data(), words * 8, cudaMemcpyHostToDevice, ctx->stream));
  LIFEB200_CUDA(cudaMemcpyAsync(d_tab + words, packed.data(), packed_words * 4, cudaMemcpyHostToDevice, ctx->stream));
  LIFEB200_CUDA(cudaStreamSynchronize(ctx->stream));
  ctx->tab_key = key; ctx->tab_ncat = n_catalysts; ctx->tab_nproj = n_projectiles; ctx->tab_valid = true;
  return LIFEB200_OK;
}

int enumerate_to_device(lifeb200_ctx* ctx, const uint64_t* catalysts, unsigned n_catalysts, const uint64_t* projectiles,
                        unsigned n_projectiles, uint64_t id_begin, uint64_t id_end, unsigned gens, unsigned stable_interval,
                        lifeb200_hit* d_hits, size_t cap) {
  const uint64_t n = id_end - id_begin;
  if (int rc = ensure_tables(ctx, catalysts, n_catalysts, projectiles, n_projectiles)) return rc;
  if (int rc = ensure_dense(ctx, n)) return rc;
  uint64_t* dense = static_cast<uint64_t*>(ctx->d_dense);
  const size_t words = (size_t)(n_catalysts + n_projectiles) * 2 * 64;
  const uint64_t* d_cat = static_cast<const uint64_t*>(ctx->d_tabbuf);
  const uint64_t* d_cz = d_cat + (size_t)n_catalysts * 64;
  const uint64_t* d_pr = d_cz + (size_t)n_catalysts * 64;
  const uint64_t* d_pz = d_pr + (size_t)n_projectiles * 64;
  const uint32_t* d_packed = reinterpret_cast<const uint32_t*>(d_cat + words);
  LIFEB200_CUDA(cudaMemsetAsync(ctx->d_counters, 0, 4 * sizeof(unsigned long long), ctx->stream));
  // many generations: 4 placements per lane group in the row-class layout (enumerate_kernel_x), unless the
  // caller pinned the lane grouping or another layout
  const bool multi = ctx->lpp == 0 && (ctx->step_layout == 8 || (ctx->step_layout == 0 && gens >= kEnumerateMultiFrom));
  if (multi) {
    const uint64_t groups = (id_end - (id_begin & ~3ull) + 3) / 4;
    LIFEB200_CUDA(cudaFuncSetAttribute(enumerate_kernel_x, cudaFuncAttributeMaxDynamicSharedMemorySize, kEnumXSmem));
    enumerate_kernel_x<<<(unsigned)((groups * 8 + 127) / 128), 128, kEnumXSmem, ctx->stream>>>(
        d_cz, d_pr, d_pz, d_packed, d_packed + (size_t)n_catalysts * 64 * 8, n_projectiles, id_begin, id_end, gens,
        stable_interval, dense, ctx->d_counters + 1);
  } else {
    int lpp = match_lpp(ctx);
    LIFEB200_DISPATCH_LPP_MIN4(lpp, (enumerate_kernel<LPP><<<grid_m(n, LPP), kBlockM, 0, ctx->stream>>>(
                                        d_cat, d_cz, d_pr, d_pz, n_projectiles, id_begin, id_end, gens, stable_interval, dense,
                                        ctx->d_counters + 1)));
  }
  LIFEB200_LAUNCHED(ctx);
  if (int rc = compact_hits(ctx, n, id_begin, d_hits, cap, ctx->d_counters + 0)) return rc;
  LIFEB200_CUDA(cudaMemcpyAsync(ctx->h_counters, ctx->d_counters, 4 * sizeof(unsigned long long), cudaMemcpyDeviceToHost,
                                ctx->stream));
  LIFEB200_CUDA(cudaStreamSynchronize(ctx->stream));
  return LIFEB200_OK;
}

int survey_to_device(lifeb200_ctx* ctx, uint64_t first_id, size_t n, uint64_t seed, int dist, unsigned gens,
                     lifeb200_hit* d_hits, size_t cap) {
  static const bool trace = std::getenv("LIFEB200_TRACE_GATHER") != nullptr;
  const auto t0 = std::chrono::steady_clock::now();
  auto stamp = [&](const char* what) {
    if (trace)
      std::fprintf(stderr, "[lifeb200 survey dev %d] %-20s +%.3f ms\n", ctx->device, what,
                   std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count());
  };
  if (int rc = ensure_dense(ctx, n)) return rc;
  stamp("scratch ready");
  uint64_t* dense = static_cast<uint64_t*>(ctx->d_dense);
  LIFEB200_CUDA(cudaMemsetAsync(ctx->d_counters, 0, 4 * sizeof(unsigned long long), ctx->stream));
  // many generations: 4 soups per lane group in the row-class layout (life_multi.cuh), unless the caller
  // pinned the lane grouping or the layout
  const bool multi = ctx->lpp == 0 && (ctx->step_layout == 8 || (ctx->step_layout == 0 && gens >= kSurveyMultiFrom));
  if (multi) {
    const size_t groups = (n + 3) / 4;
    survey_kernel_x<8><<<(unsigned)((groups * 8 + 127) / 128), 128, 0, ctx->stream>>>(
        first_id, n, seed, dist, gens, dense, ctx->d_counters + 2, ctx->d_counters + 3);
  } else {
    int lpp = survey_lpp(ctx);
    LIFEB200_DISPATCH_LPP_MIN4(lpp, (survey_kernel<LPP><<<grid_m(n, LPP), kBlockM, 0, ctx->stream>>>(
                                        first_id, n, seed, dist, gens, dense, ctx->d_counters + 2, ctx->d_counters + 3)));
  }
  LIFEB200_LAUNCHED(ctx);
  stamp("kernel launched");
  if (int rc = compact_hits(ctx, n, first_id, d_hits, cap, ctx->d_counters + 0)) return rc;
  LIFEB200_CUDA(cudaMemcpyAsync(ctx->h_counters, ctx->d_counters, 4 * sizeof(unsigned long long), cudaMemcpyDeviceToHost,
                                ctx->stream));
  stamp("all enqueued");
  LIFEB200_CUDA(cudaStreamSynchronize(ctx->stream));
  stamp("stream drained");
  return LIFEB200_OK;
}

// hits already sorted on the device: report the count, download what fits
static int download_hits(lifeb200_ctx* ctx, const lifeb200_hit* d_hits, size_t cap, lifeb200_hit* hits, uint64_t* n_hits) {
  const uint64_t found = ctx->h_counters[0];
  if (n_hits) *n_hits = found;
  const uint64_t k = std::min<uint64_t>(found, cap);
  if (k && hits) {
    LIFEB200_CUDA(cudaMemcpyAsync(hits, d_hits, k * sizeof(lifeb200_hit), cudaMemcpyDeviceToHost, ctx->stream));
    LIFEB200_CUDA(cudaStreamSynchronize(ctx->stream));
  }
  if (found > cap) return set_error(LIFEB200_E_OVERFLOW, "hit buffer too small: %llu hits, capacity %llu",
                                    (unsigned long long)found, (unsigned long long)cap);
  return LIFEB200_OK;
}

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
  if (int rc = ensure_dense(ctx, n)) return rc;
  uint64_t* dense = static_cast<uint64_t*>(ctx->d_dense);
  // many generations: 4 patterns per lane group in the row-class layout (step_match_kernel_x), unless the caller
  // pinned the lane grouping or another layout; measured 14-18 % faster from 8 generations up (profiles/r02_thresholds.txt)
  const bool multi = ctx->lpp == 0 && (ctx->step_layout == 8 || (ctx->step_layout == 0 && gens >= kStepMatchMultiFrom));
  if (multi) {
    const size_t groups = (n + 3) / 4;
    LIFEB200_CUDA(cudaFuncSetAttribute(step_match_kernel_x, cudaFuncAttributeMaxDynamicSharedMemorySize, kEnumXSmem));
    step_match_kernel_x<<<(unsigned)((groups * 8 + 127) / 128), 128, kEnumXSmem, ctx->stream>>>(
        patterns, targets, target_index, n, gens, stable_interval, dense);
  } else {
    int lpp = match_lpp(ctx);
    LIFEB200_DISPATCH_LPP_MIN4(lpp, (step_match_kernel<LPP><<<grid_m(n, LPP), kBlockM, 0, ctx->stream>>>(
                                        patterns, targets, target_index, n, gens, stable_interval, dense)));
  }
  LIFEB200_LAUNCHED(ctx);
  return compact_hits(ctx, n, first_id, hits, cap, n_hits);
}

// Host form of the call above: patterns, targets, the optional index and the hit records are HOST arrays; hits come
// back in id order.  One upload, one launch, one compaction, one download of the records that were found.
int lifeb200_step_match_host(lifeb200_ctx* ctx, const uint64_t* patterns, const uint64_t* targets,
                             const uint32_t* target_index, size_t n, size_t n_targets, uint64_t first_id, unsigned gens,
                             unsigned stable_interval, lifeb200_hit* hits, size_t cap, uint64_t* n_hits) {
  LIFEB200_CHECK_CTX(ctx);
  if (n_hits) *n_hits = 0;
  if (stable_interval == 0) return set_error(LIFEB200_E_ARG, "stable_interval must be >= 1");
  if (n == 0) return LIFEB200_OK;
  if (!patterns || !targets || n_targets == 0 || (!hits && cap)) return set_error(LIFEB200_E_ARG, "null pointer");
  const size_t pat_b = n * LIFEB200_PATTERN_BYTES, tgt_b = n_targets * 2 * LIFEB200_PATTERN_BYTES;
  const size_t idx_b = target_index ? ((n * sizeof(uint32_t) + 15) & ~(size_t)15) : 0;
  if (int rc = ensure_slot(ctx, 0, pat_b + tgt_b + idx_b, 0)) return rc;
  if (int rc = ensure_scratch(ctx->d_hitbuf, ctx->hitbuf_cap, std::max<size_t>(cap, 1) * sizeof(lifeb200_hit))) return rc;
  char* d = static_cast<char*>(ctx->slots[0].dbuf);
  LIFEB200_CUDA(cudaMemcpyAsync(d, patterns, pat_b, cudaMemcpyHostToDevice, ctx->stream));
  LIFEB200_CUDA(cudaMemcpyAsync(d + pat_b, targets, tgt_b, cudaMemcpyHostToDevice, ctx->stream));
  if (target_index)
    LIFEB200_CUDA(cudaMemcpyAsync(d + pat_b + tgt_b, target_index, n * sizeof(uint32_t), cudaMemcpyHostToDevice, ctx->stream));
  lifeb200_hit* d_hits = static_cast<lifeb200_hit*>(ctx->d_hitbuf);
  int rc = lifeb200_step_match_dev(ctx, reinterpret_cast<const uint64_t*>(d), reinterpret_cast<const uint64_t*>(d + pat_b),
                                   target_index ? reinterpret_cast<const uint32_t*>(d + pat_b + tgt_b) : nullptr, n, first_id,
                                   gens, stable_interval, d_hits, cap, ctx->d_counters + 0);
  if (rc) { cudaStreamSynchronize(ctx->stream); return rc; }
  LIFEB200_CUDA(cudaMemcpyAsync(ctx->h_counters, ctx->d_counters, sizeof(unsigned long long), cudaMemcpyDeviceToHost, ctx->stream));
  LIFEB200_CUDA(cudaStreamSynchronize(ctx->stream));
  return download_hits(ctx, d_hits, cap, hits, n_hits);
}

// ---- allocation-free variant of lifeb200_step_match_dev (no counterpart in the reference; added
// for callers that cannot call cudaMalloc: code under CUDA graph capture, tight serving loops) ---
// The dense per-id results and the per-tile counts of the ordered compaction live in a file-scope __device__ array
// instead of the context's grow-only cudaMalloc scratch.  Every GPU that loads the library holds the array
// (2 MiB + 1 KiB) until exit.  In host code the array's NAME is a host-side stand-in, not a device address: the
// address is taken once per device, with that device current, by cudaGetSymbolAddress below.
constexpr size_t kStaticMaxIds = LIFEB200_STEP_MATCH_STATIC_MAX;
constexpr size_t kStaticMaxTiles = (kStaticMaxIds + kCompactTile - 1) / kCompactTile;
static __device__ __align__(256) unsigned long long g_static_match_scratch[kStaticMaxIds + kStaticMaxTiles];

constexpr int kStaticMaxDevices = 64;
static std::mutex g_static_mu;
static unsigned long long* g_static_addr[kStaticMaxDevices];  // per device, nullptr until that device's first call

static int static_scratch_for(const lifeb200_ctx* ctx, unsigned long long** out) {
  if (ctx->device < 0 || ctx->device >= kStaticMaxDevices)
    return set_error(LIFEB200_E_ARG, "static scratch: device ordinal %d outside 0..%d", ctx->device, kStaticMaxDevices - 1);
  std::lock_guard<std::mutex> lock(g_static_mu);
  if (!g_static_addr[ctx->device]) {
    void* p = nullptr;  // the context's device is current here (LIFEB200_CHECK_CTX)
    LIFEB200_CUDA(cudaGetSymbolAddress(&p, g_static_match_scratch));
    if (!p) return set_error(LIFEB200_E_CUDA, "cudaGetSymbolAddress returned a null address");
    g_static_addr[ctx->device] = static_cast<unsigned long long*>(p);
  }
  *out = g_static_addr[ctx->device];
  return LIFEB200_OK;
}

int lifeb200_step_match_static_dev(lifeb200_ctx* ctx, const uint64_t* patterns, const uint64_t* targets,
                                   const uint32_t* target_index, size_t n, uint64_t first_id, unsigned gens,
                                   unsigned stable_interval, lifeb200_hit* hits, size_t cap, unsigned long long* n_hits) {
  LIFEB200_CHECK_CTX(ctx);
  if (!n_hits) return set_error(LIFEB200_E_ARG, "null n_hits");
  if (stable_interval == 0) return set_error(LIFEB200_E_ARG, "stable_interval must be >= 1");
  if (n > kStaticMaxIds)
    return set_error(LIFEB200_E_ARG, "lifeb200_step_match_static_dev: n = %llu patterns exceeds the static scratch "
                     "(LIFEB200_STEP_MATCH_STATIC_MAX = %llu); split the batch or use lifeb200_step_match_dev",
                     (unsigned long long)n, (unsigned long long)kStaticMaxIds);
  LIFEB200_CUDA(cudaMemsetAsync(n_hits, 0, sizeof(unsigned long long), ctx->stream));
  if (n == 0) return LIFEB200_OK;
  if (!patterns || !targets || (!hits && cap)) return set_error(LIFEB200_E_ARG, "null pointer");
  if (!aligned16(patterns) || !aligned16(targets)) return set_error(LIFEB200_E_ARG, "pointers must be 16-byte aligned");
  unsigned long long* scratch = nullptr;
  if (int rc = static_scratch_for(ctx, &scratch)) return rc;
  // nothing here must start at zero: every id in [0, n) writes its dense word, every tile its count, and the scan
  // writes *n_hits; words beyond n keep the previous call's contents and are never read
  uint64_t* dense = reinterpret_cast<uint64_t*>(scratch);
  unsigned long long* tiles = scratch + kStaticMaxIds;
  const bool multi = ctx->lpp == 0 && (ctx->step_layout == 8 || (ctx->step_layout == 0 && gens >= kStepMatchMultiFrom));
  if (multi) {
    const size_t groups = (n + 3) / 4;
    LIFEB200_CUDA(cudaFuncSetAttribute(step_match_kernel_x, cudaFuncAttributeMaxDynamicSharedMemorySize, kEnumXSmem));
    step_match_kernel_x<<<(unsigned)((groups * 8 + 127) / 128), 128, kEnumXSmem, ctx->stream>>>(
        patterns, targets, target_index, n, gens, stable_interval, dense);
  } else {
    int lpp = match_lpp(ctx);
    LIFEB200_DISPATCH_LPP_MIN4(lpp, (step_match_kernel<LPP><<<grid_m(n, LPP), kBlockM, 0, ctx->stream>>>(
                                        patterns, targets, target_index, n, gens, stable_interval, dense)));
  }
  LIFEB200_LAUNCHED(ctx);
  const size_t n_tiles = (n + kCompactTile - 1) / kCompactTile;
  compact_count_kernel<<<(unsigned)n_tiles, kCompactThreads, 0, ctx->stream>>>(dense, n, tiles);
  LIFEB200_LAUNCHED(ctx);
  compact_scan_kernel<<<1, 1024, 0, ctx->stream>>>(tiles, n_tiles, n_hits);
  LIFEB200_LAUNCHED(ctx);
  compact_scatter_kernel<<<(unsigned)n_tiles, kCompactThreads, 0, ctx->stream>>>(dense, n, tiles, first_id, hits, cap);
  LIFEB200_LAUNCHED(ctx);
  return LIFEB200_OK;
}

static int enumerate_args(const uint64_t* catalysts, unsigned n_catalysts, const uint64_t* projectiles,
                          unsigned n_projectiles, uint64_t id_begin, uint64_t* id_end, unsigned stable_interval,
                          const lifeb200_hit* hits, size_t cap, uint64_t* n_hits, uint64_t* n_skipped, bool* empty) {
  if (!catalysts || !projectiles || n_catalysts == 0 || n_projectiles == 0)
    return set_error(LIFEB200_E_ARG, "empty catalyst / projectile table");
  if (stable_interval == 0) return set_error(LIFEB200_E_ARG, "stable_interval must be >= 1");
  const uint64_t total = (uint64_t)n_catalysts * n_projectiles * 4096ull;
  if (*id_end > total) *id_end = total;
  if (n_hits) *n_hits = 0;
  if (n_skipped) *n_skipped = 0;
  *empty = id_begin >= *id_end;
  if (!*empty && !hits && cap) return set_error(LIFEB200_E_ARG, "null hits with non-zero capacity");
  return LIFEB200_OK;
}

int lifeb200_enumerate_collisions_dev(lifeb200_ctx* ctx, const uint64_t* catalysts, unsigned n_catalysts,
                                      const uint64_t* projectiles, unsigned n_projectiles, uint64_t id_begin,
                                      uint64_t id_end, unsigned gens, unsigned stable_interval, lifeb200_hit* d_hits,
                                      size_t cap, uint64_t* n_hits, uint64_t* n_skipped) {
  LIFEB200_CHECK_CTX(ctx);
  bool empty = false;
  if (int rc = enumerate_args(catalysts, n_catalysts, projectiles, n_projectiles, id_begin, &id_end, stable_interval,
                              d_hits, cap, n_hits, n_skipped, &empty)) return rc;
  if (empty) return LIFEB200_OK;
  if (int rc = enumerate_to_device(ctx, catalysts, n_catalysts, projectiles, n_projectiles, id_begin, id_end, gens,
                                   stable_interval, d_hits, cap)) return rc;
  if (n_skipped) *n_skipped = ctx->h_counters[1];
  return download_hits(ctx, d_hits, cap, nullptr, n_hits);
}

int lifeb200_enumerate_collisions(lifeb200_ctx* ctx, const uint64_t* catalysts, unsigned n_catalysts,
                                  const uint64_t* projectiles, unsigned n_projectiles, uint64_t id_begin,
                                  uint64_t id_end, unsigned gens, unsigned stable_interval, lifeb200_hit* hits,
                                  size_t cap, uint64_t* n_hits, uint64_t* n_skipped) {
  LIFEB200_CHECK_CTX(ctx);
  bool empty = false;
  if (int rc = enumerate_args(catalysts, n_catalysts, projectiles, n_projectiles, id_begin, &id_end, stable_interval,
                              hits, cap, n_hits, n_skipped, &empty)) return rc;
  if (empty) return LIFEB200_OK;
  // context-owned scratch (ctx.h): no allocation after the first call of a given size
  if (int rc = ensure_scratch(ctx->d_hitbuf, ctx->hitbuf_cap, std::max<size_t>(cap, 1) * sizeof(lifeb200_hit))) return rc;
  lifeb200_hit* d_hits = static_cast<lifeb200_hit*>(ctx->d_hitbuf);
  if (int rc = enumerate_to_device(ctx, catalysts, n_catalysts, projectiles, n_projectiles, id_begin, id_end, gens,
                                   stable_interval, d_hits, cap)) return rc;
  if (n_skipped) *n_skipped = ctx->h_counters[1];
  return download_hits(ctx, d_hits, cap, hits, n_hits);
}

static int survey_args(uint64_t* n_hits, uint64_t* pop_sum, uint64_t* hash_xor, int dist, unsigned gens) {
  if (gens < 2) return set_error(LIFEB200_E_ARG, "survey needs gens >= 2");
  if (dist != 0 && dist != 1) return set_error(LIFEB200_E_ARG, "dist must be 0 or 1");
  if (n_hits) *n_hits = 0;
  if (pop_sum) *pop_sum = 0;
  if (hash_xor) *hash_xor = 0;
  return LIFEB200_OK;
}

int lifeb200_soup_survey_dev(lifeb200_ctx* ctx, uint64_t first_id, size_t n, uint64_t seed, int dist, unsigned gens,
                             lifeb200_hit* d_hits, size_t cap, uint64_t* n_hits, uint64_t* pop_sum, uint64_t* hash_xor) {
  LIFEB200_CHECK_CTX(ctx);
  if (int rc = survey_args(n_hits, pop_sum, hash_xor, dist, gens)) return rc;
  if (n == 0) return LIFEB200_OK;
  if (!d_hits && cap) return set_error(LIFEB200_E_ARG, "null hits with non-zero capacity");
  if (int rc = survey_to_device(ctx, first_id, n, seed, dist, gens, d_hits, cap)) return rc;
  if (pop_sum) *pop_sum = ctx->h_counters[2];
  if (hash_xor) *hash_xor = ctx->h_counters[3];
  return download_hits(ctx, d_hits, cap, nullptr, n_hits);
}

int lifeb200_soup_survey(lifeb200_ctx* ctx, uint64_t first_id, size_t n, uint64_t seed, int dist, unsigned gens,
                         lifeb200_hit* hits, size_t cap, uint64_t* n_hits, uint64_t* pop_sum, uint64_t* hash_xor) {
  LIFEB200_CHECK_CTX(ctx);
  if (int rc = survey_args(n_hits, pop_sum, hash_xor, dist, gens)) return rc;
  if (n == 0) return LIFEB200_OK;
  if (!hits && cap) return set_error(LIFEB200_E_ARG, "null hits with non-zero capacity");
  if (int rc = ensure_scratch(ctx->d_hitbuf, ctx->hitbuf_cap, std::max<size_t>(cap, 1) * sizeof(lifeb200_hit))) return rc;
  lifeb200_hit* d_hits = static_cast<lifeb200_hit*>(ctx->d_hitbuf);
  if (int rc = survey_to_device(ctx, first_id, n, seed, dist, gens, d_hits, cap)) return rc;
  if (pop_sum) *pop_sum = ctx->h_counters[2];
  if (hash_xor) *hash_xor = ctx->h_counters[3];
  return download_hits(ctx, d_hits, cap, hits, n_hits);
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
