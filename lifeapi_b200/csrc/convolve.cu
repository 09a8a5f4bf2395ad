// convolve.cu -- the step AFTER the stepping path: locating patterns on the torus.
//   LifeState::Convolve            LifeAPI.hpp:1293-1370 (+ Bits.hpp:132-143)
//   MatchLive / MatchLiveAndDead   LifeAPI.hpp:427-435;  Match(LifeState) :440-442;  Match(LifeTarget) LifeTarget.hpp:53-54
//   LifeState::InteractionOffsets  LifeAPI.hpp:1066-1095
//   LifeState::Mirrored            LifeAPI.hpp:789-795
// One warp owns one problem: lane L keeps words 2L and 2L+1 of every plane.  Convolve is the torus
// Minkowski sum  result = OR over ON cells (j,k) of X of Y.Moved(j,k); it is commutative, so the warp
// walks whichever operand has fewer ON cells.  The walk is warp-uniform (the walked word is broadcast
// with a shuffle, every lane takes the same branches): empty columns are skipped with one ballot,
// and a run of r consecutive ON bits costs O(log r) rotate-ORs (the doubling the reference applies to
// runs, LifeAPI.hpp:1309-1366) instead of r.  The shifted operand's column j arrives by shuffle from
// lane L - j/2, so nothing touches shared or global memory between the load and the store.
#include <algorithm>

#include "ctx.h"
#include "host_pipeline.cuh"
#include "life_core.cuh"

namespace lifeb200 {

constexpr int kBlockC = 256;
constexpr unsigned kFull = 0xffffffffu;
static inline unsigned grid_warps(size_t n) { return (unsigned)((n * 32 + kBlockC - 1) / kBlockC); }

struct Plane {
  uint64_t w0, w1;  // words 2L, 2L+1 of this lane
};

// 64-bit rotate as two 32-bit funnel shifts (SHF.L.W) after a half swap for s >= 32: 2 SHF + 2 SEL, where
// (v << s) | (v >> (64 - s)) on a 64-bit value compiles to two variable 64-bit shifts and two ORs
__device__ __forceinline__ uint64_t rotl64(uint64_t v, int s) {
  uint32_t lo = (uint32_t)v, hi = (uint32_t)(v >> 32);
  if (s & 32) { const uint32_t t = lo; lo = hi; hi = t; }
  const uint32_t nlo = __funnelshift_l(hi, lo, s), nhi = __funnelshift_l(lo, hi, s);  // shift taken mod 32
  return ((uint64_t)nhi << 32) | nlo;
}
__device__ __forceinline__ uint64_t shfl64(uint64_t v, int src) { return __shfl_sync(kFull, v, src & 31); }

__device__ __forceinline__ Plane load_plane(const uint64_t* __restrict__ p, int lane) {
  const uint4 v = __ldg(reinterpret_cast<const uint4*>(p) + lane);
  Plane r;
  r.w0 = ((uint64_t)v.y << 32) | v.x;
  r.w1 = ((uint64_t)v.w << 32) | v.z;
  return r;
}
__device__ __forceinline__ void store_plane(uint64_t* __restrict__ p, int lane, const Plane& r) {
  reinterpret_cast<uint4*>(p)[lane] = make_uint4((uint32_t)r.w0, (uint32_t)(r.w0 >> 32), (uint32_t)r.w1, (uint32_t)(r.w1 >> 32));
}
__device__ __forceinline__ Plane zero_plane() { Plane r; r.w0 = r.w1 = 0; return r; }
__device__ __forceinline__ Plane operator~(const Plane& a) { Plane r; r.w0 = ~a.w0; r.w1 = ~a.w1; return r; }
__device__ __forceinline__ Plane operator&(const Plane& a, const Plane& b) { Plane r; r.w0 = a.w0 & b.w0; r.w1 = a.w1 & b.w1; return r; }
__device__ __forceinline__ Plane operator|(const Plane& a, const Plane& b) { Plane r; r.w0 = a.w0 | b.w0; r.w1 = a.w1 | b.w1; return r; }

__device__ __forceinline__ unsigned warp_pop(const Plane& a) {
  return __reduce_add_sync(kFull, (unsigned)(__popcll(a.w0) + __popcll(a.w1)));  // one REDUX instead of five shuffles
}

// (x, y) -> (-x, -y): word i = rotl(brev(s[(64 - i) & 63]), 1)   (FlipX, FlipY, Move(1,1))
__device__ __forceinline__ Plane mirrored(const Plane& s, int lane) {
  Plane r;
  r.w0 = rotl64(__brevll(shfl64(s.w0, 32 - lane)), 1);  // word (64 - 2L) & 63 = word 0 of lane (32 - L) & 31
  r.w1 = rotl64(__brevll(shfl64(s.w1, 31 - lane)), 1);  // word 63 - 2L = word 1 of lane 31 - L
  return r;
}

// acc |= OR over ON bits k of x of rotl(y, k), for this lane's two result words (x is warp-uniform)
// Sparse columns (the usual case: the walked operand is a catalyst, a target or an 8x8 object) go bit by bit on
// 32-bit halves: find-first-set, clear, and a rotate that is two funnel shifts because the half says whether the
// words' halves swap -- about 14 instructions per ON cell, where the run logic on 64-bit values costs about 50 per
// run.  Denser columns keep the doubling, which is what bounds the cost of solid blobs.
constexpr int kBitwiseMax = 12;
__device__ __forceinline__ void spread(uint64_t x, uint64_t y0, uint64_t y1, Plane& acc) {
  if (__popcll(x) <= kBitwiseMax) {  // warp-uniform
    const uint32_t y0l = (uint32_t)y0, y0h = (uint32_t)(y0 >> 32), y1l = (uint32_t)y1, y1h = (uint32_t)(y1 >> 32);
    uint32_t a0l = 0, a0h = 0, a1l = 0, a1h = 0;
    for (uint32_t xl = (uint32_t)x; xl; xl &= xl - 1) {  // rotl by k < 32
      const int k = __ffs(xl) - 1;
      a0l |= __funnelshift_l(y0h, y0l, k); a0h |= __funnelshift_l(y0l, y0h, k);
      a1l |= __funnelshift_l(y1h, y1l, k); a1h |= __funnelshift_l(y1l, y1h, k);
    }
    for (uint32_t xh = (uint32_t)(x >> 32); xh; xh &= xh - 1) {  // rotl by 32 + k: halves swapped
      const int k = __ffs(xh) - 1;
      a0l |= __funnelshift_l(y0l, y0h, k); a0h |= __funnelshift_l(y0h, y0l, k);
      a1l |= __funnelshift_l(y1l, y1h, k); a1h |= __funnelshift_l(y1h, y1l, k);
    }
    acc.w0 |= ((uint64_t)a0h << 32) | a0l;
    acc.w1 |= ((uint64_t)a1h << 32) | a1l;
    return;
  }
  while (x) {
    const int k = __ffsll((long long)x) - 1;
    const uint64_t sh = x >> k;
    const int r = (~sh == 0) ? 64 : __ffsll((long long)~sh) - 1;  // run length starting at bit k
    uint64_t t0 = y0, t1 = y1;
    int s = 1;
    while (2 * s <= r) { t0 |= rotl64(t0, s); t1 |= rotl64(t1, s); s *= 2; }
    if (r > s) { t0 |= rotl64(t0, r - s); t1 |= rotl64(t1, r - s); }
    acc.w0 |= rotl64(t0, k);
    acc.w1 |= rotl64(t1, k);
    x = (r == 64) ? 0 : x & ~(((1ULL << r) - 1) << k);
  }
}

// acc |= X (*) Y, walking X
__device__ __forceinline__ void conv_walk(const Plane& x, const Plane& y, Plane& acc, int lane) {
  unsigned nz0 = __ballot_sync(kFull, x.w0 != 0);  // columns 2*lane
  unsigned nz1 = __ballot_sync(kFull, x.w1 != 0);  // columns 2*lane + 1
  while (nz0) {
    const int jj = __ffs(nz0) - 1;
    nz0 &= nz0 - 1;
    const uint64_t xa = shfl64(x.w0, jj);  // column j = 2jj: result words 2L, 2L+1 take y[2L-2jj], y[2L+1-2jj]
    spread(xa, shfl64(y.w0, lane - jj), shfl64(y.w1, lane - jj), acc);
  }
  while (nz1) {
    const int jj = __ffs(nz1) - 1;
    nz1 &= nz1 - 1;
    const uint64_t xb = shfl64(x.w1, jj);  // column j = 2jj+1: y[2L-2jj-1], y[2L-2jj]
    spread(xb, shfl64(y.w1, lane - jj - 1), shfl64(y.w0, lane - jj), acc);
  }
}

__device__ __forceinline__ Plane convolve(const Plane& a, const Plane& b, int lane) {
  Plane acc = zero_plane();
  if (warp_pop(a) <= warp_pop(b)) conv_walk(a, b, acc, lane);
  else conv_walk(b, a, acc, lane);
  return acc;
}

__device__ __forceinline__ Plane to_plane(const Tile<32>& t) {
  Plane r;
  r.w0 = ((uint64_t)t.hi[0] << 32) | t.lo[0];
  r.w1 = ((uint64_t)t.hi[1] << 32) | t.lo[1];
  return r;
}
__device__ __forceinline__ Tile<32> to_tile(const Plane& p) {
  Tile<32> t;
  t.lo[0] = (uint32_t)p.w0; t.hi[0] = (uint32_t)(p.w0 >> 32);
  t.lo[1] = (uint32_t)p.w1; t.hi[1] = (uint32_t)(p.w1 >> 32);
  return t;
}

__global__ void __launch_bounds__(kBlockC) convolve_kernel(const uint64_t* __restrict__ a, size_t n_a,
                                                           const uint64_t* __restrict__ b, size_t n_b,
                                                           uint64_t* __restrict__ out, size_t n) {
  const size_t p = ((size_t)blockIdx.x * kBlockC + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (p >= n) return;  // whole warps leave together
  const Plane A = load_plane(a + (n_a == 1 ? 0 : p) * kN, lane);
  const Plane B = load_plane(b + (n_b == 1 ? 0 : p) * kN, lane);
  store_plane(out + p * kN, lane, convolve(A, B, lane));
}

// mode 0: live + dead planes given; 1: live only (MatchLive); 2: dead = live.GetBoundary() (Match(LifeState))
__global__ void __launch_bounds__(kBlockC) match_kernel(const uint64_t* __restrict__ states, size_t n,
                                                        const uint64_t* __restrict__ live, const uint64_t* __restrict__ dead,
                                                        size_t n_t, size_t t_stride, int mode, uint64_t* __restrict__ out) {
  const size_t p = ((size_t)blockIdx.x * kBlockC + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (p >= n) return;
  const size_t ti = (n_t == 1 ? 0 : p) * t_stride;
  const Plane S = load_plane(states + p * kN, lane);
  const Plane L = load_plane(live + ti, lane);
  // ~((~this).Convolve(live.Mirrored()))
  Plane r = ~convolve(~S, mirrored(L, lane), lane);
  if (mode != 1) {
    Plane D;
    if (mode == 0) D = load_plane(dead + ti, lane);
    else D = to_plane(dilate<32>(to_tile(L), lane, false)) & ~L;  // ZOI() & ~live (LifeAPI.hpp:538)
    r = r & ~convolve(S, mirrored(D, lane), lane);
  }
  store_plane(out + p * kN, lane, r);
}

__global__ void __launch_bounds__(kBlockC) interaction_offsets_kernel(const uint64_t* __restrict__ a, size_t n_a,
                                                                      const uint64_t* __restrict__ b, size_t n_b,
                                                                      uint64_t* __restrict__ out, size_t n) {
  const size_t p = ((size_t)blockIdx.x * kBlockC + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (p >= n) return;
  const Plane A = load_plane(a + (n_a == 1 ? 0 : p) * kN, lane);
  const Plane B = mirrored(load_plane(b + (n_b == 1 ? 0 : p) * kN, lane), lane);
  Tile<32> t3, t2, t1, t0;
  to_tile(A).neighbour_count(lane, t3, t2, t1, t0);
  const Plane a3 = to_plane(t3), a2 = to_plane(t2), a1 = to_plane(t1), a0 = to_plane(t0);
  to_tile(B).neighbour_count(lane, t3, t2, t1, t0);
  const Plane b3 = to_plane(t3), b2 = to_plane(t2), b1 = to_plane(t1), b0 = to_plane(t0);
  const Plane a_lo = ~a3 & ~a2, b_lo = ~b3 & ~b2;  // count < 4
  Plane r = convolve(A, B, lane);                                                  // overlaps
  r = r | convolve(a_lo & ~a1 & a0 & ~A, b_lo & b1 & ~b0 & ~B, lane);              // births: 1 next to 2
  r = r | convolve(b_lo & ~b1 & b0 & ~B, a_lo & a1 & ~a0 & ~A, lane);
  r = r | convolve(a_lo & a1 & a0 & A, (b3 | b2 | b1) & ~B, lane);                 // overcrowding
  r = r | convolve((a2 | a3) & A, (b3 | b2 | b1 | b0) & ~B, lane);
  r = r | convolve(b_lo & b1 & b0 & B, (a3 | a2 | a1) & ~A, lane);
  r = r | convolve((b2 | b3) & B, (a3 | a2 | a1 | a0) & ~A, lane);
  store_plane(out + p * kN, lane, r);
}

// LifeWeld::InteractionOffsets (LifeWeld.hpp:217-262): InteractionOffsets of the two states with the interactions that
// involve frozen cells left out -- every operand is masked by the zone of influence of its weld's NON-frozen cells.
// The expressions follow the reference member term by term, including which `ignored` mask each operand carries
// (its second and third terms mask the OTHER weld's operand with the first weld's zone: kept, parity is bit-exact).
// Records are LifeWeld layouts: state, frozen2, frozen1, frozen0 (4 x 64 words).
__global__ void __launch_bounds__(kBlockC, 2) weld_interaction_offsets_kernel(const uint64_t* __restrict__ a, size_t n_a,
                                                                           const uint64_t* __restrict__ b, size_t n_b,
                                                                           uint64_t* __restrict__ out, size_t n) {
  const size_t p = ((size_t)blockIdx.x * kBlockC + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (p >= n) return;
  const uint64_t* ra = a + (n_a == 1 ? 0 : p) * 4 * kN;
  const uint64_t* rb = b + (n_b == 1 ? 0 : p) * 4 * kN;
  const Plane A = load_plane(ra, lane);
  const Plane a_free = A & ~load_plane(ra + kN, lane) & ~load_plane(ra + 2 * kN, lane) & ~load_plane(ra + 3 * kN, lane);
  const Plane a_zone = to_plane(dilate<32>(to_tile(a_free), lane, false));  // ~a_ignored
  const Plane Bs = load_plane(rb, lane);
  const Plane b_free = Bs & ~load_plane(rb + kN, lane) & ~load_plane(rb + 2 * kN, lane) & ~load_plane(rb + 3 * kN, lane);
  const Plane B = mirrored(Bs, lane);
  const Plane b_zone = mirrored(to_plane(dilate<32>(to_tile(b_free), lane, false)), lane);  // ~b_ignored
  Tile<32> t3, t2, t1, t0;
  to_tile(A).neighbour_count(lane, t3, t2, t1, t0);
  const Plane a3 = to_plane(t3), a2 = to_plane(t2), a1 = to_plane(t1), a0 = to_plane(t0);
  to_tile(B).neighbour_count(lane, t3, t2, t1, t0);
  const Plane b3 = to_plane(t3), b2 = to_plane(t2), b1 = to_plane(t1), b0 = to_plane(t0);
  const Plane a_lo = ~a3 & ~a2, b_lo = ~b3 & ~b2;
  Plane r = convolve(A, B, lane);                                                                        // overlaps
  r = r | convolve(a_lo & ~a1 & a0 & ~A & a_zone, b_lo & b1 & ~b0 & ~B & a_zone, lane);                  // births
  r = r | convolve(b_lo & ~b1 & b0 & ~B & b_zone, a_lo & a1 & ~a0 & ~A & b_zone, lane);
  r = r | convolve(a_lo & a1 & a0 & A & a_zone, (b3 | b2 | b1) & ~B & b_zone, lane);                      // overcrowding
  r = r | convolve((a2 | a3) & A & a_zone, (b3 | b2 | b1 | b0) & ~B & b_zone, lane);
  r = r | convolve(b_lo & b1 & b0 & B & b_zone, (a3 | a2 | a1) & ~A & a_zone, lane);
  r = r | convolve((b2 | b3) & B & b_zone, (a3 | a2 | a1 | a0) & ~A & a_zone, lane);
  store_plane(out + p * kN, lane, r);
}

// ---- host-pointer variants: up to three plane inputs, each either per-problem or broadcast ----
struct HostIn {
  const uint64_t* ptr;
  size_t count;   // 1 (broadcast) or n
  size_t words;   // words per item (64, or 128 for interleaved LifeTarget records)
};

template <class LaunchFn>
int planes_host(lifeb200_ctx* ctx, const HostIn* ins, int n_ins, uint64_t* out, size_t n, LaunchFn launch) {
  const size_t chunk = std::min<size_t>(n, (size_t)1 << 15);
  const size_t nchunks = (n + chunk - 1) / chunk;
  const int used = (int)std::min<size_t>(kSlots, nchunks);
  size_t in_bytes = 0;
  for (int i = 0; i < n_ins; i++) in_bytes += (ins[i].count == 1 ? 1 : chunk) * ins[i].words * 8;
  for (int s = 0; s < used; s++) {
    int rc = ensure_slot(ctx, s, in_bytes, chunk * LIFEB200_PATTERN_BYTES);
    if (rc) return rc;
  }
  auto body = [&]() -> int {
  for (size_t c = 0; c < nchunks; c++) {
    Slot& sl = ctx->slots[c % kSlots];
    const size_t off = c * chunk, cnt = std::min(chunk, n - off);
    const uint64_t* dptr[3] = {nullptr, nullptr, nullptr};
    uint64_t* d = static_cast<uint64_t*>(sl.dbuf);
    for (int i = 0; i < n_ins; i++) {
      if (!ins[i].ptr) continue;
      const bool bc = ins[i].count == 1;
      const size_t items = bc ? 1 : cnt;
      LIFEB200_CUDA(cudaMemcpyAsync(d, ins[i].ptr + (bc ? 0 : off * ins[i].words), items * ins[i].words * 8,
                                    cudaMemcpyHostToDevice, sl.stream));
      dptr[i] = d;
      d += (bc ? 1 : chunk) * ins[i].words;
    }
    uint64_t* dout = static_cast<uint64_t*>(sl.dbuf2);
    launch(dptr, dout, cnt, sl.stream);
    LIFEB200_LAUNCHED(ctx);
    LIFEB200_CUDA(cudaMemcpyAsync(out + off * kN, dout, cnt * LIFEB200_PATTERN_BYTES, cudaMemcpyDeviceToHost, sl.stream));
  }
  for (int s = 0; s < used; s++) LIFEB200_CUDA(cudaStreamSynchronize(ctx->slots[s].stream));
  return LIFEB200_OK;
  };
  const int rc = body();
  if (rc) drain_slots(ctx, used);  // nothing may still be copying towards the caller's buffers after the call returns
  return rc;
}

// LifeState::Contains(const LifeTarget&, int dx, int dy) (LifeTarget.hpp:38-42 -> Contains(pat, dx, dy) and
// AreDisjoint(pat, dx, dy), LifeAPI.hpp:399-422): the target sits at offset (dx, dy) of the state -- for every target
// word i the state word (i + dx) & 63 rotated right by dy must hold every wanted cell and no unwanted one.  A warp
// per state: lane L tests target words 2L, 2L + 1 (one coalesced 512-byte request per plane; the state words arrive
// by the shifted index straight from global memory), one ballot folds the verdict.  HBM-bound: 512 B in, 1 B out.
__global__ void __launch_bounds__(kBlockC) contains_shifted_kernel(const uint64_t* __restrict__ states, size_t n,
                                                                   const uint64_t* __restrict__ targets, size_t n_targets,
                                                                   int dx, int dy, uint8_t* __restrict__ out) {
  const size_t p = ((size_t)blockIdx.x * kBlockC + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (p >= n) return;
  const uint64_t* st = states + p * kN;
  const uint64_t* tg = targets + (n_targets == 1 ? 0 : p) * 2 * kN;
  const unsigned sy = (unsigned)dy & 63u;
  uint64_t bad = 0;
#pragma unroll
  for (int k = 0; k < 2; k++) {
    const int i = 2 * lane + k;
    const uint64_t w = tg[i], u = tg[kN + i];
    const uint64_t v = st[(i + dx) & 63];
    const uint64_t r = sy ? (v >> sy) | (v << (64u - sy)) : v;
    bad |= (w & ~r) | (u & r);
  }
  const unsigned any = __ballot_sync(0xffffffffu, bad != 0);
  if (lane == 0) out[p] = any ? 0 : 1;
}

static int check_pair(const void* a, size_t n_a, const void* b, size_t n_b, const void* out, size_t n, bool need_align) {
  if (!a || !b || !out) return set_error(LIFEB200_E_ARG, "null pointer");
  if ((n_a != 1 && n_a != n) || (n_b != 1 && n_b != n)) return set_error(LIFEB200_E_ARG, "operand counts must be 1 (broadcast) or n");
  if (need_align && (!aligned16(a) || !aligned16(b) || !aligned16(out))) return set_error(LIFEB200_E_ARG, "misaligned device pointer");
  return LIFEB200_OK;
}

}  // namespace lifeb200

using namespace lifeb200;

extern "C" {

int lifeb200_convolve_batch_dev(lifeb200_ctx* ctx, const uint64_t* a, size_t n_a, const uint64_t* b, size_t n_b,
                                uint64_t* out, size_t n) {
  LIFEB200_CHECK_CTX(ctx);
  if (n == 0) return LIFEB200_OK;
  if (int rc = check_pair(a, n_a, b, n_b, out, n, true)) return rc;
  convolve_kernel<<<grid_warps(n), kBlockC, 0, ctx->stream>>>(a, n_a, b, n_b, out, n);
  LIFEB200_LAUNCHED(ctx);
  return LIFEB200_OK;
}

int lifeb200_convolve_batch_host(lifeb200_ctx* ctx, const uint64_t* a, size_t n_a, const uint64_t* b, size_t n_b,
                                 uint64_t* out, size_t n) {
  LIFEB200_CHECK_CTX(ctx);
  if (n == 0) return LIFEB200_OK;
  if (int rc = check_pair(a, n_a, b, n_b, out, n, false)) return rc;
  const HostIn ins[2] = {{a, n_a, (size_t)kN}, {b, n_b, (size_t)kN}};
  return planes_host(ctx, ins, 2, out, n, [&](const uint64_t* const* d, uint64_t* dout, size_t cnt, cudaStream_t st) {
    convolve_kernel<<<grid_warps(cnt), kBlockC, 0, st>>>(d[0], n_a == 1 ? 1 : cnt, d[1], n_b == 1 ? 1 : cnt, dout, cnt);
  });
}

int lifeb200_match_batch_dev(lifeb200_ctx* ctx, const uint64_t* states, size_t n, const uint64_t* live,
                             const uint64_t* dead, size_t n_t, int mode, uint64_t* out) {
  LIFEB200_CHECK_CTX(ctx);
  if (n == 0) return LIFEB200_OK;
  if (mode < 0 || mode > 2) return set_error(LIFEB200_E_ARG, "match mode must be 0, 1 or 2");
  if (int rc = check_pair(states, n, live, n_t, out, n, true)) return rc;
  if (mode == 0 && (!dead || !aligned16(dead))) return set_error(LIFEB200_E_ARG, "mode 0 needs a dead plane");
  match_kernel<<<grid_warps(n), kBlockC, 0, ctx->stream>>>(states, n, live, dead, n_t, kN, mode, out);
  LIFEB200_LAUNCHED(ctx);
  return LIFEB200_OK;
}

int lifeb200_match_batch_host(lifeb200_ctx* ctx, const uint64_t* states, size_t n, const uint64_t* live,
                              const uint64_t* dead, size_t n_t, int mode, uint64_t* out) {
  LIFEB200_CHECK_CTX(ctx);
  if (n == 0) return LIFEB200_OK;
  if (mode < 0 || mode > 2) return set_error(LIFEB200_E_ARG, "match mode must be 0, 1 or 2");
  if (int rc = check_pair(states, n, live, n_t, out, n, false)) return rc;
  if (mode == 0 && !dead) return set_error(LIFEB200_E_ARG, "mode 0 needs a dead plane");
  const HostIn ins[3] = {{states, n, (size_t)kN}, {live, n_t, (size_t)kN}, {mode == 0 ? dead : nullptr, n_t, (size_t)kN}};
  return planes_host(ctx, ins, 3, out, n, [&](const uint64_t* const* d, uint64_t* dout, size_t cnt, cudaStream_t st) {
    match_kernel<<<grid_warps(cnt), kBlockC, 0, st>>>(d[0], cnt, d[1], d[2], n_t == 1 ? 1 : cnt, kN, mode, dout);
  });
}

int lifeb200_match_target_batch_dev(lifeb200_ctx* ctx, const uint64_t* states, size_t n, const uint64_t* targets,
                                    size_t n_targets, uint64_t* out) {
  LIFEB200_CHECK_CTX(ctx);
  if (n == 0) return LIFEB200_OK;
  if (int rc = check_pair(states, n, targets, n_targets, out, n, true)) return rc;
  match_kernel<<<grid_warps(n), kBlockC, 0, ctx->stream>>>(states, n, targets, targets + kN, n_targets, 2 * kN, 0, out);
  LIFEB200_LAUNCHED(ctx);
  return LIFEB200_OK;
}

int lifeb200_match_target_batch_host(lifeb200_ctx* ctx, const uint64_t* states, size_t n, const uint64_t* targets,
                                     size_t n_targets, uint64_t* out) {
  LIFEB200_CHECK_CTX(ctx);
  if (n == 0) return LIFEB200_OK;
  if (int rc = check_pair(states, n, targets, n_targets, out, n, false)) return rc;
  const HostIn ins[2] = {{states, n, (size_t)kN}, {targets, n_targets, (size_t)(2 * kN)}};
  return planes_host(ctx, ins, 2, out, n, [&](const uint64_t* const* d, uint64_t* dout, size_t cnt, cudaStream_t st) {
    match_kernel<<<grid_warps(cnt), kBlockC, 0, st>>>(d[0], cnt, d[1], d[1] + kN, n_targets == 1 ? 1 : cnt, 2 * kN, 0, dout);
  });
}

int lifeb200_interaction_offsets_dev(lifeb200_ctx* ctx, const uint64_t* a, size_t n_a, const uint64_t* b, size_t n_b,
                                     uint64_t* out, size_t n) {
  LIFEB200_CHECK_CTX(ctx);
  if (n == 0) return LIFEB200_OK;
  if (int rc = check_pair(a, n_a, b, n_b, out, n, true)) return rc;
  interaction_offsets_kernel<<<grid_warps(n), kBlockC, 0, ctx->stream>>>(a, n_a, b, n_b, out, n);
  LIFEB200_LAUNCHED(ctx);
  return LIFEB200_OK;
}

int lifeb200_interaction_offsets_host(lifeb200_ctx* ctx, const uint64_t* a, size_t n_a, const uint64_t* b, size_t n_b,
                                      uint64_t* out, size_t n) {
  LIFEB200_CHECK_CTX(ctx);
  if (n == 0) return LIFEB200_OK;
  if (int rc = check_pair(a, n_a, b, n_b, out, n, false)) return rc;
  const HostIn ins[2] = {{a, n_a, (size_t)kN}, {b, n_b, (size_t)kN}};
  return planes_host(ctx, ins, 2, out, n, [&](const uint64_t* const* d, uint64_t* dout, size_t cnt, cudaStream_t st) {
    interaction_offsets_kernel<<<grid_warps(cnt), kBlockC, 0, st>>>(d[0], n_a == 1 ? 1 : cnt, d[1], n_b == 1 ? 1 : cnt, dout, cnt);
  });
}

int lifeb200_contains_target_batch_dev(lifeb200_ctx* ctx, const uint64_t* states, size_t n, const uint64_t* targets,
                                       size_t n_targets, int dx, int dy, uint8_t* contains) {
  LIFEB200_CHECK_CTX(ctx);
  if (n == 0) return LIFEB200_OK;
  if (!states || !targets || !contains) return set_error(LIFEB200_E_ARG, "null pointer");
  if (n_targets != 1 && n_targets != n) return set_error(LIFEB200_E_ARG, "target count must be 1 (broadcast) or n");
  contains_shifted_kernel<<<grid_warps(n), kBlockC, 0, ctx->stream>>>(states, n, targets, n_targets, dx, dy, contains);
  LIFEB200_LAUNCHED(ctx);
  return LIFEB200_OK;
}

int lifeb200_contains_target_batch_host(lifeb200_ctx* ctx, const uint64_t* states, size_t n, const uint64_t* targets,
                                        size_t n_targets, int dx, int dy, uint8_t* contains) {
  LIFEB200_CHECK_CTX(ctx);
  if (n == 0) return LIFEB200_OK;
  if (!states || !targets || !contains) return set_error(LIFEB200_E_ARG, "null pointer");
  if (n_targets != 1 && n_targets != n) return set_error(LIFEB200_E_ARG, "target count must be 1 (broadcast) or n");
  // item = state (512 B) for a shared target, state + its target (1536 B staged side by side) otherwise
  if (n_targets == 1) {
    if (int rc = ensure_scratch(ctx->d_tabbuf, ctx->tabbuf_cap, 2 * LIFEB200_PATTERN_BYTES)) return rc;
    ctx->tab_valid = false;  // the table scratch is reused: the enumeration's cached tables are gone
    LIFEB200_CUDA(cudaMemcpyAsync(ctx->d_tabbuf, targets, 2 * LIFEB200_PATTERN_BYTES, cudaMemcpyHostToDevice, ctx->stream));
    LIFEB200_CUDA(cudaStreamSynchronize(ctx->stream));
    const uint64_t* d_t = static_cast<const uint64_t*>(ctx->d_tabbuf);
    return host_pipeline_impl(ctx, states, contains, n, LIFEB200_PATTERN_BYTES, 1, true, (size_t)1 << 16,
                              [&](cudaStream_t st, const uint64_t* din, uint64_t* dout, size_t cnt) -> int {
                                contains_shifted_kernel<<<grid_warps(cnt), kBlockC, 0, st>>>(din, cnt, d_t, 1, dx, dy,
                                                                                             reinterpret_cast<uint8_t*>(dout));
                                LIFEB200_LAUNCHED(ctx);
                                return LIFEB200_OK;
                              });
  }
  const size_t chunk = std::min<size_t>(n, (size_t)1 << 15);
  const size_t nchunks = (n + chunk - 1) / chunk;
  const int used = (int)std::min<size_t>(kSlots, nchunks);
  for (int s = 0; s < used; s++)
    if (int rc = ensure_slot(ctx, s, chunk * 3 * LIFEB200_PATTERN_BYTES, chunk)) return rc;
  auto body = [&]() -> int {
    for (size_t c = 0; c < nchunks; c++) {
      Slot& sl = ctx->slots[c % kSlots];
      const size_t off = c * chunk, cnt = std::min(chunk, n - off);
      uint64_t* d_s = static_cast<uint64_t*>(sl.dbuf);
      uint64_t* d_t = d_s + chunk * kN;
      LIFEB200_CUDA(cudaMemcpyAsync(d_s, states + off * kN, cnt * LIFEB200_PATTERN_BYTES, cudaMemcpyHostToDevice, sl.stream));
      LIFEB200_CUDA(cudaMemcpyAsync(d_t, targets + off * 2 * kN, cnt * 2 * LIFEB200_PATTERN_BYTES, cudaMemcpyHostToDevice, sl.stream));
      contains_shifted_kernel<<<grid_warps(cnt), kBlockC, 0, sl.stream>>>(d_s, cnt, d_t, cnt, dx, dy, static_cast<uint8_t*>(sl.dbuf2));
      LIFEB200_LAUNCHED(ctx);
      LIFEB200_CUDA(cudaMemcpyAsync(contains + off, sl.dbuf2, cnt, cudaMemcpyDeviceToHost, sl.stream));
    }
    for (int s = 0; s < used; s++) LIFEB200_CUDA(cudaStreamSynchronize(ctx->slots[s].stream));
    return LIFEB200_OK;
  };
  const int rc = body();
  if (rc) drain_slots(ctx, used);
  return rc;
}

int lifeb200_weld_interaction_offsets_dev(lifeb200_ctx* ctx, const uint64_t* a, size_t n_a, const uint64_t* b, size_t n_b,
                                          uint64_t* out, size_t n) {
  LIFEB200_CHECK_CTX(ctx);
  if (n == 0) return LIFEB200_OK;
  if (int rc = check_pair(a, n_a, b, n_b, out, n, true)) return rc;
  weld_interaction_offsets_kernel<<<grid_warps(n), kBlockC, 0, ctx->stream>>>(a, n_a, b, n_b, out, n);
  LIFEB200_LAUNCHED(ctx);
  return LIFEB200_OK;
}

int lifeb200_weld_interaction_offsets_host(lifeb200_ctx* ctx, const uint64_t* a, size_t n_a, const uint64_t* b, size_t n_b,
                                           uint64_t* out, size_t n) {
  LIFEB200_CHECK_CTX(ctx);
  if (n == 0) return LIFEB200_OK;
  if (int rc = check_pair(a, n_a, b, n_b, out, n, false)) return rc;
  const HostIn ins[2] = {{a, n_a, (size_t)(4 * kN)}, {b, n_b, (size_t)(4 * kN)}};
  return planes_host(ctx, ins, 2, out, n, [&](const uint64_t* const* d, uint64_t* dout, size_t cnt, cudaStream_t st) {
    weld_interaction_offsets_kernel<<<grid_warps(cnt), kBlockC, 0, st>>>(d[0], n_a == 1 ? 1 : cnt, d[1], n_b == 1 ? 1 : cnt, dout, cnt);
  });
}

}  // extern "C"
