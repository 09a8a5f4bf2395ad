// step.cu -- batched LifeState::Step(n) (LifeAPI.hpp:877-881,1196-1216), the soup source,
// NeighbourCount planes and the integer-pipe microbenchmark, with their C-ABI entry points.
#include <algorithm>
#include <vector>

#include "ctx.h"
#include "life_core.cuh"
#include "life_multi.cuh"
#include "host_pipeline.cuh"

namespace lifeb200 {

constexpr int kBlock = 256;
// The step kernel is bound by the integer pipe, not by occupancy: with the register budget of one
// small CTA per SM (min blocks = 1) ptxas keeps every column count of a generation in flight instead of
// squeezing to 64 registers, and the ALU pipe goes from 95-96% to 99% busy (tools/ubench_step.cu,
// profiles/r01h_ubench_step.txt).
constexpr int kStepBlock = 128;
// Register layouts of the step kernel (lifeb200_set_step_layout; life_core.cuh / life_multi.cuh):
//   1 = the reference's (rows 0..31 | 32..63): 2 shifts per register and generation, no conversion
//   2 = even | odd rows: 1 shift per register; 4, 8 = rows by class mod 4 / 8 with 2 / 4 patterns
//   bit-interleaved per register: 1/2, 1/4 shift per register.  The conversions after the load and before
//   the store cost about 4 (layout 2) or 3 (layout 8: four register <-> position bit exchanges, life_multi.cuh)
//   generations' worth of work, so auto picks by gens.  Measured crossovers (profiles/r02x_step_sweep.txt): layout 8
//   beats the reference layout from about 15 generations (1.099 against 1.119 ms at 16) and layout 2 everywhere; layout 2 is what a pinned lane
//   grouping (lifeb200_set_lanes_per_pattern) falls back to.
constexpr unsigned kInterleaveFrom = 36, kMultiFrom = 16;

// Each warp owns 32/LPP whole patterns; `gens` generations run in registers between one
// 128-bit vectorised load and one 128-bit vectorised store per 16 bytes of state.
template <int LPP>
__global__ void __launch_bounds__(kStepBlock, 1) step_kernel(const uint64_t* in, uint64_t* out,
                                                              size_t n, unsigned gens, bool interleaved) {
  const size_t gid = (size_t)blockIdx.x * kStepBlock + threadIdx.x;
  const size_t p = gid / LPP;
  const int sub = threadIdx.x & (LPP - 1);
  const bool valid = p < n;
  Tile<LPP> t;
  if (valid) {
    t.load(in + p * kN, sub);
  } else {
#pragma unroll
    for (int w = 0; w < Tile<LPP>::W; w++) t.lo[w] = t.hi[w] = 0;
  }
  if (interleaved) {  // uniform over the launch: many generations pay for the two layout conversions
    t.interleave();
#pragma unroll 1
    for (unsigned g = 0; g < gens; g++) t.template step<true>(sub);
    t.deinterleave();
  } else {
#pragma unroll 1
    for (unsigned g = 0; g < gens; g++) t.template step<false>(sub);
  }
  if (valid) t.store(out + p * kN, sub);
}

// P = M/2 patterns per group of 8 lanes in the row-class layout of life_multi.cuh.
template <int M>
__global__ void __launch_bounds__(kStepBlock, 1) step_kernel_x(const uint64_t* in, uint64_t* out,
                                                                size_t n, unsigned gens) {
  constexpr int P = M / 2;
  const size_t gid = (size_t)blockIdx.x * kStepBlock + threadIdx.x;
  const size_t p0 = (gid >> 3) * P;  // first pattern of this lane group
  const int sub = threadIdx.x & 7;   // this lane owns columns 8*sub .. 8*sub+7 of all P patterns
  TileX<M> t;
#pragma unroll
  for (int c = 0; c < 8; c += 2) {
    uint64_t wa[P], wb[P];
#pragma unroll
    for (int s = 0; s < P; s++) {
      uint4 v = make_uint4(0, 0, 0, 0);  // groups past the end step empty patterns (they still shuffle)
      if (p0 + s < n) v = *reinterpret_cast<const uint4*>(in + (p0 + s) * kN + sub * 8 + c);  // coherent: in may be out
      wa[s] = ((uint64_t)v.y << 32) | v.x;
      wb[s] = ((uint64_t)v.w << 32) | v.z;
    }
    t.pack(c, wa);
    t.pack(c + 1, wb);
  }
#pragma unroll 1
  for (unsigned g = 0; g < gens; g++) t.step(sub);
#pragma unroll
  for (int c = 0; c < 8; c += 2) {
    uint64_t wa[P], wb[P];
    t.unpack(c, wa);
    t.unpack(c + 1, wb);
#pragma unroll
    for (int s = 0; s < P; s++)
      if (p0 + s < n)
        *reinterpret_cast<uint4*>(out + (p0 + s) * kN + sub * 8 + c) =
            make_uint4((uint32_t)wa[s], (uint32_t)(wa[s] >> 32), (uint32_t)wb[s], (uint32_t)(wb[s] >> 32));
  }
}

template <int LPP>
__global__ void __launch_bounds__(kBlock) soup_kernel(uint64_t* __restrict__ out, uint64_t first_id, size_t n,
                                                      uint64_t seed, int dist) {
  const size_t gid = (size_t)blockIdx.x * kBlock + threadIdx.x;
  const size_t p = gid / LPP;
  const int sub = threadIdx.x & (LPP - 1);
  if (p >= n) return;
  constexpr int W = Tile<LPP>::W;
  uint64_t* dst = out + p * kN + sub * W;
#pragma unroll
  for (int w = 0; w < W; w += 2) {
    uint64_t a = soup_word(seed, first_id + p, sub * W + w, dist);
    uint64_t b = soup_word(seed, first_id + p, sub * W + w + 1, dist);
    *reinterpret_cast<uint4*>(dst + w) =
        make_uint4((uint32_t)a, (uint32_t)(a >> 32), (uint32_t)b, (uint32_t)(b >> 32));
  }
}

template <int LPP>
__global__ void __launch_bounds__(kBlock) ncount_kernel(const uint64_t* __restrict__ in, uint64_t* __restrict__ planes,
                                                        size_t n) {
  const size_t gid = (size_t)blockIdx.x * kBlock + threadIdx.x;
  const size_t p = gid / LPP;
  const int sub = threadIdx.x & (LPP - 1);
  const bool valid = p < n;
  Tile<LPP> t, b3, b2, b1, b0;
  if (valid) {
    t.load(in + p * kN, sub);
  } else {
#pragma unroll
    for (int w = 0; w < Tile<LPP>::W; w++) t.lo[w] = t.hi[w] = 0;
  }
  t.neighbour_count(sub, b3, b2, b1, b0);
  if (valid) {
    uint64_t* o = planes + p * 4 * kN;
    b3.store(o, sub); b2.store(o + kN, sub); b1.store(o + 2 * kN, sub); b0.store(o + 3 * kN, sub);
  }
}

// Integer-pipe peak: 16 independent LOP3 chains per thread, nothing else in the loop body.
constexpr int kPeakChains = 16;
constexpr int kPeakInner = 64;
__global__ void __launch_bounds__(1024) int_peak_kernel(uint32_t* sink, unsigned iters) {
  uint32_t v[kPeakChains];
#pragma unroll
  for (int i = 0; i < kPeakChains; i++) v[i] = threadIdx.x * 2654435761u + i * 40503u + blockIdx.x;
  uint32_t k1 = threadIdx.x | 0x55u, k2 = ~threadIdx.x;
#pragma unroll 1
  for (unsigned it = 0; it < iters; it++) {
#pragma unroll
    for (int j = 0; j < kPeakInner; j++) {
#pragma unroll
      for (int i = 0; i < kPeakChains; i++) v[i] = lop3<0x96>(v[i], v[(i + 1) % kPeakChains], k1 + 0 * k2);
    }
  }
  uint32_t acc = 0;
#pragma unroll
  for (int i = 0; i < kPeakChains; i++) acc ^= v[i];
  if (acc == 0x12345u) sink[0] = acc;  // keep the chains alive
}

static inline unsigned grid_for(size_t n, int lpp) {
  size_t threads = n * (size_t)lpp;
  return (unsigned)((threads + kBlock - 1) / kBlock);
}

int launch_step(lifeb200_ctx* ctx, cudaStream_t st, const uint64_t* in, uint64_t* out, size_t n, unsigned gens) {
  if (n == 0) return LIFEB200_OK;
  // Measured on B200 (profiles/r01h_ubench_step.txt): 8 lanes per pattern (8 words per lane) keeps the ALU
  // pipe busiest when many generations run per launch; with few generations the launch is HBM-bound and
  // 16 lanes per pattern (32 contiguous bytes per lane) coalesces best.
  int layout = ctx->step_layout;
  if (layout == 0) layout = (gens >= kMultiFrom && ctx->lpp == 0) ? 8 : gens >= kInterleaveFrom ? 2 : 1;
  if (layout >= 4) {
    const size_t groups = (n + layout / 2 - 1) / (layout / 2);
    const unsigned grid = (unsigned)((groups * 8 + kStepBlock - 1) / kStepBlock);
    if (layout == 8) step_kernel_x<8><<<grid, kStepBlock, 0, st>>>(in, out, n, gens);
    else step_kernel_x<4><<<grid, kStepBlock, 0, st>>>(in, out, n, gens);
  } else {
    int lpp = ctx->lpp ? ctx->lpp : (gens >= 8 ? 8 : 16);
    LIFEB200_DISPATCH_LPP(lpp, (step_kernel<LPP><<<(unsigned)((n * (size_t)LPP + kStepBlock - 1) / kStepBlock), kStepBlock, 0, st>>>(
                                   in, out, n, gens, layout == 2)));
  }
  LIFEB200_LAUNCHED(ctx);
  return LIFEB200_OK;
}

int launch_ncount(lifeb200_ctx* ctx, cudaStream_t st, const uint64_t* in, uint64_t* planes, size_t n) {
  if (n == 0) return LIFEB200_OK;
  // 5 tiles in registers: keep W <= 16.  A single pass with 2 KiB out per pattern: a warp per pattern makes every
  // plane store one contiguous 512-byte request (see interaction_kernel, csrc/next.cu)
  int lpp = ctx->lpp ? std::max(4, ctx->lpp) : 32;
  LIFEB200_DISPATCH_LPP_MIN4(lpp, (ncount_kernel<LPP><<<grid_for(n, LPP), kBlock, 0, st>>>(in, planes, n)));
  LIFEB200_LAUNCHED(ctx);
  return LIFEB200_OK;
}

}  // namespace lifeb200

using namespace lifeb200;

extern "C" {

int lifeb200_step_batch_dev(lifeb200_ctx* ctx, const uint64_t* in, uint64_t* out, size_t n, unsigned gens) {
  LIFEB200_CHECK_CTX(ctx);
  if (n == 0) return LIFEB200_OK;
  if (!in || !out) return set_error(LIFEB200_E_ARG, "null pattern pointer");
  if (!aligned16(in) || !aligned16(out)) return set_error(LIFEB200_E_ARG, "pattern pointers must be 16-byte aligned");
  if (in != out && in < out + n * kN && out < in + n * kN)
    return set_error(LIFEB200_E_ARG, "in and out overlap partially (in == out is the only aliasing allowed)");
  if (gens == 0) {
    if (in != out)
      LIFEB200_CUDA(cudaMemcpyAsync(out, in, n * LIFEB200_PATTERN_BYTES, cudaMemcpyDeviceToDevice, ctx->stream));
    return LIFEB200_OK;
  }
  return launch_step(ctx, ctx->stream, in, out, n, gens);
}

int lifeb200_step_batch_host(lifeb200_ctx* ctx, const uint64_t* in, uint64_t* out, size_t n, unsigned gens) {
  LIFEB200_CHECK_CTX(ctx);
  if (n == 0) return LIFEB200_OK;
  if (!in || !out) return set_error(LIFEB200_E_ARG, "null pattern pointer");
  return host_pipeline(ctx, in, out, n, LIFEB200_PATTERN_BYTES, LIFEB200_PATTERN_BYTES, false,
                       [&](cudaStream_t st, const uint64_t* din, uint64_t* dout, size_t cnt) -> int {
                         if (gens == 0) return LIFEB200_OK;
                         return launch_step(ctx, st, din, dout, cnt, gens);
                       });
}

int lifeb200_soup_dev(lifeb200_ctx* ctx, uint64_t* out, uint64_t first_id, size_t n, uint64_t seed, int dist) {
  LIFEB200_CHECK_CTX(ctx);
  if (n == 0) return LIFEB200_OK;
  if (!out || !aligned16(out)) return set_error(LIFEB200_E_ARG, "bad output pointer");
  if (dist != 0 && dist != 1) return set_error(LIFEB200_E_ARG, "dist must be 0 or 1");
  soup_kernel<8><<<grid_for(n, 8), kBlock, 0, ctx->stream>>>(out, first_id, n, seed, dist);
  LIFEB200_LAUNCHED(ctx);
  return LIFEB200_OK;
}

int lifeb200_neighbour_count_dev(lifeb200_ctx* ctx, const uint64_t* in, uint64_t* planes, size_t n) {
  LIFEB200_CHECK_CTX(ctx);
  if (n == 0) return LIFEB200_OK;
  if (!in || !planes) return set_error(LIFEB200_E_ARG, "null pointer");
  if (!aligned16(in) || !aligned16(planes)) return set_error(LIFEB200_E_ARG, "pointers must be 16-byte aligned");
  return launch_ncount(ctx, ctx->stream, in, planes, n);
}

int lifeb200_neighbour_count_host(lifeb200_ctx* ctx, const uint64_t* in, uint64_t* planes, size_t n) {
  LIFEB200_CHECK_CTX(ctx);
  if (n == 0) return LIFEB200_OK;
  if (!in || !planes) return set_error(LIFEB200_E_ARG, "null pointer");
  return host_pipeline(ctx, in, planes, n, LIFEB200_PATTERN_BYTES, 4 * LIFEB200_PATTERN_BYTES, true,
                       [&](cudaStream_t st, const uint64_t* din, uint64_t* dout, size_t cnt) -> int {
                         return launch_ncount(ctx, st, din, dout, cnt);
                       });
}

int lifeb200_measure_int_peak(lifeb200_ctx* ctx, double* lop3_lane_instr_per_s, double* sm_mhz_effective) {
  LIFEB200_CHECK_CTX(ctx);
  const unsigned iters = 256;
  const int blocks = ctx->sm_count * 2, threads = 1024;
  cudaEvent_t a, b;
  LIFEB200_CUDA(cudaEventCreate(&a));
  LIFEB200_CUDA(cudaEventCreate(&b));
  uint32_t* sink = reinterpret_cast<uint32_t*>(ctx->d_counters + 7);
  double best = 0;
  for (int rep = 0; rep < 5; rep++) {
    LIFEB200_CUDA(cudaEventRecord(a, ctx->stream));
    int_peak_kernel<<<blocks, threads, 0, ctx->stream>>>(sink, iters);
    LIFEB200_LAUNCHED(ctx);
    LIFEB200_CUDA(cudaEventRecord(b, ctx->stream));
    LIFEB200_CUDA(cudaEventSynchronize(b));
    float ms = 0;
    LIFEB200_CUDA(cudaEventElapsedTime(&ms, a, b));
    double instr = (double)blocks * threads * (double)iters * kPeakInner * kPeakChains;
    double rate = instr / (ms * 1e-3);
    if (rep > 0 && rate > best) best = rate;
  }
  cudaEventDestroy(a);
  cudaEventDestroy(b);
  if (lop3_lane_instr_per_s) *lop3_lane_instr_per_s = best;
  if (sm_mhz_effective) *sm_mhz_effective = best / ((double)ctx->sm_count * 64.0) / 1e6;
  return LIFEB200_OK;
}

}  // extern "C"
