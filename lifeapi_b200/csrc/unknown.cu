// unknown.cu -- batched three-valued stepping through the generated circuits
// and the history-accumulating step, with their C-ABI entry points.
//
// The reference ships bitslicing/unknown_step*.hpp without an in-repo consumer;
// the loop shape here is the one LifeStable.hpp:429-469 uses for stable_simple.hpp: two
// NeighbourCounts (NeighbourCount.hpp:40-70) feed one generated circuit per word.  On the GPU the
// counts and the circuit run on 32-bit half-words held in registers, many generations per
// HBM round trip.
#include <algorithm>

#include "ctx.h"
#include "host_pipeline.cuh"
#include "generated/circuits.cuh"
#include "life_core.cuh"
#include "life_three.cuh"

namespace lifeb200 {

constexpr int kBlockU = 256;

// One three-valued generation in place: (state, unknown) -> unknown_step.hpp outputs.
template <int LPP>
__device__ __forceinline__ void unknown_step_tile(Tile<LPP>& s, Tile<LPP>& u, int sub) {
  constexpr int W = Tile<LPP>::W;
  Tile<LPP> on3, on2, on1, on0, k3, k2, k1, k0;
  s.neighbour_count(sub, on3, on2, on1, on0);  // on3 is declared but unused by the cover (dead code)
  u.neighbour_count(sub, k3, k2, k1, k0);
#pragma unroll
  for (int k = 0; k < 2 * W; k++) {
    uint32_t nu, ns;
    circuits::unknown_step(reg(u, k), reg(s, k), reg(on2, k), reg(on1, k), reg(on0, k), reg(k3, k), reg(k2, k),
                           reg(k1, k), reg(k0, k), nu, ns);
    reg(s, k) = ns;
    reg(u, k) = nu;
  }
}

// A warp whose patterns all keep state & unknown == 0 (every reachable LifeStable / LifeHistory input)
// steps through the two-select form of life_three.cuh, row-interleaved when IL; a warp that sees
// overlapping planes is outside the generator's domain and follows the fragment text literally
// (unknown_step_tile), so the result is the cover's on EVERY input.  The choice is per warp because
// both paths shuffle with a full mask.
// Lane groups are persistent: a group walks patterns p, p + stride, ... and, with 16 or 32 lanes per pattern, issues the
// NEXT pattern's loads before it steps the current one.  One pattern per short-lived group left the HBM phase
// exposed (config 3, 16 generations: 0.816 ms = 16 x 0.047 ms of stepping + 0.067 ms of load / store): every
// CTA of a wave takes the same time, so co-resident CTAs load, step and store in lockstep and nothing overlaps.
template <int LPP, bool IL>
__global__ void __launch_bounds__(kBlockU, 1) unknown_step_kernel(const uint64_t* state_in,
                                                               const uint64_t* unknown_in,
                                                               uint64_t* state_out,
                                                               uint64_t* unknown_out, size_t n,
                                                               unsigned gens, int force_literal) {
  constexpr bool kPrefetch = LPP >= 16;  // 2 x W more registers; 8 or 4 lanes per pattern have none to spare
  const size_t gid = (size_t)blockIdx.x * kBlockU + threadIdx.x;
  const size_t stride = (size_t)gridDim.x * kBlockU / LPP;
  size_t p = gid / LPP;
  size_t pw = (gid - (threadIdx.x & 31)) / LPP;  // first pattern of this warp: the loop bound is warp-uniform
  const int sub = threadIdx.x & (LPP - 1);
  Tile<LPP> s, u, ns, nu;
  auto fetch = [&](size_t q, Tile<LPP>& ts, Tile<LPP>& tu) {
    if (q < n) {
      ts.load(state_in + q * kN, sub);
      tu.load(unknown_in + q * kN, sub);
    } else {
#pragma unroll
      for (int w = 0; w < Tile<LPP>::W; w++) ts.lo[w] = ts.hi[w] = tu.lo[w] = tu.hi[w] = 0;
    }
  };
  if (kPrefetch && pw < n) fetch(p, s, u);
  for (; pw < n; pw += stride, p += stride) {
    const bool valid = p < n;  // invalid groups still take part in the full-warp shuffles
    if constexpr (kPrefetch) {
      if (pw + stride < n) fetch(p + stride, ns, nu);
    } else {
      fetch(p, s, u);
    }
    uint32_t both = 0;
#pragma unroll
    for (int w = 0; w < Tile<LPP>::W; w++) both |= (s.lo[w] & u.lo[w]) | (s.hi[w] & u.hi[w]);
    if (__any_sync(0xffffffffu, both != 0 || force_literal)) {
#pragma unroll 1
      for (unsigned g = 0; g < gens; g++) unknown_step_tile<LPP>(s, u, sub);
    } else {
      if constexpr (IL) { s.interleave(); u.interleave(); }
#pragma unroll 1
      for (unsigned g = 0; g < gens; g++) three_valued_step<LPP, IL>(s, u, sub);
      if constexpr (IL) { s.deinterleave(); u.deinterleave(); }
    }
    if (valid) {
      s.store(state_out + p * kN, sub);
      u.store(unknown_out + p * kN, sub);
    }
    if constexpr (kPrefetch) { s = ns; u = nu; }
  }
}

// LifeHistory record = state, history, marked, original (LifeHistory.hpp:8-12).  The reference
// has no stepping member for it; the build-defined step is state.Step(); history |= state
// (LifeAPI.hpp:1196-1216 and :255-260), repeated `gens` times.  marked / original are untouched.
template <int LPP>
__global__ void __launch_bounds__(kBlockU, 1) history_step_kernel(uint64_t* __restrict__ records, size_t n, unsigned gens) {
  const size_t gid = (size_t)blockIdx.x * kBlockU + threadIdx.x;
  const size_t p = gid / LPP;
  const int sub = threadIdx.x & (LPP - 1);
  const bool valid = p < n;
  Tile<LPP> s, h;
  if (valid) {
    s.load(records + p * 4 * kN, sub);
    h.load(records + p * 4 * kN + kN, sub);
  } else {
#pragma unroll
    for (int w = 0; w < Tile<LPP>::W; w++) s.lo[w] = s.hi[w] = h.lo[w] = h.hi[w] = 0;
  }
#pragma unroll 1
  for (unsigned g = 0; g < gens; g++) {
    s.step(sub);
#pragma unroll
    for (int w = 0; w < Tile<LPP>::W; w++) { h.lo[w] |= s.lo[w]; h.hi[w] |= s.hi[w]; }
  }
  if (valid) {
    s.store(records + p * 4 * kN, sub);
    h.store(records + p * 4 * kN + kN, sub);
  }
}

static inline unsigned grid_u(size_t n, int lpp) { return (unsigned)((n * (size_t)lpp + kBlockU - 1) / kBlockU); }

// lanes per pattern: 16 by default (W = 4 words per lane; the literal path keeps 10 count/plane tiles
// live at once), 4 / 8 / 32 on request
static inline int unknown_lpp(const lifeb200_ctx* ctx) { return ctx->lpp == 0 ? 16 : std::max(4, ctx->lpp); }
// row-interleaving costs about 17 integer instructions per register each way and saves 2 of 26 per
// generation; lifeb200_set_step_layout(1) pins the reference layout, (2) the interleaved one
constexpr unsigned kUnknownInterleaveFrom = 40;

// one wave of persistent CTAs (or fewer when the batch is small)
template <typename K>
static unsigned grid_unknown(const lifeb200_ctx* ctx, K kernel, size_t n, int lpp) {
  int per_sm = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, kBlockU, 0) != cudaSuccess || per_sm < 1) per_sm = 1;
  const unsigned resident = (unsigned)(ctx->sm_count > 0 ? ctx->sm_count : 148) * (unsigned)per_sm;
  return std::min(grid_u(n, lpp), resident);
}

template <bool IL>
static int launch_unknown_step_il(const lifeb200_ctx* ctx, cudaStream_t st, int lpp, const uint64_t* si, const uint64_t* ui,
                                  uint64_t* so, uint64_t* uo, size_t n, unsigned gens, int literal) {
  switch (lpp) {
    case 4: unknown_step_kernel<4, IL><<<grid_unknown(ctx, unknown_step_kernel<4, IL>, n, 4), kBlockU, 0, st>>>(si, ui, so, uo, n, gens, literal); break;
    case 8: unknown_step_kernel<8, IL><<<grid_unknown(ctx, unknown_step_kernel<8, IL>, n, 8), kBlockU, 0, st>>>(si, ui, so, uo, n, gens, literal); break;
    case 16: unknown_step_kernel<16, IL><<<grid_unknown(ctx, unknown_step_kernel<16, IL>, n, 16), kBlockU, 0, st>>>(si, ui, so, uo, n, gens, literal); break;
    default: unknown_step_kernel<32, IL><<<grid_unknown(ctx, unknown_step_kernel<32, IL>, n, 32), kBlockU, 0, st>>>(si, ui, so, uo, n, gens, literal); break;
  }
  return LIFEB200_OK;
}

int launch_unknown_step(lifeb200_ctx* ctx, cudaStream_t st, const uint64_t* si, const uint64_t* ui, uint64_t* so,
                        uint64_t* uo, size_t n, unsigned gens) {
  if (n == 0) return LIFEB200_OK;
  const int lpp = unknown_lpp(ctx);
  const bool il = ctx->step_layout == 0 ? gens >= kUnknownInterleaveFrom : ctx->step_layout >= 2;
  if (il) launch_unknown_step_il<true>(ctx, st, lpp, si, ui, so, uo, n, gens, ctx->unknown_literal);
  else launch_unknown_step_il<false>(ctx, st, lpp, si, ui, so, uo, n, gens, ctx->unknown_literal);
  LIFEB200_LAUNCHED(ctx);
  return LIFEB200_OK;
}

}  // namespace lifeb200

using namespace lifeb200;

extern "C" {

int lifeb200_unknown_step_batch_dev(lifeb200_ctx* ctx, const uint64_t* state_in, const uint64_t* unknown_in,
                                    uint64_t* state_out, uint64_t* unknown_out, size_t n, unsigned gens) {
  LIFEB200_CHECK_CTX(ctx);
  if (n == 0) return LIFEB200_OK;
  if (!state_in || !unknown_in || !state_out || !unknown_out) return set_error(LIFEB200_E_ARG, "null plane pointer");
  if (!aligned16(state_in) || !aligned16(unknown_in) || !aligned16(state_out) || !aligned16(unknown_out))
    return set_error(LIFEB200_E_ARG, "plane pointers must be 16-byte aligned");
  if (gens == 0) {
    if (state_in != state_out)
      LIFEB200_CUDA(cudaMemcpyAsync(state_out, state_in, n * LIFEB200_PATTERN_BYTES, cudaMemcpyDeviceToDevice, ctx->stream));
    if (unknown_in != unknown_out)
      LIFEB200_CUDA(cudaMemcpyAsync(unknown_out, unknown_in, n * LIFEB200_PATTERN_BYTES, cudaMemcpyDeviceToDevice, ctx->stream));
    return LIFEB200_OK;
  }
  return launch_unknown_step(ctx, ctx->stream, state_in, unknown_in, state_out, unknown_out, n, gens);
}

int lifeb200_unknown_step_batch_host(lifeb200_ctx* ctx, const uint64_t* state_in, const uint64_t* unknown_in,
                                     uint64_t* state_out, uint64_t* unknown_out, size_t n, unsigned gens) {
  LIFEB200_CHECK_CTX(ctx);
  if (n == 0) return LIFEB200_OK;
  if (!state_in || !unknown_in || !state_out || !unknown_out) return set_error(LIFEB200_E_ARG, "null plane pointer");
  // chunked, two staging buffers per slot (state | unknown); slot streams overlap copy and compute
  const size_t chunk = std::min<size_t>(n, (size_t)1 << 15);
  const size_t nchunks = (n + chunk - 1) / chunk;
  const int used = (int)std::min<size_t>(kSlots, nchunks);
  for (int s = 0; s < used; s++) {
    int rc = ensure_slot(ctx, s, chunk * LIFEB200_PATTERN_BYTES, chunk * LIFEB200_PATTERN_BYTES);
    if (rc) return rc;
  }
  for (size_t c = 0; c < nchunks; c++) {
    Slot& sl = ctx->slots[c % kSlots];
    const size_t off = c * chunk, cnt = std::min(chunk, n - off), bytes = cnt * LIFEB200_PATTERN_BYTES;
    uint64_t* ds = static_cast<uint64_t*>(sl.dbuf);
    uint64_t* du = static_cast<uint64_t*>(sl.dbuf2);
    LIFEB200_CUDA(cudaMemcpyAsync(ds, state_in + off * kN, bytes, cudaMemcpyHostToDevice, sl.stream));
    LIFEB200_CUDA(cudaMemcpyAsync(du, unknown_in + off * kN, bytes, cudaMemcpyHostToDevice, sl.stream));
    if (gens) {
      int rc = launch_unknown_step(ctx, sl.stream, ds, du, ds, du, cnt, gens);
      if (rc) return rc;
    }
    LIFEB200_CUDA(cudaMemcpyAsync(state_out + off * kN, ds, bytes, cudaMemcpyDeviceToHost, sl.stream));
    LIFEB200_CUDA(cudaMemcpyAsync(unknown_out + off * kN, du, bytes, cudaMemcpyDeviceToHost, sl.stream));
  }
  for (int s = 0; s < used; s++) LIFEB200_CUDA(cudaStreamSynchronize(ctx->slots[s].stream));
  return LIFEB200_OK;
}

static int launch_history(lifeb200_ctx* ctx, cudaStream_t st, uint64_t* records, size_t n, unsigned gens) {
  int lpp = std::max(4, effective_lpp(ctx));
  LIFEB200_DISPATCH_LPP_MIN4(lpp, (history_step_kernel<LPP><<<grid_u(n, LPP), kBlockU, 0, st>>>(records, n, gens)));
  LIFEB200_LAUNCHED(ctx);
  return LIFEB200_OK;
}

int lifeb200_history_step_batch_dev(lifeb200_ctx* ctx, uint64_t* records, size_t n, unsigned gens) {
  LIFEB200_CHECK_CTX(ctx);
  if (n == 0 || gens == 0) return LIFEB200_OK;
  if (!records || !aligned16(records)) return set_error(LIFEB200_E_ARG, "bad record pointer");
  return launch_history(ctx, ctx->stream, records, n, gens);
}

int lifeb200_history_step_batch_host(lifeb200_ctx* ctx, uint64_t* records, size_t n, unsigned gens) {
  LIFEB200_CHECK_CTX(ctx);
  if (n == 0 || gens == 0) return LIFEB200_OK;
  if (!records) return set_error(LIFEB200_E_ARG, "null record pointer");
  return host_pipeline_impl(ctx, records, records, n, 4 * LIFEB200_PATTERN_BYTES, 4 * LIFEB200_PATTERN_BYTES, false, (size_t)1 << 14,
                            [&](cudaStream_t st, const uint64_t* din, uint64_t*, size_t cnt) -> int {
                              return launch_history(ctx, st, const_cast<uint64_t*>(din), cnt, gens);
                            });
}

}  // extern "C"
