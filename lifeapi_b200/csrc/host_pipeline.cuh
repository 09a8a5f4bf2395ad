// host_pipeline.cuh -- internal: the host-pointer pipeline shared by the *_host entry points.
//
// Chunk c uses slot c % kSlots; each slot's stream runs H2D -> kernel -> D2H, so the copies of
// neighbouring chunks overlap the kernel of the current one.  The copies go straight from / to the caller's
// buffers: with PINNED memory (lifeb200_host_alloc, cudaHostRegister) they are asynchronous and overlap; with
// pageable memory CUDA stages them itself, the calls are still correct but mostly serial.
// On any failure the slot streams already used are drained before the error is returned, so no copy is in
// flight towards the caller's buffers after the call has come back.
#pragma once
#include <algorithm>
#include <new>
#include <vector>

#include "ctx.h"

namespace lifeb200 {

// What stays exposed is the first chunk's upload and the last chunk's download, so a large batch starts and
// ends with small chunks (4 Ki, then 20 Ki items: each still uploads in less time than its predecessor
// computes) and runs `big`-item chunks in between.
inline std::vector<size_t> chunk_plan(size_t n, size_t big = (size_t)1 << 16) {
  const size_t kEdge0 = std::min<size_t>((size_t)1 << 12, big), kEdge1 = std::min<size_t>((size_t)20 << 10, big);
  std::vector<size_t> plan;
  if (n <= kEdge0) {
    plan.push_back(n);
  } else if (n < 4 * big || big < ((size_t)1 << 16)) {
    for (size_t off = 0; off < n; off += big) plan.push_back(std::min(big, n - off));
  } else {
    plan.push_back(kEdge0);
    plan.push_back(kEdge1);
    for (size_t left = n - 2 * (kEdge0 + kEdge1); left > 0;) {
      const size_t c = std::min(big, left);
      plan.push_back(c);
      left -= c;
    }
    plan.push_back(kEdge1);
    plan.push_back(kEdge0);
  }
  return plan;
}

inline void drain_slots(lifeb200_ctx* ctx, int used) {
  for (int s = 0; s < used; s++)
    if (ctx->slots[s].stream) cudaStreamSynchronize(ctx->slots[s].stream);
}

// One input array and one output array of fixed-size items.  launch(stream, d_in, d_out, count) enqueues the
// kernel(s); d_out == d_in when !out_separate (in-place kernels).  extra_out (optional): a second output of
// extra_bytes_per bytes per item that the kernel writes behind the main output in the slot's second buffer.
template <class LaunchFn>
int host_pipeline_impl(lifeb200_ctx* ctx, const void* in, void* out, size_t n, size_t in_bytes_per, size_t out_bytes_per,
                       bool out_separate, size_t big, LaunchFn launch) {
  if (n == 0) return LIFEB200_OK;
  std::vector<size_t> plan;
  try {
    plan = chunk_plan(n, big);
  } catch (const std::bad_alloc&) {
    return set_error(LIFEB200_E_ARG, "out of host memory");
  }
  const size_t chunk = *std::max_element(plan.begin(), plan.end());
  const int used = (int)std::min<size_t>(kSlots, plan.size());
  for (int s = 0; s < used; s++)
    if (int rc = ensure_slot(ctx, s, chunk * in_bytes_per, out_separate ? chunk * out_bytes_per : 0)) return rc;
  auto body = [&]() -> int {
    size_t off = 0;
    for (size_t c = 0; c < plan.size(); c++) {
      Slot& sl = ctx->slots[c % kSlots];
      const size_t cnt = plan[c];
      const char* src = static_cast<const char*>(in) + off * in_bytes_per;
      char* dst = static_cast<char*>(out) + off * out_bytes_per;
      LIFEB200_CUDA(cudaMemcpyAsync(sl.dbuf, src, cnt * in_bytes_per, cudaMemcpyHostToDevice, sl.stream));
      void* dout = out_separate ? sl.dbuf2 : sl.dbuf;
      if (int rc = launch(sl.stream, static_cast<const uint64_t*>(sl.dbuf), static_cast<uint64_t*>(dout), cnt)) return rc;
      LIFEB200_CUDA(cudaMemcpyAsync(dst, dout, cnt * out_bytes_per, cudaMemcpyDeviceToHost, sl.stream));
      off += cnt;
    }
    for (int s = 0; s < used; s++) LIFEB200_CUDA(cudaStreamSynchronize(ctx->slots[s].stream));
    return LIFEB200_OK;
  };
  const int rc = body();
  if (rc) drain_slots(ctx, used);
  return rc;
}

template <class LaunchFn>
int host_pipeline(lifeb200_ctx* ctx, const void* in, void* out, size_t n, size_t in_bytes_per, size_t out_bytes_per,
                  bool out_separate, LaunchFn launch) {
  return host_pipeline_impl(ctx, in, out, n, in_bytes_per, out_bytes_per, out_separate, (size_t)1 << 16, launch);
}

}  // namespace lifeb200
