// ctx.cu -- context, memory, error and timing entry points of the C ABI (include/lifeapi_b200.h).
#include <cstring>
#include <new>

#include "ctx.h"

namespace lifeb200 {

static thread_local char g_err[512] = "";

int set_error(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof g_err, fmt, ap);
  va_end(ap);
  return code;
}

int cuda_fail(cudaError_t e, const char* what, const char* file, int line) {
  const char* base = strrchr(file, '/');
  int code = (e == cudaErrorNoDevice || e == cudaErrorInsufficientDriver) ? LIFEB200_E_NODEVICE : LIFEB200_E_CUDA;
  return set_error(code, "CUDA error %d (%s) at %s:%d: %s", (int)e, cudaGetErrorString(e), base ? base + 1 : file,
                   line, what);
}

int ensure_slot(lifeb200_ctx* ctx, int s, size_t bytes, size_t bytes2) {
  Slot& sl = ctx->slots[s];
  if (!sl.stream) LIFEB200_CUDA(cudaStreamCreateWithFlags(&sl.stream, cudaStreamNonBlocking));
  if (sl.cap < bytes) {
    if (sl.dbuf) LIFEB200_CUDA(cudaFree(sl.dbuf));
    sl.dbuf = nullptr; sl.cap = 0;
    LIFEB200_CUDA(cudaMalloc(&sl.dbuf, bytes));
    sl.cap = bytes;
  }
  if (sl.cap2 < bytes2) {
    if (sl.dbuf2) LIFEB200_CUDA(cudaFree(sl.dbuf2));
    sl.dbuf2 = nullptr; sl.cap2 = 0;
    LIFEB200_CUDA(cudaMalloc(&sl.dbuf2, bytes2));
    sl.cap2 = bytes2;
  }
  return LIFEB200_OK;
}

int ensure_scratch(void*& buf, size_t& cap, size_t bytes) {
  if (cap >= bytes) return LIFEB200_OK;
  if (buf) LIFEB200_CUDA(cudaFree(buf));  // synchronises the device: earlier users of the block are done
  buf = nullptr; cap = 0;
  // grow by at least half so that a caller walking through slightly larger batches reallocates O(log) times
  size_t want = bytes + bytes / 2;
  if (cudaMalloc(&buf, want) != cudaSuccess) {
    (void)cudaGetLastError();
    want = bytes;
    LIFEB200_CUDA(cudaMalloc(&buf, want));
  }
  cap = want;
  return LIFEB200_OK;
}

}  // namespace lifeb200

using namespace lifeb200;

extern "C" {

int lifeb200_abi_version(void) { return LIFEB200_ABI_VERSION; }
const char* lifeb200_last_error(void) { return g_err; }

int lifeb200_device_count(int* count) {
  if (!count) return set_error(LIFEB200_E_ARG, "null count");
  *count = 0;
  LIFEB200_CUDA(cudaGetDeviceCount(count));
  return LIFEB200_OK;
}

int lifeb200_create(int device, lifeb200_ctx** out) {
  if (!out) return set_error(LIFEB200_E_ARG, "null out");
  *out = nullptr;
  int count = 0;
  LIFEB200_CUDA(cudaGetDeviceCount(&count));
  if (count <= 0) return set_error(LIFEB200_E_NODEVICE, "no CUDA device; this library has no CPU fallback");
  if (device < 0 || device >= count) return set_error(LIFEB200_E_ARG, "device %d out of range [0,%d)", device, count);
  LIFEB200_CUDA(cudaSetDevice(device));
  cudaDeviceProp prop;
  LIFEB200_CUDA(cudaGetDeviceProperties(&prop, device));
  if (prop.major != 10)
    return set_error(LIFEB200_E_NODEVICE, "device %d is sm_%d%d; kernels are built for sm_100a only", device, prop.major,
                     prop.minor);
  lifeb200_ctx* c = new (std::nothrow) lifeb200_ctx();
  if (!c) return set_error(LIFEB200_E_HOST, "out of host memory");
  c->device = device;
  c->sm_count = prop.multiProcessorCount;
  c->cc_major = prop.major;
  c->cc_minor = prop.minor;
  c->total_mem = prop.totalGlobalMem;
  int khz = 0;
  cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, device);
  c->clock_khz = khz;
  cudaError_t e;
  if ((e = cudaStreamCreateWithFlags(&c->own_stream, cudaStreamNonBlocking)) != cudaSuccess ||
      (e = cudaEventCreate(&c->ev_start)) != cudaSuccess || (e = cudaEventCreate(&c->ev_stop)) != cudaSuccess ||
      (e = cudaMalloc(&c->d_counters, 8 * sizeof(unsigned long long))) != cudaSuccess ||
      (e = cudaMallocHost(&c->h_counters, 8 * sizeof(unsigned long long))) != cudaSuccess) {
    lifeb200_destroy(c);
    return cuda_fail(e, "context resources", __FILE__, __LINE__);
  }
  c->stream = c->own_stream;
  *out = c;
  return LIFEB200_OK;
}

int lifeb200_destroy(lifeb200_ctx* c) {
  if (!c) return LIFEB200_OK;
  cudaSetDevice(c->device);
  if (c->own_stream) cudaStreamSynchronize(c->own_stream);
  for (auto& s : c->slots) {
    if (s.stream) { cudaStreamSynchronize(s.stream); cudaStreamDestroy(s.stream); }
    if (s.dbuf) cudaFree(s.dbuf);
    if (s.dbuf2) cudaFree(s.dbuf2);
  }
  if (c->d_hitbuf) cudaFree(c->d_hitbuf);
  if (c->d_tabbuf) cudaFree(c->d_tabbuf);
  if (c->d_dense) cudaFree(c->d_dense);
  if (c->d_tiles) cudaFree(c->d_tiles);
  if (c->d_counters) cudaFree(c->d_counters);
  if (c->h_counters) cudaFreeHost(c->h_counters);
  if (c->ev_start) cudaEventDestroy(c->ev_start);
  if (c->ev_stop) cudaEventDestroy(c->ev_stop);
  if (c->own_stream) cudaStreamDestroy(c->own_stream);
  delete c;
  return LIFEB200_OK;
}

int lifeb200_set_stream(lifeb200_ctx* ctx, void* cuda_stream) {
  LIFEB200_CHECK_CTX(ctx);
  ctx->stream = cuda_stream ? static_cast<cudaStream_t>(cuda_stream) : ctx->own_stream;
  return LIFEB200_OK;
}

int lifeb200_sync(lifeb200_ctx* ctx) {
  LIFEB200_CHECK_CTX(ctx);
  LIFEB200_CUDA(cudaStreamSynchronize(ctx->stream));
  return LIFEB200_OK;
}

int lifeb200_device_info(lifeb200_ctx* ctx, int* sm_count, int* cc_major, int* cc_minor, int* clock_khz,
                         size_t* total_mem) {
  if (!ctx) return set_error(LIFEB200_E_ARG, "null context");
  if (sm_count) *sm_count = ctx->sm_count;
  if (cc_major) *cc_major = ctx->cc_major;
  if (cc_minor) *cc_minor = ctx->cc_minor;
  if (clock_khz) *clock_khz = ctx->clock_khz;
  if (total_mem) *total_mem = ctx->total_mem;
  return LIFEB200_OK;
}

int lifeb200_set_step_layout(lifeb200_ctx* ctx, int layout) {
  LIFEB200_CHECK_CTX(ctx);
  if (layout != 0 && layout != 1 && layout != 2 && layout != 4 && layout != 8)
    return set_error(LIFEB200_E_ARG, "step layout must be 0 (auto), 1, 2, 4 or 8 (got %d)", layout);
  ctx->step_layout = layout;
  return LIFEB200_OK;
}

int lifeb200_set_three_valued_path(lifeb200_ctx* ctx, int path) {
  if (!ctx) return set_error(LIFEB200_E_ARG, "null context");
  if (path != 0 && path != 1) return set_error(LIFEB200_E_ARG, "three-valued path must be 0 (auto) or 1 (literal cover), got %d", path);
  ctx->unknown_literal = path;
  return LIFEB200_OK;
}

int lifeb200_set_lanes_per_pattern(lifeb200_ctx* ctx, int lpp) {
  if (!ctx) return set_error(LIFEB200_E_ARG, "null context");
  if (lpp != 0 && lpp != 1 && lpp != 2 && lpp != 4 && lpp != 8 && lpp != 16 && lpp != 32)
    return set_error(LIFEB200_E_ARG, "lanes per pattern must be 0,1,2,4,8,16,32 (got %d)", lpp);
  ctx->lpp = lpp;
  return LIFEB200_OK;
}

int lifeb200_malloc(lifeb200_ctx* ctx, void** dptr, size_t bytes) {
  LIFEB200_CHECK_CTX(ctx);
  if (!dptr) return set_error(LIFEB200_E_ARG, "null dptr");
  *dptr = nullptr;
  if (bytes == 0) return LIFEB200_OK;
  LIFEB200_CUDA(cudaMalloc(dptr, bytes));
  return LIFEB200_OK;
}
int lifeb200_free(lifeb200_ctx* ctx, void* dptr) {
  LIFEB200_CHECK_CTX(ctx);
  if (dptr) LIFEB200_CUDA(cudaFree(dptr));
  return LIFEB200_OK;
}
int lifeb200_host_alloc(lifeb200_ctx* ctx, void** hptr, size_t bytes) {
  LIFEB200_CHECK_CTX(ctx);
  if (!hptr) return set_error(LIFEB200_E_ARG, "null hptr");
  *hptr = nullptr;
  if (bytes == 0) return LIFEB200_OK;
  LIFEB200_CUDA(cudaMallocHost(hptr, bytes));
  return LIFEB200_OK;
}
int lifeb200_host_free(lifeb200_ctx* ctx, void* hptr) {
  LIFEB200_CHECK_CTX(ctx);
  if (hptr) LIFEB200_CUDA(cudaFreeHost(hptr));
  return LIFEB200_OK;
}
int lifeb200_host_register(lifeb200_ctx* ctx, void* hptr, size_t bytes) {
  LIFEB200_CHECK_CTX(ctx);
  if (!hptr || bytes == 0) return set_error(LIFEB200_E_ARG, "null or empty host range");
  LIFEB200_CUDA(cudaHostRegister(hptr, bytes, cudaHostRegisterPortable));
  return LIFEB200_OK;
}
int lifeb200_host_unregister(lifeb200_ctx* ctx, void* hptr) {
  LIFEB200_CHECK_CTX(ctx);
  if (!hptr) return set_error(LIFEB200_E_ARG, "null host pointer");
  LIFEB200_CUDA(cudaHostUnregister(hptr));
  return LIFEB200_OK;
}
int lifeb200_memcpy_h2d(lifeb200_ctx* ctx, void* dst, const void* src, size_t bytes) {
  LIFEB200_CHECK_CTX(ctx);
  if (bytes) LIFEB200_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, ctx->stream));
  return LIFEB200_OK;
}
int lifeb200_memcpy_d2h(lifeb200_ctx* ctx, void* dst, const void* src, size_t bytes) {
  LIFEB200_CHECK_CTX(ctx);
  if (bytes) LIFEB200_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, ctx->stream));
  return LIFEB200_OK;
}
int lifeb200_memset(lifeb200_ctx* ctx, void* dst, int value, size_t bytes) {
  LIFEB200_CHECK_CTX(ctx);
  if (bytes) LIFEB200_CUDA(cudaMemsetAsync(dst, value, bytes, ctx->stream));
  return LIFEB200_OK;
}

int lifeb200_timer_start(lifeb200_ctx* ctx) {
  LIFEB200_CHECK_CTX(ctx);
  LIFEB200_CUDA(cudaEventRecord(ctx->ev_start, ctx->stream));
  return LIFEB200_OK;
}
int lifeb200_timer_stop(lifeb200_ctx* ctx, float* elapsed_ms) {
  LIFEB200_CHECK_CTX(ctx);
  LIFEB200_CUDA(cudaEventRecord(ctx->ev_stop, ctx->stream));
  LIFEB200_CUDA(cudaEventSynchronize(ctx->ev_stop));
  float ms = 0.f;
  LIFEB200_CUDA(cudaEventElapsedTime(&ms, ctx->ev_start, ctx->ev_stop));
  if (elapsed_ms) *elapsed_ms = ms;
  return LIFEB200_OK;
}
int lifeb200_launch_count(lifeb200_ctx* ctx, uint64_t* launches) {
  if (!ctx || !launches) return set_error(LIFEB200_E_ARG, "null argument");
  *launches = ctx->launches;
  return LIFEB200_OK;
}

}  // extern "C"
