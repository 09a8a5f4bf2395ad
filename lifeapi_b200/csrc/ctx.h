// ctx.h -- internal: context object, error plumbing and launch helpers shared by the .cu files.
#pragma once
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

#include "../../include/lifeapi_b200.h"

namespace lifeb200 {

constexpr int kSlots = 3;  // host-path pipeline depth (H2D | kernel | D2H in flight).  6 slots were tried for the 8-rank
                           // e2e case and lost (64.6 against 63.3 ms on the GPUs behind the shared PCIe uplink)

struct Slot {
  cudaStream_t stream = nullptr;
  void* dbuf = nullptr;   // device staging
  void* dbuf2 = nullptr;  // second device staging (outputs larger than inputs)
  size_t cap = 0, cap2 = 0;
};

}  // namespace lifeb200

struct lifeb200_ctx {
  int device = 0;
  int sm_count = 0;
  int cc_major = 0, cc_minor = 0;
  int clock_khz = 0;
  size_t total_mem = 0;
  int lpp = 0;  // 0 = default
  int step_layout = 0;  // 0 = auto, 1 = reference, 2 / 4 / 8 = row classes per column (lifeb200_set_step_layout)
  int unknown_literal = 0;  // 1 = three-valued step always through the literal cover (lifeb200_set_three_valued_path)
  cudaStream_t own_stream = nullptr;
  cudaStream_t stream = nullptr;  // == own_stream unless lifeb200_set_stream
  cudaEvent_t ev_start = nullptr, ev_stop = nullptr;
  uint64_t launches = 0;
  lifeb200::Slot slots[lifeb200::kSlots];
  // small device scratch for counters: [0]=n_hits [1]=n_skipped [2]=pop_sum [3]=hash_xor
  unsigned long long* d_counters = nullptr;
  unsigned long long* h_counters = nullptr;  // pinned mirror
  // grow-only device scratch of the host-facing enumeration / survey calls (hit records, catalyst and projectile
  // tables): allocated on first use, reused by every later call, freed by lifeb200_destroy -- cudaMalloc / cudaFree
  // cost milliseconds and cudaFree synchronises the device, so they stay out of the per-batch path
  void* d_hitbuf = nullptr;
  size_t hitbuf_cap = 0;
  void* d_tabbuf = nullptr;
  size_t tabbuf_cap = 0;
};

namespace lifeb200 {

int set_error(int code, const char* fmt, ...);
int cuda_fail(cudaError_t e, const char* what, const char* file, int line);
int ensure_slot(lifeb200_ctx* ctx, int slot, size_t bytes, size_t bytes2);
int ensure_scratch(void*& buf, size_t& cap, size_t bytes);  // grow-only; the old block is released after a stream sync

#define LIFEB200_CUDA(expr)                                                        \
  do {                                                                             \
    cudaError_t _e = (expr);                                                       \
    if (_e != cudaSuccess) return lifeb200::cuda_fail(_e, #expr, __FILE__, __LINE__); \
  } while (0)

#define LIFEB200_CHECK_CTX(ctx)                                              \
  do {                                                                       \
    if (!(ctx)) return lifeb200::set_error(LIFEB200_E_ARG, "null context"); \
    LIFEB200_CUDA(cudaSetDevice((ctx)->device));                             \
  } while (0)

#define LIFEB200_LAUNCHED(ctx)       \
  do {                               \
    (ctx)->launches++;               \
    LIFEB200_CUDA(cudaGetLastError()); \
  } while (0)

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

constexpr int kDefaultLpp = 8;
inline int effective_lpp(const lifeb200_ctx* ctx) { return ctx->lpp ? ctx->lpp : kDefaultLpp; }

// dispatch a functor templated on lanes-per-pattern
#define LIFEB200_DISPATCH_LPP(lpp, CALL) \
  switch (lpp) {                         \
    case 1: { constexpr int LPP = 1; CALL; } break;   \
    case 2: { constexpr int LPP = 2; CALL; } break;   \
    case 4: { constexpr int LPP = 4; CALL; } break;   \
    case 8: { constexpr int LPP = 8; CALL; } break;   \
    case 16: { constexpr int LPP = 16; CALL; } break; \
    case 32: { constexpr int LPP = 32; CALL; } break; \
    default: return lifeb200::set_error(LIFEB200_E_ARG, "lanes per pattern must be 1,2,4,8,16,32"); \
  }

}  // namespace lifeb200
