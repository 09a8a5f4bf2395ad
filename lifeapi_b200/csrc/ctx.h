// ctx.h -- internal: context object, error plumbing and launch helpers shared by the .cu files.
#pragma once
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <exception>
#include <new>
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
  // the enumeration tables in d_tabbuf are kept across calls, keyed by a hash of their words (csrc/match.cu)
  bool tab_valid = false;
  uint64_t tab_key = 0;
  unsigned tab_ncat = 0, tab_nproj = 0;
  // dense per-id results (8 B per id) and per-tile hit counts of the ordered compaction (csrc/match.cu)
  void* d_dense = nullptr;
  size_t dense_cap = 0;
  void* d_tiles = nullptr;
  size_t tiles_cap = 0;
};

namespace lifeb200 {

int set_error(int code, const char* fmt, ...);
int cuda_fail(cudaError_t e, const char* what, const char* file, int line);
int ensure_slot(lifeb200_ctx* ctx, int slot, size_t bytes, size_t bytes2);
int ensure_scratch(void*& buf, size_t& cap, size_t bytes);  // grow-only; the old block is released after a stream sync

// csrc/match.cu: run one shard, leave the hit records sorted by id in d_hits (capacity cap) and the counters
// n_hits, n_skipped, pop_sum, hash_xor in ctx->d_counters[0..3] AND ctx->h_counters[0..3]; the stream is synchronised
int survey_to_device(lifeb200_ctx* ctx, uint64_t first_id, size_t n, uint64_t seed, int dist, unsigned gens,
                     lifeb200_hit* d_hits, size_t cap);
int enumerate_to_device(lifeb200_ctx* ctx, const uint64_t* catalysts, unsigned n_catalysts, const uint64_t* projectiles,
                        unsigned n_projectiles, uint64_t id_begin, uint64_t id_end, unsigned gens, unsigned stable_interval,
                        lifeb200_hit* d_hits, size_t cap);

int reserve_ids(lifeb200_ctx* ctx, size_t n);  // size the dense-result / tile scratch for calls over n ids

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

// the same for kernels that keep three or more tiles in registers: with 1 or 2 lanes per pattern (64 / 32 words
// per lane and tile) they only measured register spills (1.2 k LDL/STL per kernel), so those instantiations
// are not built; callers clamp the lane grouping to 4..32 first
#define LIFEB200_DISPATCH_LPP_MIN4(lpp, CALL) \
  switch (lpp) {                              \
    case 4: { constexpr int LPP = 4; CALL; } break;   \
    case 8: { constexpr int LPP = 8; CALL; } break;   \
    case 16: { constexpr int LPP = 16; CALL; } break; \
    case 32: { constexpr int LPP = 32; CALL; } break; \
    default: return lifeb200::set_error(LIFEB200_E_ARG, "lanes per pattern must be 4,8,16,32 for this kernel"); \
  }

// No C++ exception may cross the extern "C" boundary: entry points that allocate host memory or start threads run
// their body through this (std::bad_alloc, std::system_error from std::thread, ...) and report LIFEB200_E_HOST.
template <class F>
inline int guarded(F body) {
  try {
    return body();
  } catch (const std::bad_alloc&) {
    return set_error(LIFEB200_E_HOST, "out of host memory");
  } catch (const std::exception& e) {
    return set_error(LIFEB200_E_HOST, "host-side failure: %s", e.what());
  } catch (...) {
    return set_error(LIFEB200_E_HOST, "host-side failure");
  }
}

}  // namespace lifeb200
