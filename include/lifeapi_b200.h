/* lifeapi_b200.h -- C ABI of the B200-native batched stepping engine for LifeAPI.
 *
 * The reference (scorbiclife/LifeAPI @ 4e9bb99) has no FFI layer: its API is the header
 * LifeAPI.hpp.  These entry points are what a drop-in replacement header binds for the stepping
 * path (see include/LifeAPI_b200.hpp and INTEGRATION.md).  Each function cites the reference
 * interface it replaces.  Plain pointers and sizes only; no torch / C++ types.
 *
 * Conventions
 *   pattern   = 64 little-endian uint64_t = 512 bytes; word i = column x=i, bit j = row y=j
 *               (LifeAPI.hpp:39-40,131).  A batch is a flat array of patterns (== LifeState[]).
 *   return    = 0 on success, non-zero on failure (LIFEB200_E_*); lifeb200_last_error() gives
 *               the message for the calling thread.  The reference has no recoverable errors
 *               (SURVEY.md 8b), so header wrappers abort on non-zero.
 *   *_dev     = pointers are DEVICE pointers on the context's GPU; work is enqueued on the
 *               context's stream and is asynchronous; call lifeb200_sync().
 *   *_host    = pointers are HOST pointers; the batch is cut into chunks that go H2D -> kernel -> D2H on
 *               three streams, straight from / to the caller's buffers, and the call returns when the
 *               result is in the caller's buffer.  With PINNED buffers (lifeb200_host_alloc or
 *               cudaHostRegister) the copies overlap the kernels; pageable buffers are correct but the
 *               copies are staged by the driver and mostly serial.  On failure every copy has finished
 *               (or been abandoned) before the call returns.
 *   There is no CPU fallback: without a usable CUDA device every compute call fails.
 */
#ifndef LIFEAPI_B200_H
#define LIFEAPI_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define LIFEB200_ABI_VERSION 3
#define LIFEB200_WORDS 64          /* LifeAPI.hpp:12  const int N = 64 */
#define LIFEB200_PATTERN_BYTES 512 /* sizeof(LifeState) */

enum {
  LIFEB200_OK = 0,
  LIFEB200_E_CUDA = 1,      /* CUDA runtime error (message has the cudaError string) */
  LIFEB200_E_ARG = 2,       /* bad argument (null / misaligned pointer, bad enum) */
  LIFEB200_E_OVERFLOW = 3,  /* hit buffer too small: *n_hits holds the required count */
  LIFEB200_E_NODEVICE = 4,  /* no CUDA device / wrong architecture */
  LIFEB200_E_NCCL = 5,      /* NCCL could not be loaded or a collective failed (multi-GPU groups only) */
  LIFEB200_E_HOST = 6       /* host-side failure inside the library (out of host memory, thread creation) */
};

typedef struct lifeb200_ctx lifeb200_ctx; /* opaque: device, stream, scratch, pinned staging */

/* ---- context / device management ------------------------------------------------------- */
int lifeb200_abi_version(void);
const char* lifeb200_last_error(void);
int lifeb200_device_count(int* count);
int lifeb200_create(int device, lifeb200_ctx** out);
int lifeb200_destroy(lifeb200_ctx* ctx);
/* Use an externally owned cudaStream_t (e.g. torch's current stream); NULL restores the own one. */
int lifeb200_set_stream(lifeb200_ctx* ctx, void* cuda_stream);
int lifeb200_sync(lifeb200_ctx* ctx);
int lifeb200_device_info(lifeb200_ctx* ctx, int* sm_count, int* cc_major, int* cc_minor, int* clock_khz,
                         size_t* total_mem);
/* Choose the lane-group width of the step kernels: lanes per pattern in {1,2,4,8,16,32};
 * 0 restores the default.  Results never depend on it. */
int lifeb200_set_lanes_per_pattern(lifeb200_ctx* ctx, int lpp);
/* Choose the REGISTER layout lifeb200_step_batch_* steps in (HBM layout and results never change):
 * 1 = the reference's, 2 = even/odd rows per register, 4 / 8 = rows by class mod 4 / 8 with 2 / 4
 * patterns bit-interleaved per register (fewer shifts per generation, a conversion per launch);
 * 0 restores the default (chosen from the generation count). */
int lifeb200_set_step_layout(lifeb200_ctx* ctx, int layout);

/* ---- memory ---------------------------------------------------------------------------- */
int lifeb200_malloc(lifeb200_ctx* ctx, void** dptr, size_t bytes);
int lifeb200_free(lifeb200_ctx* ctx, void* dptr);
int lifeb200_host_alloc(lifeb200_ctx* ctx, void** hptr, size_t bytes); /* pinned */
int lifeb200_host_free(lifeb200_ctx* ctx, void* hptr);
/* Page-lock a host range the CALLER owns (a std::vector<LifeState>'s storage, a numpy array) so that the *_host calls
 * copy from / to it asynchronously and overlap with the kernels, as they do for lifeb200_host_alloc memory; pinning
 * costs about as much as one pageable copy of the range, so it pays for buffers that are used more than once.
 * Unregister before the memory is freed. */
int lifeb200_host_register(lifeb200_ctx* ctx, void* hptr, size_t bytes);
int lifeb200_host_unregister(lifeb200_ctx* ctx, void* hptr);
int lifeb200_memcpy_h2d(lifeb200_ctx* ctx, void* dst, const void* src, size_t bytes); /* async */
int lifeb200_memcpy_d2h(lifeb200_ctx* ctx, void* dst, const void* src, size_t bytes); /* async */
int lifeb200_memset(lifeb200_ctx* ctx, void* dst, int value, size_t bytes);           /* async */

/* ---- timing on the context's stream (CUDA events) -------------------------------------- */
int lifeb200_timer_start(lifeb200_ctx* ctx);
int lifeb200_timer_stop(lifeb200_ctx* ctx, float* elapsed_ms); /* records, syncs, returns ms */
/* Number of kernels this library launched on ctx since creation. */
int lifeb200_launch_count(lifeb200_ctx* ctx, uint64_t* launches);

/* ---- A4/A5  LifeState::Step(), Step(unsigned)  (LifeAPI.hpp:866,877-881,1196-1216) ------ */
/* out[p] = in[p] stepped `gens` generations, for p < n.  in == out is allowed. gens may be 0. */
int lifeb200_step_batch_dev(lifeb200_ctx* ctx, const uint64_t* in, uint64_t* out, size_t n, unsigned gens);
int lifeb200_step_batch_host(lifeb200_ctx* ctx, const uint64_t* in, uint64_t* out, size_t n, unsigned gens);

/* ---- A14  soup source (replaces RandomState, LifeAPI.hpp:63-69, with a seeded generator) - */
/* word(p,i) = splitmix64(seed + 64*(first_id+p) + i); dist 0 = full 64-bit words,
 * dist 1 = the reference-shaped draw uniform in [2^61, 2^62] (LifeAPI.hpp:21-22). */
int lifeb200_soup_dev(lifeb200_ctx* ctx, uint64_t* out, uint64_t first_id, size_t n, uint64_t seed, int dist);

/* ---- A8  NeighbourCount(const LifeState&) / CountNeighbourhood --------------------------
 * (NeighbourCount.hpp:40-70, LifeAPI.hpp:909-952).  planes[p] = 4 x 64 words in the member
 * order bit3, bit2, bit1, bit0 (NeighbourCount.hpp:7-11): 3x3 sum INCLUDING the centre. */
int lifeb200_neighbour_count_dev(lifeb200_ctx* ctx, const uint64_t* in, uint64_t* planes, size_t n);
int lifeb200_neighbour_count_host(lifeb200_ctx* ctx, const uint64_t* in, uint64_t* planes, size_t n);

/* NeighbourCount arithmetic on batches of count records (NeighbourCount.hpp:71-102; a record = bit3, bit2, bit1, bit0
 * = 2048 B, as the constructor above writes them): out[p] = a[p].Add(b[p], incarry[p]) -- a 4-stage ripple with a
 * carry-in plane (n x 512 B; NULL = the empty plane, i.e. operator+), carry-out discarded; Subtract(o) = Add(~o,
 * all-ones) (operator-); WithExactly(k) = AND of the planes, negated where k has a 0 bit (out: n x 512 B). */
int lifeb200_nc_add_dev(lifeb200_ctx* ctx, const uint64_t* a, const uint64_t* b, const uint64_t* incarry, uint64_t* out, size_t n);
int lifeb200_nc_subtract_dev(lifeb200_ctx* ctx, const uint64_t* a, const uint64_t* b, uint64_t* out, size_t n);
int lifeb200_nc_with_exactly_dev(lifeb200_ctx* ctx, const uint64_t* counts, unsigned k, uint64_t* out, size_t n);

/* ---- A11/A12  fused step + LifeState::Contains(const LifeTarget&) (LifeTarget.hpp:44-51) - */
typedef struct lifeb200_hit {
  uint64_t id;             /* pattern / placement id */
  uint32_t g_first_false;  /* first generation g>=1 at which Contains(target) was false */
  uint32_t g_recovered;    /* first generation of the run of `stable_interval` matches */
} lifeb200_hit;

/* For each pattern p (device array `patterns`, n x 512 B) with target targets[target_index[p]]
 * (targets: n_targets x {wanted[64], unwanted[64]} = 1024 B each, LifeTarget.hpp:5-7; a NULL
 * target_index means target 0 for all): step up to `gens` generations evaluating Contains at every
 * generation; a HIT is a pattern whose predicate was false at some generation >= 1 and later
 * true for `stable_interval` consecutive generations.  Hits are compacted (dense per-id results + scan,
 * in id order, no sort: csrc/match.cu) into `hits` (device, capacity `cap`); *n_hits (device) receives the
 * count.  Patterns are not written back.  first_id is added to p in the reported id. */
int lifeb200_step_match_dev(lifeb200_ctx* ctx, const uint64_t* patterns, const uint64_t* targets,
                            const uint32_t* target_index, size_t n, uint64_t first_id, unsigned gens,
                            unsigned stable_interval, lifeb200_hit* hits, size_t cap,
                            unsigned long long* n_hits);

/* Host form: patterns (n x 512 B), targets (n_targets x 1024 B), target_index (n, may be NULL) and hits are HOST
 * arrays; *n_hits (host) receives the count; LIFEB200_E_OVERFLOW if cap is too small (the first cap records, in id
 * order, are still delivered). */
int lifeb200_step_match_host(lifeb200_ctx* ctx, const uint64_t* patterns, const uint64_t* targets,
                             const uint32_t* target_index, size_t n, size_t n_targets, uint64_t first_id, unsigned gens,
                             unsigned stable_interval, lifeb200_hit* hits, size_t cap, uint64_t* n_hits);

/* Allocation-free variant of lifeb200_step_match_dev.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  Same arguments, same hit records in the same order; the call (and everything it
 * enqueues) performs no cudaMalloc / cudaFree / allocating library call, on any invocation including the first.
 * Its scratch (8 bytes per pattern of dense results + the per-tile counts of the ordered compaction) is a file-scope
 * `static __device__ __align__(256)` array in csrc/match.cu of LIFEB200_STEP_MATCH_STATIC_MAX patterns (2 MiB + 1 KiB),
 * which every GPU that loads the library holds until exit; its device address is taken once per device with
 * cudaGetSymbolAddress.  n > LIFEB200_STEP_MATCH_STATIC_MAX is rejected with LIFEB200_E_ARG (split the batch, or
 * use lifeb200_step_match_dev).  EVERY call on a device shares that one array and it keeps the previous call's
 * contents (nothing in it has to start at zero: each call overwrites the words it reads): two calls that overlap on
 * the same GPU -- different contexts or streams -- RACE; serialise them (one stream, or events). */
#define LIFEB200_STEP_MATCH_STATIC_MAX 262144u
int lifeb200_step_match_static_dev(lifeb200_ctx* ctx, const uint64_t* patterns, const uint64_t* targets,
                                   const uint32_t* target_index, size_t n, uint64_t first_id, unsigned gens,
                                   unsigned stable_interval, lifeb200_hit* hits, size_t cap,
                                   unsigned long long* n_hits);

/* ---- A11  LifeState::Contains(const LifeTarget&, int dx, int dy) (LifeTarget.hpp:38-42 -> LifeAPI.hpp:399-422):
 * contains[p] = 1 iff targets[p] (or the one shared target, n_targets == 1) placed at offset (dx, dy) fits
 * states[p]: every wanted cell ON and every unwanted cell OFF in state word (i + dx) & 63 rotated right by dy. */
int lifeb200_contains_target_batch_dev(lifeb200_ctx* ctx, const uint64_t* states, size_t n, const uint64_t* targets,
                                       size_t n_targets, int dx, int dy, uint8_t* contains);
int lifeb200_contains_target_batch_host(lifeb200_ctx* ctx, const uint64_t* states, size_t n, const uint64_t* targets,
                                        size_t n_targets, int dx, int dy, uint8_t* contains);

/* ---- A13 + A11  collision enumeration --------------------------
 * Placement id = ((c * n_projectiles + g) * 64 + dx) * 64 + dy.  pattern = catalyst[c] |
 * projectile[g].Moved(dx,dy) (LifeAPI.hpp:699-717, :247-253).  Placements whose ZOIs
 * (LifeAPI.hpp:521-536) intersect at generation 0 are counted in *n_skipped and not stepped.
 * Target of a placement = LifeTarget(catalyst[c]) (LifeTarget.hpp:10-13).  Everything is built on
 * the device from the two small tables (HOST pointers: n_catalysts / n_projectiles x 512 B).
 * Ids [id_begin, id_end) are processed (a rank's shard).  hits (HOST, capacity cap) come back
 * sorted by id; returns LIFEB200_E_OVERFLOW with *n_hits = required if cap is too small. */
int lifeb200_enumerate_collisions(lifeb200_ctx* ctx, const uint64_t* catalysts, unsigned n_catalysts,
                                  const uint64_t* projectiles, unsigned n_projectiles, uint64_t id_begin,
                                  uint64_t id_end, unsigned gens, unsigned stable_interval,
                                  lifeb200_hit* hits, size_t cap, uint64_t* n_hits, uint64_t* n_skipped);

/* ---- config 5  on-device soup survey ---------------------------------------------------
 * Soups first_id..first_id+n-1 are generated in registers (never touch HBM), stepped `gens`
 * generations (gens >= 2); a soup is ACTIVE iff state(gens) != state(gens-2)
 * (operator== LifeAPI.hpp:213-220).  Active ids with their final population (GetPop,
 * LifeAPI.hpp:290-298) are compacted into `hits` (HOST, sorted by id; hit.g_first_false = pop,
 * hit.g_recovered = 0).  *pop_sum = sum of the final populations of ALL soups, *hash_xor = XOR of
 * a 64-bit fold of every final state (checksum for size-independent parity). */
int lifeb200_soup_survey(lifeb200_ctx* ctx, uint64_t first_id, size_t n, uint64_t seed, int dist,
                         unsigned gens, lifeb200_hit* hits, size_t cap, uint64_t* n_hits,
                         uint64_t* pop_sum, uint64_t* hash_xor);
/* Device-resident forms of the two calls above: `d_hits` is a DEVICE buffer of `cap` records on the context's
 * GPU; the records are left there sorted by id (for a gather over NVLink, lifeb200_multi_* below, or a caller's
 * own collective) and only the counters come back to the host.  The calls synchronise the stream. */
int lifeb200_enumerate_collisions_dev(lifeb200_ctx* ctx, const uint64_t* catalysts, unsigned n_catalysts,
                                      const uint64_t* projectiles, unsigned n_projectiles, uint64_t id_begin,
                                      uint64_t id_end, unsigned gens, unsigned stable_interval, lifeb200_hit* d_hits,
                                      size_t cap, uint64_t* n_hits, uint64_t* n_skipped);
int lifeb200_soup_survey_dev(lifeb200_ctx* ctx, uint64_t first_id, size_t n, uint64_t seed, int dist, unsigned gens,
                             lifeb200_hit* d_hits, size_t cap, uint64_t* n_hits, uint64_t* pop_sum,
                             uint64_t* hash_xor);
/* The same 64-bit fold of a resident batch: XOR over patterns of fold(p). */
int lifeb200_fold_dev(lifeb200_ctx* ctx, const uint64_t* patterns, size_t n, uint64_t first_id,
                      uint64_t* pop_sum, uint64_t* hash_xor);

/* ---- A9  three-valued step through bitslicing/unknown_step.hpp --------------------------
 * The reference ships the generated cover without a consumer; the loop is the one
 * LifeStable.hpp:429-469 uses for stable_simple.hpp: on = NeighbourCount(state), unk =
 * NeighbourCount(unknown) (3x3 sums INCLUDING the centre), per word the cover maps
 * (current_unknown, current_on, on2..on0, unk3..unk0) -> (naive_next_unknown, naive_next_on).
 * state / unknown are separate plane arrays of n x 512 B; out may alias in.  gens may be 0.
 * The result is the cover's on every input.  Patterns that keep the LifeStable invariant
 * state & unknown == 0 (LifeStable.hpp:220-221) are stepped through an equivalent two-select form
 * (22 instead of 36 LUTs per 32 cells, csrc/life_three.cuh); overlapping planes -- the cover's
 * don't-care rows -- go through the fragment text itself.  lifeb200_set_three_valued_path(ctx, 1)
 * sends everything through the fragment text (0 restores the default); results never depend on it. */
int lifeb200_set_three_valued_path(lifeb200_ctx* ctx, int path);
int lifeb200_unknown_step_batch_dev(lifeb200_ctx* ctx, const uint64_t* state_in, const uint64_t* unknown_in,
                                    uint64_t* state_out, uint64_t* unknown_out, size_t n, unsigned gens);
int lifeb200_unknown_step_batch_host(lifeb200_ctx* ctx, const uint64_t* state_in, const uint64_t* unknown_in,
                                     uint64_t* state_out, uint64_t* unknown_out, size_t n, unsigned gens);

/* ---- LifeHistory (LifeHistory.hpp:8-12): record = state, history, marked, original = 2048 B.
 * The reference has no stepping member; the build-defined step is, `gens` times,
 * state.Step() (LifeAPI.hpp:1196-1216) then history |= state (:255-260).  In place. */
int lifeb200_history_step_batch_dev(lifeb200_ctx* ctx, uint64_t* records, size_t n, unsigned gens);
int lifeb200_history_step_batch_host(lifeb200_ctx* ctx, uint64_t* records, size_t n, unsigned gens);

/* ---- A10  LifeStable full-grid passes (LifeStable.hpp:366-729) ---------------------------
 * record = the 10 member planes in declaration order state, unknown, live2, live3, dead0, dead1,
 * dead2, dead4, dead5, dead6 (LifeStable.hpp:39-53) = 640 words = 5120 B == sizeof(LifeStable).
 * Records are updated in place exactly as the reference member leaves *this (including what has
 * already been written when a pass reports an inconsistency); flags[i] (may be NULL) receives
 * PropagateResult (LifeStable.hpp:123-126) as CONSISTENT | CHANGED bits. */
#define LIFEB200_STABLE_WORDS 640
#define LIFEB200_STABLE_BYTES 5120
enum {
  LIFEB200_STABLE_PROPAGATE_SIMPLE_STEP = 0, /* LifeStable.hpp:414-503 + stable_simple.hpp */
  LIFEB200_STABLE_SYNCHRONISE = 1,           /* SynchroniseStateKnown :526-556 */
  LIFEB200_STABLE_UPDATE_OPTIONS = 2,        /* :558-615 + stable_count.hpp */
  LIFEB200_STABLE_SIGNAL_NEIGHBOURS = 3,     /* :617-675 + stable_signal.hpp */
  LIFEB200_STABLE_PROPAGATE_STEP = 4,        /* :695-716 */
  LIFEB200_STABLE_PROPAGATE = 5,             /* :718-729, to the fixed point */
  LIFEB200_STABLE_STABILISE_OPTIONS = 6,     /* :677-693 */
  LIFEB200_STABLE_PROPAGATE_SIMPLE = 7       /* :505-524 */
};
#define LIFEB200_STABLE_CONSISTENT 1u
#define LIFEB200_STABLE_CHANGED 2u
int lifeb200_stable_pass_dev(lifeb200_ctx* ctx, uint64_t* records, size_t n, int pass, uint8_t* flags);
int lifeb200_stable_pass_host(lifeb200_ctx* ctx, uint64_t* records, size_t n, int pass, uint8_t* flags);
/* LifeStable::Vulnerable() (LifeStable.hpp:366-412 + stable_vulnerable.hpp): out = n x 512 B. */
int lifeb200_stable_vulnerable_dev(lifeb200_ctx* ctx, const uint64_t* records, uint64_t* out, size_t n);
int lifeb200_stable_vulnerable_host(lifeb200_ctx* ctx, const uint64_t* records, uint64_t* out, size_t n);

/* ---- A9b  refined three-valued step (bitslicing/unknown_step_refined_m.hpp, variant 0; the
 * older unknown_step_refined.hpp cover, variant 1).  No consumer in the reference; inputs follow
 * the generator (unknown_step_refined.py:93-118): option planes l2..d6 of the stable record,
 * s2..s0 = NeighbourCount(stable.state), live = NeighbourCount(current_on), m3..m0 =
 * NeighbourCount::Subtract(s, live) (NeighbourCount.hpp:85-91), on2..on0 = live for variant 1.
 * Outputs: three planes of n x 512 B each. */
int lifeb200_unknown_step_refined_dev(lifeb200_ctx* ctx, const uint64_t* records, const uint64_t* current_on,
                                      const uint64_t* current_unknown, size_t n, int variant, uint64_t* next_on,
                                      uint64_t* next_unknown, uint64_t* next_unknown_stable);
int lifeb200_unknown_step_refined_host(lifeb200_ctx* ctx, const uint64_t* records, const uint64_t* current_on,
                                       const uint64_t* current_unknown, size_t n, int variant, uint64_t* next_on,
                                       uint64_t* next_unknown, uint64_t* next_unknown_stable);

/* ---- 8f "next" rows: the callers and formats either side of the stepping path ------------
 * LifeWeld::Step (LifeWeld.hpp:169-186): record = state, frozen2, frozen1, frozen0 = 2048 B
 * (LifeWeld.hpp:18-22); per generation the 3x3 count (bit3 dropped, as the reference does) plus
 * the frozen count feeds the B3/S23 rule.  In place, `gens` generations per launch. */
int lifeb200_weld_step_batch_dev(lifeb200_ctx* ctx, uint64_t* records, size_t n, unsigned gens);
int lifeb200_weld_step_batch_host(lifeb200_ctx* ctx, uint64_t* records, size_t n, unsigned gens);
/* LifeState::InteractionCountsAndNext (LifeAPI.hpp:997-1040; the first three planes are
 * InteractionCounts, :956-995): planes[p] = out1, out2, outMore, next (4 x 512 B). */
int lifeb200_interaction_counts_dev(lifeb200_ctx* ctx, const uint64_t* in, uint64_t* planes, size_t n);
int lifeb200_interaction_counts_host(lifeb200_ctx* ctx, const uint64_t* in, uint64_t* planes, size_t n);
/* LifeWeld::InteractionCounts (LifeWeld.hpp:193-215): InteractionCounts of the record's state with every plane
 * masked by the zone of influence of its NON-frozen cells; planes[p] = out1, out2, outMore (3 x 512 B). */
int lifeb200_weld_interaction_counts_dev(lifeb200_ctx* ctx, const uint64_t* records, uint64_t* planes, size_t n);
int lifeb200_weld_interaction_counts_host(lifeb200_ctx* ctx, const uint64_t* records, uint64_t* planes, size_t n);
/* LifeState::GetHash (LifeAPI.hpp:373) = XXH3_64bits(state, 512) with the default secret
 * (xxHash v0.8.3, the reference's vendored submodule): hashes[p] for p < n. */
int lifeb200_hash_batch_dev(lifeb200_ctx* ctx, const uint64_t* patterns, uint64_t* hashes, size_t n);
int lifeb200_hash_batch_host(lifeb200_ctx* ctx, const uint64_t* patterns, uint64_t* hashes, size_t n);

/* ---- 8f rank 4: locating patterns on the torus, the step AFTER the stepping path ----------
 * Operands are plane arrays of 512 B; every count argument (n_a, n_b, n_t, n_targets) is either n
 * (one operand per problem) or 1 (the same operand for all problems).
 *
 * LifeState::Convolve (LifeAPI.hpp:1293-1370): out[p] = a[p].Convolve(b[p]), the torus Minkowski sum
 * (OR over ON cells (j,k) of a of b.Moved(j,k)). */
int lifeb200_convolve_batch_dev(lifeb200_ctx* ctx, const uint64_t* a, size_t n_a, const uint64_t* b, size_t n_b,
                                uint64_t* out, size_t n);
int lifeb200_convolve_batch_host(lifeb200_ctx* ctx, const uint64_t* a, size_t n_a, const uint64_t* b, size_t n_b,
                                 uint64_t* out, size_t n);
/* LifeState::MatchLiveAndDead / MatchLive / Match(const LifeState&) (LifeAPI.hpp:427-442): out[p] has a
 * cell ON at every offset where `live` fits inside states[p] (and `dead` fits inside its complement).
 * mode LIFEB200_MATCH_LIVE_AND_DEAD uses both planes; LIFEB200_MATCH_LIVE ignores `dead` (may be NULL);
 * LIFEB200_MATCH_BOUNDARY takes dead = live.GetBoundary() (LifeAPI.hpp:538), i.e. Match(live). */
enum { LIFEB200_MATCH_LIVE_AND_DEAD = 0, LIFEB200_MATCH_LIVE = 1, LIFEB200_MATCH_BOUNDARY = 2 };
int lifeb200_match_batch_dev(lifeb200_ctx* ctx, const uint64_t* states, size_t n, const uint64_t* live,
                             const uint64_t* dead, size_t n_t, int mode, uint64_t* out);
int lifeb200_match_batch_host(lifeb200_ctx* ctx, const uint64_t* states, size_t n, const uint64_t* live,
                              const uint64_t* dead, size_t n_t, int mode, uint64_t* out);
/* LifeState::Match(const LifeTarget&) (LifeTarget.hpp:53-54); targets are LifeTarget records
 * {wanted[64], unwanted[64]} = 1024 B (LifeTarget.hpp:5-7). */
int lifeb200_match_target_batch_dev(lifeb200_ctx* ctx, const uint64_t* states, size_t n, const uint64_t* targets,
                                    size_t n_targets, uint64_t* out);
int lifeb200_match_target_batch_host(lifeb200_ctx* ctx, const uint64_t* states, size_t n, const uint64_t* targets,
                                     size_t n_targets, uint64_t* out);
/* LifeState::InteractionOffsets (LifeAPI.hpp:1066-1095): out[p] = a[p].InteractionOffsets(b[p]), the
 * offsets at which b placed on a interacts in the next generation (overlap, birth or overcrowding);
 * the step BEFORE the stepping path in a search. */
int lifeb200_interaction_offsets_dev(lifeb200_ctx* ctx, const uint64_t* a, size_t n_a, const uint64_t* b, size_t n_b,
                                     uint64_t* out, size_t n);
int lifeb200_interaction_offsets_host(lifeb200_ctx* ctx, const uint64_t* a, size_t n_a, const uint64_t* b, size_t n_b,
                                      uint64_t* out, size_t n);

/* LifeWeld::InteractionOffsets (LifeWeld.hpp:217-262): the same for two LifeWeld records (state, frozen2, frozen1,
 * frozen0 = 2048 B each), interactions that involve frozen cells left out. */
int lifeb200_weld_interaction_offsets_dev(lifeb200_ctx* ctx, const uint64_t* a, size_t n_a, const uint64_t* b, size_t n_b,
                                          uint64_t* out, size_t n);
int lifeb200_weld_interaction_offsets_host(lifeb200_ctx* ctx, const uint64_t* a, size_t n_a, const uint64_t* b, size_t n_b,
                                           uint64_t* out, size_t n);

/* ---- 8f rank 4: RLE in and out (host functions; strings are a host format) ---------------
 * LifeState::RLE (Parsing.hpp:8-66,198-202) / LifeHistory::RLE (LifeHistory.hpp:25-44,62-68): writes
 * the NUL-terminated string into buf (capacity cap) and its length into *len; LIFEB200_E_OVERFLOW
 * with *len = required length if it does not fit.  LifeState::Parse / LifeHistory::Parse
 * (Parsing.hpp:141-196, LifeHistory.hpp:70-93): state = 512 B, record = state, history, marked,
 * original = 2048 B. */
int lifeb200_rle(const uint64_t* state, char* buf, size_t cap, size_t* len);
int lifeb200_history_rle(const uint64_t* record, char* buf, size_t cap, size_t* len);
int lifeb200_parse_rle(const char* rle, uint64_t* state);
int lifeb200_history_parse_rle(const char* rle, uint64_t* record);

/* ---- multi-GPU groups (SURVEY.md 8b "device management (init with device list)", "multi-GPU run + gather";
 * 8e).  The reference is single-threaded; the loop a caller shards is LifeAPI.hpp:877-881.  A group is `world`
 * ranks, one GPU each; rank r owns the contiguous id range [first + n*r/world, first + n*(r+1)/world) of every
 * call.  The path has no exchange step, so the only collective is the gather of hit lists and counters over
 * NCCL (csrc/multi.cu).  Every process of the group makes the same calls with the same GLOBAL arguments;
 * counters come back on every rank, the merged hit list (sorted by id) on the process that owns rank 0.
 *   lifeb200_multi_create        all ranks in THIS process: devices[0..n) (NULL = 0..n-1; n <= 0 with NULL = every
 *                                visible GPU), ncclCommInitAll, one host thread per GPU inside the calls
 *   lifeb200_multi_unique_id /   one process per GPU (torchrun / MPI): rank 0 makes a 128-byte id, the caller
 *   lifeb200_multi_create_rank   distributes it, every process joins with its device, rank and world
 * NCCL is loaded at run time (libnccl.so.2 or $LIFEB200_NCCL_LIB); a group of one needs none. */
#define LIFEB200_UNIQUE_ID_BYTES 128
typedef struct lifeb200_multi lifeb200_multi;
int lifeb200_multi_create(const int* devices, int n_devices, lifeb200_multi** out);
int lifeb200_multi_unique_id(char* id128);
int lifeb200_multi_create_rank(int device, int rank, int world, const char* id128, lifeb200_multi** out);
int lifeb200_multi_destroy(lifeb200_multi* m);
int lifeb200_multi_info(lifeb200_multi* m, int* world, int* first_rank, int* n_local, int* nccl_version);
/* Borrow the context of local GPU `local_index` (owned by the group) for per-device calls. */
int lifeb200_multi_ctx(lifeb200_multi* m, int local_index, lifeb200_ctx** ctx);
int lifeb200_multi_sync(lifeb200_multi* m);
/* The id range rank `rank` owns of a call over [first_id, first_id + n). */
int lifeb200_multi_shard(lifeb200_multi* m, uint64_t first_id, uint64_t n, int rank, uint64_t* begin, uint64_t* end);
/* How hit lists reach the host: 0 = NCCL send/recv over NVLink into rank 0's GPU, one download (default);
 * 1 = every GPU downloads its list straight to its place in `hits` (groups created by lifeb200_multi_create only). */
int lifeb200_multi_set_gather(lifeb200_multi* m, int mode);
/* Device time of the last call's gather (counters + hit lists + download), CUDA events on local GPU 0. */
int lifeb200_multi_last_gather_ms(lifeb200_multi* m, float* ms);
/* Size every local GPU's scratch (dense results, hit records, rank 0's gather room) for calls over up to `n` ids with
 * hit capacity `cap`, so that no later call of that size allocates: growing a device buffer once NCCL has mapped the
 * peers costs milliseconds per GPU and serialises across the GPUs of a process.  Optional: calls grow on demand. */
int lifeb200_multi_reserve(lifeb200_multi* m, uint64_t n, size_t cap);
/* lifeb200_soup_survey / lifeb200_enumerate_collisions over the whole group (same arguments, HOST hit buffer). */
int lifeb200_multi_soup_survey(lifeb200_multi* m, uint64_t first_id, uint64_t n, uint64_t seed, int dist, unsigned gens,
                               lifeb200_hit* hits, size_t cap, uint64_t* n_hits, uint64_t* pop_sum, uint64_t* hash_xor);
int lifeb200_multi_enumerate_collisions(lifeb200_multi* m, const uint64_t* catalysts, unsigned n_catalysts,
                                        const uint64_t* projectiles, unsigned n_projectiles, uint64_t id_begin,
                                        uint64_t id_end, unsigned gens, unsigned stable_interval, lifeb200_hit* hits,
                                        size_t cap, uint64_t* n_hits, uint64_t* n_skipped);
/* lifeb200_step_batch_host with THIS process's n patterns split over its local GPUs (one host thread each). */
int lifeb200_multi_step_batch_host(lifeb200_multi* m, const uint64_t* in, uint64_t* out, size_t n, unsigned gens);

/* ---- roofline denominator: measured LOP3 lane-instructions per second ------------------- */
int lifeb200_measure_int_peak(lifeb200_ctx* ctx, double* lop3_lane_instr_per_s, double* sm_mhz_effective);

#ifdef __cplusplus
}
#endif
#endif /* LIFEAPI_B200_H */
