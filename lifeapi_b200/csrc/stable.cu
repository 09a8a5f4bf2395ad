// stable.cu -- batched LifeStable full-grid propagation passes and the refined
// three-valued steps (A9b), with their C-ABI entry points.
//
// One warp owns one LifeStable record: 10 planes x 64 words (LifeStable.hpp:39-53, 5120 bytes);
// lane L holds words 2L, 2L+1 of every plane as (lo,hi) register pairs, so a plane is loaded with
// one fully coalesced 512-byte request per warp.  Each pass is the reference's per-word loop
// (two NeighbourCounts + one generated circuit + plane algebra) on 32-bit half-words; the
// fixed-point drivers (Propagate, StabiliseOptions, PropagateSimple) iterate in registers, so a
// record crosses HBM once per call however many passes it needs.  Because a whole warp works on
// one record, the reference's early returns are warp-uniform branches and the record is left
// exactly as the reference leaves it when a pass reports an inconsistency.
#include <algorithm>

#include "ctx.h"
#include "host_pipeline.cuh"
#include "generated/circuits.cuh"
#include "life_core.cuh"

namespace lifeb200 {

constexpr int kBlockS = 128;  // 4 records per CTA
constexpr int SL = 32;        // lanes per record
using T32 = Tile<SL>;
constexpr int kRegs = 2 * T32::W;  // half-words per lane per plane (4)

enum { S_STATE, S_UNKNOWN, S_LIVE2, S_LIVE3, S_DEAD0, S_DEAD1, S_DEAD2, S_DEAD4, S_DEAD5, S_DEAD6, S_PLANES };
constexpr unsigned kConsistent = LIFEB200_STABLE_CONSISTENT, kChanged = LIFEB200_STABLE_CHANGED;

struct Stable {
  T32 p[S_PLANES];
  __device__ __forceinline__ void load(const uint64_t* rec, int lane) {
#pragma unroll
    for (int k = 0; k < S_PLANES; k++) p[k].load(rec + k * kN, lane);
  }
  __device__ __forceinline__ void store(uint64_t* rec, int lane) const {
#pragma unroll
    for (int k = 0; k < S_PLANES; k++) p[k].store(rec + k * kN, lane);
  }
};

__device__ __forceinline__ bool warp_any(uint32_t bits) { return __any_sync(0xffffffffu, bits != 0); }

// LifeStable.hpp:320-335 SetOn / SetOff on whole planes
__device__ __forceinline__ void set_on(Stable& s, const T32& which) {
#pragma unroll
  for (int k = 0; k < kRegs; k++) {
    const uint32_t w = reg(which, k);
    reg(s.p[S_STATE], k) |= w;
    reg(s.p[S_UNKNOWN], k) &= ~w;
#pragma unroll
    for (int d = S_DEAD0; d <= S_DEAD6; d++) reg(s.p[d], k) |= w;
  }
}
__device__ __forceinline__ void set_off(Stable& s, const T32& which) {
#pragma unroll
  for (int k = 0; k < kRegs; k++) {
    const uint32_t w = reg(which, k);
    reg(s.p[S_STATE], k) &= ~w;
    reg(s.p[S_UNKNOWN], k) &= ~w;
    reg(s.p[S_LIVE2], k) |= w;
    reg(s.p[S_LIVE3], k) |= w;
  }
}

// LifeStable.hpp:414-503
__device__ __forceinline__ unsigned propagate_simple_step(Stable& s, int lane) {
  T32& state = s.p[S_STATE];
  T32& unknown = s.p[S_UNKNOWN];
  const T32 start_unknown = unknown;
  T32 on3, on2, on1, on0, u3, u2, u1, u0;
  state.neighbour_count(lane, on3, on2, on1, on0);
  unknown.neighbour_count(lane, u3, u2, u1, u0);
  T32 new_off, new_on, sig_off, sig_on;
  uint32_t has_set_off = 0, has_set_on = 0, has_sig_off = 0, has_sig_on = 0, has_abort = 0;
#pragma unroll
  for (int k = 0; k < kRegs; k++) {
    const uint32_t sat = reg(u2, k) | reg(u3, k);
    const uint32_t unk1 = reg(u1, k) | sat, unk0 = reg(u0, k) | sat;  // :439-440
    uint32_t set_off_, set_on_, signal_off, signal_on, abort_;
    circuits::stable_simple(reg(unknown, k), reg(state, k), reg(on2, k), reg(on1, k), reg(on0, k), unk1, unk0, set_off_,
                            set_on_, signal_off, signal_on, abort_);
    signal_off &= unk0 | unk1;
    signal_on &= unk0 | unk1;
    reg(new_off, k) = set_off_ & reg(unknown, k);
    reg(new_on, k) = set_on_ & reg(unknown, k);
    reg(sig_off, k) = signal_off;
    reg(sig_on, k) = signal_on;
    has_set_off |= set_off_; has_set_on |= set_on_; has_sig_off |= signal_off; has_sig_on |= signal_on;
    has_abort |= abort_;
  }
  if (warp_any(has_abort)) return 0;
  // the reference guards each update with has_*; an all-zero plane makes the update a no-op, so
  // only the signal dilations (which cost shuffles) keep their guards
#pragma unroll
  for (int k = 0; k < kRegs; k++) {
    reg(state, k) |= reg(new_on, k);
    reg(unknown, k) &= ~reg(new_on, k);
    reg(unknown, k) &= ~reg(new_off, k);
  }
  const bool any_off = warp_any(has_sig_off), any_on = warp_any(has_sig_on);
  T32 off_zoi, on_zoi;
  if (any_off) {
    off_zoi = dilate<SL>(sig_off, lane, false);
#pragma unroll
    for (int k = 0; k < kRegs; k++) reg(unknown, k) &= ~reg(off_zoi, k);
  }
  if (any_on) {
    on_zoi = dilate<SL>(sig_on, lane, false);
#pragma unroll
    for (int k = 0; k < kRegs; k++) {
      reg(state, k) |= reg(on_zoi, k) & reg(unknown, k);
      reg(unknown, k) &= ~reg(on_zoi, k);
    }
  }
  if (any_on && any_off) {
    uint32_t bad = 0;
#pragma unroll
    for (int k = 0; k < kRegs; k++) bad |= reg(on_zoi, k) & reg(off_zoi, k) & reg(unknown, k);
    if (warp_any(bad)) return 0;
  }
  return kConsistent | (warp_any(unknown.xor_diff(start_unknown)) ? kChanged : 0u);
}

// LifeStable.hpp:526-556
__device__ __forceinline__ unsigned synchronise(Stable& s) {
  uint32_t changes = 0, abort_ = 0;
  T32 maybe_live, maybe_dead;
#pragma unroll
  for (int k = 0; k < kRegs; k++) {
    const uint32_t st = reg(s.p[S_STATE], k), un = reg(s.p[S_UNKNOWN], k);
    const uint32_t known_on = ~un & st, known_off = ~un & ~st;
    uint32_t all_dead = reg(s.p[S_DEAD0], k);
#pragma unroll
    for (int d = S_DEAD1; d <= S_DEAD6; d++) all_dead &= reg(s.p[d], k);
    const uint32_t md = ~all_dead;
    changes |= md & known_on;
#pragma unroll
    for (int d = S_DEAD0; d <= S_DEAD6; d++) reg(s.p[d], k) |= known_on;
    const uint32_t ml = ~(reg(s.p[S_LIVE2], k) & reg(s.p[S_LIVE3], k));
    changes |= ml & known_off;
    reg(s.p[S_LIVE2], k) |= known_off;
    reg(s.p[S_LIVE3], k) |= known_off;
    abort_ |= ~ml & ~md;
    reg(maybe_live, k) = ml;
    reg(maybe_dead, k) = md;
  }
  if (warp_any(abort_)) return 0;
#pragma unroll
  for (int k = 0; k < kRegs; k++) {
    const uint32_t only_live = reg(maybe_live, k) & ~reg(maybe_dead, k), both = reg(maybe_live, k) & reg(maybe_dead, k);
    changes |= ~reg(s.p[S_STATE], k) & only_live;
    reg(s.p[S_STATE], k) |= only_live;
    changes |= ~reg(s.p[S_UNKNOWN], k) & both;
    reg(s.p[S_UNKNOWN], k) &= both;
  }
  return kConsistent | (warp_any(changes) ? kChanged : 0u);
}

// LifeStable.hpp:558-615
// (on2, on1, on0) = the low three planes of NeighbourCount(state); the pass does not write `state`, so
// PropagateStep hands the same planes on to SignalNeighbours
// and returns (f3..f0) = NeighbourCount(known off), from which it derives NeighbourCount(state | unknown)
__device__ __forceinline__ unsigned update_options(Stable& s, int lane, const T32& on2, const T32& on1, const T32& on0,
                                                   T32& f3, T32& f2, T32& f1, T32& f0) {
  T32 known_off;
#pragma unroll
  for (int k = 0; k < kRegs; k++) reg(known_off, k) = ~reg(s.p[S_UNKNOWN], k) & ~reg(s.p[S_STATE], k);
  known_off.neighbour_count(lane, f3, f2, f1, f0);
  uint32_t has_abort = 0, changes = 0;
#pragma unroll
  for (int k = 0; k < kRegs; k++) {
    uint32_t o[8], abort_;
    circuits::stable_count(reg(known_off, k), reg(s.p[S_STATE], k), reg(on2, k), reg(on1, k), reg(on0, k), reg(f3, k),
                           reg(f2, k), reg(f1, k), reg(f0, k), o[0], o[1], o[2], o[3], o[4], o[5], o[6], o[7], abort_);
#pragma unroll
    for (int j = 0; j < 8; j++) {
      changes |= o[j] & ~reg(s.p[S_LIVE2 + j], k);
      reg(s.p[S_LIVE2 + j], k) |= o[j];
    }
    has_abort |= abort_;
  }
  return (warp_any(has_abort) ? 0u : kConsistent) | (warp_any(changes) ? kChanged : 0u);
}
__device__ __forceinline__ unsigned update_options(Stable& s, int lane) {
  T32 on3, on2, on1, on0, f3, f2, f1, f0;
  s.p[S_STATE].neighbour_count(lane, on3, on2, on1, on0);
  return update_options(s, lane, on2, on1, on0, f3, f2, f1, f0);
}

// LifeStable.hpp:617-675; (s2, s1, s0) = low planes of NeighbourCount(state), (m3..m0) = NeighbourCount(state | unknown)
__device__ __forceinline__ unsigned signal_neighbours(Stable& s, int lane, const T32& s2, const T32& s1, const T32& s0,
                                                      const T32& m3, const T32& m2, const T32& m1, const T32& m0) {
  T32 sig_off, sig_on, cen_off, cen_on;
#pragma unroll
  for (int k = 0; k < kRegs; k++) {
    circuits::stable_signal(reg(s.p[S_UNKNOWN], k), reg(s.p[S_STATE], k), reg(s.p[S_LIVE2], k), reg(s.p[S_LIVE3], k),
                            reg(s.p[S_DEAD0], k), reg(s.p[S_DEAD1], k), reg(s.p[S_DEAD2], k), reg(s.p[S_DEAD4], k),
                            reg(s.p[S_DEAD5], k), reg(s.p[S_DEAD6], k), reg(s2, k), reg(s1, k), reg(s0, k), reg(m3, k),
                            reg(m2, k), reg(m1, k), reg(m0, k), reg(sig_on, k), reg(sig_off, k), reg(cen_on, k),
                            reg(cen_off, k));
  }
  T32 off_zoi = dilate<SL>(sig_off, lane, true), on_zoi = dilate<SL>(sig_on, lane, true);
  uint32_t bad = 0, changes = 0;
#pragma unroll
  for (int k = 0; k < kRegs; k++) {
    reg(off_zoi, k) = (reg(off_zoi, k) | reg(cen_off, k)) & reg(s.p[S_UNKNOWN], k);
    reg(on_zoi, k) = (reg(on_zoi, k) | reg(cen_on, k)) & reg(s.p[S_UNKNOWN], k);
    bad |= reg(off_zoi, k) & reg(on_zoi, k);
    changes |= reg(off_zoi, k) | reg(on_zoi, k);
  }
  if (warp_any(bad)) return 0;
  set_off(s, off_zoi);
  set_on(s, on_zoi);
  return kConsistent | (warp_any(changes) ? kChanged : 0u);
}
__device__ __forceinline__ unsigned signal_neighbours(Stable& s, int lane) {
  T32 maxp;
#pragma unroll
  for (int k = 0; k < kRegs; k++) reg(maxp, k) = reg(s.p[S_STATE], k) | reg(s.p[S_UNKNOWN], k);
  T32 s3, s2, s1, s0, m3, m2, m1, m0;
  s.p[S_STATE].neighbour_count(lane, s3, s2, s1, s0);
  maxp.neighbour_count(lane, m3, m2, m1, m0);
  return signal_neighbours(s, lane, s2, s1, s0, m3, m2, m1, m0);
}

// LifeStable.hpp:677-693
__device__ __forceinline__ unsigned stabilise_options(Stable& s, int lane) {
  bool changed_ever = false;
  for (;;) {
    const unsigned k = synchronise(s);
    if (!(k & kConsistent)) return 0;
    const unsigned o = update_options(s, lane);
    if (!(o & kConsistent)) return 0;
    if (!((k | o) & kChanged)) break;
    changed_ever = true;
  }
  return kConsistent | (changed_ever ? kChanged : 0u);
}

// LifeStable.hpp:505-524
__device__ __forceinline__ unsigned propagate_simple(Stable& s, int lane) {
  bool changed = false;
  for (;;) {
    const unsigned r = propagate_simple_step(s, lane);
    if (!(r & kConsistent)) return 0;
    if (!(r & kChanged)) break;
    changed = true;
  }
  if (changed && !(stabilise_options(s, lane) & kConsistent)) return 0;
  return kConsistent | (changed ? kChanged : 0u);
}

// LifeStable.hpp:695-716 (the PropagateSimpleStep call there is commented out in the reference)
__device__ __forceinline__ unsigned propagate_step(Stable& s, int lane) {
  const unsigned k = synchronise(s);
  if (!(k & kConsistent)) return 0;
  // UpdateOptions writes neither `state` nor `unknown`, so of the four NeighbourCounts the two passes take, two
  // are needed: NeighbourCount(state) serves both, and state | unknown is the complement of known-off, whose
  // 3x3 count f (0..9, exact) gives NeighbourCount(state | unknown) = 9 - f plane by plane.
  T32 on3, on2, on1, on0, f3, f2, f1, f0;
  s.p[S_STATE].neighbour_count(lane, on3, on2, on1, on0);
  const unsigned o = update_options(s, lane, on2, on1, on0, f3, f2, f1, f0);
  if (!(o & kConsistent)) return 0;
  T32 m3, m2, m1, m0;
#pragma unroll
  for (int k = 0; k < kRegs; k++) {
    reg(m0, k) = ~reg(f0, k);                                 // 9 - f: 9..0 = 1001 1000 0111 0110 0101 0100 0011 0010 0001 0000
    reg(m1, k) = reg(f1, k);
    reg(m2, k) = reg(f2, k) ^ reg(f1, k);
    reg(m3, k) = ~(reg(f3, k) | reg(f2, k) | reg(f1, k));
  }
  const unsigned g = signal_neighbours(s, lane, on2, on1, on0, m3, m2, m1, m0);
  if (!(g & kConsistent)) return 0;
  return kConsistent | ((k | o | g) & kChanged);
}

// LifeStable.hpp:718-729
__device__ __forceinline__ unsigned propagate(Stable& s, int lane) {
  bool changed_ever = false;
  for (;;) {
    const unsigned r = propagate_step(s, lane);
    if (!(r & kConsistent)) return 0;
    if (!(r & kChanged)) break;
    changed_ever = true;
  }
  return kConsistent | (changed_ever ? kChanged : 0u);
}

__global__ void __launch_bounds__(kBlockS) stable_pass_kernel(uint64_t* __restrict__ records, size_t n, int pass,
                                                              uint8_t* __restrict__ flags) {
  const size_t rec = ((size_t)blockIdx.x * kBlockS + threadIdx.x) / SL;
  const int lane = threadIdx.x & 31;
  if (rec >= n) return;  // whole warp
  uint64_t* base = records + rec * LIFEB200_STABLE_WORDS;
  Stable s;
  s.load(base, lane);
  unsigned r;
  switch (pass) {  // uniform across the grid
    case LIFEB200_STABLE_PROPAGATE_SIMPLE_STEP: r = propagate_simple_step(s, lane); break;
    case LIFEB200_STABLE_SYNCHRONISE: r = synchronise(s); break;
    case LIFEB200_STABLE_UPDATE_OPTIONS: r = update_options(s, lane); break;
    case LIFEB200_STABLE_SIGNAL_NEIGHBOURS: r = signal_neighbours(s, lane); break;
    case LIFEB200_STABLE_PROPAGATE_STEP: r = propagate_step(s, lane); break;
    case LIFEB200_STABLE_PROPAGATE: r = propagate(s, lane); break;
    case LIFEB200_STABLE_STABILISE_OPTIONS: r = stabilise_options(s, lane); break;
    default: r = propagate_simple(s, lane); break;
  }
  s.store(base, lane);
  if (lane == 0 && flags) flags[rec] = (uint8_t)r;
}

// LifeStable.hpp:366-412 Vulnerable()
__global__ void __launch_bounds__(kBlockS) stable_vulnerable_kernel(const uint64_t* __restrict__ records, size_t n,
                                                                    uint64_t* __restrict__ out) {
  const size_t rec = ((size_t)blockIdx.x * kBlockS + threadIdx.x) / SL;
  const int lane = threadIdx.x & 31;
  if (rec >= n) return;
  Stable s;
  s.load(records + rec * LIFEB200_STABLE_WORDS, lane);
  T32 s3, s2, s1, s0, u3, u2, u1, u0;
  s.p[S_STATE].neighbour_count(lane, s3, s2, s1, s0);
  s.p[S_UNKNOWN].neighbour_count(lane, u3, u2, u1, u0);
  T32 von, voff, vcon, vcoff;
#pragma unroll
  for (int k = 0; k < kRegs; k++) {
    circuits::stable_vulnerable(reg(s.p[S_LIVE2], k), reg(s.p[S_LIVE3], k), reg(s.p[S_DEAD0], k), reg(s.p[S_DEAD1], k),
                                reg(s.p[S_DEAD2], k), reg(s.p[S_DEAD4], k), reg(s.p[S_DEAD5], k), reg(s.p[S_DEAD6], k),
                                reg(s2, k), reg(s1, k), reg(s0, k), reg(u3, k), reg(u2, k), reg(u1, k), reg(u0, k),
                                reg(von, k), reg(voff, k), reg(vcon, k), reg(vcoff, k));
  }
  T32 zon = dilate<SL>(von, lane, true), zoff = dilate<SL>(voff, lane, true), res;
#pragma unroll
  for (int k = 0; k < kRegs; k++) reg(res, k) = (reg(zon, k) | reg(vcon, k)) & (reg(zoff, k) | reg(vcoff, k));
  res.store(out + rec * kN, lane);
}

// A9b refined three-valued step (bitslicing/unknown_step_refined_m.hpp, or the older
// unknown_step_refined.hpp cover with variant = 1).  No consumer exists in the reference; the
// harness is the generator's own input definition (unknown_step_refined.py:93-118): option planes
// from the stable record, s = NeighbourCount(stable.state), live = NeighbourCount(current_on),
// m = s - live as a 4-bit two's-complement NeighbourCount::Subtract (NeighbourCount.hpp:85-91),
// on = live for the older cover.
__global__ void __launch_bounds__(kBlockS) refined_step_kernel(const uint64_t* __restrict__ records,
                                                               const uint64_t* __restrict__ cur_on,
                                                               const uint64_t* __restrict__ cur_unknown, size_t n,
                                                               int variant, uint64_t* __restrict__ next_on,
                                                               uint64_t* __restrict__ next_unknown,
                                                               uint64_t* __restrict__ next_unknown_stable) {
  const size_t rec = ((size_t)blockIdx.x * kBlockS + threadIdx.x) / SL;
  const int lane = threadIdx.x & 31;
  if (rec >= n) return;
  Stable s;
  s.load(records + rec * LIFEB200_STABLE_WORDS, lane);
  T32 con, cunk;
  con.load(cur_on + rec * kN, lane);
  cunk.load(cur_unknown + rec * kN, lane);
  T32 s3, s2, s1, s0, l3, l2, l1, l0;
  s.p[S_STATE].neighbour_count(lane, s3, s2, s1, s0);
  con.neighbour_count(lane, l3, l2, l1, l0);
  T32 o_on, o_unk, o_stab;
#pragma unroll
  for (int k = 0; k < kRegs; k++) {
    uint32_t a, b, c;
    if (variant == 0) {
      // m = s + ~live + 1, ripple over four planes, carry out dropped
      uint32_t x0 = reg(s0, k), y0 = ~reg(l0, k), x1 = reg(s1, k), y1 = ~reg(l1, k);
      uint32_t x2 = reg(s2, k), y2 = ~reg(l2, k), x3 = reg(s3, k), y3 = ~reg(l3, k);
      uint32_t m0 = ~(x0 ^ y0), c0 = x0 | y0;  // carry in = 1
      uint32_t m1 = lop3<kXor3>(x1, y1, c0), c1 = lop3<kMaj>(x1, y1, c0);
      uint32_t m2 = lop3<kXor3>(x2, y2, c1), c2 = lop3<kMaj>(x2, y2, c1);
      uint32_t m3 = lop3<kXor3>(x3, y3, c2);
      circuits::unknown_step_refined_m(reg(s.p[S_LIVE2], k), reg(s.p[S_LIVE3], k), reg(s.p[S_DEAD0], k),
                                       reg(s.p[S_DEAD1], k), reg(s.p[S_DEAD2], k), reg(s.p[S_DEAD4], k),
                                       reg(s.p[S_DEAD5], k), reg(s.p[S_DEAD6], k), reg(con, k), m3, m2, m1, m0, a, b, c);
    } else {
      circuits::unknown_step_refined(reg(s.p[S_LIVE2], k), reg(s.p[S_LIVE3], k), reg(s.p[S_DEAD0], k), reg(s.p[S_DEAD1], k),
                                     reg(s.p[S_DEAD2], k), reg(s.p[S_DEAD4], k), reg(s.p[S_DEAD5], k), reg(s.p[S_DEAD6], k),
                                     reg(cunk, k), reg(con, k), reg(s2, k), reg(s1, k), reg(s0, k), reg(l2, k), reg(l1, k),
                                     reg(l0, k), a, b, c);
    }
    reg(o_on, k) = a; reg(o_unk, k) = b; reg(o_stab, k) = c;
  }
  o_on.store(next_on + rec * kN, lane);
  o_unk.store(next_unknown + rec * kN, lane);
  o_stab.store(next_unknown_stable + rec * kN, lane);
}

static inline unsigned grid_s(size_t n) { return (unsigned)((n * SL + kBlockS - 1) / kBlockS); }

}  // namespace lifeb200

using namespace lifeb200;

extern "C" {

int lifeb200_stable_pass_dev(lifeb200_ctx* ctx, uint64_t* records, size_t n, int pass, uint8_t* flags) {
  LIFEB200_CHECK_CTX(ctx);
  if (pass < 0 || pass > LIFEB200_STABLE_PROPAGATE_SIMPLE) return set_error(LIFEB200_E_ARG, "unknown LifeStable pass %d", pass);
  if (n == 0) return LIFEB200_OK;
  if (!records || !aligned16(records)) return set_error(LIFEB200_E_ARG, "bad record pointer");
  stable_pass_kernel<<<grid_s(n), kBlockS, 0, ctx->stream>>>(records, n, pass, flags);
  LIFEB200_LAUNCHED(ctx);
  return LIFEB200_OK;
}

int lifeb200_stable_pass_host(lifeb200_ctx* ctx, uint64_t* records, size_t n, int pass, uint8_t* flags) {
  LIFEB200_CHECK_CTX(ctx);
  if (pass < 0 || pass > LIFEB200_STABLE_PROPAGATE_SIMPLE) return set_error(LIFEB200_E_ARG, "unknown LifeStable pass %d", pass);
  if (n == 0) return LIFEB200_OK;
  if (!records) return set_error(LIFEB200_E_ARG, "null record pointer");
  const size_t chunk = std::min<size_t>(n, (size_t)1 << 13);  // 8 Ki records = 40 MiB
  const size_t nchunks = (n + chunk - 1) / chunk;
  const int used = (int)std::min<size_t>(kSlots, nchunks);
  for (int s = 0; s < used; s++) {
    int rc = ensure_slot(ctx, s, chunk * LIFEB200_STABLE_BYTES, chunk);
    if (rc) return rc;
  }
  for (size_t c = 0; c < nchunks; c++) {
    Slot& sl = ctx->slots[c % kSlots];
    const size_t off = c * chunk, cnt = std::min(chunk, n - off);
    uint64_t* d = static_cast<uint64_t*>(sl.dbuf);
    uint8_t* df = static_cast<uint8_t*>(sl.dbuf2);
    LIFEB200_CUDA(cudaMemcpyAsync(d, records + off * LIFEB200_STABLE_WORDS, cnt * LIFEB200_STABLE_BYTES,
                                  cudaMemcpyHostToDevice, sl.stream));
    stable_pass_kernel<<<grid_s(cnt), kBlockS, 0, sl.stream>>>(d, cnt, pass, df);
    LIFEB200_LAUNCHED(ctx);
    LIFEB200_CUDA(cudaMemcpyAsync(records + off * LIFEB200_STABLE_WORDS, d, cnt * LIFEB200_STABLE_BYTES,
                                  cudaMemcpyDeviceToHost, sl.stream));
    if (flags) LIFEB200_CUDA(cudaMemcpyAsync(flags + off, df, cnt, cudaMemcpyDeviceToHost, sl.stream));
  }
  for (int s = 0; s < used; s++) LIFEB200_CUDA(cudaStreamSynchronize(ctx->slots[s].stream));
  return LIFEB200_OK;
}

int lifeb200_stable_vulnerable_dev(lifeb200_ctx* ctx, const uint64_t* records, uint64_t* out, size_t n) {
  LIFEB200_CHECK_CTX(ctx);
  if (n == 0) return LIFEB200_OK;
  if (!records || !out || !aligned16(records) || !aligned16(out)) return set_error(LIFEB200_E_ARG, "bad pointer");
  stable_vulnerable_kernel<<<grid_s(n), kBlockS, 0, ctx->stream>>>(records, n, out);
  LIFEB200_LAUNCHED(ctx);
  return LIFEB200_OK;
}

int lifeb200_stable_vulnerable_host(lifeb200_ctx* ctx, const uint64_t* records, uint64_t* out, size_t n) {
  LIFEB200_CHECK_CTX(ctx);
  if (n == 0) return LIFEB200_OK;
  if (!records || !out) return set_error(LIFEB200_E_ARG, "null pointer");
  return host_pipeline_impl(ctx, records, out, n, LIFEB200_STABLE_BYTES, LIFEB200_PATTERN_BYTES, true, (size_t)1 << 13,
                            [&](cudaStream_t st, const uint64_t* din, uint64_t* dout, size_t cnt) -> int {
                              stable_vulnerable_kernel<<<grid_s(cnt), kBlockS, 0, st>>>(din, cnt, dout);
                              LIFEB200_LAUNCHED(ctx);
                              return LIFEB200_OK;
                            });
}

int lifeb200_unknown_step_refined_dev(lifeb200_ctx* ctx, const uint64_t* records, const uint64_t* current_on,
                                      const uint64_t* current_unknown, size_t n, int variant, uint64_t* next_on,
                                      uint64_t* next_unknown, uint64_t* next_unknown_stable) {
  LIFEB200_CHECK_CTX(ctx);
  if (variant != 0 && variant != 1) return set_error(LIFEB200_E_ARG, "variant must be 0 (refined_m) or 1 (refined)");
  if (n == 0) return LIFEB200_OK;
  const void* ptrs[] = {records, current_on, current_unknown, next_on, next_unknown, next_unknown_stable};
  for (const void* p : ptrs)
    if (!p || !aligned16(p)) return set_error(LIFEB200_E_ARG, "null or misaligned pointer");
  refined_step_kernel<<<grid_s(n), kBlockS, 0, ctx->stream>>>(records, current_on, current_unknown, n, variant, next_on,
                                                               next_unknown, next_unknown_stable);
  LIFEB200_LAUNCHED(ctx);
  return LIFEB200_OK;
}

// Host form: the three inputs of a chunk are staged side by side in the slot's first buffer, the three output
// planes in its second (records | current_on | current_unknown  ->  next_on | next_unknown | next_unknown_stable).
int lifeb200_unknown_step_refined_host(lifeb200_ctx* ctx, const uint64_t* records, const uint64_t* current_on,
                                       const uint64_t* current_unknown, size_t n, int variant, uint64_t* next_on,
                                       uint64_t* next_unknown, uint64_t* next_unknown_stable) {
  LIFEB200_CHECK_CTX(ctx);
  if (variant != 0 && variant != 1) return set_error(LIFEB200_E_ARG, "variant must be 0 (refined_m) or 1 (refined)");
  if (n == 0) return LIFEB200_OK;
  const void* ptrs[] = {records, current_on, current_unknown, next_on, next_unknown, next_unknown_stable};
  for (const void* p : ptrs)
    if (!p) return set_error(LIFEB200_E_ARG, "null pointer");
  const size_t chunk = std::min<size_t>(n, (size_t)1 << 13);
  const size_t nchunks = (n + chunk - 1) / chunk;
  const int used = (int)std::min<size_t>(kSlots, nchunks);
  for (int s = 0; s < used; s++)
    if (int rc = ensure_slot(ctx, s, chunk * (LIFEB200_STABLE_BYTES + 2 * LIFEB200_PATTERN_BYTES), chunk * 3 * LIFEB200_PATTERN_BYTES)) return rc;
  auto body = [&]() -> int {
    for (size_t c = 0; c < nchunks; c++) {
      Slot& sl = ctx->slots[c % kSlots];
      const size_t off = c * chunk, cnt = std::min(chunk, n - off), pb = cnt * LIFEB200_PATTERN_BYTES;
      uint64_t* d_rec = static_cast<uint64_t*>(sl.dbuf);
      uint64_t* d_on = d_rec + chunk * LIFEB200_STABLE_WORDS;
      uint64_t* d_unk = d_on + chunk * kN;
      uint64_t* o_on = static_cast<uint64_t*>(sl.dbuf2);
      uint64_t* o_unk = o_on + chunk * kN;
      uint64_t* o_st = o_unk + chunk * kN;
      LIFEB200_CUDA(cudaMemcpyAsync(d_rec, records + off * LIFEB200_STABLE_WORDS, cnt * LIFEB200_STABLE_BYTES, cudaMemcpyHostToDevice, sl.stream));
      LIFEB200_CUDA(cudaMemcpyAsync(d_on, current_on + off * kN, pb, cudaMemcpyHostToDevice, sl.stream));
      LIFEB200_CUDA(cudaMemcpyAsync(d_unk, current_unknown + off * kN, pb, cudaMemcpyHostToDevice, sl.stream));
      refined_step_kernel<<<grid_s(cnt), kBlockS, 0, sl.stream>>>(d_rec, d_on, d_unk, cnt, variant, o_on, o_unk, o_st);
      LIFEB200_LAUNCHED(ctx);
      LIFEB200_CUDA(cudaMemcpyAsync(next_on + off * kN, o_on, pb, cudaMemcpyDeviceToHost, sl.stream));
      LIFEB200_CUDA(cudaMemcpyAsync(next_unknown + off * kN, o_unk, pb, cudaMemcpyDeviceToHost, sl.stream));
      LIFEB200_CUDA(cudaMemcpyAsync(next_unknown_stable + off * kN, o_st, pb, cudaMemcpyDeviceToHost, sl.stream));
    }
    for (int s = 0; s < used; s++) LIFEB200_CUDA(cudaStreamSynchronize(ctx->slots[s].stream));
    return LIFEB200_OK;
  };
  const int rc = body();
  if (rc) drain_slots(ctx, used);
  return rc;
}

}  // extern "C"
