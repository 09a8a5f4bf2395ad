/*
 * stable_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * CPU restatement of the bitsliced-circuit consumers of the reference:
 *   - a generic evaluator of the committed sum-of-products fragments (bitslicing/[name].hpp), driven by
 *     the cube tables in circuit_tables.h (oracle/make_tables.py, from the fragments' text as parsed
 *     by the emitter), so don't-care behaviour is the fragments' own;
 *   - the three-valued step through unknown_step.hpp in the loop shape of LifeStable.hpp:429-469;
 *   - LifeStable's full-grid passes (LifeStable.hpp:366-729) on a 10-plane record in the member
 *     order state, unknown, live2, live3, dead0, dead1, dead2, dead4, dead5, dead6 (LifeStable.hpp:39-53),
 *     including their early-return behaviour (what has and has not been written when a pass
 *     reports an inconsistency).
 *
 * Parity status: the reference has no tests for any of this ("parity unpinned" by its own tests,
 * SURVEY.md 8c).  The pin is the unmodified reference compiled into oracle/_ref (ref_circuits.cpp,
 * ref_stable.cpp): tests/test_oracle.py checks every function here against it on seeded inputs,
 * and against golden vectors generated from it (tests/golden/make_golden.py).
 */
#include <stdint.h>
#include <stddef.h>
#include <string.h>

#include "circuit_tables.h"

#define LN 64
enum { P_STATE, P_UNKNOWN, P_LIVE2, P_LIVE3, P_DEAD0, P_DEAD1, P_DEAD2, P_DEAD4, P_DEAD5, P_DEAD6, P_COUNT };
#define RES_CONSISTENT 1u
#define RES_CHANGED 2u

void lo_neighbour_count(const uint64_t *s, uint64_t *planes); /* life_oracle.c: bit3,bit2,bit1,bit0 */
void lo_zoi(const uint64_t *s, uint64_t *out);
void lo_parallel_for(size_t n, int threads, void (*fn)(size_t, size_t, void *), void *arg);

static inline uint64_t rotl1(uint64_t v) { return (v << 1) | (v >> 63); }
static inline uint64_t rotr1(uint64_t v) { return (v >> 1) | (v << 63); }

/* One word of one fragment: out[o] = OR of matching cubes, then the phase inversions. */
static void eval_circuit(const lo_circuit *c, const uint64_t *in, uint64_t *out) {
  for (int o = 0; o < c->n_out; o++) out[o] = 0;
  for (int t = 0; t < c->n_terms; t++) {
    const lo_cube *q = &c->terms[t];
    uint64_t v = ~0ULL;
    for (int i = 0; i < c->n_in; i++)
      if (q->care >> i & 1) v &= (q->value >> i & 1) ? in[i] : ~in[i];
    for (int o = 0; o < c->n_out; o++)
      if (q->outs >> o & 1) out[o] |= v;
  }
  for (int o = 0; o < c->n_out; o++)
    if (c->phase >> o & 1) out[o] = ~out[o];
}

/* Exposed for the emitter's cross-check: evaluate fragment `which` on one word per input. */
int lo_circuit_eval(int which, const uint64_t *in, uint64_t *out) {
  static const lo_circuit *all[] = { &lo_circuit_unknown_step, &lo_circuit_stable_simple, &lo_circuit_stable_count,
                                     &lo_circuit_stable_signal, &lo_circuit_stable_vulnerable,
                                     &lo_circuit_unknown_step_refined, &lo_circuit_unknown_step_refined_m };
  if (which < 0 || which >= (int)(sizeof all / sizeof all[0])) return -1;
  eval_circuit(all[which], in, out);
  return all[which]->n_out;
}

/* LifeAPI.hpp:541-562 ZOIHollow: convolve with 3o$obo$3o. */
static void zoi_hollow(const uint64_t *s, uint64_t *out) {
  uint64_t full[LN], mid[LN];
  for (int i = 0; i < LN; i++) { mid[i] = rotl1(s[i]) | rotr1(s[i]); full[i] = mid[i] | s[i]; }
  for (int i = 0; i < LN; i++) out[i] = full[(i + LN - 1) & 63] | mid[i] | full[(i + 1) & 63];
}
void lo_zoi_hollow(const uint64_t *s, uint64_t *out) { zoi_hollow(s, out); }

/* ---- three-valued step through bitslicing/unknown_step.hpp ------------------ */
void lo_unknown_step_sop(uint64_t *state, uint64_t *unknown) {
  uint64_t on[4 * LN], unk[4 * LN], ns[LN], nu[LN];
  lo_neighbour_count(state, on);
  lo_neighbour_count(unknown, unk);
  for (int i = 0; i < LN; i++) {
    /* input order: current_unknown current_on on2 on1 on0 unk3 unk2 unk1 unk0 */
    uint64_t in[9] = { unknown[i], state[i], on[LN + i], on[2 * LN + i], on[3 * LN + i],
                       unk[i], unk[LN + i], unk[2 * LN + i], unk[3 * LN + i] };
    uint64_t out[2];
    eval_circuit(&lo_circuit_unknown_step, in, out);
    nu[i] = out[0];
    ns[i] = out[1];
  }
  memcpy(state, ns, sizeof ns);
  memcpy(unknown, nu, sizeof nu);
}

/* ---- LifeStable passes ------------------------------------------------------------------------ */
#define PL(p, k) ((p) + (size_t)(k) * LN)

/* LifeStable.hpp:320-335 */
static void set_on(uint64_t *p, const uint64_t *which) {
  for (int i = 0; i < LN; i++) {
    uint64_t w = which[i];
    PL(p, P_STATE)[i] |= w; PL(p, P_UNKNOWN)[i] &= ~w;
    PL(p, P_DEAD0)[i] |= w; PL(p, P_DEAD1)[i] |= w; PL(p, P_DEAD2)[i] |= w;
    PL(p, P_DEAD4)[i] |= w; PL(p, P_DEAD5)[i] |= w; PL(p, P_DEAD6)[i] |= w;
  }
}
static void set_off(uint64_t *p, const uint64_t *which) {
  for (int i = 0; i < LN; i++) {
    uint64_t w = which[i];
    PL(p, P_STATE)[i] &= ~w; PL(p, P_UNKNOWN)[i] &= ~w;
    PL(p, P_LIVE2)[i] |= w; PL(p, P_LIVE3)[i] |= w;
  }
}

/* LifeStable.hpp:414-503 PropagateSimpleStep */
unsigned lo_stable_propagate_simple_step(uint64_t *p) {
  uint64_t *state = PL(p, P_STATE), *unknown = PL(p, P_UNKNOWN);
  uint64_t start_unknown[LN], on[4 * LN], unk[4 * LN];
  uint64_t new_off[LN], new_on[LN], new_sig_off[LN], new_sig_on[LN];
  memcpy(start_unknown, unknown, sizeof start_unknown);
  lo_neighbour_count(state, on);
  lo_neighbour_count(unknown, unk);
  uint64_t has_set_off = 0, has_set_on = 0, has_sig_off = 0, has_sig_on = 0, has_abort = 0;
  for (int i = 0; i < LN; i++) {
    uint64_t unk3 = unk[i], unk2 = unk[LN + i], unk1 = unk[2 * LN + i], unk0 = unk[3 * LN + i];
    unk1 |= unk2 | unk3;
    unk0 |= unk2 | unk3;
    /* stateunk stateon on2 on1 on0 unk1 unk0 -> set_off set_on signal_off signal_on abort */
    uint64_t in[7] = { unknown[i], state[i], on[LN + i], on[2 * LN + i], on[3 * LN + i], unk1, unk0 };
    uint64_t out[5];
    eval_circuit(&lo_circuit_stable_simple, in, out);
    uint64_t sig_off = out[2] & (unk0 | unk1), sig_on = out[3] & (unk0 | unk1);
    new_off[i] = out[0] & unknown[i];
    new_on[i] = out[1] & unknown[i];
    new_sig_off[i] = sig_off;
    new_sig_on[i] = sig_on;
    has_set_off |= out[0]; has_set_on |= out[1]; has_sig_off |= sig_off; has_sig_on |= sig_on; has_abort |= out[4];
  }
  if (has_abort) return 0;
  if (has_set_on) for (int i = 0; i < LN; i++) { state[i] |= new_on[i]; unknown[i] &= ~new_on[i]; }
  if (has_set_off) for (int i = 0; i < LN; i++) unknown[i] &= ~new_off[i];
  uint64_t off_zoi[LN], on_zoi[LN];
  if (has_sig_off) { lo_zoi(new_sig_off, off_zoi); for (int i = 0; i < LN; i++) unknown[i] &= ~off_zoi[i]; }
  if (has_sig_on) {
    lo_zoi(new_sig_on, on_zoi);
    for (int i = 0; i < LN; i++) { state[i] |= on_zoi[i] & unknown[i]; unknown[i] &= ~on_zoi[i]; }
  }
  if (has_sig_on && has_sig_off) {
    uint64_t bad = 0;
    for (int i = 0; i < LN; i++) bad |= on_zoi[i] & off_zoi[i] & unknown[i];
    if (bad) return 0;
  }
  return RES_CONSISTENT | (memcmp(unknown, start_unknown, sizeof start_unknown) ? RES_CHANGED : 0);
}

/* LifeStable.hpp:526-556 SynchroniseStateKnown */
unsigned lo_stable_synchronise(uint64_t *p) {
  uint64_t changes = 0, abort = 0;
  uint64_t maybe_live[LN], maybe_dead[LN];
  for (int i = 0; i < LN; i++) {
    uint64_t st = PL(p, P_STATE)[i], un = PL(p, P_UNKNOWN)[i];
    uint64_t known_on = ~un & st;
    maybe_dead[i] = ~(PL(p, P_DEAD0)[i] & PL(p, P_DEAD1)[i] & PL(p, P_DEAD2)[i] & PL(p, P_DEAD4)[i] &
                      PL(p, P_DEAD5)[i] & PL(p, P_DEAD6)[i]);
    changes |= maybe_dead[i] & known_on;
    for (int k = P_DEAD0; k <= P_DEAD6; k++) PL(p, k)[i] |= known_on;
    uint64_t known_off = ~un & ~st;
    maybe_live[i] = ~(PL(p, P_LIVE2)[i] & PL(p, P_LIVE3)[i]);
    changes |= maybe_live[i] & known_off;
    PL(p, P_LIVE2)[i] |= known_off;
    PL(p, P_LIVE3)[i] |= known_off;
    abort |= ~maybe_live[i] & ~maybe_dead[i];
  }
  if (abort) return 0;
  for (int i = 0; i < LN; i++) {
    uint64_t only_live = maybe_live[i] & ~maybe_dead[i], both = maybe_live[i] & maybe_dead[i];
    changes |= ~PL(p, P_STATE)[i] & only_live;
    PL(p, P_STATE)[i] |= only_live;
    changes |= ~PL(p, P_UNKNOWN)[i] & both;
    PL(p, P_UNKNOWN)[i] &= both;
  }
  return RES_CONSISTENT | (changes ? RES_CHANGED : 0);
}

/* LifeStable.hpp:558-615 UpdateOptions */
unsigned lo_stable_update_options(uint64_t *p) {
  uint64_t on[4 * LN], off[4 * LN], known_off_plane[LN];
  for (int i = 0; i < LN; i++) known_off_plane[i] = ~PL(p, P_UNKNOWN)[i] & ~PL(p, P_STATE)[i];
  lo_neighbour_count(PL(p, P_STATE), on);
  lo_neighbour_count(known_off_plane, off);
  uint64_t has_abort = 0, changes = 0;
  for (int i = 0; i < LN; i++) {
    /* known_off known_on on2 on1 on0 off3 off2 off1 off0 -> l2 l3 d0 d1 d2 d4 d5 d6 abort */
    uint64_t in[9] = { known_off_plane[i], PL(p, P_STATE)[i], on[LN + i], on[2 * LN + i], on[3 * LN + i],
                       off[i], off[LN + i], off[2 * LN + i], off[3 * LN + i] };
    uint64_t out[9];
    eval_circuit(&lo_circuit_stable_count, in, out);
    for (int k = 0; k < 8; k++) {
      changes |= out[k] & ~PL(p, P_LIVE2 + k)[i];
      PL(p, P_LIVE2 + k)[i] |= out[k];
    }
    has_abort |= out[8];
  }
  return (has_abort == 0 ? RES_CONSISTENT : 0) | (changes ? RES_CHANGED : 0);
}

/* LifeStable.hpp:617-675 SignalNeighbours */
unsigned lo_stable_signal_neighbours(uint64_t *p) {
  uint64_t s[4 * LN], m[4 * LN], maxp[LN];
  for (int i = 0; i < LN; i++) maxp[i] = PL(p, P_STATE)[i] | PL(p, P_UNKNOWN)[i];
  lo_neighbour_count(PL(p, P_STATE), s);
  lo_neighbour_count(maxp, m);
  uint64_t sig_off[LN], sig_on[LN], cen_off[LN], cen_on[LN];
  for (int i = 0; i < LN; i++) {
    /* stateunk stateon l2 l3 d0 d1 d2 d4 d5 d6 s2 s1 s0 m3 m2 m1 m0 -> signalon signaloff centeron centeroff */
    uint64_t in[17] = { PL(p, P_UNKNOWN)[i], PL(p, P_STATE)[i], PL(p, P_LIVE2)[i], PL(p, P_LIVE3)[i], PL(p, P_DEAD0)[i],
                        PL(p, P_DEAD1)[i], PL(p, P_DEAD2)[i], PL(p, P_DEAD4)[i], PL(p, P_DEAD5)[i], PL(p, P_DEAD6)[i],
                        s[LN + i], s[2 * LN + i], s[3 * LN + i], m[i], m[LN + i], m[2 * LN + i], m[3 * LN + i] };
    uint64_t out[4];
    eval_circuit(&lo_circuit_stable_signal, in, out);
    sig_on[i] = out[0]; sig_off[i] = out[1]; cen_on[i] = out[2]; cen_off[i] = out[3];
  }
  uint64_t off_zoi[LN], on_zoi[LN], bad = 0, changes = 0;
  zoi_hollow(sig_off, off_zoi);
  zoi_hollow(sig_on, on_zoi);
  for (int i = 0; i < LN; i++) {
    off_zoi[i] = (off_zoi[i] | cen_off[i]) & PL(p, P_UNKNOWN)[i];
    on_zoi[i] = (on_zoi[i] | cen_on[i]) & PL(p, P_UNKNOWN)[i];
    bad |= off_zoi[i] & on_zoi[i];
    changes |= off_zoi[i] | on_zoi[i];
  }
  if (bad) return 0;
  set_off(p, off_zoi);
  set_on(p, on_zoi);
  return RES_CONSISTENT | (changes ? RES_CHANGED : 0);
}

/* LifeStable.hpp:677-693 StabiliseOptions */
unsigned lo_stable_stabilise_options(uint64_t *p) {
  int changed_ever = 0;
  for (;;) {
    unsigned k = lo_stable_synchronise(p);
    if (!(k & RES_CONSISTENT)) return 0;
    unsigned o = lo_stable_update_options(p);
    if (!(o & RES_CONSISTENT)) return 0;
    int done = !(o & RES_CHANGED) && !(k & RES_CHANGED);
    if (!done) changed_ever = 1;
    if (done) break;
  }
  return RES_CONSISTENT | (changed_ever ? RES_CHANGED : 0);
}

/* LifeStable.hpp:505-524 PropagateSimple */
unsigned lo_stable_propagate_simple(uint64_t *p) {
  int changed = 0;
  for (;;) {
    unsigned r = lo_stable_propagate_simple_step(p);
    if (!(r & RES_CONSISTENT)) return 0;
    if (r & RES_CHANGED) changed = 1; else break;
  }
  if (changed) {
    unsigned r = lo_stable_stabilise_options(p);
    if (!(r & RES_CONSISTENT)) return 0;
  }
  return RES_CONSISTENT | (changed ? RES_CHANGED : 0);
}

/* LifeStable.hpp:695-716 PropagateStep */
unsigned lo_stable_propagate_step(uint64_t *p) {
  unsigned k = lo_stable_synchronise(p);
  if (!(k & RES_CONSISTENT)) return 0;
  unsigned o = lo_stable_update_options(p);
  if (!(o & RES_CONSISTENT)) return 0;
  unsigned s = lo_stable_signal_neighbours(p);
  if (!(s & RES_CONSISTENT)) return 0;
  return RES_CONSISTENT | (((k | o | s) & RES_CHANGED) ? RES_CHANGED : 0);
}

/* LifeStable.hpp:718-729 Propagate */
unsigned lo_stable_propagate(uint64_t *p) {
  int changed_ever = 0;
  for (;;) {
    unsigned r = lo_stable_propagate_step(p);
    if (!(r & RES_CONSISTENT)) return 0;
    if (r & RES_CHANGED) changed_ever = 1; else break;
  }
  return RES_CONSISTENT | (changed_ever ? RES_CHANGED : 0);
}

/* LifeStable.hpp:366-412 Vulnerable */
void lo_stable_vulnerable(const uint64_t *p, uint64_t *out) {
  uint64_t s[4 * LN], u[4 * LN];
  lo_neighbour_count(PL(p, P_STATE), s);
  lo_neighbour_count(PL(p, P_UNKNOWN), u);
  uint64_t von[LN], voff[LN], vcon[LN], vcoff[LN];
  for (int i = 0; i < LN; i++) {
    /* l2 l3 d0 d1 d2 d4 d5 d6 s2 s1 s0 unk3 unk2 unk1 unk0 */
    uint64_t in[15] = { PL(p, P_LIVE2)[i], PL(p, P_LIVE3)[i], PL(p, P_DEAD0)[i], PL(p, P_DEAD1)[i], PL(p, P_DEAD2)[i],
                        PL(p, P_DEAD4)[i], PL(p, P_DEAD5)[i], PL(p, P_DEAD6)[i], s[LN + i], s[2 * LN + i], s[3 * LN + i],
                        u[i], u[LN + i], u[2 * LN + i], u[3 * LN + i] };
    uint64_t o[4];
    eval_circuit(&lo_circuit_stable_vulnerable, in, o);
    von[i] = o[0]; voff[i] = o[1]; vcon[i] = o[2]; vcoff[i] = o[3];
  }
  uint64_t zon[LN], zoff[LN];
  zoi_hollow(von, zon);
  zoi_hollow(voff, zoff);
  for (int i = 0; i < LN; i++) out[i] = (zon[i] | vcon[i]) & (zoff[i] | vcoff[i]);
}

/* pass selector shared with the C ABI (include/lifeapi_b200.h LIFEB200_STABLE_*) */
unsigned lo_stable_pass(uint64_t *p, int pass) {
  switch (pass) {
    case 0: return lo_stable_propagate_simple_step(p);
    case 1: return lo_stable_synchronise(p);
    case 2: return lo_stable_update_options(p);
    case 3: return lo_stable_signal_neighbours(p);
    case 4: return lo_stable_propagate_step(p);
    case 5: return lo_stable_propagate(p);
    case 6: return lo_stable_stabilise_options(p);
    case 7: return lo_stable_propagate_simple(p);
    default: return 0xffu;
  }
}

typedef struct { uint64_t *planes; uint8_t *flags; int pass; } sp_args;
static void sp_range(size_t lo, size_t hi, void *q) {
  sp_args *a = (sp_args *)q;
  for (size_t i = lo; i < hi; i++) a->flags[i] = (uint8_t)lo_stable_pass(a->planes + i * P_COUNT * LN, a->pass);
}
void lo_stable_pass_batch(uint64_t *planes, size_t n, int pass, uint8_t *flags, int threads) {
  sp_args a = { planes, flags, pass };
  lo_parallel_for(n, threads, sp_range, &a);
}

typedef struct { uint64_t *state, *unknown; unsigned gens; } us_args;
static void us_range(size_t lo, size_t hi, void *q) {
  us_args *a = (us_args *)q;
  for (size_t i = lo; i < hi; i++)
    for (unsigned g = 0; g < a->gens; g++) lo_unknown_step_sop(a->state + i * LN, a->unknown + i * LN);
}
void lo_unknown_step_sop_batch(uint64_t *state, uint64_t *unknown, size_t n, unsigned gens, int threads) {
  us_args a = { state, unknown, gens };
  lo_parallel_for(n, threads, us_range, &a);
}

/* ---- A9b refined three-valued step (no consumer in the reference; the harness is the generator's
 * own input definition, bitslicing/unknown_step_refined.py:93-118).  variant 0 = _m cover
 * (m = stable count - live count, NeighbourCount.hpp:85-91 two's complement), 1 = older cover. */
void lo_nc_subtract(const uint64_t *a, const uint64_t *b, uint64_t *out); /* life_oracle.c */
void lo_refined_step(const uint64_t *p, const uint64_t *cur_on, const uint64_t *cur_unk, int variant, uint64_t *next_on,
                     uint64_t *next_unk, uint64_t *next_unk_stable) {
  uint64_t s[4 * LN], live[4 * LN], m[4 * LN];
  lo_neighbour_count(PL(p, P_STATE), s);
  lo_neighbour_count(cur_on, live);
  lo_nc_subtract(s, live, m);
  for (int i = 0; i < LN; i++) {
    uint64_t o[3];
    if (variant == 0) {
      /* l2 l3 d0 d1 d2 d4 d5 d6 current_on m3 m2 m1 m0 */
      uint64_t in[13] = { PL(p, P_LIVE2)[i], PL(p, P_LIVE3)[i], PL(p, P_DEAD0)[i], PL(p, P_DEAD1)[i], PL(p, P_DEAD2)[i],
                          PL(p, P_DEAD4)[i], PL(p, P_DEAD5)[i], PL(p, P_DEAD6)[i], cur_on[i],
                          m[i], m[LN + i], m[2 * LN + i], m[3 * LN + i] };
      eval_circuit(&lo_circuit_unknown_step_refined_m, in, o);
    } else {
      /* l2 l3 d0 d1 d2 d4 d5 d6 current_unknown current_on s2 s1 s0 on2 on1 on0 */
      uint64_t in[16] = { PL(p, P_LIVE2)[i], PL(p, P_LIVE3)[i], PL(p, P_DEAD0)[i], PL(p, P_DEAD1)[i], PL(p, P_DEAD2)[i],
                          PL(p, P_DEAD4)[i], PL(p, P_DEAD5)[i], PL(p, P_DEAD6)[i], cur_unk[i], cur_on[i],
                          s[LN + i], s[2 * LN + i], s[3 * LN + i], live[LN + i], live[2 * LN + i], live[3 * LN + i] };
      eval_circuit(&lo_circuit_unknown_step_refined, in, o);
    }
    next_on[i] = o[0]; next_unk[i] = o[1]; next_unk_stable[i] = o[2];
  }
}
