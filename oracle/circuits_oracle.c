/*
 * circuits_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * Per-cell (non-bitsliced) CPU restatement of the three-valued semantics behind the reference's
 * generated circuits, and of the build-defined compositions the reference does not implement
 * itself: collision enumeration, soup survey, state fold.
 *
 * Parity status of the three-valued step: the reference has NO test for bitslicing/unknown_step.hpp
 * and no in-repo consumer ("parity unpinned" by the reference's own tests).  The pin
 * used here is the committed .hpp text itself, compiled unmodified into oracle/_ref
 * (ref_circuits.cpp): tests/test_oracle.py checks this semantic port against it and against
 * golden vectors generated from it.
 */
#include <stdint.h>
#include <stddef.h>
#include <string.h>

#define LN 64

/* from life_oracle.c */
void lo_step(uint64_t *s);
void lo_zoi(const uint64_t *s, uint64_t *out);
void lo_moved(const uint64_t *s, int x, int y, uint64_t *out);
void lo_target_from_state(const uint64_t *s, uint64_t *target);
int lo_step_match(uint64_t *s, const uint64_t *target, unsigned gens, unsigned stable_interval,
                  unsigned *g_first_false, unsigned *g_recovered);
void lo_soup(uint64_t *out, uint64_t first_id, size_t n, uint64_t seed, int dist);
unsigned lo_pop(const uint64_t *s);
int lo_equal(const uint64_t *a, const uint64_t *b);
void lo_parallel_for(size_t n, int threads, void (*fn)(size_t, size_t, void *), void *arg);

enum { C_OFF = 0, C_ON = 1, C_UNK = 2 };

/* bitslicing/common.py life_rule: B3/S23 on the count of the 8 neighbours. */
static int life_rule(int centre_on, int neighbours) {
  if (centre_on) return neighbours == 2 || neighbours == 3;
  return neighbours == 3;
}

/* bitslicing/unknown_step.py:3-38 stepactive_function: the cell is ON/OFF next generation only if
 * every neighbour count in [on, on+unk] gives that outcome; an UNKNOWN centre stays UNKNOWN. */
static int step_active(int centre, int on, int unk) {
  if (centre == C_UNK) return C_UNK;
  int maybe_on = 0, maybe_off = 0;
  for (int i = on; i <= on + unk; i++) {
    if (life_rule(centre == C_ON, i)) maybe_on = 1; else maybe_off = 1;
  }
  if (maybe_on && maybe_off) return C_UNK;
  return maybe_on ? C_ON : C_OFF;
}

static inline int cell(const uint64_t *p, int x, int y) { return (int)((p[x & 63] >> (y & 63)) & 1); }

/* One three-valued generation.  state/unknown are disjoint planes (LifeStable.hpp:220-221). */
void lo_unknown_step(uint64_t *state, uint64_t *unknown) {
  uint64_t ns[LN], nu[LN];
  memset(ns, 0, sizeof ns);
  memset(nu, 0, sizeof nu);
  for (int x = 0; x < LN; x++)
    for (int y = 0; y < LN; y++) {
      int on = 0, unk = 0;
      for (int dx = -1; dx <= 1; dx++)
        for (int dy = -1; dy <= 1; dy++)
          if (dx || dy) { on += cell(state, x + dx, y + dy); unk += cell(unknown, x + dx, y + dy); }
      int centre = cell(unknown, x, y) ? C_UNK : (cell(state, x, y) ? C_ON : C_OFF);
      int r = step_active(centre, on, unk);
      if (r == C_ON) ns[x] |= 1ULL << y;
      if (r == C_UNK) nu[x] |= 1ULL << y;
    }
  memcpy(state, ns, sizeof ns);
  memcpy(unknown, nu, sizeof nu);
}

typedef struct { uint64_t *state, *unknown; unsigned gens; } us_args;
static void us_range(size_t lo, size_t hi, void *p) {
  us_args *a = (us_args *)p;
  for (size_t i = lo; i < hi; i++)
    for (unsigned g = 0; g < a->gens; g++) lo_unknown_step(a->state + i * LN, a->unknown + i * LN);
}
void lo_unknown_step_batch(uint64_t *state, uint64_t *unknown, size_t n, unsigned gens, int threads) {
  us_args a = { state, unknown, gens };
  lo_parallel_for(n, threads, us_range, &a);
}

/* ---- build-defined checksum: fold(p) = XOR_i splitmix64(state[i] + i*K);
 *      batch = XOR_p splitmix64(fold(p) + id_p).  (include/lifeapi_b200.h lifeb200_fold_dev) */
static inline uint64_t splitmix64(uint64_t z) {
  z += 0x9E3779B97F4A7C15ULL;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
  return z ^ (z >> 31);
}
uint64_t lo_fold_one(const uint64_t *s) {
  uint64_t f = 0;
  for (int i = 0; i < LN; i++) f ^= splitmix64(s[i] + (uint64_t)i * 0xD6E8FEB86659FD93ULL);
  return f;
}
void lo_fold_batch(const uint64_t *states, size_t n, uint64_t first_id, uint64_t *pop_sum, uint64_t *hash_xor) {
  uint64_t ps = 0, hx = 0;
  for (size_t p = 0; p < n; p++) {
    ps += lo_pop(states + p * LN);
    hx ^= splitmix64(lo_fold_one(states + p * LN) + first_id + p);
  }
  *pop_sum = ps;
  *hash_xor = hx;
}

/* ---- collision enumeration: composition of ZOI / Moved / | / Step /
 * Contains(LifeTarget).  hits: id, g_first_false, g_recovered as three parallel arrays, ascending
 * id.  Returns the number of hits (writes at most cap). */
size_t lo_enumerate(const uint64_t *catalysts, unsigned n_cat, const uint64_t *projectiles, unsigned n_proj,
                    uint64_t id_begin, uint64_t id_end, unsigned gens, unsigned stable_interval, uint64_t *hit_id,
                    uint32_t *hit_first_false, uint32_t *hit_recovered, size_t cap, uint64_t *n_skipped) {
  uint64_t total = (uint64_t)n_cat * n_proj * 4096ULL;
  if (id_end > total) id_end = total;
  size_t nh = 0;
  uint64_t skipped = 0;
  for (uint64_t id = id_begin; id < id_end; id++) {
    int dy = (int)(id & 63), dx = (int)((id >> 6) & 63);
    uint64_t cg = id >> 12;
    unsigned g = (unsigned)(cg % n_proj);
    uint64_t c = cg / n_proj;
    const uint64_t *cat = catalysts + c * LN;
    uint64_t moved[LN], cz[LN], pz[LN], s[LN], target[2 * LN];
    lo_moved(projectiles + (size_t)g * LN, dx, dy, moved);
    lo_zoi(cat, cz);
    lo_zoi(moved, pz);
    uint64_t ov = 0;
    for (int i = 0; i < LN; i++) ov |= cz[i] & pz[i];
    if (ov) { skipped++; continue; }
    for (int i = 0; i < LN; i++) s[i] = cat[i] | moved[i];
    lo_target_from_state(cat, target);
    unsigned ff = 0, rec = 0;
    if (lo_step_match(s, target, gens, stable_interval, &ff, &rec)) {
      if (nh < cap) { hit_id[nh] = id; hit_first_false[nh] = ff; hit_recovered[nh] = rec; }
      nh++;
    }
  }
  if (n_skipped) *n_skipped = skipped;
  return nh;
}

/* ---- soup survey: active iff state(G) != state(G-2). */
size_t lo_survey(uint64_t first_id, size_t n, uint64_t seed, int dist, unsigned gens, uint64_t *hit_id,
                 uint32_t *hit_pop, size_t cap, uint64_t *pop_sum, uint64_t *hash_xor) {
  size_t nh = 0;
  uint64_t ps = 0, hx = 0;
  for (size_t p = 0; p < n; p++) {
    uint64_t s[LN], before[LN];
    lo_soup(s, first_id + p, 1, seed, dist);
    for (unsigned g = 0; g + 2 < gens; g++) lo_step(s);
    memcpy(before, s, sizeof s);
    lo_step(s);
    lo_step(s);
    unsigned pop = lo_pop(s);
    ps += pop;
    hx ^= splitmix64(lo_fold_one(s) + first_id + p);
    if (!lo_equal(s, before)) {
      if (nh < cap) { hit_id[nh] = first_id + p; hit_pop[nh] = pop; }
      nh++;
    }
  }
  *pop_sum = ps;
  *hash_xor = hx;
  return nh;
}
