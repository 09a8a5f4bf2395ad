// ref_circuits.cpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// Harnesses around the UNMODIFIED reference for the things the reference ships only as building
// blocks: the committed bitslicing/*.hpp circuit text is #included from
// /root/reference at compile time inside the per-word loop shape the reference itself uses
// (LifeStable.hpp:429-469), and the enumeration / survey compositions are written only in terms
// of reference member functions.  Compiled only where /root/reference exists; see ref_harness.cpp.
#include <bit>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <string>
#include <thread>
#include <vector>

#include "LifeAPI.hpp"
#include "Parsing.hpp"
#include "Symmetry.hpp"
#include "LifeStrip.hpp"
#include "LifeTarget.hpp"
#include "NeighbourCount.hpp"

static inline LifeState load(const uint64_t *p) {
  LifeState s;
  std::memcpy(s.state, p, 512);
  return s;
}
static inline void store(const LifeState &s, uint64_t *p) { std::memcpy(p, s.state, 512); }

// One three-valued generation through bitslicing/unknown_step.hpp (inputs: SURVEY.md 8a A9).
static void unknown_step_once(LifeState &state, LifeState &unknown) {
  NeighbourCount onCount(state);
  NeighbourCount unkCount(unknown);
  LifeState next_on, next_unknown;
  for (int i = 0; i < N; i++) {
    uint64_t current_on = state[i];
    uint64_t current_unknown = unknown[i];
    [[maybe_unused]] uint64_t on3 = onCount.bit3[i];
    uint64_t on2 = onCount.bit2[i];
    uint64_t on1 = onCount.bit1[i];
    uint64_t on0 = onCount.bit0[i];
    uint64_t unk3 = unkCount.bit3[i];
    uint64_t unk2 = unkCount.bit2[i];
    uint64_t unk1 = unkCount.bit1[i];
    uint64_t unk0 = unkCount.bit0[i];
    uint64_t naive_next_unknown = 0;
    uint64_t naive_next_on = 0;
#include "bitslicing/unknown_step.hpp"
    next_on[i] = naive_next_on;
    next_unknown[i] = naive_next_unknown;
  }
  state = next_on;
  unknown = next_unknown;
}

extern "C" {

void ref_unknown_step(uint64_t *state, uint64_t *unknown) {
  LifeState s = load(state), u = load(unknown);
  unknown_step_once(s, u);
  store(s, state);
  store(u, unknown);
}

void ref_unknown_step_batch(uint64_t *state, uint64_t *unknown, size_t n, unsigned gens, int threads) {
  if (threads <= 0) threads = (int)std::thread::hardware_concurrency();
  if (threads < 1) threads = 1;
  std::vector<std::thread> pool;
  for (int t = 0; t < threads; t++) {
    size_t lo = n * t / threads, hi = n * (t + 1) / threads;
    pool.emplace_back([=] {
      for (size_t p = lo; p < hi; p++) {
        LifeState s = load(state + p * 64), u = load(unknown + p * 64);
        for (unsigned g = 0; g < gens; g++) unknown_step_once(s, u);
        store(s, state + p * 64);
        store(u, unknown + p * 64);
      }
    });
  }
  for (auto &th : pool) th.join();
}

// Collision enumeration, reference calls only.
size_t ref_enumerate(const uint64_t *catalysts, unsigned n_cat, const uint64_t *projectiles, unsigned n_proj,
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
    LifeState cat = load(catalysts + c * 64);
    LifeState moved = load(projectiles + (size_t)g * 64).Moved(dx, dy);
    if (!(cat.ZOI() & moved.ZOI()).IsEmpty()) { skipped++; continue; }
    LifeState st = cat | moved;
    LifeTarget target(cat);
    unsigned first_false = 0, run = 0;
    bool seen_false = false;
    for (unsigned gen = 1; gen <= gens; gen++) {
      st.Step();
      bool ok = st.Contains(target);
      if (!ok) { if (!seen_false) { seen_false = true; first_false = gen; } run = 0; }
      else if (seen_false && ++run == stable_interval) {
        if (nh < cap) { hit_id[nh] = id; hit_first_false[nh] = first_false; hit_recovered[nh] = gen + 1 - stable_interval; }
        nh++;
        break;
      }
    }
  }
  if (n_skipped) *n_skipped = skipped;
  return nh;
}

// Soup survey predicate on caller-provided soups: active iff
// state(G) != state(G-2); pops via GetPop.  active[p] in {0,1}.
void ref_survey_states(uint64_t *states, size_t n, unsigned gens, uint8_t *active, uint32_t *pop) {
  for (size_t p = 0; p < n; p++) {
    LifeState s = load(states + p * 64);
    if (gens >= 2) s.Step(gens - 2);
    LifeState before = s;
    s.Step(2);
    active[p] = (s == before) ? 0 : 1;
    pop[p] = s.GetPop();
    store(s, states + p * 64);
  }
}

}  // extern "C"
