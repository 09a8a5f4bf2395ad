// ref_stable.cpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// The UNMODIFIED reference LifeStable (LifeStable.hpp, with its #included bitslicing/stable_*.hpp
// fragments) behind a C interface, compiled from /root/reference where it lies into oracle/_ref.
// A record is the 10 member planes in declaration order (LifeStable.hpp:39-53) = 5120 bytes.
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
#include "LifeStable.hpp"

static_assert(sizeof(LifeStable) == 5120, "LifeStable layout");

static inline LifeStable load(const uint64_t *p) {
  LifeStable s;
  std::memcpy(static_cast<void *>(&s), p, sizeof s);
  return s;
}
static inline void store(const LifeStable &s, uint64_t *p) { std::memcpy(p, static_cast<const void *>(&s), sizeof s); }

static unsigned flags(const LifeStable::PropagateResult &r) { return (r.consistent ? 1u : 0u) | (r.changed ? 2u : 0u); }

static unsigned run_pass(LifeStable &s, int pass) {
  switch (pass) {
    case 0: return flags(s.PropagateSimpleStep());
    case 1: return flags(s.SynchroniseStateKnown());
    case 2: return flags(s.UpdateOptions());
    case 3: return flags(s.SignalNeighbours());
    case 4: return flags(s.PropagateStep());
    case 5: return flags(s.Propagate());
    case 6: return flags(s.StabiliseOptions());
    case 7: return flags(s.PropagateSimple());
    default: return 0xffu;
  }
}

extern "C" {

unsigned ref_stable_pass(uint64_t *planes, int pass) {
  LifeStable s = load(planes);
  unsigned r = run_pass(s, pass);
  store(s, planes);
  return r;
}

void ref_stable_pass_batch(uint64_t *planes, size_t n, int pass, uint8_t *out_flags, int threads) {
  if (threads <= 0) threads = (int)std::thread::hardware_concurrency();
  if (threads < 1) threads = 1;
  std::vector<std::thread> pool;
  for (int t = 0; t < threads; t++) {
    size_t lo = n * t / threads, hi = n * (t + 1) / threads;
    pool.emplace_back([=] {
      for (size_t i = lo; i < hi; i++) {
        LifeStable s = load(planes + i * 640);
        out_flags[i] = (uint8_t)run_pass(s, pass);
        store(s, planes + i * 640);
      }
    });
  }
  for (auto &th : pool) th.join();
}

void ref_stable_vulnerable(const uint64_t *planes, uint64_t *out) {
  LifeStable s = load(planes);
  LifeState v = s.Vulnerable();
  std::memcpy(out, v.state, 512);
}

void ref_zoi_hollow(const uint64_t *in, uint64_t *out) {
  LifeState s;
  std::memcpy(s.state, in, 512);
  LifeState z = s.ZOIHollow();
  std::memcpy(out, z.state, 512);
}

}  // extern "C"

// A9b: the committed refined covers inside the reference's per-word loop shape, fed by the
// reference's own NeighbourCount / Subtract.
extern "C" void ref_refined_step(const uint64_t *planes, const uint64_t *cur_on, const uint64_t *cur_unk, int variant,
                                 uint64_t *out_on, uint64_t *out_unk, uint64_t *out_unk_stable) {
  LifeStable st = load(planes);
  LifeState con, cunk;
  std::memcpy(con.state, cur_on, 512);
  std::memcpy(cunk.state, cur_unk, 512);
  NeighbourCount sc(st.state), lc(con);
  NeighbourCount mc = sc - lc;
  for (int i = 0; i < N; i++) {
    uint64_t l2 = st.live2[i], l3 = st.live3[i], d0 = st.dead0[i], d1 = st.dead1[i], d2 = st.dead2[i];
    uint64_t d4 = st.dead4[i], d5 = st.dead5[i], d6 = st.dead6[i];
    [[maybe_unused]] uint64_t current_unknown = cunk[i];
    uint64_t current_on = con[i];
    [[maybe_unused]] uint64_t s2 = sc.bit2[i], s1 = sc.bit1[i], s0 = sc.bit0[i];
    uint64_t next_on = 0, next_unknown = 0, next_unknown_stable = 0;
    if (variant == 0) {
      uint64_t m3 = mc.bit3[i], m2 = mc.bit2[i], m1 = mc.bit1[i], m0 = mc.bit0[i];
#include "bitslicing/unknown_step_refined_m.hpp"
    } else {
      uint64_t on2 = lc.bit2[i], on1 = lc.bit1[i], on0 = lc.bit0[i];
#include "bitslicing/unknown_step_refined.hpp"
    }
    out_on[i] = next_on; out_unk[i] = next_unknown; out_unk_stable[i] = next_unknown_stable;
  }
}

// SURVEY.md 8(f) rows through the reference's own members.
#include "LifeWeld.hpp"
static_assert(sizeof(LifeWeld) == 2048, "LifeWeld layout");
extern "C" void ref_weld_step_n(uint64_t *rec, unsigned gens) {
  LifeWeld w;
  std::memcpy(static_cast<void *>(&w), rec, sizeof w);
  for (unsigned g = 0; g < gens; g++) w.Step();
  std::memcpy(rec, static_cast<const void *>(&w), sizeof w);
}
extern "C" void ref_interaction_counts(const uint64_t *s, uint64_t *out) {
  LifeState st, o1, o2, om, nx, p1, p2, pm;
  std::memcpy(st.state, s, 512);
  st.InteractionCountsAndNext(o1, o2, om, nx);
  st.InteractionCounts(p1, p2, pm);
  if (!(p1 == o1 && p2 == o2 && pm == om)) std::abort();  // the two members agree on the shared planes
  std::memcpy(out, o1.state, 512); std::memcpy(out + 64, o2.state, 512);
  std::memcpy(out + 128, om.state, 512); std::memcpy(out + 192, nx.state, 512);
}
