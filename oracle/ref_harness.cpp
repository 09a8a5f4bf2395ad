// ref_harness.cpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// Thin extern "C" shims around the UNMODIFIED reference headers.  This file contains no
// LifeAPI logic of its own: every function body is a call into /root/reference (found via
// -I at compile time, see oracle/Makefile).  It is compiled only where /root/reference
// exists (this container); the resulting oracle/_ref/libliferef_*.so travels to the GPU box
// as a prebuilt binary.  Used (a) to pin oracle/life_oracle.c, (b) to generate
// tests/golden/*, (c) as the CPU baseline "reference" arm of bench.py.
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
#include "LifeHistory.hpp"

static_assert(sizeof(LifeState) == 512, "LifeState layout");
static_assert(sizeof(LifeTarget) == 1024, "LifeTarget layout");
static_assert(sizeof(NeighbourCount) == 2048, "NeighbourCount layout");

static inline LifeState load(const uint64_t *p) {
  LifeState s;
  std::memcpy(s.state, p, 512);
  return s;
}
static inline void store(const LifeState &s, uint64_t *p) { std::memcpy(p, s.state, 512); }

extern "C" {

const char *ref_isa(void) {
#if defined(__AVX512F__)
  return "avx512";
#elif defined(__AVX2__)
  return "avx2";
#else
  return "sse";
#endif
}

// LifeState::Step(unsigned) on a batch, contiguous chunks over `threads` std::threads.
void ref_step_batch(uint64_t *states, size_t n, unsigned gens, int threads) {
  LifeState *b = reinterpret_cast<LifeState *>(states);  // caller guarantees 64-B alignment
  if (threads <= 0) threads = (int)std::thread::hardware_concurrency();
  if (threads <= 1 || n < 2) {
    for (size_t i = 0; i < n; i++) b[i].Step(gens);
    return;
  }
  std::vector<std::thread> pool;
  for (int t = 0; t < threads; t++) {
    size_t lo = n * t / threads, hi = n * (t + 1) / threads;
    pool.emplace_back([=] { for (size_t i = lo; i < hi; i++) b[i].Step(gens); });
  }
  for (auto &th : pool) th.join();
}

int ref_hardware_threads(void) { return (int)std::thread::hardware_concurrency(); }

void ref_step(uint64_t *s) { LifeState t = load(s); t.Step(); store(t, s); }
void ref_step_alt(uint64_t *s) { LifeState t = load(s); t.StepAlt(); store(t, s); }
void ref_step_n(uint64_t *s, unsigned gens) { LifeState t = load(s); t.Step(gens); store(t, s); }

uint64_t ref_hash(const uint64_t *s) { return load(s).GetHash(); }
unsigned ref_pop(const uint64_t *s) { return load(s).GetPop(); }

// LifeState::Parse (NOT ConstantParse: bare-'$' bug).
void ref_parse(const char *rle, uint64_t *out) { store(LifeState::Parse(rle), out); }
void ref_parse_at(const char *rle, int dx, int dy, uint64_t *out) {
  LifeState s = LifeState::Parse(rle);
  s.Move(dx, dy);
  store(s, out);
}

void ref_moved(const uint64_t *s, int x, int y, uint64_t *out) { store(load(s).Moved(x, y), out); }
void ref_zoi(const uint64_t *s, uint64_t *out) { store(load(s).ZOI(), out); }

// planes out = [bit3|bit2|bit1|bit0] (NeighbourCount member order).
void ref_neighbour_count(const uint64_t *s, uint64_t *planes) {
  NeighbourCount nc(load(s));
  std::memcpy(planes, &nc, 2048);
}
void ref_count_neighbourhood(const uint64_t *s, uint64_t *planes) {
  LifeState b3, b2, b1, b0;
  load(s).CountNeighbourhood(b3, b2, b1, b0);
  store(b3, planes); store(b2, planes + 64); store(b1, planes + 128); store(b0, planes + 192);
}
void ref_nc_add(const uint64_t *a, const uint64_t *b, const uint64_t *incarry, uint64_t *out) {
  NeighbourCount x, y;
  std::memcpy(&x, a, 2048); std::memcpy(&y, b, 2048);
  NeighbourCount r = x.Add(y, load(incarry));
  std::memcpy(out, &r, 2048);
}
void ref_nc_subtract(const uint64_t *a, const uint64_t *b, uint64_t *out) {
  NeighbourCount x, y;
  std::memcpy(&x, a, 2048); std::memcpy(&y, b, 2048);
  NeighbourCount r = x - y;
  std::memcpy(out, &r, 2048);
}
void ref_nc_with_exactly(const uint64_t *planes, unsigned n, uint64_t *out) {
  NeighbourCount x;
  std::memcpy(&x, planes, 2048);
  store(x.WithExactly(n), out);
}

// LifeTarget(state): wanted | unwanted.
void ref_target_from_state(const uint64_t *s, uint64_t *target) {
  LifeTarget t(load(s));
  std::memcpy(target, &t, 1024);
}
int ref_contains_target(const uint64_t *s, const uint64_t *target) {
  LifeTarget t;
  std::memcpy(&t, target, 1024);
  return load(s).Contains(t) ? 1 : 0;
}

// Same composition as lo_step_match, built only from reference calls.
int ref_step_match(uint64_t *s, const uint64_t *target, unsigned gens, unsigned stable_interval,
                   unsigned *g_first_false, unsigned *g_recovered) {
  LifeState st = load(s);
  LifeTarget t;
  std::memcpy(&t, target, 1024);
  unsigned first_false = 0, run = 0;
  bool seen_false = false;
  int hit = 0;
  for (unsigned g = 1; g <= gens; g++) {
    st.Step();
    bool c = st.Contains(t);
    if (!c) { if (!seen_false) { seen_false = true; first_false = g; } run = 0; }
    else if (seen_false) {
      if (++run == stable_interval) { *g_first_false = first_false; *g_recovered = g + 1 - stable_interval; hit = 1; break; }
    }
  }
  store(st, s);
  return hit;
}

// ---- Convolve / Match / InteractionOffsets / RLE, the reference's own members ----
void ref_mirrored(const uint64_t *s, uint64_t *out) { store(load(s).Mirrored(), out); }
void ref_convolve(const uint64_t *a, const uint64_t *b, uint64_t *out) { store(load(a).Convolve(load(b)), out); }
void ref_match_live(const uint64_t *s, const uint64_t *live, uint64_t *out) { store(load(s).MatchLive(load(live)), out); }
void ref_match_live_and_dead(const uint64_t *s, const uint64_t *live, const uint64_t *dead, uint64_t *out) {
  store(load(s).MatchLiveAndDead(load(live), load(dead)), out);
}
void ref_match_state(const uint64_t *s, const uint64_t *live, uint64_t *out) { store(load(s).Match(load(live)), out); }
void ref_match_target(const uint64_t *s, const uint64_t *target, uint64_t *out) {
  LifeTarget t;
  std::memcpy(&t, target, 1024);
  store(load(s).Match(t), out);
}
void ref_interaction_offsets(const uint64_t *a, const uint64_t *b, uint64_t *out) {
  store(load(a).InteractionOffsets(load(b)), out);
}
static size_t copy_out(const std::string &r, char *buf, size_t cap) {
  if (cap) { size_t k = r.size() < cap - 1 ? r.size() : cap - 1; std::memcpy(buf, r.data(), k); buf[k] = 0; }
  return r.size();
}
size_t ref_rle(const uint64_t *s, char *buf, size_t cap) { return copy_out(load(s).RLE(), buf, cap); }
size_t ref_history_rle(const uint64_t *rec, char *buf, size_t cap) {
  LifeHistory h(load(rec), load(rec + 64), load(rec + 128), load(rec + 192));
  return copy_out(h.RLE(), buf, cap);
}
void ref_history_parse(const char *rle, uint64_t *rec) {
  LifeHistory h = LifeHistory::Parse(rle);
  store(h.state, rec); store(h.history, rec + 64); store(h.marked, rec + 128); store(h.original, rec + 192);
}
// batch timing helpers for bench_extras (CPU baseline of the same rows), contiguous chunks over threads
void ref_match_target_batch(const uint64_t *states, size_t n, const uint64_t *target, uint64_t *out, int threads) {
  LifeTarget t;
  std::memcpy(&t, target, 1024);
  if (threads <= 0) threads = (int)std::thread::hardware_concurrency();
  std::vector<std::thread> pool;
  for (int k = 0; k < threads; k++) {
    size_t lo = n * k / threads, hi = n * (k + 1) / threads;
    pool.emplace_back([=] { for (size_t i = lo; i < hi; i++) store(load(states + 64 * i).Match(t), out + 64 * i); });
  }
  for (auto &th : pool) th.join();
}
void ref_interaction_offsets_batch(const uint64_t *a, const uint64_t *b, size_t n, uint64_t *out, int threads) {
  if (threads <= 0) threads = (int)std::thread::hardware_concurrency();
  std::vector<std::thread> pool;
  LifeState as = load(a);
  for (int k = 0; k < threads; k++) {
    size_t lo = n * k / threads, hi = n * (k + 1) / threads;
    pool.emplace_back([=] { for (size_t i = lo; i < hi; i++) store(as.InteractionOffsets(load(b + 64 * i)), out + 64 * i); });
  }
  for (auto &th : pool) th.join();
}

}  // extern "C"
