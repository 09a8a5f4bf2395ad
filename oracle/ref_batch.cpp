// ref_batch.cpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// Native batch loops over the UNMODIFIED reference members, spread over all host threads, so that the
// CPU baselines bench.py prints beside the side rows are the reference itself on every core (BASELINE.md
// section 3) and not one thread or one ctypes call per state.  Every body is a call into /root/reference;
// compiled only where /root/reference exists (see oracle/Makefile, ref_harness.cpp).
#include <algorithm>
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
#include "LifeWeld.hpp"

static inline LifeState load(const uint64_t *p) {
  LifeState s;
  std::memcpy(s.state, p, 512);
  return s;
}
static inline void store(const LifeState &s, uint64_t *p) { std::memcpy(p, s.state, 512); }

template <class F>
static void parallel_chunks(size_t n, int threads, F body) {  // body(lo, hi, thread index)
  if (threads <= 0) threads = (int)std::thread::hardware_concurrency();
  if (threads < 1) threads = 1;
  if ((size_t)threads > n) threads = n ? (int)n : 1;
  std::vector<std::thread> pool;
  for (int t = 0; t < threads; t++) {
    size_t lo = n * t / threads, hi = n * (t + 1) / threads;
    pool.emplace_back([=] { body(lo, hi, t); });
  }
  for (auto &th : pool) th.join();
}

extern "C" {

// LifeState::GetHash (LifeAPI.hpp:373) over a batch.
void ref_hash_batch(const uint64_t *states, size_t n, uint64_t *hashes, int threads) {
  parallel_chunks(n, threads, [=](size_t lo, size_t hi, int) {
    for (size_t i = lo; i < hi; i++) hashes[i] = load(states + 64 * i).GetHash();
  });
}

// LifeWeld::Step (LifeWeld.hpp:169-186) `gens` times on a batch of 2048-byte records, in place.
void ref_weld_step_batch(uint64_t *recs, size_t n, unsigned gens, int threads) {
  parallel_chunks(n, threads, [=](size_t lo, size_t hi, int) {
    for (size_t i = lo; i < hi; i++) {
      LifeWeld w;
      std::memcpy(static_cast<void *>(&w), recs + 256 * i, sizeof w);
      for (unsigned g = 0; g < gens; g++) w.Step();
      std::memcpy(recs + 256 * i, static_cast<const void *>(&w), sizeof w);
    }
  });
}

// LifeState::InteractionCountsAndNext (LifeAPI.hpp:997-1040) over a batch: out1, out2, outMore, next.
void ref_interaction_counts_batch(const uint64_t *states, size_t n, uint64_t *planes, int threads) {
  parallel_chunks(n, threads, [=](size_t lo, size_t hi, int) {
    for (size_t i = lo; i < hi; i++) {
      LifeState o1, o2, om, nx;
      load(states + 64 * i).InteractionCountsAndNext(o1, o2, om, nx);
      store(o1, planes + 256 * i); store(o2, planes + 256 * i + 64);
      store(om, planes + 256 * i + 128); store(nx, planes + 256 * i + 192);
    }
  });
}

// LifeWeld::InteractionCounts (LifeWeld.hpp:193-215) over a batch of 2048-byte records: out1, out2, outMore.
void ref_weld_interaction_counts_batch(const uint64_t *recs, size_t n, uint64_t *planes, int threads) {
  parallel_chunks(n, threads, [=](size_t lo, size_t hi, int) {
    for (size_t i = lo; i < hi; i++) {
      LifeWeld w;
      std::memcpy(static_cast<void *>(&w), recs + 256 * i, sizeof w);
      LifeState o1, o2, om;
      w.InteractionCounts(o1, o2, om);
      store(o1, planes + 192 * i); store(o2, planes + 192 * i + 64); store(om, planes + 192 * i + 128);
    }
  });
}

// LifeWeld::InteractionOffsets (LifeWeld.hpp:217-262): out[i] = a[i or 0].InteractionOffsets(b[i or 0]).
void ref_weld_interaction_offsets_batch(const uint64_t *a, size_t n_a, const uint64_t *b, size_t n_b, size_t n, uint64_t *out,
                                        int threads) {
  parallel_chunks(n, threads, [=](size_t lo, size_t hi, int) {
    for (size_t i = lo; i < hi; i++) {
      LifeWeld x, y;
      std::memcpy(static_cast<void *>(&x), a + 256 * (n_a == 1 ? 0 : i), sizeof x);
      std::memcpy(static_cast<void *>(&y), b + 256 * (n_b == 1 ? 0 : i), sizeof y);
      store(x.InteractionOffsets(y), out + 64 * i);
    }
  });
}

// NeighbourCount arithmetic over batches of 2048-byte records (NeighbourCount.hpp:71-102).
void ref_nc_add_batch(const uint64_t *a, const uint64_t *b, const uint64_t *incarry, size_t n, uint64_t *out, int threads) {
  parallel_chunks(n, threads, [=](size_t lo, size_t hi, int) {
    for (size_t i = lo; i < hi; i++) {
      NeighbourCount x, y;
      std::memcpy(&x, a + 256 * i, 2048); std::memcpy(&y, b + 256 * i, 2048);
      NeighbourCount r = incarry ? x.Add(y, load(incarry + 64 * i)) : x + y;
      std::memcpy(out + 256 * i, &r, 2048);
    }
  });
}
void ref_nc_subtract_batch(const uint64_t *a, const uint64_t *b, size_t n, uint64_t *out, int threads) {
  parallel_chunks(n, threads, [=](size_t lo, size_t hi, int) {
    for (size_t i = lo; i < hi; i++) {
      NeighbourCount x, y;
      std::memcpy(&x, a + 256 * i, 2048); std::memcpy(&y, b + 256 * i, 2048);
      NeighbourCount r = x - y;
      std::memcpy(out + 256 * i, &r, 2048);
    }
  });
}
void ref_nc_with_exactly_batch(const uint64_t *a, size_t n, unsigned k, uint64_t *out, int threads) {
  parallel_chunks(n, threads, [=](size_t lo, size_t hi, int) {
    for (size_t i = lo; i < hi; i++) {
      NeighbourCount x;
      std::memcpy(&x, a + 256 * i, 2048);
      store(x.WithExactly(k), out + 64 * i);
    }
  });
}

// LifeState::Contains(const LifeTarget&, int dx, int dy) (LifeTarget.hpp:38-42 -> LifeAPI.hpp:399-422).
int ref_contains_target_shifted(const uint64_t *s, const uint64_t *target, int dx, int dy) {
  LifeTarget t;
  std::memcpy(&t, target, 1024);
  return load(s).Contains(t, dx, dy) ? 1 : 0;
}

// Collision enumeration over [id_begin, id_end) on `threads` host threads: the id range
// is cut into contiguous chunks, each thread runs ref_enumerate's loop on its chunk, the per-thread lists are
// concatenated in id order.  Same composition as ref_enumerate (ref_circuits.cpp), reference calls only.
size_t ref_enumerate(const uint64_t *catalysts, unsigned n_cat, const uint64_t *projectiles, unsigned n_proj,
                     uint64_t id_begin, uint64_t id_end, unsigned gens, unsigned stable_interval, uint64_t *hit_id,
                     uint32_t *hit_first_false, uint32_t *hit_recovered, size_t cap, uint64_t *n_skipped);
size_t ref_enumerate_mt(const uint64_t *catalysts, unsigned n_cat, const uint64_t *projectiles, unsigned n_proj,
                        uint64_t id_begin, uint64_t id_end, unsigned gens, unsigned stable_interval, uint64_t *hit_id,
                        uint32_t *hit_first_false, uint32_t *hit_recovered, size_t cap, uint64_t *n_skipped, int threads) {
  uint64_t total = (uint64_t)n_cat * n_proj * 4096ULL;
  if (id_end > total) id_end = total;
  if (id_begin >= id_end) { if (n_skipped) *n_skipped = 0; return 0; }
  const uint64_t n = id_end - id_begin;
  if (threads <= 0) threads = (int)std::thread::hardware_concurrency();
  // many small chunks: placements differ a lot in cost (skipped ones are free)
  const uint64_t chunk = 1024;
  const size_t n_chunks = (size_t)((n + chunk - 1) / chunk);
  struct Part { std::vector<uint64_t> id; std::vector<uint32_t> ff, rec; uint64_t skipped = 0; };
  std::vector<Part> parts(n_chunks);
  parallel_chunks(n_chunks, threads, [&](size_t lo, size_t hi, int) {
    for (size_t c = lo; c < hi; c++) {
      const uint64_t b = id_begin + c * chunk, e = std::min(id_end, b + chunk);
      Part &p = parts[c];
      p.id.resize(e - b); p.ff.resize(e - b); p.rec.resize(e - b);
      size_t nh = ref_enumerate(catalysts, n_cat, projectiles, n_proj, b, e, gens, stable_interval, p.id.data(),
                                p.ff.data(), p.rec.data(), e - b, &p.skipped);
      p.id.resize(nh); p.ff.resize(nh); p.rec.resize(nh);
    }
  });
  size_t nh = 0;
  uint64_t skipped = 0;
  for (const Part &p : parts) {
    skipped += p.skipped;
    for (size_t k = 0; k < p.id.size(); k++, nh++)
      if (nh < cap) { hit_id[nh] = p.id[k]; hit_first_false[nh] = p.ff[k]; hit_recovered[nh] = p.rec[k]; }
  }
  if (n_skipped) *n_skipped = skipped;
  return nh;
}

// Soup survey predicate on caller-provided soups, all threads: active iff
// state(G) != state(G-2) (operator== LifeAPI.hpp:213-220), pop = GetPop (:290-298); final states in place.
void ref_survey_states_mt(uint64_t *states, size_t n, unsigned gens, uint8_t *active, uint32_t *pop, int threads) {
  parallel_chunks(n, threads, [=](size_t lo, size_t hi, int) {
    for (size_t p = lo; p < hi; p++) {
      LifeState s = load(states + p * 64);
      if (gens >= 2) s.Step(gens - 2);
      LifeState before = s;
      s.Step(2);
      active[p] = (s == before) ? 0 : 1;
      pop[p] = s.GetPop();
      store(s, states + p * 64);
    }
  });
}

}  // extern "C"
