// LifeAPI_b200.hpp -- drop-in host header for the STEPPING PATH of scorbiclife/LifeAPI, backed by
// the B200 CUDA library (include/lifeapi_b200.h).  Same type names, member signatures and memory
// layout as the reference for everything on the path:
//
//   LifeState   { uint64_t state[64]; }  alignas(64)          LifeAPI.hpp:39-40
//       Step(), Step(unsigned), Stepped(), Stepped(unsigned), StepAlt()   :866-886
//       Contains(const LifeTarget&)                                      LifeTarget.hpp:44-51
//       HalfAdd / FullAdd (sum, carry by reference first)                 :822-833,850-864
//   LifeTarget  { wanted, unwanted }                                     LifeTarget.hpp:5-18
//   NeighbourCount { bit3, bit2, bit1, bit0 }                            NeighbourCount.hpp:7-102
//   LifeHistory { state, history, marked, original }                     LifeHistory.hpp:8-12
//   LifeStable  { state, unknown, live2, live3, dead0..dead6 } + full-grid passes  LifeStable.hpp:39-53,366-729
//
// Every stepping member is a batch of ONE through the same sm_100a kernels the batched API uses;
// there is NO CPU implementation of Step in this header and no fallback: if the library cannot
// create a context (no B200) the process aborts with the library's error message, which is the
// reference's error model (it has no recoverable errors).
//
// Throughput comes from the batched API (LifeBatch below / the C ABI), not from single calls.
// Plane algebra (operator&,|,^,~, ==, IsEmpty, GetPop, Moved, ZOI, NeighbourCount::Add ...) is
// host-side value arithmetic exactly as in the reference; it is not part of the accelerated path.
#pragma once

#include <bit>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <ostream>
#include <random>
#include <string>
#include <utility>
#include <vector>

#include "lifeapi_b200.h"

constexpr int N = 64;  // LifeAPI.hpp:12

constexpr unsigned torus_wrap(int x) { return x & (N - 1); }  // LifeAPI.hpp:14-16 (valid for negative x)

// LifeAPI.hpp:18-23: the soup source of RandomState -- 64 draws uniform in [2^61, 2^62] from a mt19937_64 seeded by
// random_device.  PRNG::Seed(s) makes a run reproducible (the reference has no such knob; tests use it).
namespace PRNG {
inline std::mt19937_64& Engine() {
  static std::mt19937_64 e2{std::random_device{}()};
  return e2;
}
inline void Seed(uint64_t s) { Engine().seed(s); }
inline uint64_t Draw() {
  static std::uniform_int_distribution<uint64_t> dist(std::llround(std::pow(2, 61)), std::llround(std::pow(2, 62)));
  return dist(Engine());
}
}  // namespace PRNG

enum class InitializedTag { UNINITIALIZED };  // LifeAPI.hpp:37

namespace lifeb200 {

[[noreturn]] inline void Die(const char* what) {
  std::fprintf(stderr, "LifeAPI_b200: %s: %s\n", what, lifeb200_last_error());
  std::abort();
}
inline void Check(int rc, const char* what) {
  if (rc != LIFEB200_OK) Die(what);
}

// Process-wide default context on device $LIFEB200_DEVICE (default 0), created on first use.
inline lifeb200_ctx* DefaultContext() {
  static lifeb200_ctx* ctx = [] {
    const char* e = std::getenv("LIFEB200_DEVICE");
    lifeb200_ctx* c = nullptr;
    Check(lifeb200_create(e ? std::atoi(e) : 0, &c), "lifeb200_create");
    return c;
  }();
  return ctx;
}

}  // namespace lifeb200

struct LifeTarget;

struct alignas(64) LifeState {
  uint64_t state[N];

  constexpr LifeState() : state{0} {}
  explicit constexpr LifeState(InitializedTag) {}  // LifeAPI.hpp:47: planes every word of which is about to be written
  LifeState& operator=(bool) = delete;             // :55 avoid accidentally coercing from a bool
  LifeState(const LifeState&) = default;
  LifeState& operator=(const LifeState&) = default;

  uint64_t& operator[](unsigned i) { return state[i]; }
  uint64_t operator[](unsigned i) const { return state[i]; }

  static constexpr LifeState Cell(std::pair<int, int> cell) {  // LifeAPI.hpp:57-61
    LifeState result;
    result.Set(cell.first, cell.second);
    return result;
  }
  static LifeState RandomState() {  // LifeAPI.hpp:63-69
    LifeState result;
    for (unsigned i = 0; i < (unsigned)N; i++) result[i] = PRNG::Draw();
    return result;
  }

  // ---- cell access (LifeAPI.hpp:131-146): word = column x, bit = row y ----------------------
  constexpr void Set(unsigned x, unsigned y) { state[x & 63] |= 1ULL << (y & 63); }
  constexpr void Erase(unsigned x, unsigned y) { state[x & 63] &= ~(1ULL << (y & 63)); }
  constexpr bool Get(unsigned x, unsigned y) const { return (state[x & 63] >> (y & 63)) & 1ULL; }
  void Set(std::pair<int, int> c) { Set((unsigned)c.first, (unsigned)c.second); }
  void Erase(std::pair<int, int> c) { Erase((unsigned)c.first, (unsigned)c.second); }
  bool Get(std::pair<int, int> c) const { return Get((unsigned)c.first, (unsigned)c.second); }

  // ---- plane algebra (LifeAPI.hpp:213-298) --------------------------------------------------
  bool operator==(const LifeState& o) const { return std::memcmp(state, o.state, sizeof state) == 0; }
  bool operator!=(const LifeState& o) const { return !(*this == o); }
  LifeState operator~() const { LifeState r; for (int i = 0; i < N; i++) r.state[i] = ~state[i]; return r; }
  LifeState operator&(const LifeState& o) const { LifeState r; for (int i = 0; i < N; i++) r.state[i] = state[i] & o.state[i]; return r; }
  LifeState operator|(const LifeState& o) const { LifeState r; for (int i = 0; i < N; i++) r.state[i] = state[i] | o.state[i]; return r; }
  LifeState operator^(const LifeState& o) const { LifeState r; for (int i = 0; i < N; i++) r.state[i] = state[i] ^ o.state[i]; return r; }
  LifeState& operator&=(const LifeState& o) { for (int i = 0; i < N; i++) state[i] &= o.state[i]; return *this; }
  LifeState& operator|=(const LifeState& o) { for (int i = 0; i < N; i++) state[i] |= o.state[i]; return *this; }
  LifeState& operator^=(const LifeState& o) { for (int i = 0; i < N; i++) state[i] ^= o.state[i]; return *this; }
  bool IsEmpty() const { uint64_t a = 0; for (int i = 0; i < N; i++) a |= state[i]; return a == 0; }
  unsigned GetPop() const { unsigned p = 0; for (int i = 0; i < N; i++) p += (unsigned)std::popcount(state[i]); return p; }

  // ---- geometry used to build placements and targets (LifeAPI.hpp:521-538,699-717) -----------
  LifeState Moved(int x, int y) const {
    LifeState r;
    const unsigned sx = (unsigned)x & 63u;
    for (unsigned i = 0; i < (unsigned)N; i++) r.state[(i + sx) & 63u] = std::rotl(state[i], y);
    return r;
  }
  void Move(int x, int y) { *this = Moved(x, y); }
  void Move(std::pair<int, int> vec) { Move(vec.first, vec.second); }                            // LifeAPI.hpp:697
  LifeState Moved(std::pair<int, int> vec) const { return Moved(vec.first, vec.second); }        // :719-721
  LifeState ZOI() const {
    uint64_t col[N];
    for (int i = 0; i < N; i++) col[i] = state[i] | std::rotl(state[i], 1) | std::rotr(state[i], 1);
    LifeState r;
    for (int i = 0; i < N; i++) r.state[i] = col[(i + N - 1) & 63] | col[i] | col[(i + 1) & 63];
    return r;
  }
  LifeState GetBoundary() const { return ZOI() & ~*this; }

  // ---- RLE input: LifeState::Parse (Parsing.hpp:141-198) through the library's parser, which follows the
  // reference's GenericParse character for character (header lines dropped, counts carried over spaces and
  // newlines, only 'o' sets a cell, unknown characters advance x, the 129-row abort); then Move(dx, dy)
  static LifeState Parse(const std::string& rle) {
    LifeState r;
    lifeb200::Check(lifeb200_parse_rle(rle.c_str(), r.state), "LifeState::Parse");
    return r;
  }
  static LifeState Parse(const std::string& rle, int dx, int dy) { return Parse(rle).Moved(dx, dy); }

  // ---- per-cell queries (LifeAPI.hpp:889-895,1042-1064) --------------------------------------
  unsigned CountNeighboursWithCenter(std::pair<int, int> cell) const {
    unsigned result = 0;
    for (unsigned i = 0; i <= 2; i++) {
      const uint64_t column = std::rotr(state[torus_wrap(cell.first + (int)i - 1)], cell.second - 1);
      result += (unsigned)std::popcount(column & 0b111);
    }
    return result;
  }
  unsigned CountNeighbours(std::pair<int, int> cell) const { return CountNeighboursWithCenter(cell) - (Get(cell) ? 1 : 0); }
  bool StepFor(std::pair<int, int> cell) const {
    const unsigned count = CountNeighbours(cell);
    return Get(cell) ? (count == 2 || count == 3) : count == 3;
  }

  // ---- adder conventions (LifeAPI.hpp:822-833,850-864): (sum, carry) first, by reference ------
  static inline void HalfAdd(uint64_t& out0, uint64_t& out1, uint64_t a, uint64_t b) { out0 = a ^ b; out1 = a & b; }
  static inline void FullAdd(uint64_t& out0, uint64_t& out1, uint64_t a, uint64_t b, uint64_t c) {
    const uint64_t t = a ^ b;
    out0 = t ^ c;
    out1 = (a & b) | (t & c);
  }
  static inline void HalfAdd(LifeState& out0, LifeState& out1, const LifeState& a, const LifeState& b) {
    for (int i = 0; i < N; i++) HalfAdd(out0.state[i], out1.state[i], a.state[i], b.state[i]);
  }
  static inline void FullAdd(LifeState& out0, LifeState& out1, const LifeState& a, const LifeState& b, const LifeState& c) {
    for (int i = 0; i < N; i++) FullAdd(out0.state[i], out1.state[i], a.state[i], b.state[i], c.state[i]);
  }

  // ---- THE PATH: stepping on the B200 (LifeAPI.hpp:866-886,1196-1254) ------------------------
  void Step() { Step(1u); }
  void Step(unsigned numIters) {
    lifeb200::Check(lifeb200_step_batch_host(lifeb200::DefaultContext(), state, state, 1, numIters), "LifeState::Step");
  }
  // StepAlt (LifeAPI.hpp:1218-1254): the reference's SECOND formulation of the same function -- two full adders
  // over the column counts, cc ^= fc & cs, (fs ^ cc) & (fc ^ cs ^ cc) & (a | fs).  Here too it is a second circuit:
  // the `next` plane of the interaction kernel, which evaluates exactly that form (csrc/next.cu), whereas Step()
  // runs the pair-shared select of the step kernels.  tests/StepAltTest.cpp compares the two, as in the reference.
  void StepAlt() {
    alignas(64) uint64_t planes[4 * N];
    lifeb200::Check(lifeb200_interaction_counts_host(lifeb200::DefaultContext(), state, planes, 1), "LifeState::StepAlt");
    std::memcpy(state, planes + 3 * N, sizeof state);
  }
  [[nodiscard]] LifeState Stepped() const { LifeState r = *this; r.Step(); return r; }
  [[nodiscard]] LifeState Stepped(unsigned numIters) const { LifeState r = *this; r.Step(numIters); return r; }

  // CountNeighbourhood (LifeAPI.hpp:909-952): 3x3 sum including the centre, four planes.
  void CountNeighbourhood(LifeState& bit3, LifeState& bit2, LifeState& bit1, LifeState& bit0) const {
    alignas(64) uint64_t planes[4 * N];
    lifeb200::Check(lifeb200_neighbour_count_host(lifeb200::DefaultContext(), state, planes, 1), "CountNeighbourhood");
    std::memcpy(bit3.state, planes, 512); std::memcpy(bit2.state, planes + N, 512);
    std::memcpy(bit1.state, planes + 2 * N, 512); std::memcpy(bit0.state, planes + 3 * N, 512);
  }

  // GetHash (LifeAPI.hpp:373): XXH3_64bits of the 512-byte state, computed on the GPU
  uint64_t GetHash() const {
    uint64_t h = 0;
    lifeb200::Check(lifeb200_hash_batch_host(lifeb200::DefaultContext(), state, &h, 1), "GetHash");
    return h;
  }

  // InteractionCounts / InteractionCountsAndNext (LifeAPI.hpp:956-1040)
  void InteractionCountsAndNext(LifeState& out1, LifeState& out2, LifeState& outMore, LifeState& next) const {
    alignas(64) uint64_t planes[4 * N];
    lifeb200::Check(lifeb200_interaction_counts_host(lifeb200::DefaultContext(), state, planes, 1), "InteractionCounts");
    std::memcpy(out1.state, planes, 512); std::memcpy(out2.state, planes + N, 512);
    std::memcpy(outMore.state, planes + 2 * N, 512); std::memcpy(next.state, planes + 3 * N, 512);
  }
  void InteractionCounts(LifeState& out1, LifeState& out2, LifeState& outMore) const {
    LifeState next;
    InteractionCountsAndNext(out1, out2, outMore, next);
  }

  // ---- locating patterns on the torus (LifeAPI.hpp:427-442,789-795,1066-1095,1293-1370), on the GPU ----
  [[nodiscard]] LifeState Mirrored() const {  // (x, y) -> (-x, -y): FlipX, FlipY, Move(1, 1)
    LifeState r;
    for (int i = 0; i < N; i++) {
      uint64_t v = state[(N - i) & 63], b = 0;
      for (int k = 0; k < 64; k++) b |= ((v >> k) & 1ULL) << (63 - k);
      r.state[i] = std::rotl(b, 1);
    }
    return r;
  }
  [[nodiscard]] LifeState Convolve(const LifeState& other) const {
    LifeState r;
    lifeb200::Check(lifeb200_convolve_batch_host(lifeb200::DefaultContext(), state, 1, other.state, 1, r.state, 1), "Convolve");
    return r;
  }
  [[nodiscard]] LifeState MatchLive(const LifeState& live) const {
    LifeState r;
    lifeb200::Check(lifeb200_match_batch_host(lifeb200::DefaultContext(), state, 1, live.state, nullptr, 1, LIFEB200_MATCH_LIVE, r.state), "MatchLive");
    return r;
  }
  [[nodiscard]] LifeState MatchLiveAndDead(const LifeState& live, const LifeState& dead) const {
    LifeState r;
    lifeb200::Check(lifeb200_match_batch_host(lifeb200::DefaultContext(), state, 1, live.state, dead.state, 1, LIFEB200_MATCH_LIVE_AND_DEAD, r.state), "MatchLiveAndDead");
    return r;
  }
  [[nodiscard]] LifeState Match(const LifeState& live) const {  // dead = live.GetBoundary()
    LifeState r;
    lifeb200::Check(lifeb200_match_batch_host(lifeb200::DefaultContext(), state, 1, live.state, nullptr, 1, LIFEB200_MATCH_BOUNDARY, r.state), "Match");
    return r;
  }
  [[nodiscard]] LifeState Match(const LifeTarget& target) const;  // LifeTarget.hpp:53-54 (defined below)
  [[nodiscard]] LifeState InteractionOffsets(const LifeState& other) const {
    LifeState r;
    lifeb200::Check(lifeb200_interaction_offsets_host(lifeb200::DefaultContext(), state, 1, other.state, 1, r.state, 1), "InteractionOffsets");
    return r;
  }

  // ---- RLE output (Parsing.hpp:8-66,198-202); written from cell (32, 32), as the reference does ----
  [[nodiscard]] std::string RLE() const {
    std::string s(1024, '\0');
    size_t len = 0;
    int rc = lifeb200_rle(state, s.data(), s.size(), &len);
    if (rc == LIFEB200_E_OVERFLOW) {
      s.assign(len + 1, '\0');
      rc = lifeb200_rle(state, s.data(), s.size(), &len);
    }
    lifeb200::Check(rc, "RLE");
    s.resize(len);
    return s;
  }

  bool Contains(const LifeTarget& target) const;  // LifeTarget.hpp:44-51 (defined below)
  bool Contains(const LifeTarget& target, int dx, int dy) const;  // LifeTarget.hpp:38-42 (defined below)
  bool Contains(const LifeState& pat) const { return (*this & pat) == pat; }                 // :388-397
  bool AreDisjoint(const LifeState& pat) const { return (*this & pat).IsEmpty(); }          // :377-386
  bool Contains(const LifeState& pat, int targetDx, int targetDy) const {                    // :399-409
    const int dy = (int)torus_wrap(targetDy);
    for (unsigned i = 0; i < (unsigned)N; i++)
      if ((std::rotr(state[torus_wrap((int)i + targetDx)], dy) & pat[i]) != pat[i]) return false;
    return true;
  }
  bool AreDisjoint(const LifeState& pat, int targetDx, int targetDy) const {                 // :411-422
    const int dy = (int)torus_wrap(targetDy);
    for (unsigned i = 0; i < (unsigned)N; i++)
      if (((~std::rotr(state[torus_wrap((int)i + targetDx)], dy)) & pat[i]) != pat[i]) return false;
    return true;
  }

  friend std::ostream& operator<<(std::ostream& os, LifeState const& self) { return os << self.RLE(); }  // :1176-1178
};
static_assert(sizeof(LifeState) == 512 && alignof(LifeState) == 64, "LifeState layout (LifeAPI.hpp:39-40)");

struct LifeTarget {
  LifeState wanted;
  LifeState unwanted;
  LifeTarget() = default;
  explicit LifeTarget(const LifeState& s) : wanted{s}, unwanted{s.GetBoundary()} {}
  LifeTarget(const LifeState& w, const LifeState& u) : wanted{w}, unwanted{u} {}
  [[nodiscard]] LifeTarget Moved(std::pair<int, int> v) const { return LifeTarget(wanted.Moved(v.first, v.second), unwanted.Moved(v.first, v.second)); }
};
static_assert(sizeof(LifeTarget) == 1024, "LifeTarget layout (LifeTarget.hpp:5-7)");

inline LifeState LifeState::Match(const LifeTarget& t) const {
  LifeState r;
  lifeb200::Check(lifeb200_match_target_batch_host(lifeb200::DefaultContext(), state, 1, t.wanted.state, 1, r.state), "Match(LifeTarget)");
  return r;
}

inline bool LifeState::Contains(const LifeTarget& t, int dx, int dy) const {
  return Contains(t.wanted, dx, dy) && AreDisjoint(t.unwanted, dx, dy);
}

inline bool LifeState::Contains(const LifeTarget& t) const {
  uint64_t d = 0;
  for (int i = 0; i < N; i++) d |= (state[i] ^ t.wanted.state[i]) & (t.wanted.state[i] | t.unwanted.state[i]);
  return d == 0;
}

// NeighbourCount.hpp:7-102 -- construction runs on the GPU; Add/Subtract/WithExactly are the
// reference's ripple plane algebra.
struct NeighbourCount {
  LifeState bit3, bit2, bit1, bit0;
  NeighbourCount() = default;
  explicit NeighbourCount(const LifeState& s) { s.CountNeighbourhood(bit3, bit2, bit1, bit0); }
  NeighbourCount Add(const NeighbourCount& o, const LifeState& incarry) const {
    NeighbourCount r;
    LifeState carry = incarry;
    LifeState::FullAdd(r.bit0, carry, bit0, o.bit0, carry);
    LifeState::FullAdd(r.bit1, carry, bit1, o.bit1, carry);
    LifeState::FullAdd(r.bit2, carry, bit2, o.bit2, carry);
    LifeState::FullAdd(r.bit3, carry, bit3, o.bit3, carry);
    return r;
  }
  NeighbourCount operator~() const { NeighbourCount r; r.bit3 = ~bit3; r.bit2 = ~bit2; r.bit1 = ~bit1; r.bit0 = ~bit0; return r; }
  NeighbourCount operator+(const NeighbourCount& o) const { return Add(o, LifeState()); }
  NeighbourCount Subtract(const NeighbourCount& o) const { return Add(~o, ~LifeState()); }
  NeighbourCount operator-(const NeighbourCount& o) const { return Subtract(o); }
  LifeState WithExactly(unsigned n) const {
    LifeState r = ~LifeState();
    r &= (n & 1) ? bit0 : ~bit0;
    r &= (n & 2) ? bit1 : ~bit1;
    r &= (n & 4) ? bit2 : ~bit2;
    r &= (n & 8) ? bit3 : ~bit3;
    return r;
  }
};
static_assert(sizeof(NeighbourCount) == 2048, "NeighbourCount layout (NeighbourCount.hpp:7-11)");

// LifeHistory.hpp:8-12: passive 4-plane container (no Step in the reference).
struct LifeHistory {
  LifeState state, history, marked, original;
  // RLE / Parse in the reference's LifeHistory alphabet (LifeHistory.hpp:25-44,62-93)
  [[nodiscard]] std::string RLE() const {
    std::string s(4096, '\0');
    size_t len = 0;
    int rc = lifeb200_history_rle(state.state, s.data(), s.size(), &len);
    if (rc == LIFEB200_E_OVERFLOW) {
      s.assign(len + 1, '\0');
      rc = lifeb200_history_rle(state.state, s.data(), s.size(), &len);
    }
    lifeb200::Check(rc, "LifeHistory::RLE");
    s.resize(len);
    return s;
  }
  [[nodiscard]] std::string RLEWHeader() const { return "x = 0, y = 0, rule = LifeHistory\n" + RLE(); }
  static LifeHistory Parse(const std::string& rle) {
    LifeHistory h;
    lifeb200::Check(lifeb200_history_parse_rle(rle.c_str(), h.state.state), "LifeHistory::Parse");
    return h;
  }
  // build-defined accumulate step: n times { state.Step(); history |= state; } in one launch
  void StepAccumulate(unsigned n = 1) {
    lifeb200::Check(lifeb200_history_step_batch_host(lifeb200::DefaultContext(), state.state, 1, n), "StepAccumulate");
  }
};
static_assert(sizeof(LifeHistory) == 2048, "LifeHistory layout (LifeHistory.hpp:8-12)");

// LifeWeld.hpp:18-63,169-191: state + frozen neighbour counts; Step adds the frozen count to the true
// 3x3 count before the rule.  Runs on the GPU (lifeb200_weld_step_batch_dev).
struct LifeWeld {
  LifeState state, frozen2, frozen1, frozen0;
  LifeWeld() = default;
  LifeWeld(const LifeState& s) : state{s} {}
  LifeWeld(const LifeState& s, const LifeState& f2, const LifeState& f1, const LifeState& f0)
      : state{s}, frozen2{f2}, frozen1{f1}, frozen0{f0} {}
  bool operator==(const LifeWeld& o) const { return std::memcmp(this, &o, sizeof *this) == 0; }
  LifeState AllFrozen() const { return frozen2 | frozen1 | frozen0; }
  bool Contains(const LifeTarget& target) const { return state.Contains(target); }
  uint64_t GetHash() const {  // LifeWeld.hpp:61-63
    return 3 * frozen2.GetHash() + (3 * frozen1.GetHash() + (3 * frozen0.GetHash() + state.GetHash()));
  }
  void Step(unsigned numIters = 1) {  // LifeWeld.hpp:169-186
    lifeb200::Check(lifeb200_weld_step_batch_host(lifeb200::DefaultContext(), state.state, 1, numIters), "LifeWeld::Step");
  }
  LifeTarget ToTarget() const;  // LifeWeld.hpp:188-191 (defined after LifeTarget)
  LifeState InteractionOffsets(const LifeWeld& b) const {  // LifeWeld.hpp:217-262
    LifeState r;
    lifeb200::Check(lifeb200_weld_interaction_offsets_host(lifeb200::DefaultContext(), state.state, 1, b.state.state, 1, r.state, 1),
                    "LifeWeld::InteractionOffsets");
    return r;
  }
  void InteractionCounts(LifeState& out1, LifeState& out2, LifeState& outMore) const {  // LifeWeld.hpp:193-215
    alignas(64) uint64_t planes[3 * N];
    lifeb200::Check(lifeb200_weld_interaction_counts_host(lifeb200::DefaultContext(), state.state, planes, 1), "LifeWeld::InteractionCounts");
    std::memcpy(out1.state, planes, 512); std::memcpy(out2.state, planes + N, 512); std::memcpy(outMore.state, planes + 2 * N, 512);
  }
};
static_assert(sizeof(LifeWeld) == 2048, "LifeWeld layout (LifeWeld.hpp:18-22)");
inline LifeTarget LifeWeld::ToTarget() const {
  const LifeState nonFrozen = state & ~frozen2 & ~frozen1 & ~frozen0;
  return LifeTarget(state, nonFrozen.ZOI() & ~state);
}

// LifeStable.hpp:36-131: the 10 planes in the reference's declaration order; option planes are
// stored inverted (1 = ruled out, LifeStable.hpp:44-46).  The full-grid propagation passes run on
// the GPU and leave *this exactly as the reference members do; the strip variants and the
// CompleteStable search are not part of this header.
struct LifeStable {
  LifeState state, unknown, live2, live3, dead0, dead1, dead2, dead4, dead5, dead6;

  struct PropagateResult {
    bool consistent;
    bool changed;
  };

  bool operator==(const LifeStable& o) const { return std::memcmp(this, &o, sizeof *this) == 0; }

  void SetOn(const LifeState& which) {  // LifeStable.hpp:320-328
    state |= which; unknown &= ~which;
    dead0 |= which; dead1 |= which; dead2 |= which; dead4 |= which; dead5 |= which; dead6 |= which;
  }
  void SetOff(const LifeState& which) {  // LifeStable.hpp:329-335
    state &= ~which; unknown &= ~which; live2 |= which; live3 |= which;
  }

  PropagateResult PropagateSimpleStep() { return Run(LIFEB200_STABLE_PROPAGATE_SIMPLE_STEP); }  // :414-503
  PropagateResult PropagateSimple() { return Run(LIFEB200_STABLE_PROPAGATE_SIMPLE); }           // :505-524
  PropagateResult SynchroniseStateKnown() { return Run(LIFEB200_STABLE_SYNCHRONISE); }          // :526-556
  PropagateResult UpdateOptions() { return Run(LIFEB200_STABLE_UPDATE_OPTIONS); }               // :558-615
  PropagateResult SignalNeighbours() { return Run(LIFEB200_STABLE_SIGNAL_NEIGHBOURS); }         // :617-675
  PropagateResult StabiliseOptions() { return Run(LIFEB200_STABLE_STABILISE_OPTIONS); }         // :677-693
  PropagateResult PropagateStep() { return Run(LIFEB200_STABLE_PROPAGATE_STEP); }               // :695-716
  PropagateResult Propagate() { return Run(LIFEB200_STABLE_PROPAGATE); }                        // :718-729

  LifeState Vulnerable() const {  // :366-412
    LifeState out;
    lifeb200::Check(lifeb200_stable_vulnerable_host(lifeb200::DefaultContext(), state.state, out.state, 1), "Vulnerable");
    return out;
  }

  // batched form: one pass (LIFEB200_STABLE_*) over an array of records, results[i] per record
  static void PassBatch(LifeStable* recs, size_t n, int pass, PropagateResult* results, lifeb200_ctx* ctx = nullptr) {
    std::vector<uint8_t> flags(n);
    lifeb200::Check(lifeb200_stable_pass_host(ctx ? ctx : lifeb200::DefaultContext(), reinterpret_cast<uint64_t*>(recs), n,
                                              pass, flags.data()), "LifeStable::PassBatch");
    if (results)
      for (size_t i = 0; i < n; i++)
        results[i] = {(flags[i] & LIFEB200_STABLE_CONSISTENT) != 0, (flags[i] & LIFEB200_STABLE_CHANGED) != 0};
  }
  static void PropagateBatch(LifeStable* recs, size_t n, PropagateResult* results, lifeb200_ctx* ctx = nullptr) {
    PassBatch(recs, n, LIFEB200_STABLE_PROPAGATE, results, ctx);
  }

 private:
  PropagateResult Run(int pass) {
    PropagateResult r;
    PassBatch(this, 1, pass, &r);
    return r;
  }
};
static_assert(sizeof(LifeStable) == LIFEB200_STABLE_BYTES, "LifeStable layout (LifeStable.hpp:39-53)");

// ---------------------------------------------------------------------------------------------
// Batched API: a std::vector<LifeState> is already the flat AoS batch.
// ---------------------------------------------------------------------------------------------
struct LifeHit {
  uint64_t id;
  uint32_t g_first_false, g_recovered;
};
static_assert(sizeof(LifeHit) == sizeof(lifeb200_hit), "hit record layout");

// RAII page-lock of a caller-owned batch: LifeBatch::Step(batch, n) on a std::vector copies through pageable memory
// (correct, but staged by the driver and not overlapped); pinning the vector's storage once makes every later host
// call on it as fast as one on lifeb200_host_alloc memory.  The vector must not reallocate while pinned.
class LifePinned {
 public:
  explicit LifePinned(std::vector<LifeState>& batch, lifeb200_ctx* ctx = nullptr)
      : ctx_(ctx ? ctx : lifeb200::DefaultContext()), ptr_(batch.empty() ? nullptr : batch.data()) {
    if (ptr_) lifeb200::Check(lifeb200_host_register(ctx_, ptr_, batch.size() * sizeof(LifeState)), "lifeb200_host_register");
  }
  ~LifePinned() { if (ptr_) lifeb200_host_unregister(ctx_, ptr_); }
  LifePinned(const LifePinned&) = delete;
  LifePinned& operator=(const LifePinned&) = delete;

 private:
  lifeb200_ctx* ctx_;
  void* ptr_;
};

class LifeBatch {
 public:
  explicit LifeBatch(lifeb200_ctx* ctx = nullptr) : ctx_(ctx ? ctx : lifeb200::DefaultContext()) {}
  ~LifeBatch() { if (d_) lifeb200_free(ctx_, d_); }
  LifeBatch(const LifeBatch&) = delete;
  LifeBatch& operator=(const LifeBatch&) = delete;

  // batch[i].Step(gens) for every i (LifeAPI.hpp:877-881), host in / host out, pipelined copies
  static void Step(std::vector<LifeState>& batch, unsigned gens, lifeb200_ctx* ctx = nullptr) {
    if (batch.empty()) return;
    lifeb200::Check(lifeb200_step_batch_host(ctx ? ctx : lifeb200::DefaultContext(), batch[0].state, batch[0].state,
                                             batch.size(), gens), "LifeBatch::Step");
  }

  // resident batch: upload once, step many times, download when needed
  void Upload(const std::vector<LifeState>& batch) {
    Reserve(batch.size());
    n_ = batch.size();
    if (n_) lifeb200::Check(lifeb200_memcpy_h2d(ctx_, d_, batch.data(), n_ * 512), "Upload");
  }
  void Soups(uint64_t first_id, size_t n, uint64_t seed, int dist = 0) {
    Reserve(n);
    n_ = n;
    lifeb200::Check(lifeb200_soup_dev(ctx_, static_cast<uint64_t*>(d_), first_id, n, seed, dist), "Soups");
  }
  void Step(unsigned gens) {
    lifeb200::Check(lifeb200_step_batch_dev(ctx_, static_cast<uint64_t*>(d_), static_cast<uint64_t*>(d_), n_, gens), "Step");
  }
  std::vector<LifeState> Download() {
    std::vector<LifeState> out(n_);
    if (n_) lifeb200::Check(lifeb200_memcpy_d2h(ctx_, out.data(), d_, n_ * 512), "Download");
    lifeb200::Check(lifeb200_sync(ctx_), "sync");
    return out;
  }
  size_t size() const { return n_; }

  // catalyst x projectile collision enumeration with fused Contains(LifeTarget(catalyst)) and
  // stabilisation match; hits sorted by placement id
  static std::vector<LifeHit> EnumerateCollisions(const std::vector<LifeState>& catalysts,
                                                  const std::vector<LifeState>& projectiles, unsigned gens,
                                                  unsigned stable_interval, uint64_t* n_skipped = nullptr,
                                                  lifeb200_ctx* ctx = nullptr) {
    ctx = ctx ? ctx : lifeb200::DefaultContext();
    const uint64_t total = (uint64_t)catalysts.size() * projectiles.size() * 4096ull;
    std::vector<LifeHit> hits(4096);
    for (;;) {
      uint64_t nh = 0, ns = 0;
      int rc = lifeb200_enumerate_collisions(ctx, catalysts[0].state, (unsigned)catalysts.size(), projectiles[0].state,
                                             (unsigned)projectiles.size(), 0, total, gens, stable_interval,
                                             reinterpret_cast<lifeb200_hit*>(hits.data()), hits.size(), &nh, &ns);
      if (rc == LIFEB200_E_OVERFLOW) { hits.resize(nh); continue; }
      lifeb200::Check(rc, "EnumerateCollisions");
      hits.resize(nh);
      if (n_skipped) *n_skipped = ns;
      return hits;
    }
  }

 private:
  void Reserve(size_t n) {
    if (n <= cap_) return;
    if (d_) lifeb200::Check(lifeb200_free(ctx_, d_), "free");
    d_ = nullptr;
    lifeb200::Check(lifeb200_malloc(ctx_, &d_, n * 512), "malloc");
    cap_ = n;
  }
  lifeb200_ctx* ctx_;
  void* d_ = nullptr;
  size_t n_ = 0, cap_ = 0;
};

// ---------------------------------------------------------------------------------------------
// Multi-GPU batches (SURVEY.md 8b "init with device list", "multi-GPU run + gather"; 8e): a group of GPUs in this
// process, every call sharded by contiguous id range, hit lists gathered over NCCL into this process, sorted by id.
// The reference is single-threaded; this is the loop a caller shards (LifeAPI.hpp:877-881) done by the library.
// ---------------------------------------------------------------------------------------------
class LifeMultiBatch {
 public:
  // the first n_devices GPUs (0 = every visible GPU), or an explicit device list
  explicit LifeMultiBatch(int n_devices = 0) { lifeb200::Check(lifeb200_multi_create(nullptr, n_devices, &m_), "lifeb200_multi_create"); }
  explicit LifeMultiBatch(const std::vector<int>& devices) {
    lifeb200::Check(lifeb200_multi_create(devices.data(), (int)devices.size(), &m_), "lifeb200_multi_create");
  }
  ~LifeMultiBatch() { lifeb200_multi_destroy(m_); }
  LifeMultiBatch(const LifeMultiBatch&) = delete;
  LifeMultiBatch& operator=(const LifeMultiBatch&) = delete;

  int World() const { int w = 0; lifeb200_multi_info(m_, &w, nullptr, nullptr, nullptr); return w; }
  int NcclVersion() const { int v = 0; lifeb200_multi_info(m_, nullptr, nullptr, nullptr, &v); return v; }
  // size the scratch of every GPU for calls over up to n ids / cap hit records (no allocation in later calls of that size)
  void Reserve(uint64_t n, size_t cap) { lifeb200::Check(lifeb200_multi_reserve(m_, n, cap), "lifeb200_multi_reserve"); }
  float LastGatherMs() const { float ms = 0; lifeb200_multi_last_gather_ms(m_, &ms); return ms; }
  lifeb200_ctx* Context(int local_index) const { lifeb200_ctx* c = nullptr; lifeb200::Check(lifeb200_multi_ctx(m_, local_index, &c), "lifeb200_multi_ctx"); return c; }

  // batch[i].Step(gens) for every i, the batch split over the GPUs (one host thread and one copy pipeline per GPU)
  void Step(std::vector<LifeState>& batch, unsigned gens) {
    if (batch.empty()) return;
    lifeb200::Check(lifeb200_multi_step_batch_host(m_, batch[0].state, batch[0].state, batch.size(), gens), "LifeMultiBatch::Step");
  }
  // config 5: on-device soups, active iff state(G) != state(G - 2); returns the number of hits (hits resized to fit)
  size_t SoupSurvey(uint64_t first_id, uint64_t n, uint64_t seed, int dist, unsigned gens, std::vector<LifeHit>& hits,
                    uint64_t* pop_sum = nullptr, uint64_t* hash_xor = nullptr) {
    if (hits.size() < 4096) hits.resize(4096);
    for (;;) {
      uint64_t nh = 0;
      const int rc = lifeb200_multi_soup_survey(m_, first_id, n, seed, dist, gens, reinterpret_cast<lifeb200_hit*>(hits.data()),
                                                hits.size(), &nh, pop_sum, hash_xor);
      if (rc == LIFEB200_E_OVERFLOW) { hits.resize(nh); continue; }
      lifeb200::Check(rc, "LifeMultiBatch::SoupSurvey");
      return (size_t)nh;
    }
  }
  // config 4 over the group; hits sorted by placement id
  std::vector<LifeHit> EnumerateCollisions(const std::vector<LifeState>& catalysts, const std::vector<LifeState>& projectiles,
                                           unsigned gens, unsigned stable_interval, uint64_t* n_skipped = nullptr) {
    const uint64_t total = (uint64_t)catalysts.size() * projectiles.size() * 4096ull;
    std::vector<LifeHit> hits(4096);
    for (;;) {
      uint64_t nh = 0, ns = 0;
      const int rc = lifeb200_multi_enumerate_collisions(m_, catalysts[0].state, (unsigned)catalysts.size(), projectiles[0].state,
                                                         (unsigned)projectiles.size(), 0, total, gens, stable_interval,
                                                         reinterpret_cast<lifeb200_hit*>(hits.data()), hits.size(), &nh, &ns);
      if (rc == LIFEB200_E_OVERFLOW) { hits.resize(nh); continue; }
      lifeb200::Check(rc, "LifeMultiBatch::EnumerateCollisions");
      hits.resize(nh);
      if (n_skipped) *n_skipped = ns;
      return hits;
    }
  }

 private:
  lifeb200_multi* m_ = nullptr;
};
