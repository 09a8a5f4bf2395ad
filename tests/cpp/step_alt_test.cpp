// step_alt_test.cpp -- the reference's tests/StepAltTest.cpp:5-13 pattern (randomised differential,
// bit-exact ==) written against the drop-in header include/LifeAPI_b200.hpp.  Where the reference
// compares Step() with StepAlt(), this compares the B200 Step() with the oracle (oracle/ C port,
// linked here as the checker) -- same call sites, same EXPECT_EQ discipline, no gtest needed.
//
// Build (tests/test_cpp_header.py does this):
//   g++ -std=c++20 -O2 -Iinclude tests/cpp/step_alt_test.cpp -o step_alt_test
//       lifeapi_b200/_lib/liblifeapi_b200.so oracle/_build/liblifeoracle.so
#include <cstdio>
#include <random>
#include <sstream>
#include <vector>

#include "LifeAPI_b200.hpp"

extern "C" {
void lo_step(uint64_t* s);
void lo_step_alt(uint64_t* s);
void lo_step_n(uint64_t* s, unsigned gens);
void lo_neighbour_count(const uint64_t* s, uint64_t* planes);
int lo_contains_target(const uint64_t* s, const uint64_t* target);
unsigned lo_stable_pass(uint64_t* planes, int pass);
void lo_stable_vulnerable(const uint64_t* planes, uint64_t* out);
void lo_weld_step_n(uint64_t* rec, unsigned gens);
void lo_interaction_counts(const uint64_t* s, uint64_t* out);
uint64_t lo_xxh3_512(const uint64_t* s);
void lo_mirrored(const uint64_t* s, uint64_t* out);
void lo_convolve(const uint64_t* a, const uint64_t* b, uint64_t* out);
void lo_match_state(const uint64_t* s, const uint64_t* live, uint64_t* out);
void lo_match_target(const uint64_t* s, const uint64_t* target, uint64_t* out);
void lo_match_live_and_dead(const uint64_t* s, const uint64_t* live, const uint64_t* dead, uint64_t* out);
void lo_interaction_offsets(const uint64_t* a, const uint64_t* b, uint64_t* out);
size_t lo_rle(const uint64_t* state, char* buf, size_t cap);
size_t lo_history_rle(const uint64_t* rec, char* buf, size_t cap);
}

static int failures = 0;
#define EXPECT_TRUE(c)                                                      \
  do {                                                                      \
    if (!(c)) { std::printf("FAIL %s:%d: %s\n", __FILE__, __LINE__, #c); failures++; } \
  } while (0)
#define EXPECT_EQ(a, b) EXPECT_TRUE((a) == (b))

// Seeded stand-in for LifeState::RandomState (LifeAPI.hpp:63-69): rows uniform in [2^61, 2^62].
static LifeState RandomState(std::mt19937_64& rng) {
  std::uniform_int_distribution<uint64_t> dist(1ULL << 61, 1ULL << 62);
  LifeState s;
  for (int i = 0; i < N; i++) s[i] = dist(rng);
  return s;
}

static void StepAltTest_Random() {
  std::mt19937_64 rng(20240611);
  const unsigned numIters = 10000;
  for (unsigned i = 0; i < numIters; i++) {
    LifeState a = RandomState(rng);
    LifeState b = a;
    a.Step();
    b.StepAlt();
    EXPECT_EQ(a, b);
    if (i < 2000) {
      LifeState c = RandomState(rng);
      LifeState want = c;
      lo_step(want.state);
      LifeState want_alt = c;
      lo_step_alt(want_alt.state);
      c.Step();
      EXPECT_EQ(c, want);
      EXPECT_EQ(c, want_alt);
    }
  }
}

static void StepN_RPentomino() {
  LifeState rp = LifeState::Parse("b2o$2o$bo!");
  EXPECT_EQ(rp.GetPop(), 5u);
  LifeState gpu = rp, cpu = rp;
  for (unsigned g = 1; g <= 1000; g++) {
    gpu.Step();
    lo_step(cpu.state);
    EXPECT_EQ(gpu, cpu);
  }
  EXPECT_EQ(gpu.GetPop(), 113u);                  // SURVEY.md 8c known answer
  EXPECT_EQ(rp.Stepped(1000), gpu);               // Step(n) == n x Step()
  EXPECT_EQ(rp.Stepped(100).GetPop(), 121u);
  EXPECT_EQ(rp.Stepped(0), rp);
}

static void Batch_MatchesSingles() {
  std::mt19937_64 rng(7);
  std::vector<LifeState> batch(1000);
  for (auto& s : batch) s = RandomState(rng);
  std::vector<LifeState> want = batch;
  for (auto& s : want) lo_step_n(s.state, 64);
  std::vector<LifeState> host = batch;
  LifeBatch::Step(host, 64);
  EXPECT_TRUE(host == want);
  {
    std::vector<LifeState> pinned = batch;  // the same call on a page-locked vector (asynchronous, overlapped copies)
    LifePinned lock(pinned);
    LifeBatch::Step(pinned, 32);
    LifeBatch::Step(pinned, 32);
    EXPECT_TRUE(pinned == want);
  }
  LifeBatch resident;
  resident.Upload(batch);
  resident.Step(32);
  resident.Step(32);
  EXPECT_TRUE(resident.Download() == want);
}

static void NeighbourCount_And_Target() {
  std::mt19937_64 rng(99);
  for (int it = 0; it < 50; it++) {
    LifeState s = RandomState(rng);
    NeighbourCount nc(s);
    uint64_t planes[4 * N];
    lo_neighbour_count(s.state, planes);
    EXPECT_TRUE(std::memcmp(&nc, planes, sizeof planes) == 0);
    // Subtract(o) = Add(~o, ones): nc - nc == 0 on every plane (NeighbourCount.hpp:85-91)
    NeighbourCount z = nc - nc;
    EXPECT_TRUE(z.bit0.IsEmpty() && z.bit1.IsEmpty() && z.bit2.IsEmpty() && z.bit3.IsEmpty());
    LifeState all;
    for (unsigned n = 0; n <= 9; n++) all |= nc.WithExactly(n);
    EXPECT_EQ(all, ~LifeState());
    LifeTarget t(s & s.Moved(1, 0));
    EXPECT_TRUE((s & s.Moved(1, 0)).Contains(t));
    EXPECT_EQ(s.Contains(t), (bool)lo_contains_target(s.state, reinterpret_cast<const uint64_t*>(&t)));
  }
}

// LifeStable members (LifeStable.hpp:414-729) and the LifeHistory accumulate step through the header.
static void Stable_And_History() {
  LifeStable st;
  // a block and a beehive with a few cells unknown (state &= ~unknown), all options open
  st.state = LifeState::Parse("2o$2o!", 10, 10) | LifeState::Parse("b2o$o2bo$b2o!", 30, 20);
  st.unknown = LifeState::Parse("3o$3o!", 9, 9) | LifeState::Parse("2o!", 31, 20);
  st.state &= ~st.unknown;
  for (int pass : {LIFEB200_STABLE_SYNCHRONISE, LIFEB200_STABLE_UPDATE_OPTIONS, LIFEB200_STABLE_SIGNAL_NEIGHBOURS,
                   LIFEB200_STABLE_PROPAGATE_SIMPLE_STEP, LIFEB200_STABLE_PROPAGATE}) {
    LifeStable want = st;
    unsigned wf = lo_stable_pass(reinterpret_cast<uint64_t*>(&want), pass);
    LifeStable got = st;
    LifeStable::PropagateResult r{};
    LifeStable::PassBatch(&got, 1, pass, &r);
    EXPECT_EQ(got, want);
    EXPECT_EQ(r.consistent, (wf & 1u) != 0);
    EXPECT_EQ(r.changed, (wf & 2u) != 0);
  }
  LifeStable done = st;
  auto res = done.Propagate();
  EXPECT_TRUE(res.consistent);
  auto again = done.Propagate();
  EXPECT_TRUE(again.consistent && !again.changed);
  LifeState v = done.Vulnerable(), vw;
  lo_stable_vulnerable(reinterpret_cast<const uint64_t*>(&done), vw.state);
  EXPECT_EQ(v, vw);

  LifeHistory h;
  h.state = LifeState::Parse("b2o$2o$bo!", 20, 20);
  h.history = h.state;
  LifeState cur = h.state, acc = h.state;
  for (int g = 0; g < 40; g++) { lo_step(cur.state); acc |= cur; }
  h.StepAccumulate(40);
  EXPECT_EQ(h.state, cur);
  EXPECT_EQ(h.history, acc);
}

// LifeWeld::Step, GetHash, InteractionCountsAndNext through the header.
static void Weld_Hash_Interaction() {
  LifeState rp = LifeState::Parse("b2o$2o$bo!");
  EXPECT_EQ(rp.Stepped(1000).GetHash(), 0xdc35d1573487dc91ULL);  // SURVEY.md 8c known answer
  std::mt19937_64 rng(5);
  for (int it = 0; it < 20; it++) {
    LifeState s = RandomState(rng);
    EXPECT_EQ(s.GetHash(), lo_xxh3_512(s.state));
    LifeState o1, o2, om, nx;
    s.InteractionCountsAndNext(o1, o2, om, nx);
    uint64_t want[4 * N];
    lo_interaction_counts(s.state, want);
    EXPECT_TRUE(std::memcmp(o1.state, want, 512) == 0 && std::memcmp(o2.state, want + N, 512) == 0);
    EXPECT_TRUE(std::memcmp(om.state, want + 2 * N, 512) == 0 && std::memcmp(nx.state, want + 3 * N, 512) == 0);
    EXPECT_EQ(nx, s.Stepped());
    LifeWeld w(s, RandomState(rng) & RandomState(rng) & ~s, RandomState(rng) & RandomState(rng) & ~s, RandomState(rng) & ~s);
    LifeWeld want_w = w;
    lo_weld_step_n(reinterpret_cast<uint64_t*>(&want_w), 3);
    w.Step(3);
    EXPECT_EQ(w, want_w);
    LifeWeld plain(s);
    plain.Step();
    EXPECT_EQ(plain.state, s.Stepped());  // frozen = 0 is LifeState::Step
  }
  // a block welded in place: ToTarget is matched by the weld's own state
  LifeWeld blk(LifeState::Parse("2o$2o!", 5, 5));
  EXPECT_TRUE(blk.Contains(blk.ToTarget()));
}

// SURVEY.md 8f rank 4 through the header: Mirrored / Convolve / Match / InteractionOffsets / RLE.
// Call sites read like the reference's users of these members (LifeAPI.hpp:747-750 AlignWith,
// tests/InteractionTest.cpp:16-24).
static void Locate_And_RLE() {
  std::mt19937_64 rng(77);
  const LifeState eater = LifeState::Parse("2b2o$bobo$bo$2o!");
  const LifeState glider = LifeState::Parse("bo$2bo$3o!");
  for (int it = 0; it < 12; it++) {
    LifeState soup = RandomState(rng) & RandomState(rng) & RandomState(rng);
    const int dx = (int)(rng() & 63), dy = (int)(rng() & 63);
    const LifeState placed = eater.Moved(dx, dy);
    LifeState s = (soup & ~placed.ZOI()) | placed;
    LifeState want;
    lo_mirrored(s.state, want.state);
    EXPECT_EQ(s.Mirrored(), want);
    lo_convolve(s.state, glider.state, want.state);
    EXPECT_EQ(s.Convolve(glider), want);
    EXPECT_EQ(glider.Convolve(s), want);
    lo_match_state(s.state, eater.state, want.state);
    const LifeState found = s.Match(eater);
    EXPECT_EQ(found, want);
    EXPECT_TRUE(found.Get((unsigned)dx, (unsigned)dy));  // the planted eater is found where it was put
    const LifeTarget target(eater);
    EXPECT_EQ(s.Match(target), found);
    lo_match_target(s.state, target.wanted.state, want.state);
    EXPECT_EQ(s.Match(target), want);
    lo_match_live_and_dead(s.state, eater.state, glider.state, want.state);
    EXPECT_EQ(s.MatchLiveAndDead(eater, glider), want);
    lo_interaction_offsets(eater.state, soup.state, want.state);
    EXPECT_EQ(eater.InteractionOffsets(soup), want);
    char buf[1 << 15];
    EXPECT_EQ(s.RLE(), std::string(buf, lo_rle(s.state, buf, sizeof buf)));
    LifeHistory h;
    h.state = s; h.history = soup.Moved(1, 0) & ~s; h.marked = placed;  // letters A, B, C only: all of them parse back
    EXPECT_EQ(h.RLE(), std::string(buf, lo_history_rle(h.state.state, buf, sizeof buf)));
    const LifeHistory back = LifeHistory::Parse(h.RLE());
    EXPECT_EQ(back.state, s.Moved(32, 32));  // RLE is written from cell (32, 32): a half-torus shift
    EXPECT_EQ(back.history, h.history.Moved(32, 32));
    EXPECT_EQ(back.marked, placed.Moved(32, 32));
  }
  // glider x eater: overlapping offsets are always reported, offsets whose zones of influence are disjoint never
  const LifeState offs = eater.InteractionOffsets(glider);
  int checked = 0;
  for (int x = 0; x < 64; x += 3)
    for (int y = 0; y < 64; y += 5) {
      const LifeState moved = glider.Moved(x, y);
      if (!(eater & moved).IsEmpty()) EXPECT_TRUE(offs.Get((unsigned)x, (unsigned)y));
      if ((eater.ZOI() & moved.ZOI()).IsEmpty()) {
        EXPECT_TRUE(!offs.Get((unsigned)x, (unsigned)y));
        EXPECT_EQ((eater | moved).Stepped(), eater.Stepped() | moved.Stepped());  // and they really evolve independently
      }
      checked++;
    }
  EXPECT_TRUE(checked > 200);
}

// Call sites of the reference that the round-1 header did not compile: RandomState, InitializedTag,
// operator<<, Move(pair), Cell, shifted Contains, CountNeighbours / StepFor, Parse of an RLE with a header line.
static void Reference_CallSites() {
  PRNG::Seed(12345);
  const LifeState a = LifeState::RandomState();
  PRNG::Seed(12345);
  EXPECT_EQ(a, LifeState::RandomState());  // seedable
  uint64_t all_or = 0, all_and = ~0ULL;
  for (int i = 0; i < N; i++) { all_or |= a[i]; all_and &= a[i]; }
  EXPECT_EQ(all_or >> 62, 0ULL);           // LifeAPI.hpp:21-22: draws in [2^61, 2^62] -> bits 63, 62 clear, bit 61 set
  EXPECT_EQ((all_and >> 61) & 1ULL, 1ULL);
  LifeState u(InitializedTag::UNINITIALIZED);
  u = a;
  EXPECT_EQ(u, a);
  LifeState m = a;
  m.Move(std::make_pair(3, -5));
  EXPECT_EQ(m, a.Moved(3, -5));
  EXPECT_EQ(m, a.Moved(std::make_pair(3, 59)));
  EXPECT_EQ(LifeState::Cell({7, 9}).GetPop(), 1u);
  EXPECT_TRUE(LifeState::Cell({7, 9}).Get(7, 9));
  std::ostringstream os;
  os << a;
  EXPECT_EQ(os.str(), a.RLE());
  // StepFor / CountNeighbours (LifeAPI.hpp:889-895,1042-1064) agree with the GPU Step on every cell of a soup
  const LifeState next = a.Stepped();
  int bad = 0;
  for (int x = 0; x < N; x++)
    for (int y = 0; y < N; y++) {
      bad += a.StepFor({x, y}) != next.Get((unsigned)x, (unsigned)y);
      bad += a.CountNeighbours({x, y}) + (a.Get({x, y}) ? 1u : 0u) != a.CountNeighboursWithCenter({x, y});
    }
  EXPECT_EQ(bad, 0);
  // shifted Contains (LifeTarget.hpp:38-42): a target built at the origin is found at the offset its pattern was moved by
  const LifeState eater = LifeState::Parse("2b2o$bobo$bo$2o!");
  const LifeTarget t(eater);
  for (int dx : {0, 1, 31, 63})
    for (int dy : {0, 5, 62}) {
      const LifeState s = eater.Moved(dx, dy);
      EXPECT_TRUE(s.Contains(t, dx, dy));
      EXPECT_TRUE(!s.Contains(t, dx + 1, dy));
      EXPECT_EQ(s.Contains(t, dx, dy), s.Contains(t.Moved({dx, dy})));
      EXPECT_TRUE(s.Contains(eater, dx, dy) && s.AreDisjoint(t.unwanted, dx, dy));
    }
  // Parse follows GenericParse (Parsing.hpp:141-190): the header line is dropped, so RLEWHeader round-trips
  LifeHistory h;
  h.state = eater;
  EXPECT_EQ(LifeHistory::Parse(h.RLEWHeader()).state, eater.Moved(32, 32));
  EXPECT_EQ(LifeState::Parse("x = 4, y = 4, rule = B3/S23\n2b2o$bobo$bo$2o!"), eater);
  EXPECT_EQ(LifeState::Parse("2b2o$bobo$bo$2o!", 3, 4), eater.Moved(3, 4));
  EXPECT_EQ(LifeState::Parse("2A!").GetPop(), 0u);  // only 'o' sets a cell in the two-state alphabet
}

int main() {
  Reference_CallSites();
  Locate_And_RLE();
  Weld_Hash_Interaction();
  Stable_And_History();
  StepAltTest_Random();
  StepN_RPentomino();
  Batch_MatchesSingles();
  NeighbourCount_And_Target();
  if (failures) { std::printf("%d FAILURES\n", failures); return 1; }
  std::printf("step_alt_test: all passed\n");
  return 0;
}
