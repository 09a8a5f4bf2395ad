// multi_survey.cpp -- a C++ caller (no Python, no torch) of the multi-GPU C ABI: surveys `n` on-device soups for
// `gens` generations on every visible GPU (or the first `g` of them), gathers the hit list over NCCL into this
// process and prints count / counters / a checksum of the records, then repeats the call on ONE GPU through the
// plain context API and checks that the two lists are byte-identical (BASELINE config 5 with defaults on 8 GPUs:
//   multi_survey 16777216 4096).
// Build: g++ -std=c++17 -O2 -Iinclude tests/cpp/multi_survey.cpp -o multi_survey lifeapi_b200/_lib/liblifeapi_b200.so
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "LifeAPI_b200.hpp"

static uint64_t fnv(const std::vector<LifeHit>& h, size_t n) {
  uint64_t x = 0xCBF29CE484222325ull;
  const unsigned char* p = reinterpret_cast<const unsigned char*>(h.data());
  for (size_t i = 0; i < n * sizeof(LifeHit); i++) { x ^= p[i]; x *= 0x100000001B3ull; }
  return x;
}

int main(int argc, char** argv) {
  const uint64_t n = argc > 1 ? std::strtoull(argv[1], nullptr, 0) : 1u << 16;
  const unsigned gens = argc > 2 ? (unsigned)std::atoi(argv[2]) : 256;
  const int want_gpus = argc > 3 ? std::atoi(argv[3]) : 0;   // 0 = all
  const bool check = !(argc > 4 && std::strcmp(argv[4], "nocheck") == 0);
  const uint64_t seed = 0x4C696665415049ull;

  LifeMultiBatch group(want_gpus);  // lifeb200_multi_create over the first want_gpus devices (0: every visible one)
  std::vector<LifeHit> hits(n / 8 + 4096);  // room for one hit in eight ids: a list that does not fit costs a second full run
  uint64_t pop_sum = 0, hash_xor = 0;
  group.Reserve(n, hits.size());  // buffers at their final size: nothing is allocated inside the timed call
  group.SoupSurvey(0, std::min<uint64_t>(n, hits.size()), seed, 0, 2, hits, &pop_sum, &hash_xor);  // warm-up: kernel image
  const auto t0 = std::chrono::steady_clock::now();
  const size_t nh = group.SoupSurvey(0, n, seed, 0, gens, hits, &pop_sum, &hash_xor);
  const double s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  std::printf("gpus %d nccl %d soups %llu gens %u hits %zu pop_sum %llu hash_xor %016llx records_fnv %016llx seconds %.4f gather_ms %.3f Tcell/s %.2f\n",
              group.World(), group.NcclVersion(), (unsigned long long)n, gens, nh, (unsigned long long)pop_sum,
              (unsigned long long)hash_xor, (unsigned long long)fnv(hits, nh), s, group.LastGatherMs(),
              (double)n * 4096.0 * gens / s / 1e12);
  bool sorted = true;
  for (size_t i = 1; i < nh; i++) sorted = sorted && hits[i - 1].id < hits[i].id;
  if (!sorted) { std::printf("FAIL: hit list not in id order\n"); return 1; }
  if (check) {
    lifeb200_ctx* ctx = lifeb200::DefaultContext();
    std::vector<LifeHit> one(n ? n : 1);
    uint64_t nh1 = 0, ps1 = 0, hx1 = 0;
    lifeb200::Check(lifeb200_soup_survey(ctx, 0, n, seed, 0, gens, reinterpret_cast<lifeb200_hit*>(one.data()), one.size(), &nh1,
                                         &ps1, &hx1), "lifeb200_soup_survey");
    const bool same = nh1 == nh && ps1 == pop_sum && hx1 == hash_xor && std::memcmp(one.data(), hits.data(), nh * sizeof(LifeHit)) == 0;
    std::printf("single-GPU list: hits %llu records_fnv %016llx -> %s\n", (unsigned long long)nh1, (unsigned long long)fnv(one, nh1),
                same ? "identical" : "DIFFERENT");
    if (!same) return 1;
  }
  std::printf("multi_survey: all passed\n");
  return 0;
}
