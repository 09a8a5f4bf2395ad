// gtest/gtest.h -- a minimal stand-in for GoogleTest (absent from this image), just enough for the reference's
// own tests/StepAltTest.cpp to compile and run UNCHANGED: TEST(suite, name) registers a function, EXPECT_EQ /
// EXPECT_TRUE count failures, main() runs every registered test and exits non-zero if anything failed.
#pragma once
#include <cstdio>
#include <vector>

namespace gtest_shim {
struct Case { const char* suite; const char* name; void (*fn)(); };
inline std::vector<Case>& cases() { static std::vector<Case> c; return c; }
inline int& failures() { static int f = 0; return f; }
struct Registrar { Registrar(const char* s, const char* n, void (*fn)()) { cases().push_back({s, n, fn}); } };
}  // namespace gtest_shim

#define TEST(suite, name)                                                                   \
  static void suite##_##name##_body();                                                      \
  static gtest_shim::Registrar suite##_##name##_reg(#suite, #name, &suite##_##name##_body); \
  static void suite##_##name##_body()
#define EXPECT_TRUE(c)                                                                                          \
  do {                                                                                                          \
    if (!(c)) { if (gtest_shim::failures()++ < 10) std::printf("FAIL %s:%d: %s\n", __FILE__, __LINE__, #c); }    \
  } while (0)
#define EXPECT_EQ(a, b) EXPECT_TRUE((a) == (b))
#define EXPECT_NE(a, b) EXPECT_TRUE(!((a) == (b)))
#define ASSERT_TRUE(c) do { if (!(c)) { gtest_shim::failures()++; std::printf("FAIL %s:%d: %s\n", __FILE__, __LINE__, #c); return; } } while (0)
#define ASSERT_EQ(a, b) ASSERT_TRUE((a) == (b))

int main() {
  for (const auto& c : gtest_shim::cases()) {
    const int before = gtest_shim::failures();
    c.fn();
    std::printf("[%s] %s.%s\n", gtest_shim::failures() == before ? "  OK  " : "FAILED", c.suite, c.name);
  }
  std::printf("%zu test(s), %d failure(s)\n", gtest_shim::cases().size(), gtest_shim::failures());
  if (gtest_shim::failures() == 0) std::printf("all passed\n");
  return gtest_shim::failures() ? 1 : 0;
}
