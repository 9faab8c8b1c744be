// mini_test.h -- a few gtest-shaped macros (gtest is not installed here) so that the adapter tests read like the
// reference's own unit tests.
#pragma once
#include <cstdio>
#include <cstdlib>
#include <functional>
#include <string>
#include <vector>

namespace mini_test
{
struct Case
{
  std::string name;
  std::function<void()> fn;
};
inline std::vector<Case>& cases()
{
  static std::vector<Case> c;
  return c;
}
inline int& failures()
{
  static int f = 0;
  return f;
}
struct Registrar
{
  Registrar(const char* n, std::function<void()> f) { cases().push_back({ n, std::move(f) }); }
};
inline int run_all()
{
  for (auto& c : cases())
  {
    const int before = failures();
    std::printf("[ RUN      ] %s\n", c.name.c_str());
    try
    {
      c.fn();
    }
    catch (const std::exception& e)
    {
      std::printf("  unexpected exception: %s\n", e.what());
      failures()++;
    }
    std::printf("[ %s ] %s\n", failures() == before ? "      OK" : " FAILED ", c.name.c_str());
  }
  std::printf("%zu tests, %d failures\n", cases().size(), failures());
  return failures() ? 1 : 0;
}
}  // namespace mini_test

#define TEST(suite, name)                                                                                              \
  static void suite##_##name();                                                                                        \
  static mini_test::Registrar reg_##suite##_##name(#suite "." #name, suite##_##name);                                  \
  static void suite##_##name()

#define EXPECT_TRUE(x)                                                                                                 \
  do                                                                                                                   \
  {                                                                                                                    \
    if (!(x))                                                                                                          \
    {                                                                                                                  \
      std::printf("  %s:%d: EXPECT_TRUE(%s) failed\n", __FILE__, __LINE__, #x);                                        \
      mini_test::failures()++;                                                                                         \
    }                                                                                                                  \
  } while (0)
#define EXPECT_FALSE(x) EXPECT_TRUE(!(x))
#define EXPECT_EQ(a, b)                                                                                                \
  do                                                                                                                   \
  {                                                                                                                    \
    if (!((a) == (b)))                                                                                                 \
    {                                                                                                                  \
      std::printf("  %s:%d: EXPECT_EQ(%s, %s) failed: %lld vs %lld\n", __FILE__, __LINE__, #a, #b, (long long)(a),     \
                  (long long)(b));                                                                                     \
      mini_test::failures()++;                                                                                         \
    }                                                                                                                  \
  } while (0)
#define EXPECT_LE(a, b) EXPECT_TRUE((a) <= (b))
#define EXPECT_GE(a, b) EXPECT_TRUE((a) >= (b))
