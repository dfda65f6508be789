// mini_boost_test.h -- the handful of Boost.Test macros the reference's suites use, so that the facade tests read
// like GV/test/testing_*.cu (Boost is not installed on this image).
#pragma once
#include <cstdio>
#include <iostream>
#include <string>
#include <vector>

namespace mini_test {
struct Case { const char* suite; const char* name; void (*fn)(); };
inline std::vector<Case>& cases() { static std::vector<Case> c; return c; }
inline int& failures() { static int f = 0; return f; }
inline int& checks() { static int c = 0; return c; }
struct Reg { Reg(const char* s, const char* n, void (*f)()) { Case c = {s, n, f}; cases().push_back(c); } };
inline const char*& current_suite() { static const char* s = ""; return s; }
inline int run_all(const char* filter)
{
  for (size_t i = 0; i < cases().size(); ++i)
  {
    const std::string full = std::string(cases()[i].suite) + "/" + cases()[i].name;
    // "name": only cases whose suite/name contains it; "!name": all cases except those
    if (filter && filter[0] != '!' && full.find(filter) == std::string::npos) continue;
    if (filter && filter[0] == '!' && full.find(filter + 1) != std::string::npos) continue;
    const int before = failures();
    cases()[i].fn();
    std::printf("[%s] %s\n", failures() == before ? " OK " : "FAIL", full.c_str());
  }
  std::printf("%d checks, %d failures\n", checks(), failures());
  return failures() ? 1 : 0;
}
}  // namespace mini_test

#define BOOST_AUTO_TEST_SUITE(name) namespace name { static const char* const gvx_suite_name = #name;
#define BOOST_AUTO_TEST_SUITE_END() }
#define BOOST_AUTO_TEST_CASE(name)                                        \
  static void name();                                                      \
  static mini_test::Reg reg_##name(gvx_suite_name, #name, &name);          \
  static void name()
#define BOOST_CHECK_MESSAGE(cond, msg)                                                                            \
  do { ++mini_test::checks();                                                                                     \
       if (!(cond)) { ++mini_test::failures(); std::cerr << __FILE__ << ":" << __LINE__ << ": check failed: " << #cond \
                                                         << " [" << msg << "]" << std::endl; } } while (0)
#define BOOST_CHECK(cond) BOOST_CHECK_MESSAGE(cond, "")
#define BOOST_CHECK_EQUAL(a, b) BOOST_CHECK_MESSAGE((a) == (b), (a) << " != " << (b))
