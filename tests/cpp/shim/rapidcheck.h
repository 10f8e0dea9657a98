// Minimal stand-in for <rapidcheck.h> / <rapidcheck/gtest.h> (RapidCheck is not in this image): property tests of the
// reference's own test sources run as 100 accepted random cases each (RapidCheck's default), RC_PRE discards a case,
// RC_ASSERT fails the property and prints the arguments.  Test infrastructure only (see gtest/gtest.h here).
#pragma once
#include <cstdint>
#include <cstdio>
#include <random>
#include <sstream>
#include <string>
#include <tuple>
#include <type_traits>

#include "gtest/gtest.h"

namespace rc {

struct Discard {};
struct Failure {
  std::string what;
};
inline std::mt19937_64& rng() {
  static std::mt19937_64 g(0x5eed1234abcdULL);
  return g;
}
inline double unit() { return std::uniform_real_distribution<double>(0.0, 1.0)(rng()); }

// Arbitrary values: like RapidCheck, mostly "small" magnitudes with occasional large ones, so that the properties'
// RC_PRE range filters accept a useful share of the cases.
template <typename T>
T arbitrary() {
  if constexpr (std::is_same_v<T, bool>) {
    return unit() < 0.5;
  } else if constexpr (std::is_floating_point_v<T>) {
    const double u = unit();
    double scale = 10.0;
    if (u < 0.15) scale = 1.0;
    else if (u < 0.30) scale = 100.0;
    else if (u < 0.35) scale = 1.0e4;
    else if (u < 0.38) return static_cast<T>(0);
    return static_cast<T>((unit() * 2.0 - 1.0) * scale);
  } else if constexpr (std::is_integral_v<T>) {
    const double u = unit();
    double scale = 100.0;
    if (u < 0.2) scale = 10.0;
    else if (u < 0.4) scale = 1.0e4;
    else if (u < 0.5) scale = 1.0e6;
    const double v = unit() * scale;
    if constexpr (std::is_signed_v<T>) return static_cast<T>(unit() < 0.5 ? -v : v);
    else return static_cast<T>(v);
  } else {
    static_assert(sizeof(T) == 0, "rapidcheck shim: no generator for this argument type");
  }
}

namespace gen {
template <typename T>
struct Range {
  T lo, hi;  // half-open [lo, hi)
  T operator*() const {
    if constexpr (std::is_floating_point_v<T>) {
      return static_cast<T>(lo + (hi - lo) * unit());
    } else {
      return static_cast<T>(lo + static_cast<T>(unit() * static_cast<double>(hi - lo)));
    }
  }
};
template <typename T>
Range<T> inRange(T lo, T hi) {
  return {lo, hi};
}
}  // namespace gen

template <typename... Args>
void runProperty(void (*prop)(Args...)) {
  constexpr int kWanted = 100, kMaxTries = 20000;
  int ok = 0, tries = 0;
  for (; ok < kWanted && tries < kMaxTries; tries++) {
    std::tuple<std::decay_t<Args>...> args{arbitrary<std::decay_t<Args>>()...};
    try {
      std::apply(prop, args);
      ok++;
    } catch (const Discard&) {
    } catch (const Failure& f) {
      std::ostringstream os;
      os << "property failed after " << ok << " passing cases: " << f.what;
      if constexpr (sizeof...(Args) > 0) {
        os << "   arguments:";
        std::apply([&](const auto&... a) { ((os << " " << a), ...); }, args);
      }
      ::testing::AssertHelper{__FILE__, __LINE__, os.str()} = ::testing::Message();
      return;
    }
  }
  if (ok < kWanted) {
    std::ostringstream os;
    os << "gave up: only " << ok << " of " << tries << " generated cases passed the RC_PRE filters";
    ::testing::AssertHelper{__FILE__, __LINE__, os.str()} = ::testing::Message();
  }
}

}  // namespace rc

#define RC_PRE(c)                       \
  do {                                  \
    if (!(c)) throw ::rc::Discard();    \
  } while (0)
#define RC_ASSERT(c)                                                                            \
  do {                                                                                          \
    if (!(c)) throw ::rc::Failure{std::string(#c) + "  (" __FILE__ ":" + std::to_string(__LINE__) + ")"}; \
  } while (0)

#define RC_GTEST_PROP(suite, name, arglist)                                                     \
  static void NB_GT_NAME_(NB_GT_NAME_(suite, name), prop) arglist;                              \
  static ::testing::Registrar NB_GT_NAME_(NB_GT_NAME_(suite, name), reg)(#suite, #name, [] {    \
    ::rc::runProperty(&NB_GT_NAME_(NB_GT_NAME_(suite, name), prop));                            \
  });                                                                                           \
  static void NB_GT_NAME_(NB_GT_NAME_(suite, name), prop) arglist
