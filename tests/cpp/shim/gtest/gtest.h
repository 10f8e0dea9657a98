// Minimal stand-in for <gtest/gtest.h> (GoogleTest is not in this image): just enough of the macro surface for the
// reference's own test sources (tests/*.cpp of LessUp/n-body) to compile unchanged and RUN against this repo's
// drop-in headers and libnbody_b200.so.  Built by oracle/build_ref_tests.sh; test infrastructure only.
#pragma once
#include <cmath>
#include <cstdio>
#include <cstring>
#include <functional>
#include <cstdint>
#include <map>
#include <memory>
#include <sstream>
#include <string>
#include <type_traits>
#include <vector>

namespace testing {

struct TestCase {
  std::string suite, name;
  std::function<void()> fn;
};
inline std::vector<TestCase>& registry() {
  static std::vector<TestCase> r;
  return r;
}
inline bool& current_failed() {
  static bool f = false;
  return f;
}
struct Registrar {
  Registrar(const char* s, const char* n, std::function<void()> f) { registry().push_back({s, n, std::move(f)}); }
};

class Message {
 public:
  template <typename T>
  Message& operator<<(const T& v) {
    os_ << v;
    return *this;
  }
  std::string str() const { return os_.str(); }

 private:
  std::ostringstream os_;
};
struct AssertHelper {  // `AssertHelper(...) = Message() << "text"` reports a failure (GoogleTest's own trick)
  const char* file;
  int line;
  std::string what;
  void operator=(const Message& m) const {
    current_failed() = true;
    std::printf("%s:%d: Failure\n  %s\n", file, line, what.c_str());
    if (!m.str().empty()) std::printf("  %s\n", m.str().c_str());
  }
};

template <typename A, typename B>
std::string describe(const char* op, const char* ea, const char* eb, const A& a, const B& b) {
  std::ostringstream os;
  os << ea << " " << op << " " << eb;
  if constexpr (requires(std::ostream& o, const A& x, const B& y) { o << x; o << y; }) os << "   (" << a << " vs " << b << ")";
  return os.str();
}
inline bool float_almost_equal(float a, float b) {  // 4 ULPs, as EXPECT_FLOAT_EQ
  if (std::isnan(a) || std::isnan(b)) return false;
  if (a == b) return true;
  std::int32_t ia, ib;
  std::memcpy(&ia, &a, 4);
  std::memcpy(&ib, &b, 4);
  if ((ia < 0) != (ib < 0)) return false;
  return std::abs(ia - ib) <= 4;
}

class Test {
 public:
  virtual ~Test() = default;
  virtual void SetUp() {}
  virtual void TearDown() {}
  virtual void TestBody() = 0;
  void Run() {
    SetUp();
    TestBody();
    TearDown();
  }
};
template <typename T>
class TestWithParam : public Test {
 public:
  using ParamType = T;
  static T& slot() {
    static T v{};
    return v;
  }
  const T& GetParam() const { return slot(); }
};
template <typename... A>
std::vector<std::common_type_t<A...>> Values(A... a) {
  return {static_cast<std::common_type_t<A...>>(a)...};
}

// TEST_P bodies and INSTANTIATE_TEST_SUITE_P value lists, expanded when the run starts
struct ParamBody {
  std::string name;
  std::function<void()> run;  // constructs the fixture (parameter already in its slot) and runs it
};
inline std::map<std::string, std::vector<ParamBody>>& param_bodies() {
  static std::map<std::string, std::vector<ParamBody>> m;
  return m;
}
struct ParamInstance {
  std::string prefix, fixture;
  std::function<int()> count;
  std::function<void(int)> select;
};
inline std::vector<ParamInstance>& param_instances() {
  static std::vector<ParamInstance> v;
  return v;
}
struct ParamBodyRegistrar {
  ParamBodyRegistrar(const char* fixture, const char* name, std::function<void()> run) {
    param_bodies()[fixture].push_back({name, std::move(run)});
  }
};
template <typename Fixture, typename Gen>
struct ParamInstanceRegistrar {
  ParamInstanceRegistrar(const char* prefix, const char* fixture, Gen values) {
    auto held = std::make_shared<Gen>(std::move(values));
    param_instances().push_back({prefix, fixture, [held] { return (int)held->size(); },
                                 [held](int i) { Fixture::slot() = static_cast<typename Fixture::ParamType>((*held)[i]); }});
  }
};

inline void InitGoogleTest(int*, char**) {}
inline int RunAllTests() {
  for (const ParamInstance& pi : param_instances())
    for (const ParamBody& b : param_bodies()[pi.fixture])
      for (int i = 0; i < pi.count(); i++) {
        auto sel = pi.select;
        auto run = b.run;
        registry().push_back({pi.prefix + "/" + pi.fixture, b.name + "/" + std::to_string(i), [sel, run, i] {
                                sel(i);
                                run();
                              }});
      }
  int failed = 0, ran = 0;
  std::vector<std::string> bad;
  for (const TestCase& t : registry()) {
    current_failed() = false;
    std::printf("[ RUN      ] %s.%s\n", t.suite.c_str(), t.name.c_str());
    std::fflush(stdout);
    try {
      t.fn();
    } catch (const std::exception& e) {
      current_failed() = true;
      std::printf("  uncaught exception: %s\n", e.what());
    } catch (...) {
      current_failed() = true;
      std::printf("  uncaught exception of unknown type\n");
    }
    ran++;
    if (current_failed()) {
      failed++;
      bad.push_back(t.suite + "." + t.name);
      std::printf("[  FAILED  ] %s.%s\n", t.suite.c_str(), t.name.c_str());
    } else {
      std::printf("[       OK ] %s.%s\n", t.suite.c_str(), t.name.c_str());
    }
  }
  std::printf("[==========] %d tests ran, %d passed, %d failed\n", ran, ran - failed, failed);
  for (const std::string& b : bad) std::printf("[  FAILED  ] %s\n", b.c_str());
  return failed ? 1 : 0;
}

}  // namespace testing

#define RUN_ALL_TESTS() ::testing::RunAllTests()

#define NB_GT_CAT_(a, b) a##_##b
#define NB_GT_NAME_(a, b) NB_GT_CAT_(a, b)

#define TEST(suite, name)                                                                       \
  static void NB_GT_NAME_(suite, name)();                                                      \
  static ::testing::Registrar NB_GT_NAME_(NB_GT_NAME_(suite, name), reg)(#suite, #name, &NB_GT_NAME_(suite, name)); \
  static void NB_GT_NAME_(suite, name)()

#define TEST_F(fixture, name)                                                                   \
  class NB_GT_NAME_(fixture, name) : public fixture {                                          \
   public:                                                                                      \
    void TestBody() override;                                                                   \
  };                                                                                            \
  static ::testing::Registrar NB_GT_NAME_(NB_GT_NAME_(fixture, name), reg)(#fixture, #name, [] { \
    NB_GT_NAME_(fixture, name) t;                                                               \
    t.Run();                                                                                    \
  });                                                                                           \
  void NB_GT_NAME_(fixture, name)::TestBody()

#define TEST_P(fixture, name)                                                                   \
  class NB_GT_NAME_(fixture, name) : public fixture {                                          \
   public:                                                                                      \
    void TestBody() override;                                                                   \
  };                                                                                            \
  static ::testing::ParamBodyRegistrar NB_GT_NAME_(NB_GT_NAME_(fixture, name), reg)(#fixture, #name, [] { \
    NB_GT_NAME_(fixture, name) t;                                                               \
    t.Run();                                                                                    \
  });                                                                                           \
  void NB_GT_NAME_(fixture, name)::TestBody()

#define INSTANTIATE_TEST_SUITE_P(prefix, fixture, gen)                                          \
  static ::testing::ParamInstanceRegistrar<fixture, decltype(gen)> NB_GT_NAME_(NB_GT_NAME_(prefix, fixture), inst)(#prefix, #fixture, gen)

#define NB_GT_CHECK_(cond, text) \
  if (cond)                      \
    ;                            \
  else                           \
    ::testing::AssertHelper{__FILE__, __LINE__, text} = ::testing::Message()
#define NB_GT_REQUIRE_(cond, text) \
  if (cond)                        \
    ;                              \
  else                             \
    return ::testing::AssertHelper{__FILE__, __LINE__, text} = ::testing::Message()

#define NB_GT_CMP_(macro, op, a, b)                                                    \
  macro(([&] { return (a)op(b); })(), ::testing::describe(#op, #a, #b, (a), (b)))

#define EXPECT_TRUE(c) NB_GT_CHECK_(static_cast<bool>(c), "EXPECT_TRUE(" #c ")")
#define EXPECT_FALSE(c) NB_GT_CHECK_(!static_cast<bool>(c), "EXPECT_FALSE(" #c ")")
#define EXPECT_EQ(a, b) NB_GT_CMP_(NB_GT_CHECK_, ==, a, b)
#define EXPECT_NE(a, b) NB_GT_CMP_(NB_GT_CHECK_, !=, a, b)
#define EXPECT_LT(a, b) NB_GT_CMP_(NB_GT_CHECK_, <, a, b)
#define EXPECT_LE(a, b) NB_GT_CMP_(NB_GT_CHECK_, <=, a, b)
#define EXPECT_GT(a, b) NB_GT_CMP_(NB_GT_CHECK_, >, a, b)
#define EXPECT_GE(a, b) NB_GT_CMP_(NB_GT_CHECK_, >=, a, b)
#define EXPECT_NEAR(a, b, tol) \
  NB_GT_CHECK_(std::fabs(static_cast<double>(a) - static_cast<double>(b)) <= static_cast<double>(tol), \
               ::testing::describe("~=", #a, #b, (a), (b)))
#define EXPECT_FLOAT_EQ(a, b) \
  NB_GT_CHECK_(::testing::float_almost_equal(static_cast<float>(a), static_cast<float>(b)), ::testing::describe("==", #a, #b, (a), (b)))
#define ASSERT_TRUE(c) NB_GT_REQUIRE_(static_cast<bool>(c), "ASSERT_TRUE(" #c ")")
#define ASSERT_FALSE(c) NB_GT_REQUIRE_(!static_cast<bool>(c), "ASSERT_FALSE(" #c ")")
#define ASSERT_EQ(a, b) NB_GT_CMP_(NB_GT_REQUIRE_, ==, a, b)
#define ASSERT_NE(a, b) NB_GT_CMP_(NB_GT_REQUIRE_, !=, a, b)
#define ASSERT_NEAR(a, b, tol) \
  NB_GT_REQUIRE_(std::fabs(static_cast<double>(a) - static_cast<double>(b)) <= static_cast<double>(tol), \
                 ::testing::describe("~=", #a, #b, (a), (b)))

#define EXPECT_THROW(stmt, exc)                                                             \
  NB_GT_CHECK_(([&] {                                                                       \
                 try {                                                                      \
                   stmt;                                                                    \
                 } catch (const exc&) {                                                     \
                   return true;                                                             \
                 } catch (...) {                                                            \
                   return false;                                                            \
                 }                                                                          \
                 return false;                                                              \
               })(),                                                                        \
               "EXPECT_THROW(" #stmt ", " #exc ")")
#define EXPECT_NO_THROW(stmt)                                                               \
  NB_GT_CHECK_(([&] {                                                                       \
                 try {                                                                      \
                   stmt;                                                                    \
                 } catch (...) {                                                            \
                   return false;                                                            \
                 }                                                                          \
                 return true;                                                               \
               })(),                                                                        \
               "EXPECT_NO_THROW(" #stmt ")")
