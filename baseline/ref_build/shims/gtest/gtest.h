// gtest/gtest.h shim — the subset of googletest the reference uses: TEST_F fixtures with
// SetUp/TearDown, EXPECT/ASSERT_{EQ,NE,TRUE,FALSE,DOUBLE_EQ} with streamed messages,
// FRIEND_TEST, InitGoogleTest / RUN_ALL_TESTS (--gtest_filter=Suite.Name prefix matching, --gtest_list_tests).
// driver/simple_id_mapper.cpp includes this header in production code, so it must stay light.
#ifndef MINIPS_SHIM_GTEST_H_
#define MINIPS_SHIM_GTEST_H_
#include <cmath>
#include <cstring>
#include <functional>
#include <iostream>
#include <sstream>
#include <string>
#include <vector>

#define GTEST_API_
#define FRIEND_TEST(suite, name) friend class suite##_##name##_Test

namespace testing {
class Test {
 public:
  virtual ~Test() {}
  virtual void SetUp() {}
  virtual void TearDown() {}
  virtual void TestBody() = 0;
};

struct Registry {
  struct Case {
    std::string name;
    std::function<Test*()> make;
  };
  std::vector<Case> cases;
  std::string filter;
  bool list_only = false;
  int failures_in_current = 0;
  static Registry& get() {
    static Registry r;
    return r;
  }
};

struct Registrar {
  Registrar(const char* suite, const char* name, std::function<Test*()> make) {
    Registry::get().cases.push_back({std::string(suite) + "." + name, std::move(make)});
  }
};

// collects the streamed message of a failed expectation and reports it when destroyed
class Failure {
 public:
  Failure(const char* file, int line, const std::string& what) {
    os_ << file << ":" << line << ": Failure\n" << what << "\n";
  }
  ~Failure() {
    std::cerr << os_.str() << std::flush;
    Registry::get().failures_in_current++;
  }
  template <typename T>
  Failure& operator<<(const T& v) {
    os_ << v;
    return *this;
  }

 private:
  std::ostringstream os_;
};
struct Voidify {
  void operator=(const Failure&) {}
};

template <typename T>
inline auto Show(std::ostream& os, const T& v, int) -> decltype(os << v, void()) {
  os << v;
}
template <typename T>
inline void Show(std::ostream& os, const T&, long) {
  os << "<object>";
}
template <typename A, typename B>
inline std::string Describe(const char* ea, const char* eb, const A& a, const B& b, const char* op) {
  std::ostringstream os;
  os << "Expected: (" << ea << ") " << op << " (" << eb << "), actual: ";
  Show(os, a, 0);
  os << " vs ";
  Show(os, b, 0);
  return os.str();
}
template <typename A, typename B>
inline bool Eq(const A& a, const B& b) {
  return a == b;
}
inline bool DoubleEq(double a, double b) {
  if (a == b) return true;
  const double diff = std::fabs(a - b), scale = std::fmax(std::fabs(a), std::fabs(b));
  return diff <= 4 * 2.220446049250313e-16 * scale;  // ~4 ULPs, like gtest
}

inline void InitGoogleTest(int* argc, char** argv) {
  for (int i = 1; i < *argc; ++i)
    if (!std::strncmp(argv[i], "--gtest_filter=", 15)) Registry::get().filter = argv[i] + 15;
    else if (!std::strcmp(argv[i], "--gtest_list_tests")) Registry::get().list_only = true;
}
inline int RunAllTests() {
  Registry& r = Registry::get();
  int ran = 0, failed = 0;
  std::vector<std::string> failed_names;
  if (r.list_only) {
    for (auto& c : r.cases) std::cout << c.name << "\n";
    return 0;
  }
  for (auto& c : r.cases) {
    std::string f = r.filter;
    if (!f.empty() && f.back() == '*') f.pop_back();
    if (!f.empty() && c.name.compare(0, f.size(), f) != 0) continue;
    std::cerr << "[ RUN      ] " << c.name << std::endl;
    r.failures_in_current = 0;
    Test* t = c.make();
    t->SetUp();
    t->TestBody();
    t->TearDown();
    delete t;
    ++ran;
    if (r.failures_in_current) {
      ++failed;
      failed_names.push_back(c.name);
      std::cerr << "[  FAILED  ] " << c.name << std::endl;
    } else {
      std::cerr << "[       OK ] " << c.name << std::endl;
    }
  }
  std::cerr << "[==========] " << ran << " tests ran.\n[  PASSED  ] " << (ran - failed) << " tests." << std::endl;
  for (auto& n : failed_names) std::cerr << "[  FAILED  ] " << n << std::endl;
  return failed ? 1 : 0;
}
}  // namespace testing

#define RUN_ALL_TESTS() ::testing::RunAllTests()

#define TEST_F(fixture, name)                                                                        \
  class fixture##_##name##_Test : public fixture {                                                   \
   public:                                                                                           \
    void TestBody() override;                                                                        \
  };                                                                                                 \
  static ::testing::Registrar fixture##_##name##_registrar(                                         \
      #fixture, #name, []() -> ::testing::Test* { return new fixture##_##name##_Test; });            \
  void fixture##_##name##_Test::TestBody()

#define TEST(suite, name)                                                                            \
  class suite##_##name##_Test : public ::testing::Test {                                             \
   public:                                                                                           \
    void TestBody() override;                                                                        \
  };                                                                                                 \
  static ::testing::Registrar suite##_##name##_registrar(                                           \
      #suite, #name, []() -> ::testing::Test* { return new suite##_##name##_Test; });                \
  void suite##_##name##_Test::TestBody()

#define MINIPS_GTEST_FAIL_(what) ::testing::Voidify() = ::testing::Failure(__FILE__, __LINE__, what)
#define MINIPS_GTEST_CMP_(a, b, cond, op, on_fail)                                   \
  if (cond)                                                                          \
    ;                                                                                \
  else                                                                               \
    on_fail MINIPS_GTEST_FAIL_(::testing::Describe(#a, #b, (a), (b), op))
#define MINIPS_GTEST_BOOL_(c, want, on_fail) \
  if (static_cast<bool>(c) == want)          \
    ;                                        \
  else                                       \
    on_fail MINIPS_GTEST_FAIL_(std::string("Value of: " #c "\nExpected: ") + (want ? "true" : "false"))

#define EXPECT_EQ(a, b) MINIPS_GTEST_CMP_(a, b, ::testing::Eq((a), (b)), "==", )
#define EXPECT_NE(a, b) MINIPS_GTEST_CMP_(a, b, !::testing::Eq((a), (b)), "!=", )
#define EXPECT_DOUBLE_EQ(a, b) MINIPS_GTEST_CMP_(a, b, ::testing::DoubleEq((a), (b)), "~=", )
#define EXPECT_TRUE(c) MINIPS_GTEST_BOOL_(c, true, )
#define EXPECT_FALSE(c) MINIPS_GTEST_BOOL_(c, false, )
#define ASSERT_EQ(a, b) MINIPS_GTEST_CMP_(a, b, ::testing::Eq((a), (b)), "==", return)
#define ASSERT_NE(a, b) MINIPS_GTEST_CMP_(a, b, !::testing::Eq((a), (b)), "!=", return)
#define ASSERT_TRUE(c) MINIPS_GTEST_BOOL_(c, true, return)
#define ASSERT_FALSE(c) MINIPS_GTEST_BOOL_(c, false, return)
#endif
