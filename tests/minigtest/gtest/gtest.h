// Minimal GoogleTest-compatible header (TEST INFRASTRUCTURE).  GoogleTest is not installed in this
// image and there is no network, so the reference's own test/*.cpp are compiled UNMODIFIED against this
// header: it provides the macros they use (TEST, EXPECT_/ASSERT_ TRUE, FALSE, EQ, NE, LT, LE, GT, GE,
// NEAR, DOUBLE_EQ, THROW, NO_THROW, ANY_THROW, with streamed messages) and a main().
#pragma once
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <exception>
#include <functional>
#include <iostream>
#include <limits>
#include <sstream>
#include <string>
#include <vector>

namespace testing {
struct TestInfo {
    const char* suite;
    const char* name;
    std::function<void()> body;
};
inline std::vector<TestInfo>& registry() {
    static std::vector<TestInfo> r;
    return r;
}
inline int& current_failures() {
    static int f = 0;
    return f;
}
struct Registrar {
    Registrar(const char* s, const char* n, std::function<void()> b) { registry().push_back({s, n, std::move(b)}); }
};
// Collects "<< msg" and reports on destruction; `fatal` is handled by the macro (return).
class Failure {
public:
    Failure(const char* file, int line, const std::string& what) {
        ++current_failures();
        os_ << file << ":" << line << ": Failure\n" << what;
    }
    ~Failure() { std::cerr << os_.str() << std::endl; }
    template <class T> Failure& operator<<(const T& v) {
        os_ << v;
        return *this;
    }
    // lets `return Failure(...) << x;` compile in void functions (ASSERT_*)
    void operator=(const Failure&) const {}

private:
    std::ostringstream os_;
};
struct Voidify {
    void operator&(const Failure&) const {}
};
template <class T> void print_value(std::ostringstream& o, const T& v) {
    if constexpr (requires(std::ostringstream& s, const T& x) { s << x; }) o << v;
    else o << "<" << sizeof(T) << "-byte object>";
}
template <class A, class B> std::string describe(const char* ea, const char* eb, const A& a, const B& b, const char* op) {
    std::ostringstream o;
    o << "Expected: (" << ea << ") " << op << " (" << eb << "), actual: ";
    print_value(o, a);
    o << " vs ";
    print_value(o, b);
    o << "\n";
    return o.str();
}
// 4-ULP comparison, as GoogleTest's EXPECT_DOUBLE_EQ
inline bool almost_equal_ulps(double a, double b) {
    if (std::isnan(a) || std::isnan(b)) return false;
    if (a == b) return true;
    long long ia, ib;
    std::memcpy(&ia, &a, sizeof a);
    std::memcpy(&ib, &b, sizeof b);
    if (ia < 0) ia = std::numeric_limits<long long>::min() - ia;
    if (ib < 0) ib = std::numeric_limits<long long>::min() - ib;
    const long long d = ia > ib ? ia - ib : ib - ia;
    return d <= 4;
}
inline void InitGoogleTest(int*, char**) {}
inline int RunAllTests() {
    int failed = 0;
    std::printf("[==========] Running %zu tests.\n", registry().size());
    for (auto& t : registry()) {
        std::printf("[ RUN      ] %s.%s\n", t.suite, t.name);
        std::fflush(stdout);
        current_failures() = 0;
        try {
            t.body();
        } catch (const std::exception& e) {
            ++current_failures();
            std::cerr << "unexpected exception: " << e.what() << std::endl;
        } catch (...) {
            ++current_failures();
            std::cerr << "unexpected non-std exception" << std::endl;
        }
        if (current_failures()) {
            ++failed;
            std::printf("[  FAILED  ] %s.%s\n", t.suite, t.name);
        } else {
            std::printf("[       OK ] %s.%s\n", t.suite, t.name);
        }
    }
    std::printf("[==========] %zu tests ran.\n[  PASSED  ] %zu tests.\n", registry().size(), registry().size() - failed);
    if (failed) std::printf("[  FAILED  ] %d tests.\n", failed);
    return failed ? 1 : 0;
}
}  // namespace testing

#define RUN_ALL_TESTS() ::testing::RunAllTests()

#define TEST(suite, name)                                                                              \
    static void suite##_##name##_body();                                                               \
    static ::testing::Registrar suite##_##name##_reg(#suite, #name, suite##_##name##_body);            \
    static void suite##_##name##_body()

#define MINIGTEST_NONFATAL(cond, what) \
    if (cond) {                        \
    } else                             \
        ::testing::Voidify() & ::testing::Failure(__FILE__, __LINE__, what)
#define MINIGTEST_FATAL(cond, what) \
    if (cond) {                     \
    } else                          \
        return ::testing::Voidify() & ::testing::Failure(__FILE__, __LINE__, what)

#define MINIGTEST_CMP(kind, a, b, op)                                             \
    MINIGTEST_##kind(((a)op(b)), ::testing::describe(#a, #b, (a), (b), #op))

#define EXPECT_TRUE(c) MINIGTEST_NONFATAL((c), std::string("Value of: " #c "\n  Actual: false\nExpected: true\n"))
#define EXPECT_FALSE(c) MINIGTEST_NONFATAL(!(c), std::string("Value of: " #c "\n  Actual: true\nExpected: false\n"))
#define ASSERT_TRUE(c) MINIGTEST_FATAL((c), std::string("Value of: " #c "\n  Actual: false\nExpected: true\n"))
#define ASSERT_FALSE(c) MINIGTEST_FATAL(!(c), std::string("Value of: " #c "\n  Actual: true\nExpected: false\n"))
#define EXPECT_EQ(a, b) MINIGTEST_CMP(NONFATAL, a, b, ==)
#define EXPECT_NE(a, b) MINIGTEST_CMP(NONFATAL, a, b, !=)
#define EXPECT_LT(a, b) MINIGTEST_CMP(NONFATAL, a, b, <)
#define EXPECT_LE(a, b) MINIGTEST_CMP(NONFATAL, a, b, <=)
#define EXPECT_GT(a, b) MINIGTEST_CMP(NONFATAL, a, b, >)
#define EXPECT_GE(a, b) MINIGTEST_CMP(NONFATAL, a, b, >=)
#define ASSERT_EQ(a, b) MINIGTEST_CMP(FATAL, a, b, ==)
#define ASSERT_NE(a, b) MINIGTEST_CMP(FATAL, a, b, !=)
#define ASSERT_LT(a, b) MINIGTEST_CMP(FATAL, a, b, <)
#define ASSERT_LE(a, b) MINIGTEST_CMP(FATAL, a, b, <=)
#define ASSERT_GT(a, b) MINIGTEST_CMP(FATAL, a, b, >)
#define ASSERT_GE(a, b) MINIGTEST_CMP(FATAL, a, b, >=)
#define EXPECT_NEAR(a, b, tol) \
    MINIGTEST_NONFATAL((std::abs((a) - (b)) <= (tol)), ::testing::describe(#a, #b, (a), (b), "near"))
#define ASSERT_NEAR(a, b, tol) \
    MINIGTEST_FATAL((std::abs((a) - (b)) <= (tol)), ::testing::describe(#a, #b, (a), (b), "near"))
#define EXPECT_DOUBLE_EQ(a, b) \
    MINIGTEST_NONFATAL(::testing::almost_equal_ulps((a), (b)), ::testing::describe(#a, #b, (a), (b), "=="))
#define ASSERT_DOUBLE_EQ(a, b) \
    MINIGTEST_FATAL(::testing::almost_equal_ulps((a), (b)), ::testing::describe(#a, #b, (a), (b), "=="))

#define MINIGTEST_THROW_BODY(stmt, extype, flag) \
    bool flag = false;                           \
    try {                                        \
        stmt;                                    \
    } catch (const extype&) {                    \
        flag = true;                             \
    } catch (...) {                              \
    }
#define EXPECT_THROW(stmt, extype)                                                                     \
    do {                                                                                               \
        MINIGTEST_THROW_BODY(stmt, extype, minigtest_caught)                                           \
        MINIGTEST_NONFATAL(minigtest_caught, std::string("Expected: " #stmt " throws " #extype "\n")); \
    } while (0)
#define EXPECT_ANY_THROW(stmt)                                                                   \
    do {                                                                                         \
        bool minigtest_caught = false;                                                           \
        try {                                                                                    \
            stmt;                                                                                \
        } catch (...) {                                                                          \
            minigtest_caught = true;                                                             \
        }                                                                                        \
        MINIGTEST_NONFATAL(minigtest_caught, std::string("Expected: " #stmt " throws\n"));       \
    } while (0)
#define EXPECT_NO_THROW(stmt)                                                                          \
    do {                                                                                               \
        bool minigtest_threw = false;                                                                  \
        try {                                                                                          \
            stmt;                                                                                      \
        } catch (...) {                                                                                \
            minigtest_threw = true;                                                                    \
        }                                                                                              \
        MINIGTEST_NONFATAL(!minigtest_threw, std::string("Expected: " #stmt " does not throw\n"));     \
    } while (0)
