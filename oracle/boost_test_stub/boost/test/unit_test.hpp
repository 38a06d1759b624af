// TEST INFRASTRUCTURE ONLY.  A few-dozen-line stand-in for the part of Boost.Test that the reference's
// src/test/kawpow_tests.cpp uses (Boost is not installed here), so that THAT FILE - unmodified, read where it lies under
// /root/reference - can be compiled against this repository's drop-in headers and run on the GPU implementation
// (oracle/Makefile, target `ref_tests`).  Written from the macros' documented behaviour; nothing of Boost is copied.
#pragma once
#include <cstdio>
#include <functional>
#include <sstream>
#include <string>
#include <vector>

namespace nx_boost_stub
{
struct Case
{
    const char* suite;
    const char* name;
    std::function<void()> run;
};
inline std::vector<Case>& cases()
{
    static std::vector<Case> v;
    return v;
}
inline int& failures()
{
    static int n = 0;
    return n;
}
inline int& checks()
{
    static int n = 0;
    return n;
}
struct Registrar
{
    Registrar(const char* suite, const char* name, std::function<void()> f) { cases().push_back({suite, name, std::move(f)}); }
};
inline void report(bool ok, const char* file, int line, const std::string& what)
{
    ++checks();
    if (ok) return;
    ++failures();
    std::fprintf(stderr, "%s(%d): error: check %s has failed\n", file, line, what.c_str());
}
template <class A, class B>
void check_equal(const A& a, const B& b, const char* sa, const char* sb, const char* file, int line)
{
    const bool ok = (a == b);
    std::ostringstream os;
    if (!ok) os << sa << " == " << sb << " [" << a << " != " << b << "]";
    report(ok, file, line, os.str());
}
}  // namespace nx_boost_stub

#define BOOST_FIXTURE_TEST_SUITE(suite_name, Fixture) \
    namespace suite_name                              \
    {                                                 \
    typedef Fixture nx_suite_fixture;                 \
    static const char* const nx_suite_name = #suite_name;
#define BOOST_AUTO_TEST_SUITE_END() }

#define BOOST_AUTO_TEST_CASE(case_name)                                                                            \
    struct case_name : nx_suite_fixture                                                                            \
    {                                                                                                              \
        void test_method();                                                                                        \
    };                                                                                                             \
    static nx_boost_stub::Registrar nx_reg_##case_name(nx_suite_name, #case_name, [] { case_name().test_method(); }); \
    void case_name::test_method()

#define BOOST_CHECK(expr) nx_boost_stub::report(static_cast<bool>(expr), __FILE__, __LINE__, #expr)
#define BOOST_CHECK_EQUAL(a, b) nx_boost_stub::check_equal((a), (b), #a, #b, __FILE__, __LINE__)

#ifdef NX_BOOST_STUB_MAIN
int main(int argc, char** argv)
{
    const std::string only = argc > 1 ? argv[1] : "";
    int ran = 0;
    for (const auto& c : nx_boost_stub::cases()) {
        if (!only.empty() && only != c.name) continue;
        const int before = nx_boost_stub::failures();
        c.run();
        ++ran;
        std::printf("%s/%s: %s\n", c.suite, c.name, nx_boost_stub::failures() == before ? "ok" : "FAILED");
        std::fflush(stdout);
    }
    std::printf("*** %d test cases, %d checks, %d failures\n", ran, nx_boost_stub::checks(), nx_boost_stub::failures());
    return nx_boost_stub::failures() ? 1 : 0;
}
#endif
