// Tiny assertion harness for the C++ test executables (one process per test file, driven by pytest).
#pragma once
#include <cstdio>
#include <cstdlib>
#include <functional>
#include <string>
#include <vector>

struct TestCase {
    const char *name;
    std::function<void()> fn;
};
inline std::vector<TestCase> &test_registry() {
    static std::vector<TestCase> r;
    return r;
}
struct TestRegistrar {
    TestRegistrar(const char *name, std::function<void()> fn) { test_registry().push_back({name, std::move(fn)}); }
};
#define TEST(name)                                                                                                     \
    static void test_##name();                                                                                         \
    static TestRegistrar reg_##name(#name, test_##name);                                                               \
    static void test_##name()

#define CHECK(cond)                                                                                                    \
    do {                                                                                                               \
        if (!(cond)) {                                                                                                 \
            fprintf(stderr, "CHECK failed: %s at %s:%d\n", #cond, __FILE__, __LINE__);                                 \
            std::exit(1);                                                                                              \
        }                                                                                                              \
    } while (0)
#define CHECK_EQ(a, b)                                                                                                 \
    do {                                                                                                               \
        const auto va = (a);                                                                                           \
        const auto vb = (b);                                                                                           \
        if (!(va == vb)) {                                                                                             \
            fprintf(stderr, "CHECK_EQ failed: %s == %s (%lld vs %lld) at %s:%d\n", #a, #b, (long long) va,             \
                    (long long) vb, __FILE__, __LINE__);                                                               \
            std::exit(1);                                                                                              \
        }                                                                                                              \
    } while (0)

inline int run_all_tests(int argc, char **argv) {
    const std::string only = argc > 1 ? argv[1] : "";
    int ran = 0;
    for (auto &t : test_registry()) {
        if (!only.empty() && only != t.name) continue;
        fprintf(stderr, "[ RUN  ] %s\n", t.name);
        t.fn();
        fprintf(stderr, "[  OK  ] %s\n", t.name);
        ++ran;
    }
    fprintf(stderr, "%d test(s) passed\n", ran);
    return ran > 0 ? 0 : 1;
}
#define TEST_MAIN()                                                                                                    \
    int main(int argc, char **argv) { return run_all_tests(argc, argv); }
