// Minimal unit-test harness (no third-party test framework is available offline).
#pragma once

#include <cstdio>
#include <functional>
#include <sstream>
#include <string>
#include <vector>

namespace testing {

struct Case {
    const char* name;
    std::function<void()> fn;
};

inline std::vector<Case>& registry() {
    static std::vector<Case> r;
    return r;
}

struct Registrar {
    Registrar(const char* name, std::function<void()> fn) { registry().push_back(Case{name, std::move(fn)}); }
};

inline int& failures() {
    static int f = 0;
    return f;
}

template <typename T>
std::string show(const T& v) {
    std::ostringstream os;
    os << v;
    return os.str();
}
inline std::string show(const std::string& v) {
    std::string o = "\"";
    for (char c : v) {
        if (c == '\n') o += "\\n";
        else if (c == '\t') o += "\\t";
        else if (c == '\r') o += "\\r";
        else if (c == '\f') o += "\\f";
        else if (static_cast<unsigned char>(c) < 0x20) { char b[8]; std::snprintf(b, sizeof b, "\\x%02x", c); o += b; }
        else o += c;
    }
    return o + "\"";
}
inline std::string show(bool v) { return v ? "true" : "false"; }

}  // namespace testing

#define TEST(name)                                                        \
    static void test_##name();                                            \
    static ::testing::Registrar registrar_##name(#name, test_##name);     \
    static void test_##name()

#define CHECK(cond)                                                                         \
    do {                                                                                    \
        if (!(cond)) {                                                                      \
            std::fprintf(stderr, "  %s:%d: CHECK failed: %s\n", __FILE__, __LINE__, #cond); \
            ++::testing::failures();                                                        \
        }                                                                                   \
    } while (0)

#define CHECK_EQ(a, b)                                                                                  \
    do {                                                                                                \
        const auto va_ = (a);                                                                           \
        const auto vb_ = (b);                                                                           \
        if (!(va_ == vb_)) {                                                                            \
            std::fprintf(stderr, "  %s:%d: CHECK_EQ failed: %s\n    left:  %s\n    right: %s\n", __FILE__, \
                         __LINE__, #a " == " #b, ::testing::show(va_).c_str(), ::testing::show(vb_).c_str()); \
            ++::testing::failures();                                                                    \
        }                                                                                               \
    } while (0)
