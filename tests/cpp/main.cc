#include <cstdio>
#include <cstring>

#include "testing.h"

int main(int argc, char** argv) {
    const char* filter = argc > 1 ? argv[1] : nullptr;
    int ran = 0, failed_cases = 0;
    for (const auto& c : testing::registry()) {
        if (filter && !std::strstr(c.name, filter)) continue;
        const int before = testing::failures();
        c.fn();
        ++ran;
        if (testing::failures() != before) {
            ++failed_cases;
            std::fprintf(stderr, "FAIL %s\n", c.name);
        }
    }
    std::printf("%d test(s), %d failed\n", ran, failed_cases);
    return failed_cases == 0 ? 0 : 1;
}
