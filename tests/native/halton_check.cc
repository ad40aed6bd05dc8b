// Checks the closed-form base-2 Halton stream (yb::Halton2) against the generic incremental restatement of
// mcqmc.h:29-94 (yb::HaltonT<2>), bit for bit, including the counter wrap.  Built and run by tests/test_native_units.py.
#define YB_HOSTEMU 1
#include "../../core_b200/csrc/yb_math.cuh"
#include <cstdio>
#include <cstring>
int main()
{
    const uint32_t starts[] = {0u, 1u, 2u, 5u, 63u, 64u, 1000u, 123456789u, 0x7fffffffu, 0x80000000u, 0xfffffff0u, 0xfffffffeu, 0xffffffffu};
    uint64_t lcg = 88172645463325252ull;
    int bad = 0, n = 0;
    for (int rep = 0; rep < 4000; ++rep) {
        uint32_t st = rep < (int)(sizeof(starts) / sizeof(starts[0])) ? starts[rep] : (uint32_t)((lcg = lcg * 6364136223846793005ull + 1442695040888963407ull) >> 32);
        yb::Halton2 fast; yb::HaltonT<2> slow;
        fast.setStart(st); slow.setStart(st);
        for (int j = 0; j < 40; ++j, ++n) {
            const float a = fast.getNext(), b = slow.getNext();
            if (std::memcmp(&a, &b, 4)) { if (bad < 5) std::printf("start %u step %d: %.9g vs %.9g\n", st, j, a, b); ++bad; }
        }
    }
    std::printf("%d values, %d mismatches\n", n, bad);
    return bad ? 1 : 0;
}
