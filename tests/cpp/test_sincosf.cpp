// Host-side check of gf_orb_slam_b200/csrc/glibc_sincosf.h against the platform libm sinf/cosf.
// usage: test_sincosf [stride]   (stride 1 = every float in [0, 2*pi]; default 97)
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>
#include <atomic>
#include "../../gf_orb_slam_b200/csrc/glibc_sincosf.h"

int main(int argc, char** argv) {
    uint32_t stride = argc > 1 ? (uint32_t)atoi(argv[1]) : 97;
    float hi = 6.2831855f * 1.001f;
    uint32_t hib; memcpy(&hib, &hi, 4);
    int T = 8;
    std::atomic<long> bad{0}, total{0};
    std::vector<std::thread> th;
    for (int t = 0; t < T; t++) th.emplace_back([&, t]() {
        long b = 0, n = 0;
        for (uint64_t u = (uint64_t)t * stride; u <= hib; u += (uint64_t)T * stride) {
            uint32_t ub = (uint32_t)u; float x; memcpy(&x, &ub, 4);
            float s = sinf(x), c = cosf(x);
            float s2 = gforb::glibc_sinf(x), c2 = gforb::glibc_cosf(x);
            if (memcmp(&s, &s2, 4) || memcmp(&c, &c2, 4)) { if (b < 3) printf("diff x=%a sin %a/%a cos %a/%a\n", x, s, s2, c, c2); b++; }
            n++;
        }
        bad += b; total += n;
    });
    for (auto& x : th) x.join();
    printf("%s checked %ld mismatches %ld\n", bad.load() ? "FAIL" : "OK", total.load(), bad.load());
    return bad.load() ? 1 : 0;
}
