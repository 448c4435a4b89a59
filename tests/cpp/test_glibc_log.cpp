// Host-side check of gf_orb_slam_b200/csrc/glibc_log.h against the platform libm log(), bit for bit.
// usage: test_glibc_log [millions of inputs, default 12]
#include <atomic>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>
#include "../../gf_orb_slam_b200/csrc/glibc_log.h"

static inline uint64_t sm64(uint64_t& s) {
    s += 0x9E3779B97F4A7C15ull;
    uint64_t z = s;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

int main(int argc, char** argv) {
    const long millions = argc > 1 ? atol(argv[1]) : 12;
    const int T = 8;
    std::atomic<long> bad{0}, total{0};
    std::vector<std::thread> th;
    for (int t = 0; t < T; t++) th.emplace_back([&, t]() {
        uint64_t s = 0x1234567ull * (t + 1);
        long b = 0, n = 0;
        const long per = millions * 1000000 / T;
        auto check = [&](double x) {
            volatile double xv = x;
            const double a = std::log(xv), c = gforb::glibc_log(x);
            if (memcmp(&a, &c, 8) && !(a != a && c != c)) { if (b < 5) printf("diff x=%a libm=%a ours=%a\n", x, a, c); b++; }
            n++;
        };
        for (long i = 0; i < per; i++) {
            const uint64_t r = sm64(s);
            double x;
            switch (i & 3) {
            case 0: { uint64_t u = r & 0x7FEFFFFFFFFFFFFFull; memcpy(&x, &u, 8); break; }                    // any positive finite (incl. subnormal)
            case 1: x = 0.9 + (double)(r >> 11) * (0.2 / 9007199254740992.0); break;                       // the near-1 polynomial and its edges
            case 2: x = std::exp(((double)(r >> 11) / 9007199254740992.0) * 140.0 - 95.0); break;          // log-uniform 1e-41 .. 1e19 (det products)
            default: { uint64_t u = 0x3FE0000000000000ull + (r & 0x003FFFFFFFFFFFFFull); memcpy(&x, &u, 8); break; }   // [0.5, 2): every table interval
            }
            check(x);
        }
        if (t == 0) {
            const double sp[] = {0.0, -0.0, 1.0, -1.0, 1.0 / 0.0, 4.9406564584124654e-324, 2.2250738585072014e-308, 0.9375, 1.064697265625,
                                 0x1.fffffffffffffp-1, 0x1.0000000000001p0, 1.7976931348623157e308};
            for (double x : sp) check(x);
            double nanv = std::nan(""); check(nanv);
        }
        bad += b; total += n;
    });
    for (auto& x : th) x.join();
    printf("%s checked %ld mismatches %ld\n", bad.load() ? "FAIL" : "OK", total.load(), bad.load());
    return bad.load() ? 1 : 0;
}
