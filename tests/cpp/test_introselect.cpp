// Host-side check of gf_orb_slam_b200/csrc/introselect.h against the toolchain's std::nth_element
// (the permutation, not just the nth value), with heavy ties like FAST scores.
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "../../gf_orb_slam_b200/csrc/introselect.h"

static uint64_t s = 88172645463325252ull;
static uint32_t rnd() { s ^= s << 13; s ^= s >> 7; s ^= s << 17; return (uint32_t)(s >> 11); }

int main() {
    long cases = 0;
    for (int trial = 0; trial < 60000; trial++) {
        int n = 1 + rnd() % (trial % 50 == 0 ? 3000 : 300);
        int range = (trial % 3 == 0) ? 4 : (trial % 3 == 1 ? 40 : 250);
        std::vector<uint32_t> a(n);
        int mode = trial % 7;
        for (int i = 0; i < n; i++) {
            uint32_t sc = (mode == 5) ? (uint32_t)(i % range) : (mode == 6 ? (uint32_t)((n - i) % range) : rnd() % range);
            a[i] = (sc << 24) | (uint32_t)i;   // low bits make every element distinguishable
        }
        std::vector<uint32_t> b = a;
        int nth = rnd() % n;
        std::nth_element(a.begin(), a.begin() + nth, a.end(), gforb::ScoreGreater());
        gforb::nth_element_libstdcxx(b.data(), b.data() + nth, b.data() + n, gforb::ScoreGreater());
        if (a != b) { printf("MISMATCH trial %d n %d nth %d\n", trial, n, nth); return 1; }
        cases++;
    }
    // adversarial: force the heap-select fallback with a median-of-3 killer-ish pattern of many sizes
    for (int n = 4; n < 600; n++) {
        std::vector<uint32_t> a(n);
        for (int i = 0; i < n; i++) a[i] = ((uint32_t)((i * 7919) % 3) << 24) | i;
        for (int nth = 0; nth < n; nth += 1 + n / 7) {
            std::vector<uint32_t> x = a, y = a;
            std::nth_element(x.begin(), x.begin() + nth, x.end(), gforb::ScoreGreater());
            gforb::nth_element_libstdcxx(y.data(), y.data() + nth, y.data() + n, gforb::ScoreGreater());
            if (x != y) { printf("MISMATCH adversarial n %d nth %d\n", n, nth); return 1; }
            cases++;
        }
    }
    // the depth-limit fallback (heap select) checked directly against libstdc++'s internal helper
    for (int trial = 0; trial < 20000; trial++) {
        int n = 2 + rnd() % 200, m = 1 + rnd() % n;
        std::vector<uint32_t> a(n);
        for (int i = 0; i < n; i++) a[i] = ((rnd() % 6) << 24) | i;
        std::vector<uint32_t> b = a;
        std::__heap_select(a.begin(), a.begin() + m, a.end(), __gnu_cxx::__ops::__iter_comp_iter(gforb::ScoreGreater()));
        gforb::is_heap_select(b.data(), b.data() + m, b.data() + n, gforb::ScoreGreater());
        if (a != b) { printf("MISMATCH heap_select n %d m %d\n", n, m); return 1; }
        cases++;
    }
    printf("OK %ld cases\n", cases);
    return 0;
}
