// Host-side checks of tony_b200/csrc/host_math.h, built by tests/test_host_math.py with
// g++ -fsanitize=address,undefined (sanitizer build of the host-side C++).
#include "host_math.h"

#include <cstdio>
#include <cstdlib>
#include <vector>

using namespace tony;

static int failures = 0;
#define CHECK(cond)                                                       \
    do {                                                                  \
        if (!(cond)) {                                                    \
            std::printf("FAILED %s:%d: %s\n", __FILE__, __LINE__, #cond); \
            ++failures;                                                   \
        }                                                                 \
    } while (0)

int main() {
    // FastDiv: every divisor a launch can produce (splits, column blocks, tiles per row / image, pixels per image or
    // row), dividends across the whole non-negative int range, with the boundaries around multiples of d
    std::vector<int> divisors;
    for (int d = 1; d <= 4096; ++d) divisors.push_back(d);
    for (int d : {4097, 12544, 50176, 200704, 802816, 3211264, (1 << 20) + 7, (1 << 30) - 1, 2147483647}) divisors.push_back(d);
    unsigned long long checked = 0;
    uint32_t lcg = 12345u;
    for (int d : divisors) {
        const FastDiv f = make_fastdiv(d);
        std::vector<long long> ns = {0, 1, d - 1, d, static_cast<long long>(d) + 1, 2147483647LL,
                                     2147483647LL - d, 1LL << 30};
        for (long long k : {2LL, 3LL, 7LL, 1000LL, 65536LL}) {
            ns.push_back(k * d - 1);
            ns.push_back(k * d);
        }
        for (int i = 0; i < 64; ++i) {
            lcg = lcg * 1664525u + 1013904223u;
            ns.push_back(lcg & 0x7fffffffu);
        }
        for (long long n : ns) {
            if (n < 0 || n > 2147483647LL) continue;
            CHECK(fdiv(static_cast<int>(n), f) == static_cast<int>(n / d));
            ++checked;
        }
    }
    CHECK(make_fastdiv(0).d == 1u);                      // guarded: a zero extent never divides

    // fold sizing: groups of ~sqrt(slabs); every launch shape the reductions use stays inside 1024 counter words
    for (int slabs = 1; slabs <= 1024; ++slabs) {
        const int gs = fold_group_size(slabs);
        CHECK(gs >= 8 && gs * gs >= slabs && (gs == 8 || (gs - 1) * (gs - 1) < slabs));
        const int groups = (slabs + gs - 1) / gs;
        CHECK(groups <= gs && groups <= 64);
        CHECK(fold_counter_words(1, slabs) == 1 + groups);
        CHECK(fold_group_floats(slabs, 2, 256) == static_cast<size_t>(groups) * 512);
    }
    for (int tiles : {1, 2, 4, 8, 12, 16, 120})          // column tiles x (4 CTAs per SM / tiles) slabs
        CHECK(fold_counter_words(tiles, 592 / tiles > 0 ? 592 / tiles : 1) <= 1024);

    // wave efficiency and the tiling choice on the shapes it was written for
    CHECK(wave_efficiency(148, 148) == 1.0 && wave_efficiency(149, 148) < 0.51);
    CHECK(choose_gemm_tiling(16384, 3072, 148) == 512);  // BERT-large: many tiles, the pair kernel stays
    CHECK(choose_gemm_tiling(16384, 1024, 148) == 512);
    CHECK(choose_gemm_tiling(12544, 512, 148) == 512);   // 98 pair tiles on 74 pairs: 2 waves either way, pair stays
    CHECK(choose_gemm_tiling(9728, 384, 148) == 128);    // 76 pair tiles = 2 waves, 228 tiles of 128 x 128 = 2 short ones
    CHECK(choose_gemm_tiling(2048, 512, 148) == 512);    // a handful of tiles: split-K on the pair kernel instead
    CHECK(choose_gemm_tiling(802816, 256, 148) == 512);
    for (int workers : {2, 74, 132, 148})                // never returns anything but a kernel that exists
        for (int M = 256; M <= 20000; M += 997)
            for (int N = 256; N <= 4096; N += 256) {
                const int t = choose_gemm_tiling(M, N, workers);
                CHECK(t == 512 || t == 256 || t == 128);
            }
    // split-K: the partials' traffic is part of the cost, not only the wave count
    CHECK(choose_splitk(16, 196, 74, 2048LL * 512, 16, 1.0) == 4);      // ResNet 7x7 weight gradient (measured best: 4)
    CHECK(choose_splitk(16, 256, 74, 1024LL * 1024, 16, 1.0) == 4);     // BERT attention-output weight gradient
    CHECK(choose_splitk(48, 256, 74, 3072LL * 1024, 16, 1.0) == 3);     // BERT QKV weight gradient
    CHECK(choose_splitk(1, 12544, 74, 256LL * 64, 74, 1.0) >= 32);      // one tile, K = every pixel: one split per SM pair
    CHECK(choose_splitk(64, 256, 74, 4096LL * 1024, 16, 1.0) == 1);     // already a full wave
    for (int tiles = 1; tiles <= 80; ++tiles)
        for (int nk : {8, 16, 100, 1000, 12544}) {
            const int s1 = choose_splitk(tiles, nk, 74, 1LL << 20, 16, 1.0);
            CHECK(s1 >= 1 && s1 <= 16 && (s1 == 1 || nk / s1 >= 8));
        }
    std::printf("host_math: %llu divisions checked, %d failure(s)\n", checked, failures);
    return failures ? 1 : 0;
}
