// tony_b200 - small integer helpers shared by host launch code and device code.  No CUDA types: the file also
// compiles with a plain C++ compiler, which is how tests/host/test_host_math.cpp runs it under
// -fsanitize=address,undefined on the CPU tier (sanitizer build of the host-side C++).
#pragma once

#include <stddef.h>
#include <stdint.h>

#if defined(__CUDACC__)
#define TONY_HD __host__ __device__ __forceinline__
#else
#define TONY_HD inline
#endif

namespace tony {

// Division by a launch-time constant as multiply + shift: q = (n * mul) >> shr, exact for 0 <= n < 2^31 and
// 1 <= d < 2^31 (mul = floor(2^shr / d) + 1 with shr = 31 + ceil(log2 d); mul < 2^32).  The tile -> (image, row,
// column) index math runs in every warp for every tile; as 32-bit divisions it was a third of the instructions the
// convolution epilogue warps issued.
struct FastDiv {
    uint32_t d, mul, shr;
};
TONY_HD FastDiv make_fastdiv(int d) {
    FastDiv f;
    f.d = d > 0 ? static_cast<uint32_t>(d) : 1u;
    uint32_t s = 0;
    while ((1ull << s) < f.d) ++s;
    f.shr = 31 + s;
    f.mul = static_cast<uint32_t>((1ull << f.shr) / f.d + 1);
    return f;
}
TONY_HD int fdiv(int n, const FastDiv& f) {
    return static_cast<int>((static_cast<unsigned long long>(static_cast<uint32_t>(n)) * f.mul) >> f.shr);
}

// Two-level fold of per-slab partial rows (fold_two_level_256 in common.cuh): slabs are folded in groups of
// ~sqrt(slabs), the group sums by one CTA.
TONY_HD int fold_group_size(int nparts) {
    int gs = 8;
    while (gs * gs < nparts) ++gs;
    return gs;
}
// words of arrival counters a launch with `tiles` column tiles and `slabs` slabs touches (hosts check <= 1024)
TONY_HD int fold_counter_words(int tiles, int slabs) {
    const int gs = fold_group_size(slabs);
    return tiles * (1 + (slabs + gs - 1) / gs);
}
// floats of group partials such a launch needs
TONY_HD size_t fold_group_floats(int slabs, int ns, int cols) {
    const int gs = fold_group_size(slabs);
    return static_cast<size_t>((slabs + gs - 1) / gs) * ns * cols;
}

// Fraction of the machine the last wave of `units` equal work items leaves busy, averaged over all waves.
TONY_HD double wave_efficiency(long long units, int workers) {
    const long long waves = (units + workers - 1) / workers;
    return static_cast<double>(units) / static_cast<double>(waves * workers);
}

// GEMM tiling for an M x N output on `workers` SMs when both the CTA-pair kernel (256 x 256 tiles on 2 SMs) and the
// 1-CTA kernels (128 x 256, 128 x 128) can run it: SM-time per wave x waves.  The pair kernel moves a third less
// operand traffic per SM (factor 1.12, measured on BERT-large), 128 x 128 tiles are a little less efficient (1.08).
// Returns 512 for the pair kernel, else the 1-CTA kernel's BLOCK_N.  Few-tile shapes only: with many tiles the pair
// kernel always wins and the caller does not ask.
TONY_HD int choose_gemm_tiling(int M, int N, int workers) {
    const long long m128 = (M + 127) / 128;
    const long long ta = static_cast<long long>((M + 255) / 256) * ((N + 255) / 256);
    const long long tb = m128 * ((N + 255) / 256);
    const long long tc = m128 * ((N + 127) / 128);
    const int pairs = workers / 2 > 0 ? workers / 2 : 1;
    const double cost_a = static_cast<double>((ta + pairs - 1) / pairs) * 32768.0 / 1.12;
    const double cost_b = static_cast<double>((tb + workers - 1) / workers) * 32768.0;
    const double cost_c = static_cast<double>((tc + workers - 1) / workers) * 16384.0 * 1.08;
    const double best1 = cost_b < cost_c ? cost_b : cost_c;
    if (ta * 2 > pairs && best1 < 0.9 * cost_a) return cost_c < cost_b ? 128 : 256;
    return 512;
}

// Split count for a GEMM whose `tiles` output tiles cannot fill `workers` (SMs, or SM pairs for the pair kernel).
// Cost in units of one k-block of one tile: waves x k-blocks per split, plus what the fp32 partials cost - every
// split writes and the reduce kernel re-reads out_elems x 4 bytes, at ~1.75 MB per k-block time (HBM rate x the
// ~0.27 us a 256 x 256 x 64 block takes), plus ~11 units for launching the reduce kernel.  Choosing for wave
// efficiency alone picked 9 splits where 4 run 20 % faster (profiles/r2_c91_gemm_tiling_probe.txt).
TONY_HD int choose_splitk(long long tiles, int num_k, int workers, long long out_elems, int max_s, double tile_scale) {
    int best_s = 1;
    double best = static_cast<double>((tiles + workers - 1) / workers) * num_k * tile_scale;
    for (int s = 2; s <= max_s; ++s) {
        if (num_k / s < 8) break;                       // keep >= 512 reduction elements per split
        const int kb = (num_k + s - 1) / s;
        const double mma = static_cast<double>((tiles * s + workers - 1) / workers) * kb * tile_scale;
        const double partials = static_cast<double>(s) * static_cast<double>(out_elems) * 8.0 / 1.75e6 + 11.0;
        if (mma + partials < 0.97 * best) {
            best = mma + partials;
            best_s = s;
        }
    }
    return best_s;
}

}  // namespace tony
