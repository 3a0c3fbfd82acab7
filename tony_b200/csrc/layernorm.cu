// tony_b200 — LayerNorm forward / backward for sm_100a (memory-bound: one pass over HBM).
//
// One warp per row, 16-byte vector loads (8 bf16 per lane per load), fp32 statistics.
// Backward = a dx kernel + a column-reduction kernel for dgamma/dbeta (see below).
// Replaces ATen layer_norm / native_layer_norm_backward in the launched jobs
// (the reference itself has no device code: SURVEY.md §2.4).
#include "common.cuh"
#include "tony_b200_kernels.h"

#include <algorithm>

namespace tony {

constexpr int LN_WARPS = 8;
constexpr int LN_MAX_PARTIALS = 512;
constexpr int LN_MAX_GROUPS = 32;     // second fold level: groups of ~sqrt(slabs) slabs

__device__ __forceinline__ void unpack8(const uint4& u, float (&f)[8]) {
    const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const float2 t = unpack_bf16x2(w[j]);
        f[2 * j] = t.x;
        f[2 * j + 1] = t.y;
    }
}
__device__ __forceinline__ uint4 pack8(const float (&f)[8]) {
    uint4 o;
    o.x = pack_bf16x2(f[0], f[1]);
    o.y = pack_bf16x2(f[2], f[3]);
    o.z = pack_bf16x2(f[4], f[5]);
    o.w = pack_bf16x2(f[6], f[7]);
    return o;
}

template <int NV>
__global__ void __launch_bounds__(LN_WARPS * 32)
layernorm_fwd_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ gamma,
                     const __nv_bfloat16* __restrict__ beta, __nv_bfloat16* __restrict__ y,
                     float* __restrict__ mean_out, float* __restrict__ rstd_out, int rows, float eps) {
    pdl_wait();                // launched with programmatic stream serialization: see launch_pdl_kernel
    pdl_launch_dependents();
    constexpr int H = NV * 256;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int row = blockIdx.x * LN_WARPS + warp; row < rows; row += gridDim.x * LN_WARPS) {
        const __nv_bfloat16* xr = x + static_cast<size_t>(row) * H;
        float v[NV][8];
        float sum = 0.f;
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            const uint4 u = ld_global_nc_v4(xr + (i * 32 + lane) * 8);
            unpack8(u, v[i]);
#pragma unroll
            for (int j = 0; j < 8; ++j) sum += v[i][j];
        }
        const float mean = warp_sum(sum) * (1.0f / H);
        float sq = 0.f;
#pragma unroll
        for (int i = 0; i < NV; ++i)
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const float d = v[i][j] - mean;
                sq += d * d;
            }
        const float rstd = rsqrtf(warp_sum(sq) * (1.0f / H) + eps);
        if (lane == 0) {
            mean_out[row] = mean;
            rstd_out[row] = rstd;
        }
        __nv_bfloat16* yr = y + static_cast<size_t>(row) * H;
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            float g[8], b[8], o[8];
            unpack8(__ldg(reinterpret_cast<const uint4*>(gamma + (i * 32 + lane) * 8)), g);
            unpack8(__ldg(reinterpret_cast<const uint4*>(beta + (i * 32 + lane) * 8)), b);
#pragma unroll
            for (int j = 0; j < 8; ++j) o[j] = (v[i][j] - mean) * rstd * g[j] + b[j];
            st_global_v4(yr + (i * 32 + lane) * 8, pack8(o));
        }
    }
}

// Backward is split in two light kernels instead of one register-heavy one:
//   * layernorm_bwd_dx_kernel: warp per row, no per-column state -> few registers, high
//     occupancy, streams x and dy once from HBM and writes dx;
//   * layernorm_bwd_dgdb_kernel: column reduction of dy*xhat and dy.  Each thread owns 8
//     columns and walks a slab of rows (x / dy are re-read, but 2 x 32 MB has just been
//     streamed and sits in the 126 MB L2); partial rows go to caller scratch and the last
//     CTA of each column tile folds them (counter in the first 4 KB of scratch).
// (A single kernel keeping 16*NV per-lane accumulators needed 192 registers -> one CTA per SM,
// or spilled at 128: 40-45 us for 16384 x 1024 against a 15 us HBM floor.)
template <int NV>
__global__ void __launch_bounds__(LN_WARPS * 32)
layernorm_bwd_dx_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
                        const __nv_bfloat16* __restrict__ gamma, const float* __restrict__ mean,
                        const float* __restrict__ rstd, __nv_bfloat16* __restrict__ dx, int rows,
                        const __nv_bfloat16* __restrict__ add) {
    pdl_wait();                // launched with programmatic stream serialization: see launch_pdl_kernel
    pdl_launch_dependents();
    constexpr int H = NV * 256;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int row = blockIdx.x * LN_WARPS + warp; row < rows; row += gridDim.x * LN_WARPS) {
        const size_t off = static_cast<size_t>(row) * H;
        const float mu = mean[row], rs = rstd[row];
        uint4 xq[NV], dq[NV];
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            xq[i] = ld_global_nc_v4(x + off + (i * 32 + lane) * 8);
            dq[i] = ld_global_nc_v4(dy + off + (i * 32 + lane) * 8);
        }
        float c1 = 0.f, c2 = 0.f;
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            float xv[8], dv[8], gv[8];
            unpack8(xq[i], xv);
            unpack8(dq[i], dv);
            unpack8(__ldg(reinterpret_cast<const uint4*>(gamma + (i * 32 + lane) * 8)), gv);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const float wdy = dv[j] * gv[j];
                c1 = fmaf((xv[j] - mu) * rs, wdy, c1);
                c2 += wdy;
            }
        }
        c1 = warp_sum(c1) * (1.0f / H);
        c2 = warp_sum(c2) * (1.0f / H);
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            float xv[8], dv[8], gv[8], o[8];
            unpack8(xq[i], xv);
            unpack8(dq[i], dv);
            unpack8(__ldg(reinterpret_cast<const uint4*>(gamma + (i * 32 + lane) * 8)), gv);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const float xh = (xv[j] - mu) * rs;
                o[j] = (dv[j] * gv[j] - fmaf(xh, c1, c2)) * rs;
            }
            if (add) {   // gradient of a residual branch that bypassed this LayerNorm (pre-LN blocks)
                float av[8];
                unpack8(ld_global_nc_v4(add + off + (i * 32 + lane) * 8), av);
#pragma unroll
                for (int j = 0; j < 8; ++j) o[j] += av[j];
            }
            st_global_v4(dx + off + (i * 32 + lane) * 8, pack8(o));
        }
    }
}

constexpr int LN_RED_CTAS_PER_SM = 3;   // 80 registers hold the 8 row vectors in flight; the grid is sized to match
constexpr int LN_RED_THREADS = 256;  // 8 warps x 32 lanes x 8 columns = 256 columns per CTA

__global__ void __launch_bounds__(LN_RED_THREADS, LN_RED_CTAS_PER_SM)
layernorm_bwd_dgdb_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
                          const float* __restrict__ mean, const float* __restrict__ rstd, int rows, int H,
                          int rows_per_slab, float* __restrict__ partials, unsigned int* __restrict__ counters,
                          __nv_bfloat16* __restrict__ dgamma, __nv_bfloat16* __restrict__ dbeta,
                          int accumulate) {
    pdl_wait();                // launched with programmatic stream serialization: see launch_pdl_kernel
    pdl_launch_dependents();
    __shared__ float red[2][LN_RED_THREADS / 32][256];
    __shared__ unsigned int is_last;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int col = blockIdx.x * 256 + lane * 8;
    const int r0 = blockIdx.y * rows_per_slab;
    const int r1 = min(rows, r0 + rows_per_slab);
    float ag[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    float ab[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    constexpr int W = LN_RED_THREADS / 32;
    int r = r0 + warp;
    for (; r < r1; r += 4 * W) {  // four rows (eight 16-byte loads) in flight per warp; last batch predicated
        uint4 xq[4], dq[4];
        float mu[4], rs[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int rr = r + k * W;
            if (rr < r1) {
                xq[k] = ld_global_nc_v4(x + static_cast<size_t>(rr) * H + col);
                dq[k] = ld_global_nc_v4(dy + static_cast<size_t>(rr) * H + col);
                mu[k] = mean[rr];
                rs[k] = rstd[rr];
            } else {
                xq[k] = make_uint4(0, 0, 0, 0);
                dq[k] = xq[k];
                mu[k] = 0.f;
                rs[k] = 0.f;
            }
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            float xv[8], dv[8];
            unpack8(xq[k], xv);
            unpack8(dq[k], dv);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                ag[j] = fmaf(dv[j], (xv[j] - mu[k]) * rs[k], ag[j]);
                ab[j] += dv[j];
            }
        }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        red[0][warp][lane * 8 + j] = ag[j];
        red[1][warp][lane * 8 + j] = ab[j];
    }
    __syncthreads();
    const int c = threadIdx.x;
    float sg = 0.f, sb = 0.f;
#pragma unroll
    for (int w = 0; w < W; ++w) {
        sg += red[0][w][c];
        sb += red[1][w][c];
    }
    __syncthreads();
    const int gc = blockIdx.x * 256 + c;
    const float pp[2] = {sg, sb};
    float tt[2];
    if (!fold_two_level_256<2>(pp, partials, partials + static_cast<size_t>(LN_MAX_PARTIALS) * 2 * H, counters, H, red[0],
                               &is_last, tt))
        return;
    float tg = tt[0], tb = tt[1];
    if (accumulate) {
        tg += __bfloat162float(dgamma[gc]);
        tb += __bfloat162float(dbeta[gc]);
    }
    dgamma[gc] = __float2bfloat16(tg);
    dbeta[gc] = __float2bfloat16(tb);
}

// Backward in ONE pass for H <= 1024: the dx kernel's rows, and the column sums for dgamma / dbeta kept in
// shared memory instead of registers.  Every warp owns a private [2][H] fp32 accumulator (8 warps x 8 KB at
// H = 1024); lane l owns the same 8 * NV columns it loads, so the read-modify-write needs no atomics and, with the
// [vector][half][lane][4] layout, no bank conflicts.  Per row that is 2 * H * 4 bytes read and written in shared
// memory - about 128 clk per row per SM, under the ~200 clk the row's HBM traffic takes - instead of a second
// kernel that streams x and dy again (20 us of the 39 us the two-kernel form took at 16384 x 1024).
// The CTA then folds its 8 warps and the CTAs fold through fold_two_level_256, one call per 256-column tile.
template <int NV>
__global__ void __launch_bounds__(LN_WARPS * 32, 2)
layernorm_bwd_fused_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
                           const __nv_bfloat16* __restrict__ gamma, const float* __restrict__ mean,
                           const float* __restrict__ rstd, __nv_bfloat16* __restrict__ dx, int rows,
                           const __nv_bfloat16* __restrict__ add, float* __restrict__ partials,
                           float* __restrict__ gparts, unsigned int* __restrict__ counters,
                           __nv_bfloat16* __restrict__ dgamma, __nv_bfloat16* __restrict__ dbeta, int accumulate) {
    pdl_wait();
    pdl_launch_dependents();
    constexpr int H = NV * 256;
    extern __shared__ __align__(16) float ln_acc[];          // [LN_WARPS][2][NV][2][32][4]
    __shared__ float red[8][256];
    __shared__ unsigned int is_last;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float* ag = ln_acc + static_cast<size_t>(warp) * 2 * H;  // dy * xhat sums of this warp
    float* ab = ag + H;                                      // dy sums
    for (int i = lane; i < 2 * H / 4; i += 32) reinterpret_cast<float4*>(ag)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    __syncwarp();
    float gv[NV][8];
#pragma unroll
    for (int i = 0; i < NV; ++i) unpack8(__ldg(reinterpret_cast<const uint4*>(gamma + (i * 32 + lane) * 8)), gv[i]);
    for (int row = blockIdx.x * LN_WARPS + warp; row < rows; row += gridDim.x * LN_WARPS) {
        const size_t off = static_cast<size_t>(row) * H;
        const float mu = mean[row], rs = rstd[row];
        uint4 xq[NV], dq[NV];
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            xq[i] = ld_global_nc_v4(x + off + (i * 32 + lane) * 8);
            dq[i] = ld_global_nc_v4(dy + off + (i * 32 + lane) * 8);
        }
        float c1 = 0.f, c2 = 0.f;
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            float xv[8], dv[8];
            unpack8(xq[i], xv);
            unpack8(dq[i], dv);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const float wdy = dv[j] * gv[i][j];
                c1 = fmaf((xv[j] - mu) * rs, wdy, c1);
                c2 += wdy;
            }
        }
        c1 = warp_sum(c1) * (1.0f / H);
        c2 = warp_sum(c2) * (1.0f / H);
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            float xv[8], dv[8], o[8], xh[8];
            unpack8(xq[i], xv);
            unpack8(dq[i], dv);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                xh[j] = (xv[j] - mu) * rs;
                o[j] = (dv[j] * gv[i][j] - fmaf(xh[j], c1, c2)) * rs;
            }
            if (add) {   // gradient of a residual branch that bypassed this LayerNorm (pre-LN blocks)
                float av[8];
                unpack8(ld_global_nc_v4(add + off + (i * 32 + lane) * 8), av);
#pragma unroll
                for (int j = 0; j < 8; ++j) o[j] += av[j];
            }
            st_global_v4(dx + off + (i * 32 + lane) * 8, pack8(o));
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                float4* pg = reinterpret_cast<float4*>(ag) + (i * 2 + h) * 32 + lane;
                float4* pb = reinterpret_cast<float4*>(ab) + (i * 2 + h) * 32 + lane;
                float4 a = *pg, b = *pb;
                a.x = fmaf(dv[4 * h], xh[4 * h], a.x);
                a.y = fmaf(dv[4 * h + 1], xh[4 * h + 1], a.y);
                a.z = fmaf(dv[4 * h + 2], xh[4 * h + 2], a.z);
                a.w = fmaf(dv[4 * h + 3], xh[4 * h + 3], a.w);
                b.x += dv[4 * h];
                b.y += dv[4 * h + 1];
                b.z += dv[4 * h + 2];
                b.w += dv[4 * h + 3];
                *pg = a;
                *pb = b;
            }
        }
    }
    __syncthreads();
    // column c = tile * 256 + t lives at vector i = tile, lane = t / 8, half = (t % 8) / 4, element t % 4
    const int t = threadIdx.x;
    const int sub = ((t >> 2) & 1) * 32 * 4 + (t >> 3) * 4 + (t & 3);
    // one fold for all NV column tiles: the CTA's partial row is [2 * NV][256] (thread t: columns tile * 256 + t),
    // so the cross-CTA chain of fences, counters and barriers is walked once, not once per tile
    float pp[2 * NV], tt[2 * NV];
#pragma unroll
    for (int tile = 0; tile < NV; ++tile) {
        float a = 0.f, b = 0.f;
#pragma unroll
        for (int w = 0; w < LN_WARPS; ++w) {
            const float* base = ln_acc + static_cast<size_t>(w) * 2 * H + tile * 2 * 32 * 4 + sub;
            a += base[0];
            b += base[H];
        }
        pp[tile] = a;
        pp[NV + tile] = b;
    }
    if (!fold_two_level_256<2 * NV>(pp, partials, gparts, counters, 256, red, &is_last, tt, 0, 1,
                                    static_cast<int>(blockIdx.x), static_cast<int>(gridDim.x)))
        return;
#pragma unroll
    for (int tile = 0; tile < NV; ++tile) {
        const int gc = tile * 256 + t;
        float tg = tt[tile], tb = tt[NV + tile];
        if (accumulate) {
            tg += __bfloat162float(dgamma[gc]);
            tb += __bfloat162float(dbeta[gc]);
        }
        dgamma[gc] = __float2bfloat16(tg);
        dbeta[gc] = __float2bfloat16(tb);
    }
}

// TONY_B200_LN_BWD=split keeps the two-kernel backward (dx, then the dgamma / dbeta column reduction) for H <= 1024
static bool ln_bwd_fused_enabled() {
    const char* e = getenv("TONY_B200_LN_BWD");
    return !(e && e[0] == 's');
}

template <int NV>
static int launch_fwd(const void* x, const void* gamma, const void* beta, void* y, float* mean,
                      float* rstd, int rows, float eps, cudaStream_t s) {
    const int grid = std::min((rows + LN_WARPS - 1) / LN_WARPS, device_sm_count() * 8);
    TONY_CUDA_CHECK(launch_pdl_kernel(layernorm_fwd_kernel<NV>, dim3(grid), dim3(LN_WARPS * 32), 0, s,
                                      static_cast<const __nv_bfloat16*>(x), static_cast<const __nv_bfloat16*>(gamma),
                                      static_cast<const __nv_bfloat16*>(beta), static_cast<__nv_bfloat16*>(y), mean, rstd,
                                      rows, eps));
    count_launch();
    return TONY_OK;
}

template <int NV>
static int launch_bwd(const void* dy, const void* x, const void* gamma, const float* mean,
                      const float* rstd, void* dx, void* dgamma, void* dbeta, int rows, float* scratch,
                      int accumulate, cudaStream_t s, const void* add = nullptr) {
    constexpr int H = NV * 256;
    const int sms = device_sm_count();
    if constexpr (NV <= 4) if (ln_bwd_fused_enabled()) {
        constexpr size_t SMEM = static_cast<size_t>(LN_WARPS) * 2 * H * sizeof(float);
        static bool configured[64] = {false};
        int dev = 0;
        TONY_CUDA_CHECK(cudaGetDevice(&dev));
        if (dev >= 0 && dev < 64 && !configured[dev]) {
            TONY_CUDA_CHECK(cudaFuncSetAttribute(layernorm_bwd_fused_kernel<NV>,
                                                 cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(SMEM)));
            configured[dev] = true;
        }
        const int per_sm = std::max(1, std::min(2, static_cast<int>((200 * 1024) / (SMEM + 9 * 1024))));
        const int fgrid = std::max(1, std::min({LN_MAX_PARTIALS, sms * per_sm, (rows + LN_WARPS - 1) / LN_WARPS}));
        unsigned int* fcounters = reinterpret_cast<unsigned int*>(scratch);
        if (clear_counters_each_call()) TONY_CUDA_CHECK(cudaMemsetAsync(fcounters, 0, 4096, s));
        float* fpartials = scratch + 1024;
        TONY_CUDA_CHECK(launch_pdl_kernel(
            layernorm_bwd_fused_kernel<NV>, dim3(fgrid), dim3(LN_WARPS * 32), SMEM, s,
            static_cast<const __nv_bfloat16*>(dy), static_cast<const __nv_bfloat16*>(x),
            static_cast<const __nv_bfloat16*>(gamma), mean, rstd, static_cast<__nv_bfloat16*>(dx), rows,
            static_cast<const __nv_bfloat16*>(add), fpartials, fpartials + static_cast<size_t>(LN_MAX_PARTIALS) * 2 * H,
            fcounters, static_cast<__nv_bfloat16*>(dgamma), static_cast<__nv_bfloat16*>(dbeta), accumulate));
        count_launch();
        return TONY_OK;
    }
    const int grid = std::min((rows + LN_WARPS - 1) / LN_WARPS, sms * 8);
    TONY_CUDA_CHECK(launch_pdl_kernel(layernorm_bwd_dx_kernel<NV>, dim3(grid), dim3(LN_WARPS * 32), 0, s,
                                      static_cast<const __nv_bfloat16*>(dy), static_cast<const __nv_bfloat16*>(x),
                                      static_cast<const __nv_bfloat16*>(gamma), mean, rstd,
                                      static_cast<__nv_bfloat16*>(dx), rows, static_cast<const __nv_bfloat16*>(add)));
    // scratch: [4 KB counters][partials: slabs x 2 x H floats][group partials]
    const int col_tiles = H / 256;
    int slabs = std::max(1, std::min({LN_MAX_PARTIALS, (sms * LN_RED_CTAS_PER_SM) / col_tiles, std::max(1, rows / 16)}));
    const int rows_per_slab = (rows + slabs - 1) / slabs;
    slabs = (rows + rows_per_slab - 1) / rows_per_slab;
    unsigned int* counters = reinterpret_cast<unsigned int*>(scratch);
    if (clear_counters_each_call()) TONY_CUDA_CHECK(cudaMemsetAsync(counters, 0, 4096, s));
    TONY_CUDA_CHECK(launch_pdl_kernel(layernorm_bwd_dgdb_kernel, dim3(col_tiles, slabs), dim3(LN_RED_THREADS), 0, s,
                                      static_cast<const __nv_bfloat16*>(dy), static_cast<const __nv_bfloat16*>(x), mean,
                                      rstd, rows, H, rows_per_slab, scratch + 1024, counters,
                                      static_cast<__nv_bfloat16*>(dgamma), static_cast<__nv_bfloat16*>(dbeta), accumulate));
    count_launch(2);
    return TONY_OK;
}

}  // namespace tony

using namespace tony;

extern "C" int tony_layernorm_fwd(const void* x, const void* gamma, const void* beta, void* y,
                                  float* mean, float* rstd, int rows, int H, float eps,
                                  void* stream) {
    if (!x || !gamma || !beta || !y || !mean || !rstd || rows <= 0) return TONY_ERR_INVALID;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    switch (H) {
        case 256: return launch_fwd<1>(x, gamma, beta, y, mean, rstd, rows, eps, s);
        case 512: return launch_fwd<2>(x, gamma, beta, y, mean, rstd, rows, eps, s);
        case 768: return launch_fwd<3>(x, gamma, beta, y, mean, rstd, rows, eps, s);
        case 1024: return launch_fwd<4>(x, gamma, beta, y, mean, rstd, rows, eps, s);
        case 1280: return launch_fwd<5>(x, gamma, beta, y, mean, rstd, rows, eps, s);
        case 1536: return launch_fwd<6>(x, gamma, beta, y, mean, rstd, rows, eps, s);
        case 2048: return launch_fwd<8>(x, gamma, beta, y, mean, rstd, rows, eps, s);
        case 4096: return launch_fwd<16>(x, gamma, beta, y, mean, rstd, rows, eps, s);
        default:
            set_last_error("tony_layernorm_fwd", "H must be one of 256,512,768,1024,1280,1536,2048,4096");
            return TONY_ERR_UNSUPPORTED;
    }
}

extern "C" size_t tony_layernorm_bwd_scratch_bytes(int H) {
    return 4096 + static_cast<size_t>(LN_MAX_PARTIALS + LN_MAX_GROUPS) * 2 * static_cast<size_t>(H) * sizeof(float);
}

static int layernorm_bwd_impl(const void* dy, const void* x, const void* gamma, const float* mean,
                              const float* rstd, void* dx, void* dgamma, void* dbeta, int rows, int H,
                              float* scratch, size_t scratch_bytes, int accumulate, void* stream,
                              const void* add) {
    if (!dy || !x || !gamma || !mean || !rstd || !dx || !dgamma || !dbeta || !scratch || rows <= 0)
        return TONY_ERR_INVALID;
    if (scratch_bytes < tony_layernorm_bwd_scratch_bytes(H)) {
        set_last_error("tony_layernorm_bwd", "scratch too small (see tony_layernorm_bwd_scratch_bytes)");
        return TONY_ERR_INVALID;
    }
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    switch (H) {
        case 256: return launch_bwd<1>(dy, x, gamma, mean, rstd, dx, dgamma, dbeta, rows, scratch, accumulate, s, add);
        case 512: return launch_bwd<2>(dy, x, gamma, mean, rstd, dx, dgamma, dbeta, rows, scratch, accumulate, s, add);
        case 768: return launch_bwd<3>(dy, x, gamma, mean, rstd, dx, dgamma, dbeta, rows, scratch, accumulate, s, add);
        case 1024: return launch_bwd<4>(dy, x, gamma, mean, rstd, dx, dgamma, dbeta, rows, scratch, accumulate, s, add);
        case 1280: return launch_bwd<5>(dy, x, gamma, mean, rstd, dx, dgamma, dbeta, rows, scratch, accumulate, s, add);
        case 1536: return launch_bwd<6>(dy, x, gamma, mean, rstd, dx, dgamma, dbeta, rows, scratch, accumulate, s, add);
        case 2048: return launch_bwd<8>(dy, x, gamma, mean, rstd, dx, dgamma, dbeta, rows, scratch, accumulate, s, add);
        case 4096: return launch_bwd<16>(dy, x, gamma, mean, rstd, dx, dgamma, dbeta, rows, scratch, accumulate, s, add);
        default:
            set_last_error("tony_layernorm_bwd", "H must be one of 256,512,768,1024,1280,1536,2048,4096");
            return TONY_ERR_UNSUPPORTED;
    }
}

extern "C" int tony_layernorm_bwd(const void* dy, const void* x, const void* gamma, const float* mean,
                                  const float* rstd, void* dx, void* dgamma, void* dbeta, int rows, int H,
                                  float* scratch, size_t scratch_bytes, int accumulate, void* stream) {
    return layernorm_bwd_impl(dy, x, gamma, mean, rstd, dx, dgamma, dbeta, rows, H, scratch, scratch_bytes,
                              accumulate, stream, nullptr);
}

extern "C" int tony_layernorm_bwd_add(const void* dy, const void* x, const void* gamma, const float* mean,
                                      const float* rstd, const void* add, void* dx, void* dgamma, void* dbeta,
                                      int rows, int H, float* scratch, size_t scratch_bytes, int accumulate,
                                      void* stream) {
    if (!add) return TONY_ERR_INVALID;
    return layernorm_bwd_impl(dy, x, gamma, mean, rstd, dx, dgamma, dbeta, rows, H, scratch, scratch_bytes,
                              accumulate, stream, add);
}
