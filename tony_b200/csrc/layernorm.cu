// tony_b200 — LayerNorm forward / backward for sm_100a (memory-bound: one pass over HBM).
//
// One warp per row, 16-byte vector loads (8 bf16 per lane per load), fp32 statistics.
// Backward is two-stage for dgamma/dbeta: per-lane register accumulators over a CTA's
// rows -> smem -> per-CTA partial rows in caller scratch -> a finalize kernel.
// Replaces ATen layer_norm / native_layer_norm_backward in the launched jobs
// (the reference itself has no device code: SURVEY.md §2.4).
#include "common.cuh"
#include "tony_b200_kernels.h"

#include <algorithm>

namespace tony {

constexpr int LN_WARPS = 8;
constexpr int LN_MAX_PARTIALS = 512;

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

// Registers are the occupancy limit here (per-lane dgamma/dbeta accumulators alone are
// 16*NV floats), so x / dy / gamma stay packed as bf16 and are unpacked twice (statistics
// pass, then dx pass) instead of being held as fp32: 2 CTAs x 8 warps per SM for H <= 1024.
template <int NV>
__global__ void __launch_bounds__(LN_WARPS * 32, (NV <= 4) ? 2 : 1)
layernorm_bwd_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
                     const __nv_bfloat16* __restrict__ gamma, const float* __restrict__ mean,
                     const float* __restrict__ rstd, __nv_bfloat16* __restrict__ dx,
                     float* __restrict__ partials, int rows) {
    constexpr int H = NV * 256;
    __shared__ float red[2][H];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint4 gq[NV];
    float dg[NV][8], db[NV][8];
#pragma unroll
    for (int i = 0; i < NV; ++i) {
        gq[i] = __ldg(reinterpret_cast<const uint4*>(gamma + (i * 32 + lane) * 8));
#pragma unroll
        for (int j = 0; j < 8; ++j) dg[i][j] = db[i][j] = 0.f;
    }
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
            unpack8(gq[i], gv);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const float xh = (xv[j] - mu) * rs;
                const float wdy = dv[j] * gv[j];
                c1 = fmaf(xh, wdy, c1);
                c2 += wdy;
                dg[i][j] = fmaf(dv[j], xh, dg[i][j]);
                db[i][j] += dv[j];
            }
        }
        c1 = warp_sum(c1) * (1.0f / H);
        c2 = warp_sum(c2) * (1.0f / H);
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            float xv[8], dv[8], gv[8], o[8];
            unpack8(xq[i], xv);
            unpack8(dq[i], dv);
            unpack8(gq[i], gv);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const float xh = (xv[j] - mu) * rs;
                o[j] = (dv[j] * gv[j] - fmaf(xh, c1, c2)) * rs;
            }
            st_global_v4(dx + off + (i * 32 + lane) * 8, pack8(o));
        }
    }
    // CTA reduction of the per-lane accumulators: warps take turns adding into smem
    for (int w = 0; w < LN_WARPS; ++w) {
        if (warp == w) {
#pragma unroll
            for (int i = 0; i < NV; ++i)
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const int c = (i * 32 + lane) * 8 + j;
                    if (w == 0) {
                        red[0][c] = dg[i][j];
                        red[1][c] = db[i][j];
                    } else {
                        red[0][c] += dg[i][j];
                        red[1][c] += db[i][j];
                    }
                }
        }
        __syncthreads();
    }
    float* out = partials + static_cast<size_t>(blockIdx.x) * 2 * H;
    for (int c = threadIdx.x; c < 2 * H; c += blockDim.x) out[c] = red[c / H][c % H];
}

// dgamma/dbeta[c] (+)= sum_p partials[p][which][c]
__global__ void layernorm_bwd_finalize_kernel(const float* __restrict__ partials, int num_partials,
                                              int H, __nv_bfloat16* __restrict__ dgamma,
                                              __nv_bfloat16* __restrict__ dbeta, int accumulate) {
    __shared__ float red[8][33];
    const int c = blockIdx.x * 32 + threadIdx.x;
    const int which = blockIdx.y;
    float s = 0.f;
    for (int p = threadIdx.y; p < num_partials; p += 8)
        s += partials[(static_cast<size_t>(p) * 2 + which) * H + c];
    red[threadIdx.y][threadIdx.x] = s;
    __syncthreads();
    if (threadIdx.y == 0) {
#pragma unroll
        for (int k = 1; k < 8; ++k) s += red[k][threadIdx.x];
        __nv_bfloat16* out = which == 0 ? dgamma : dbeta;
        if (accumulate) s += __bfloat162float(out[c]);
        out[c] = __float2bfloat16(s);
    }
}

template <int NV>
static int launch_fwd(const void* x, const void* gamma, const void* beta, void* y, float* mean,
                      float* rstd, int rows, float eps, cudaStream_t s) {
    const int grid = std::min((rows + LN_WARPS - 1) / LN_WARPS, device_sm_count() * 8);
    layernorm_fwd_kernel<NV><<<grid, LN_WARPS * 32, 0, s>>>(
        static_cast<const __nv_bfloat16*>(x), static_cast<const __nv_bfloat16*>(gamma),
        static_cast<const __nv_bfloat16*>(beta), static_cast<__nv_bfloat16*>(y), mean, rstd, rows,
        eps);
    count_launch();
    TONY_CUDA_CHECK(cudaGetLastError());
    return TONY_OK;
}

template <int NV>
static int launch_bwd(const void* dy, const void* x, const void* gamma, const float* mean,
                      const float* rstd, void* dx, void* dgamma, void* dbeta, int rows, float* scratch,
                      int accumulate, cudaStream_t s) {
    constexpr int H = NV * 256;
    const int grid = std::min({(rows + LN_WARPS - 1) / LN_WARPS, device_sm_count() * (NV <= 4 ? 2 : 1), LN_MAX_PARTIALS});
    layernorm_bwd_kernel<NV><<<grid, LN_WARPS * 32, 0, s>>>(
        static_cast<const __nv_bfloat16*>(dy), static_cast<const __nv_bfloat16*>(x),
        static_cast<const __nv_bfloat16*>(gamma), mean, rstd, static_cast<__nv_bfloat16*>(dx),
        scratch, rows);
    TONY_CUDA_CHECK(cudaGetLastError());
    layernorm_bwd_finalize_kernel<<<dim3(H / 32, 2), dim3(32, 8), 0, s>>>(
        scratch, grid, H, static_cast<__nv_bfloat16*>(dgamma), static_cast<__nv_bfloat16*>(dbeta),
        accumulate);
    count_launch(2);
    TONY_CUDA_CHECK(cudaGetLastError());
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
    return static_cast<size_t>(LN_MAX_PARTIALS) * 2 * static_cast<size_t>(H) * sizeof(float);
}

extern "C" int tony_layernorm_bwd(const void* dy, const void* x, const void* gamma,
                                  const float* mean, const float* rstd, void* dx, void* dgamma,
                                  void* dbeta, int rows, int H, float* scratch,
                                  size_t scratch_bytes, int accumulate, void* stream) {
    if (!dy || !x || !gamma || !mean || !rstd || !dx || !dgamma || !dbeta || !scratch || rows <= 0)
        return TONY_ERR_INVALID;
    if (scratch_bytes < tony_layernorm_bwd_scratch_bytes(H)) {
        set_last_error("tony_layernorm_bwd", "scratch too small (see tony_layernorm_bwd_scratch_bytes)");
        return TONY_ERR_INVALID;
    }
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    switch (H) {
        case 256: return launch_bwd<1>(dy, x, gamma, mean, rstd, dx, dgamma, dbeta, rows, scratch, accumulate, s);
        case 512: return launch_bwd<2>(dy, x, gamma, mean, rstd, dx, dgamma, dbeta, rows, scratch, accumulate, s);
        case 768: return launch_bwd<3>(dy, x, gamma, mean, rstd, dx, dgamma, dbeta, rows, scratch, accumulate, s);
        case 1024: return launch_bwd<4>(dy, x, gamma, mean, rstd, dx, dgamma, dbeta, rows, scratch, accumulate, s);
        case 1280: return launch_bwd<5>(dy, x, gamma, mean, rstd, dx, dgamma, dbeta, rows, scratch, accumulate, s);
        case 1536: return launch_bwd<6>(dy, x, gamma, mean, rstd, dx, dgamma, dbeta, rows, scratch, accumulate, s);
        case 2048: return launch_bwd<8>(dy, x, gamma, mean, rstd, dx, dgamma, dbeta, rows, scratch, accumulate, s);
        default:
            set_last_error("tony_layernorm_bwd", "H must be one of 256,512,768,1024,1280,1536,2048");
            return TONY_ERR_UNSUPPORTED;
    }
}
