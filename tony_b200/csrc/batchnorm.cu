// tony_b200 — fused BatchNorm (+ residual add) (+ ReLU) for channels-last bf16 activations.
//
// A channels-last [N, H, W, C] activation is a row-major [M = N*H*W, C] matrix, so BatchNorm's
// per-channel statistics are column reductions and its transform is a per-column affine map:
//   forward : stats kernel (one read of x: sum and sum of squares per channel, two-stage with the
//             last CTA folding the partial rows and producing mean / rstd / a = gamma*rstd /
//             b = beta - mean*a and the running statistics), then ONE elementwise kernel
//             y = relu(x*a + b + residual);
//   backward: reduce kernel (dbeta = sum dy_m, dgamma = sum dy_m*xhat with the ReLU mask applied on
//             the fly from a 1-bit-per-element mask the forward transform wrote), then ONE
//             elementwise kernel dx = a*(dy_m - dbeta/M - xhat*dgamma/M) that also emits the
//             residual branch's gradient dy_m.
// Narrow layers (C < 256) are processed as a [M/k, k*C] matrix (k = 256/C) so that every lane still
// moves 16 bytes per row; the k column groups are added in the fold step.
// Replaces, per BatchNorm of a stock channels-last ResNet: batch_norm_collect_statistics +
// batch_norm_transform_input + relu (+ add) forward and threshold_backward +
// batch_norm_backward_reduce + batch_norm_backward_elemt backward (ATen), which together are
// 79 % of the stock ResNet-50 step on B200 (profiles/r1_c90_*).  The reference (zuston/TonY)
// only launches such jobs; this is the BN part of worklist item K8.
#include "common.cuh"
#include "tony_b200_kernels.h"

#include <algorithm>

namespace tony {

constexpr int BN_THREADS = 256;
// Partial rows: slabs * column tiles <= 4 CTAs per SM, so every SM gets the same number of CTAs (a
// cap of 160 slabs left 12 SMs with twice the work for every layer narrower than 1024 channels).
constexpr int BN_CTAS_PER_SM = 4;
// the backward reduction keeps 8 row vectors in flight per thread and needs 80 registers: three CTAs per SM, and
// a grid of exactly that many (four per SM ran as a full wave plus a one-third wave)
constexpr int BN_BWD_CTAS_PER_SM = 3;
constexpr int BN_MAX_SMS = 256;

static size_t bn_slab_partials_floats() {   // slabs * 2 * Cw  <=  4 * SMs * 2 * 256
    return static_cast<size_t>(BN_CTAS_PER_SM) * BN_MAX_SMS * 2 * 256;
}
// second fold level: at most 64 groups of slabs x 2 x Cw, and groups ~ sqrt(slabs) with slabs * Cw <= 4 * SMs * 256
constexpr size_t BN_GROUP_PARTIALS_FLOATS = 65536;
__device__ __forceinline__ size_t bn_slab_partials_floats_dev() {
    return static_cast<size_t>(BN_CTAS_PER_SM) * BN_MAX_SMS * 2 * 256;
}        // counters[0..63]: one per column tile; then 64 group counters per tile
static size_t bn_partials_floats() { return bn_slab_partials_floats() + BN_GROUP_PARTIALS_FLOATS; }

__device__ __forceinline__ void unpack8f(const uint4& u, float (&v)[8]) {
    const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const float2 f = unpack_bf16x2(w[j]);
        v[2 * j] = f.x;
        v[2 * j + 1] = f.y;
    }
}

__device__ __forceinline__ uint4 pack8f(const float (&v)[8]) {
    uint4 o;
    o.x = pack_bf16x2(v[0], v[1]);
    o.y = pack_bf16x2(v[2], v[3]);
    o.z = pack_bf16x2(v[4], v[5]);
    o.w = pack_bf16x2(v[6], v[7]);
    return o;
}

__device__ __forceinline__ void load8f(const float* __restrict__ p, float (&v)[8]) {   // p 32-byte aligned
    const float4 a = __ldg(reinterpret_cast<const float4*>(p)), b = __ldg(reinterpret_cast<const float4*>(p) + 1);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w;
    v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
}

struct BnDims {
    long long M;      // rows of the real [M, C] matrix
    int C;            // channels
    long long Mv;     // rows of the folded [Mv, Cw] view
    int Cw;           // columns of the folded view (= k * C)
    int k;            // fold factor
    int rows_per_slab;
};

// CTA-level: add the 8 warps' per-lane column accumulators; returns the total for column threadIdx.x.
__device__ __forceinline__ float cta_colsum(const float (&acc)[8], float (*red)[256]) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
    for (int j = 0; j < 8; ++j) red[warp][lane * 8 + j] = acc[j];
    __syncthreads();
    float s = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) s += red[w][threadIdx.x];
    __syncthreads();
    return s;
}

// After the fold: thread t holds the totals of folded column (tile*256 + t).  Returns true for the
// threads that own a real channel and leaves that channel's two totals in s1 / s2.
__device__ __forceinline__ bool combine_groups(const BnDims& d, float& s1, float& s2, float (*red)[256], int& channel) {
    if (d.k == 1) {
        channel = blockIdx.x * 256 + threadIdx.x;
        return channel < d.C;
    }
    red[0][threadIdx.x] = s1;   // single column tile: Cw == 256
    red[1][threadIdx.x] = s2;
    __syncthreads();
    channel = threadIdx.x;
    if (channel >= d.C) return false;
    float a = 0.f, b = 0.f;
    for (int g = 0; g < d.k; ++g) {
        a += red[0][g * d.C + channel];
        b += red[1][g * d.C + channel];
    }
    s1 = a;
    s2 = b;
    return true;
}

__global__ void __launch_bounds__(BN_THREADS, BN_CTAS_PER_SM)
bn_stats_kernel(const __nv_bfloat16* __restrict__ x, const BnDims d, float* __restrict__ partials,
                unsigned int* __restrict__ counters, const __nv_bfloat16* __restrict__ gamma,
                const __nv_bfloat16* __restrict__ beta, float eps, float momentum, float* __restrict__ mean,
                float* __restrict__ rstd, float* __restrict__ ab, float* __restrict__ running_mean,
                float* __restrict__ running_var) {
    pdl_wait();                // launched with programmatic stream serialization: see launch_pdl_kernel
    pdl_launch_dependents();
    __shared__ float red[8][256];
    __shared__ unsigned int is_last;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int col = blockIdx.x * 256 + lane * 8;
    const long long r0 = static_cast<long long>(blockIdx.y) * d.rows_per_slab;
    const long long r1 = min(d.Mv, r0 + d.rows_per_slab);
    float a1[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    float a2[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    if (col < d.Cw) {
        for (long long r = r0 + warp; r < r1; r += 4 * 8) {   // four rows in flight per warp, last batch predicated
            uint4 u[4];
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const long long rr = r + q * 8;
                u[q] = rr < r1 ? ld_global_nc_v4(x + rr * d.Cw + col) : make_uint4(0, 0, 0, 0);
            }
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                float v[8];
                unpack8f(u[q], v);
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    a1[j] += v[j];
                    a2[j] = fmaf(v[j], v[j], a2[j]);
                }
            }
        }
    }
    const float p1 = cta_colsum(a1, red);
    const float p2 = cta_colsum(a2, red);
    const float pp[2] = {p1, p2};
    float ss[2];
    if (!fold_two_level_256<2>(pp, partials, partials + bn_slab_partials_floats_dev(), counters, d.Cw, red, &is_last, ss))
        return;
    float s1 = ss[0], s2 = ss[1];
    int c;
    if (combine_groups(d, s1, s2, red, c)) {
        const float inv_m = 1.0f / static_cast<float>(d.M);
        const float mu = s1 * inv_m;
        const float var = fmaxf(s2 * inv_m - mu * mu, 0.f);
        const float rs = rsqrtf(var + eps);
        const float a = __bfloat162float(gamma[c]) * rs;
        mean[c] = mu;
        rstd[c] = rs;
        ab[c] = a;
        ab[d.C + c] = __bfloat162float(beta[c]) - mu * a;
        if (running_mean) {
            const float unbiased = d.M > 1 ? var * static_cast<float>(d.M) / static_cast<float>(d.M - 1) : var;
            running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * mu;
            running_var[c] = (1.f - momentum) * running_var[c] + momentum * unbiased;
        }
    }
}

// Statistics that arrived as per-channel sums (accumulated by the convolution's epilogue): every thread
// derives the coefficients of its 8 channels itself, the threads of row 0 also publish mean / rstd and
// the running statistics, so the forward BatchNorm is ONE launch and one read of x.
struct BnSums {
    const float* sum;       // nullptr: coefficients come from `ab` (bn_stats_kernel ran)
    const float* sqsum;
    const __nv_bfloat16* gamma;
    const __nv_bfloat16* beta;
    float eps, momentum;
    float* mean;
    float* rstd;
    float* running_mean;
    float* running_var;
};

// y = relu(x*a + b (+ residual)); every thread keeps the 8 channels it owns in registers.
__global__ void __launch_bounds__(BN_THREADS, 4)
bn_apply_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ residual,
                const float* __restrict__ ab, __nv_bfloat16* __restrict__ y, unsigned char* __restrict__ mask,
                long long M, int C, int relu, const BnSums sums) {
    pdl_wait();                // launched with programmatic stream serialization: see launch_pdl_kernel
    pdl_launch_dependents();
    const int groups = C / 8;                                    // divides 256 for every supported C
    const long long tid = static_cast<long long>(blockIdx.x) * BN_THREADS + threadIdx.x;
    const int cg = static_cast<int>(tid % groups);
    const long long row0 = tid / groups;
    const long long row_stride = static_cast<long long>(gridDim.x) * BN_THREADS / groups;
    // per-channel coefficients with 16-byte loads: as 8 scalar loads per array, lanes 32 bytes apart, this prologue
    // was ~20 us of L1 wavefronts on the 7x7 / 14x14 layers whose whole pass moves 25-50 MB
    float a[8], b[8];
    if (sums.sum == nullptr) {
        load8f(ab + cg * 8, a);
        load8f(ab + C + cg * 8, b);
    } else {
        const float inv_m = 1.0f / static_cast<float>(M);
        float g[8], be[8], s1[8], s2[8];
        unpack8f(__ldg(reinterpret_cast<const uint4*>(sums.gamma + cg * 8)), g);
        unpack8f(__ldg(reinterpret_cast<const uint4*>(sums.beta + cg * 8)), be);
        load8f(sums.sum + cg * 8, s1);
        load8f(sums.sqsum + cg * 8, s2);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const int c = cg * 8 + j;
            const float mu = s1[j] * inv_m;
            const float var = fmaxf(s2[j] * inv_m - mu * mu, 0.f);
            const float rs = rsqrtf(var + sums.eps);
            a[j] = g[j] * rs;
            b[j] = be[j] - mu * a[j];
            if (row0 == 0) {
                sums.mean[c] = mu;
                sums.rstd[c] = rs;
                if (sums.running_mean) {
                    const float unbiased = M > 1 ? var * static_cast<float>(M) / static_cast<float>(M - 1) : var;
                    sums.running_mean[c] = (1.f - sums.momentum) * sums.running_mean[c] + sums.momentum * mu;
                    sums.running_var[c] = (1.f - sums.momentum) * sums.running_var[c] + sums.momentum * unbiased;
                }
            }
        }
    }
    for (long long r = row0; r < M; r += 2 * row_stride) {      // two rows in flight
        const long long r2 = r + row_stride;
        const bool two = r2 < M;
        const size_t off0 = static_cast<size_t>(r) * C + cg * 8;
        const size_t off1 = static_cast<size_t>(two ? r2 : r) * C + cg * 8;
        const uint4 x0 = ld_global_nc_v4(x + off0), x1 = ld_global_nc_v4(x + off1);
        uint4 q0 = make_uint4(0, 0, 0, 0), q1 = q0;
        if (residual) {
            q0 = ld_global_nc_v4(residual + off0);
            q1 = ld_global_nc_v4(residual + off1);
        }
#pragma unroll
        for (int q = 0; q < 2; ++q) {
            if (q == 1 && !two) break;
            const size_t off = q ? off1 : off0;
            float v[8];
            unpack8f(q ? x1 : x0, v);
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] = fmaf(v[j], a[j], b[j]);
            if (residual) {
                float rr[8];
                unpack8f(q ? q1 : q0, rr);
#pragma unroll
                for (int j = 0; j < 8; ++j) v[j] += rr[j];
            }
            if (relu) {
                unsigned int bits = 0;
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    bits |= (v[j] > 0.f ? 1u : 0u) << j;
                    v[j] = fmaxf(v[j], 0.f);
                }
                if (mask) mask[off >> 3] = static_cast<unsigned char>(bits);   // one bit per element for backward
            }
            st_global_v4(y + off, pack8f(v));
        }
    }
}

// Folded-view column sums of dy_m and dy_m * xhat.
__global__ void __launch_bounds__(BN_THREADS, BN_BWD_CTAS_PER_SM)
bn_bwd_reduce_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
                     const unsigned char* __restrict__ mask, const BnDims d, const float* __restrict__ mean,
                     const float* __restrict__ rstd, int relu, float* __restrict__ partials, unsigned int* __restrict__ counters,
                     __nv_bfloat16* __restrict__ dgamma, __nv_bfloat16* __restrict__ dbeta, float* __restrict__ coef,
                     const __nv_bfloat16* __restrict__ gamma) {
    pdl_wait();                // launched with programmatic stream serialization: see launch_pdl_kernel
    pdl_launch_dependents();
    __shared__ float red[8][256];
    __shared__ unsigned int is_last;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int col = blockIdx.x * 256 + lane * 8;
    const long long r0 = static_cast<long long>(blockIdx.y) * d.rows_per_slab;
    const long long r1 = min(d.Mv, r0 + d.rows_per_slab);
    float a1[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};   // sum dy_m
    float a2[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};   // sum dy_m * xhat
    if (col < d.Cw) {
        // a2 accumulates sum dy_m * x; the fold turns it into sum dy_m * xhat = rstd * (a2 - mean * a1),
        // so the streaming loop carries no per-channel statistics
        for (long long r = r0 + warp; r < r1; r += 32) {     // four rows in flight per warp
            uint4 g[4], xx[4];
            unsigned int mb[4];
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const long long rr = r + q * 8;
                if (rr < r1) {
                    const size_t o = static_cast<size_t>(rr) * d.Cw + col;
                    g[q] = ld_global_nc_v4(dy + o);
                    xx[q] = ld_global_nc_v4(x + o);
                    mb[q] = relu ? mask[o >> 3] : 0xffu;
                } else {
                    g[q] = make_uint4(0, 0, 0, 0);
                    xx[q] = g[q];
                    mb[q] = 0;
                }
            }
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                float gv[8], xv[8];
                unpack8f(g[q], gv);
                unpack8f(xx[q], xv);
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const float gm = ((mb[q] >> j) & 1u) ? gv[j] : 0.f;
                    a1[j] += gm;
                    a2[j] = fmaf(gm, xv[j], a2[j]);
                }
            }
        }
    }
    const float p1 = cta_colsum(a1, red);
    const float p2 = cta_colsum(a2, red);
    const float pp[2] = {p1, p2};
    float ss[2];
    if (!fold_two_level_256<2>(pp, partials, partials + bn_slab_partials_floats_dev(), counters, d.Cw, red, &is_last, ss))
        return;
    float s1 = ss[0], s2 = ss[1];
    int c;
    if (combine_groups(d, s1, s2, red, c)) {
        const float rs = rstd[c], mu = mean[c];
        s2 = rs * (s2 - mu * s1);
        dbeta[c] = __float2bfloat16(s1);
        dgamma[c] = __float2bfloat16(s2);
        // dx = a*(g - c1 - (x - mu)*rs*c2) = a*g + P*x + Q with c1 = dbeta/M, c2 = dgamma/M: the elementwise
        // pass reads three ready vectors instead of deriving them from five in every thread
        const float inv_m = 1.0f / static_cast<float>(d.M);
        const float c1 = s1 * inv_m, c2 = s2 * inv_m;
        const float a = __bfloat162float(gamma[c]) * rs;
        coef[c] = a;
        coef[d.C + c] = -a * rs * c2;
        coef[2 * d.C + c] = a * (rs * c2 * mu - c1);
    }
}

// dx = a * (dy_m - dbeta/M - xhat * dgamma/M); optionally dres = dy_m.
__global__ void __launch_bounds__(BN_THREADS, 4)
bn_bwd_dx_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
                 const unsigned char* __restrict__ mask, const float* __restrict__ coef, int relu,
                 __nv_bfloat16* __restrict__ dx, __nv_bfloat16* __restrict__ dres, long long M, int C) {
    pdl_wait();                // launched with programmatic stream serialization: see launch_pdl_kernel
    pdl_launch_dependents();
    const int groups = C / 8;
    const long long tid = static_cast<long long>(blockIdx.x) * BN_THREADS + threadIdx.x;
    const int cg = static_cast<int>(tid % groups);
    const long long row0 = tid / groups;
    const long long row_stride = static_cast<long long>(gridDim.x) * BN_THREADS / groups;
    float aa[8], P[8], Q[8];          // written by the reduction's closing CTA
    load8f(coef + cg * 8, aa);
    load8f(coef + C + cg * 8, P);
    load8f(coef + 2 * C + cg * 8, Q);
    for (long long r = row0; r < M; r += 2 * row_stride) {      // two rows (four 16-byte loads) in flight
        const long long r2 = r + row_stride;
        const bool two = r2 < M;
        const size_t off0 = static_cast<size_t>(r) * C + cg * 8;
        const size_t off1 = static_cast<size_t>(two ? r2 : r) * C + cg * 8;
        const uint4 g0 = ld_global_nc_v4(dy + off0), x0 = ld_global_nc_v4(x + off0);
        const uint4 g1 = ld_global_nc_v4(dy + off1), x1 = ld_global_nc_v4(x + off1);
        const unsigned int m0 = relu ? mask[off0 >> 3] : 0xffu;
        const unsigned int m1 = relu ? mask[off1 >> 3] : 0xffu;
#pragma unroll
        for (int q = 0; q < 2; ++q) {
            if (q == 1 && !two) break;
            float gv[8], xv[8], o[8];
            unpack8f(q ? g1 : g0, gv);
            unpack8f(q ? x1 : x0, xv);
            const unsigned int mb = q ? m1 : m0;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const float g = ((mb >> j) & 1u) ? gv[j] : 0.f;
                gv[j] = g;
                o[j] = fmaf(aa[j], g, fmaf(P[j], xv[j], Q[j]));
            }
            const size_t off = q ? off1 : off0;
            st_global_v4(dx + off, pack8f(o));
            if (dres) st_global_v4(dres + off, pack8f(gv));
        }
    }
}

static int make_dims(const char* what, long long M, int C, BnDims* d, int ctas_per_sm = BN_CTAS_PER_SM) {
    if (M <= 0 || C < 8 || C > 2048 || (C & (C - 1))) {
        set_last_error(what, "C must be a power of two in [8, 2048]");
        return TONY_ERR_UNSUPPORTED;
    }
    int k = 1;
    if (C < 256) {
        k = 256 / C;
        while (k > 1 && (M % k)) k /= 2;
        if (k * C != 256) {   // rows do not fold evenly: fall back to the unfolded view
            k = 1;
        }
    }
    d->M = M;
    d->C = C;
    d->k = k;
    d->Mv = M / k;
    d->Cw = C * k;
    const int col_tiles = (d->Cw + 255) / 256;
    long long slabs = std::max(1, (std::min(device_sm_count(), BN_MAX_SMS) * ctas_per_sm) / col_tiles);
    slabs = std::min<long long>(slabs, std::max<long long>(1, d->Mv / 32));
    d->rows_per_slab = static_cast<int>((d->Mv + slabs - 1) / slabs);
    return TONY_OK;
}

static int elementwise_grid(long long M, int C) {
    const int groups = C / 8;
    const long long rows_per_block = std::max(1, BN_THREADS / groups);
    long long blocks = (M + rows_per_block - 1) / rows_per_block;
    if (groups > BN_THREADS) blocks = M * ((groups + BN_THREADS - 1) / BN_THREADS);
    return static_cast<int>(std::min<long long>(blocks, device_sm_count() * 8LL));   // one wave of 8 CTAs per SM
}

}  // namespace tony

using namespace tony;

extern "C" size_t tony_bn_scratch_bytes(int C) {
    const int Cw = C < 256 ? 256 : C;
    // [4 KB counters][partials: slabs x 2 x Cw + group partials][ab: 2 x C][coef: 3 x C]
    (void)Cw;
    return 4096 + (bn_partials_floats() + 8 * static_cast<size_t>(C)) * sizeof(float);
}

extern "C" int tony_bn_fwd(const void* x, const void* residual, const void* gamma, const void* beta, void* y,
                           void* mask, float* mean, float* rstd, float* running_mean, float* running_var, long long M,
                           int C, float eps, float momentum, int relu, float* scratch, size_t scratch_bytes,
                           void* stream) {
    if (!x || !gamma || !beta || !y || !mean || !rstd || !scratch) return TONY_ERR_INVALID;
    BnDims d;
    int rc = make_dims("tony_bn_fwd", M, C, &d);
    if (rc != TONY_OK) return rc;
    if (scratch_bytes < tony_bn_scratch_bytes(C)) {
        set_last_error("tony_bn_fwd", "scratch too small (see tony_bn_scratch_bytes)");
        return TONY_ERR_INVALID;
    }
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    unsigned int* counters = reinterpret_cast<unsigned int*>(scratch);
    float* partials = scratch + 1024;
    float* ab = partials + bn_partials_floats();
    const int col_tiles = (d.Cw + 255) / 256;
    const int slabs = static_cast<int>((d.Mv + d.rows_per_slab - 1) / d.rows_per_slab);
    if (clear_counters_each_call()) TONY_CUDA_CHECK(cudaMemsetAsync(counters, 0, 4096, s));
    TONY_CUDA_CHECK(launch_pdl_kernel(bn_stats_kernel, dim3(dim3(col_tiles, slabs)), dim3(BN_THREADS), 0, s,
        static_cast<const __nv_bfloat16*>(x), d, partials, counters, static_cast<const __nv_bfloat16*>(gamma),
        static_cast<const __nv_bfloat16*>(beta), eps, momentum, mean, rstd, ab, running_mean, running_var));
    TONY_CUDA_CHECK(launch_pdl_kernel(bn_apply_kernel, dim3(elementwise_grid(M, C)), dim3(BN_THREADS), 0, s,
        static_cast<const __nv_bfloat16*>(x), static_cast<const __nv_bfloat16*>(residual), ab,
        static_cast<__nv_bfloat16*>(y), static_cast<unsigned char*>(mask), M, C, relu, BnSums{}));
    count_launch(2);
    return TONY_OK;
}

extern "C" int tony_bn_fwd_from_sums(const void* x, const void* residual, const void* gamma, const void* beta, void* y,
                                     void* mask, float* mean, float* rstd, float* running_mean, float* running_var,
                                     long long M, int C, float eps, float momentum, int relu, const float* stat_sum,
                                     const float* stat_sqsum, void* stream) {
    if (!x || !gamma || !beta || !y || !mean || !rstd || !stat_sum || !stat_sqsum) return TONY_ERR_INVALID;
    BnDims d;
    int rc = make_dims("tony_bn_fwd_from_sums", M, C, &d);
    if (rc != TONY_OK) return rc;
    BnSums sums{stat_sum, stat_sqsum, static_cast<const __nv_bfloat16*>(gamma), static_cast<const __nv_bfloat16*>(beta),
                eps, momentum, mean, rstd, running_mean, running_var};
    TONY_CUDA_CHECK(launch_pdl_kernel(bn_apply_kernel, dim3(elementwise_grid(M, C)), dim3(BN_THREADS), 0, static_cast<cudaStream_t>(stream),
        static_cast<const __nv_bfloat16*>(x), static_cast<const __nv_bfloat16*>(residual), nullptr,
        static_cast<__nv_bfloat16*>(y), static_cast<unsigned char*>(mask), M, C, relu, sums));
    count_launch(1);
    return TONY_OK;
}

extern "C" int tony_bn_bwd(const void* dy, const void* x, const void* mask, const void* gamma,
                           const float* mean, const float* rstd, void* dx, void* dres, void* dgamma, void* dbeta, long long M,
                           int C, int relu, float* scratch, size_t scratch_bytes, void* stream) {
    if (!dy || !x || !gamma || !mean || !rstd || !dx || !dgamma || !dbeta || !scratch || (relu && !mask))
        return TONY_ERR_INVALID;
    BnDims d;
    int rc = make_dims("tony_bn_bwd", M, C, &d, BN_BWD_CTAS_PER_SM);
    if (rc != TONY_OK) return rc;
    if (scratch_bytes < tony_bn_scratch_bytes(C)) {
        set_last_error("tony_bn_bwd", "scratch too small (see tony_bn_scratch_bytes)");
        return TONY_ERR_INVALID;
    }
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    unsigned int* counters = reinterpret_cast<unsigned int*>(scratch);
    float* partials = scratch + 1024;
    // a = gamma * rstd (dx scale) is rebuilt from the saved statistics in the kernel: the scratch may
    // have served other layers since the forward.  The ReLU mask comes from the forward's bit mask.
    float* coef = partials + bn_partials_floats() + 2 * C;
    const int col_tiles = (d.Cw + 255) / 256;
    const int slabs = static_cast<int>((d.Mv + d.rows_per_slab - 1) / d.rows_per_slab);
    if (clear_counters_each_call()) TONY_CUDA_CHECK(cudaMemsetAsync(counters, 0, 4096, s));
    TONY_CUDA_CHECK(launch_pdl_kernel(bn_bwd_reduce_kernel, dim3(dim3(col_tiles, slabs)), dim3(BN_THREADS), 0, s,
        static_cast<const __nv_bfloat16*>(dy), static_cast<const __nv_bfloat16*>(x),
        static_cast<const unsigned char*>(mask), d, mean, rstd, relu, partials, counters,
        static_cast<__nv_bfloat16*>(dgamma), static_cast<__nv_bfloat16*>(dbeta), coef,
        static_cast<const __nv_bfloat16*>(gamma)));
    TONY_CUDA_CHECK(launch_pdl_kernel(bn_bwd_dx_kernel, dim3(elementwise_grid(M, C)), dim3(BN_THREADS), 0, s,
        static_cast<const __nv_bfloat16*>(dy), static_cast<const __nv_bfloat16*>(x),
        static_cast<const unsigned char*>(mask), coef, relu, static_cast<__nv_bfloat16*>(dx),
        static_cast<__nv_bfloat16*>(dres), M, C));
    count_launch(2);
    return TONY_OK;
}

// ---------------------------------------------------------------------------------------
// 3x3 / stride 2 / pad 1 max pooling on channels-last bf16 (the ResNet stem).  Forward also stores
// the position (0..8) of the first maximum per output element as one byte, so backward is a pure
// gather: every INPUT element looks at the <= 4 windows that contain it and adds dy where the
// stored position is its own — no atomics, no int64 index tensor (ATen's pair costs 0.75 + 1.73 ms
// per ResNet-50 step on B200 for 0.5 GB of traffic).
namespace tony {

__global__ void __launch_bounds__(256)
maxpool3x3s2_fwd_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y,
                        unsigned char* __restrict__ idx, int N, int H, int W, int C, int OH, int OW) {
    const int groups = C / 8;
    const long long total = static_cast<long long>(N) * OH * OW * groups;
    for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        const int cg = static_cast<int>(i % groups);
        long long t = i / groups;
        const int ow = static_cast<int>(t % OW);
        t /= OW;
        const int oh = static_cast<int>(t % OH);
        const int n = static_cast<int>(t / OH);
        float best[8];
        int pos[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            best[j] = -INFINITY;
            pos[j] = 0;
        }
#pragma unroll
        for (int r = 0; r < 3; ++r) {
            const int ih = oh * 2 - 1 + r;
            if (ih < 0 || ih >= H) continue;
#pragma unroll
            for (int s = 0; s < 3; ++s) {
                const int iw = ow * 2 - 1 + s;
                if (iw < 0 || iw >= W) continue;
                float v[8];
                unpack8f(ld_global_nc_v4(x + ((static_cast<size_t>(n) * H + ih) * W + iw) * C + cg * 8), v);
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    if (v[j] > best[j]) {        // strict: the FIRST maximum keeps the gradient (ATen's rule)
                        best[j] = v[j];
                        pos[j] = r * 3 + s;
                    }
                }
            }
        }
        const size_t o = ((static_cast<size_t>(n) * OH + oh) * OW + ow) * C + cg * 8;
        st_global_v4(y + o, pack8f(best));
        uint2 packed;
        packed.x = pos[0] | (pos[1] << 8) | (pos[2] << 16) | (pos[3] << 24);
        packed.y = pos[4] | (pos[5] << 8) | (pos[6] << 16) | (pos[7] << 24);
        *reinterpret_cast<uint2*>(idx + o) = packed;
    }
}

__global__ void __launch_bounds__(256)
maxpool3x3s2_bwd_kernel(const __nv_bfloat16* __restrict__ dy, const unsigned char* __restrict__ idx,
                        __nv_bfloat16* __restrict__ dx, int N, int H, int W, int C, int OH, int OW) {
    // One thread per 2 x 2 block of input pixels (rows 2i, 2i+1; columns 2j, 2j+1) and 8 channels.  The block is
    // covered by exactly the windows (i + u, j + v), u, v in {0, 1}: pixel (a, b) of the block sits at row
    // r = a + 1 - 2u, column s = b + 1 - 2v of window (u, v), which is inside the window for u <= a, v <= b.
    // Each window's (dy, position) pair is loaded once per block - 4 loads for 4 pixels, where one thread per pixel
    // looked at 1 / 2 / 2 / 4 windows (9 loads per block, 1.4 GB through L2 for the ResNet stem).
    const int groups = C / 8;
    const int HB = (H + 1) / 2, WB = (W + 1) / 2;
    const long long total = static_cast<long long>(N) * HB * WB * groups;
    for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        const int cg = static_cast<int>(i % groups);
        long long t = i / groups;
        const int jb = static_cast<int>(t % WB);
        t /= WB;
        const int ib = static_cast<int>(t % HB);
        const int n = static_cast<int>(t / HB);
        float acc[2][2][8];
#pragma unroll
        for (int a = 0; a < 2; ++a)
#pragma unroll
            for (int b = 0; b < 2; ++b)
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[a][b][j] = 0.f;
#pragma unroll
        for (int u = 0; u < 2; ++u) {
            const int oh = ib + u;
            if (oh >= OH) continue;
#pragma unroll
            for (int v = 0; v < 2; ++v) {
                const int ow = jb + v;
                if (ow >= OW) continue;
                const size_t o = ((static_cast<size_t>(n) * OH + oh) * OW + ow) * C + cg * 8;
                const uint2 packed = *reinterpret_cast<const uint2*>(idx + o);
                float g[8];
                unpack8f(ld_global_nc_v4(dy + o), g);
#pragma unroll
                for (int a = u; a < 2; ++a)
#pragma unroll
                    for (int b = v; b < 2; ++b) {
                        const int p = (a + 1 - 2 * u) * 3 + (b + 1 - 2 * v);
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            const int pj = ((j < 4 ? packed.x : packed.y) >> ((j & 3) * 8)) & 0xff;
                            if (pj == p) acc[a][b][j] += g[j];
                        }
                    }
            }
        }
#pragma unroll
        for (int a = 0; a < 2; ++a) {
            const int ih = 2 * ib + a;
            if (ih >= H) continue;
#pragma unroll
            for (int b = 0; b < 2; ++b) {
                const int iw = 2 * jb + b;
                if (iw >= W) continue;
                st_global_v4(dx + ((static_cast<size_t>(n) * H + ih) * W + iw) * C + cg * 8, pack8f(acc[a][b]));
            }
        }
    }
}

}  // namespace tony

extern "C" int tony_maxpool3x3s2_fwd(const void* x, void* y, void* idx, int N, int H, int W, int C, void* stream) {
    if (!x || !y || !idx || N <= 0 || H <= 0 || W <= 0 || C <= 0 || (C % 8)) return TONY_ERR_INVALID;
    const int OH = (H + 2 - 3) / 2 + 1, OW = (W + 2 - 3) / 2 + 1;
    const long long total = static_cast<long long>(N) * OH * OW * (C / 8);
    const int grid = static_cast<int>(std::min<long long>((total + 255) / 256, device_sm_count() * 32LL));
    maxpool3x3s2_fwd_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        static_cast<const __nv_bfloat16*>(x), static_cast<__nv_bfloat16*>(y), static_cast<unsigned char*>(idx), N, H,
        W, C, OH, OW);
    count_launch();
    TONY_CUDA_CHECK(cudaGetLastError());
    return TONY_OK;
}

extern "C" int tony_maxpool3x3s2_bwd(const void* dy, const void* idx, void* dx, int N, int H, int W, int C,
                                     void* stream) {
    if (!dy || !dx || !idx || N <= 0 || H <= 0 || W <= 0 || C <= 0 || (C % 8)) return TONY_ERR_INVALID;
    const int OH = (H + 2 - 3) / 2 + 1, OW = (W + 2 - 3) / 2 + 1;
    const long long total = static_cast<long long>(N) * ((H + 1) / 2) * ((W + 1) / 2) * (C / 8);   // 2 x 2 pixel blocks
    const int bgrid = static_cast<int>(std::min<long long>((total + 255) / 256, device_sm_count() * 32LL));
    maxpool3x3s2_bwd_kernel<<<bgrid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        static_cast<const __nv_bfloat16*>(dy), static_cast<const unsigned char*>(idx), static_cast<__nv_bfloat16*>(dx),
        N, H, W, C, OH, OW);
    count_launch();
    TONY_CUDA_CHECK(cudaGetLastError());
    return TONY_OK;
}
