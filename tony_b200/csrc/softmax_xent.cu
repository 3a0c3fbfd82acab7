// tony_b200 — fused softmax cross-entropy forward+backward (sm_100a, memory-bound).
//
// One CTA per row.  Pass 1 streams the row once with 16-byte loads keeping an online
// (max, sum-exp) pair per thread; pass 2 re-reads the row (L2 resident: a 50k-entry
// bf16 row is 100 KB, L2 is 126 MB) and overwrites it with the gradient.  The logits
// tensor is therefore read from HBM once and written once — the stock sequence
// (log_softmax, nll_loss, then two backward kernels) moves it four to five times.
// Replaces F.cross_entropy (reference user code:
// tony-examples/mnist-pytorch/mnist_distributed.py:81) in launched jobs.
#include "common.cuh"
#include "tony_b200_kernels.h"

namespace tony {

constexpr int XENT_THREADS = 512;

__device__ __forceinline__ void online_update(float& m, float& s, float x) {
    if (x > m) {
        s = s * __expf(m - x) + 1.0f;
        m = x;
    } else {
        s += __expf(x - m);
    }
}

__global__ void __launch_bounds__(XENT_THREADS)
softmax_xent_kernel(__nv_bfloat16* __restrict__ logits, const long long* __restrict__ targets,
                    float* __restrict__ loss, int V, int V_pad, long long ld, long long ignore_index,
                    float grad_scale, const float* __restrict__ grad_scale_dev) {
    __shared__ float sm_m[XENT_THREADS / 32], sm_s[XENT_THREADS / 32];
    __shared__ float sh_max, sh_sum;
    const int row = blockIdx.x;
    __nv_bfloat16* lr = logits + static_cast<size_t>(row) * ld;
    const long long tgt = targets[row];
    const int nvec_pad = V_pad / 8;
    if (tgt == ignore_index || tgt < 0 || tgt >= V) {
        if (threadIdx.x == 0) loss[row] = 0.f;
        const uint4 z = make_uint4(0, 0, 0, 0);
        for (int i = threadIdx.x; i < nvec_pad; i += XENT_THREADS) st_global_v4(lr + i * 8, z);
        return;
    }
    const int nvec = V / 8;
    float m = -INFINITY, s = 0.f;
    for (int i = threadIdx.x; i < nvec; i += XENT_THREADS) {
        const uint4 u = ld_global_v4(lr + i * 8);
        const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float2 f = unpack_bf16x2(w[j]);
            online_update(m, s, f.x);
            online_update(m, s, f.y);
        }
    }
    for (int c = nvec * 8 + threadIdx.x; c < V; c += XENT_THREADS)
        online_update(m, s, __bfloat162float(lr[c]));
    // warp then block merge of (max, sum) pairs
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float m2 = __shfl_xor_sync(0xffffffffu, m, o);
        const float s2 = __shfl_xor_sync(0xffffffffu, s, o);
        const float mm = fmaxf(m, m2);
        s = (m == -INFINITY ? 0.f : s * __expf(m - mm)) + (m2 == -INFINITY ? 0.f : s2 * __expf(m2 - mm));
        m = mm;
    }
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0) {
        sm_m[warp] = m;
        sm_s[warp] = s;
    }
    __syncthreads();
    if (warp == 0) {
        m = lane < XENT_THREADS / 32 ? sm_m[lane] : -INFINITY;
        s = lane < XENT_THREADS / 32 ? sm_s[lane] : 0.f;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float m2 = __shfl_xor_sync(0xffffffffu, m, o);
            const float s2 = __shfl_xor_sync(0xffffffffu, s, o);
            const float mm = fmaxf(m, m2);
            s = (m == -INFINITY ? 0.f : s * __expf(m - mm)) +
                (m2 == -INFINITY ? 0.f : s2 * __expf(m2 - mm));
            m = mm;
        }
        if (lane == 0) {
            sh_max = m;
            sh_sum = s;
            loss[row] = __logf(s) + m - __bfloat162float(lr[tgt]);
        }
    }
    __syncthreads();
    const float row_max = sh_max;
    const float inv_sum = 1.0f / sh_sum;
    const float gs = grad_scale * (grad_scale_dev ? *grad_scale_dev : 1.0f);
    const int tvec = static_cast<int>(tgt >> 3), tsub = static_cast<int>(tgt & 7);
    for (int i = threadIdx.x; i < nvec_pad; i += XENT_THREADS) {
        const uint4 u = ld_global_v4(lr + i * 8);
        const uint32_t w[4] = {u.x, u.y, u.z, u.w};
        float g[8];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float2 f = unpack_bf16x2(w[j]);
            g[2 * j] = __expf(f.x - row_max) * inv_sum;
            g[2 * j + 1] = __expf(f.y - row_max) * inv_sum;
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const int c = i * 8 + j;
            if (i == tvec && j == tsub) g[j] -= 1.0f;
            g[j] = c < V ? g[j] * gs : 0.f;
        }
        uint4 o;
        o.x = pack_bf16x2(g[0], g[1]);
        o.y = pack_bf16x2(g[2], g[3]);
        o.z = pack_bf16x2(g[4], g[5]);
        o.w = pack_bf16x2(g[6], g[7]);
        st_global_v4(lr + i * 8, o);
    }
}

}  // namespace tony

using namespace tony;

extern "C" int tony_softmax_xent_fwd_bwd(void* logits, const long long* targets, float* loss,
                                         int rows, int V, long long ld, long long ignore_index,
                                         float grad_scale, const float* grad_scale_dev,
                                         void* stream) {
    if (!logits || !targets || !loss || rows <= 0 || V <= 0) return TONY_ERR_INVALID;
    // columns [V, V_pad) are alignment padding of the vocabulary: excluded from the
    // softmax, gradient forced to zero so the following GEMMs can run on the padded width
    const int V_pad = (V + 7) / 8 * 8;
    if (ld < V_pad || (ld % 8) != 0 || (reinterpret_cast<uintptr_t>(logits) & 15)) {
        set_last_error("tony_softmax_xent_fwd_bwd", "ld must be a multiple of 8 and >= V rounded up to 8");
        return TONY_ERR_INVALID;
    }
    // whole padded row width is zero-filled in the gradient
    const int V_zero = static_cast<int>(ld);
    softmax_xent_kernel<<<rows, XENT_THREADS, 0, static_cast<cudaStream_t>(stream)>>>(
        static_cast<__nv_bfloat16*>(logits), targets, loss, V, V_zero, ld, ignore_index, grad_scale,
        grad_scale_dev);
    count_launch();
    TONY_CUDA_CHECK(cudaGetLastError());
    return TONY_OK;
}
