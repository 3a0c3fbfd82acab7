// tony_b200 — optimizers on flat buffers, and the fused gradient-collective + optimizer
// kernel (K1/K2 of SURVEY.md §8) that is the heart of the DDP / Horovod / ps-worker data plane.
//
// fused_reduce_optim_kernel, one launch per gradient bucket, on every rank:
//   A. arrive/wait on per-rank flag words in peer-mapped memory (st.release.sys /
//      ld.acquire.sys) so every rank's bucket is known to be complete;
//   B. each rank reduces ITS 1/world shard directly out of all peers' bf16 gradient
//      buffers with 16-byte loads over NVLink (fp32 accumulation), applies unscale +
//      AdamW/SGD to its shard of the fp32 master weights and moments (optimizer state is
//      therefore sharded across ranks, ZeRO-1 style), and stores the new bf16 weights
//      into every rank's parameter buffer with 16-byte peer stores (or only locally for
//      the parameter-server path);
//   C. last CTA out signals "done" to every peer and waits for theirs, so kernel
//      completion on a rank implies its parameters are fully updated and its gradient
//      buffer is free to be overwritten by the next backward.
// NVLink traffic per rank: (world-1)/world * S in, (world-1)/world * S out — the
// reduce-scatter + all-gather optimum — with the optimizer's HBM traffic cut by `world`.
// Every wait is bounded (globaltimer budget) and reports TONY_ERR_TIMEOUT through a
// status word instead of hanging the GPU when a peer died.
//
// Replaces: ncclAllReduce + grad unscale + optimizer.step() + DDP bucket copies, i.e.
// what tony-examples/mnist-pytorch/mnist_distributed.py:113-126 (average_gradients) and
// Horovod's DistributedOptimizer (tony-examples/horovod-on-tony/tensorflow2_keras_mnist.py:55-56)
// do in the jobs the reference launches.
#include "common.cuh"
#include "tony_b200_kernels.h"

#include <algorithm>
#include <cstdlib>

namespace tony {

struct OptHyper {
    int optimizer;  // TONY_OPT_ADAMW, TONY_OPT_SGD, or -1 = none (plain all-reduce)
    float lr, beta1, beta2, eps, weight_decay, inv_bc1, inv_sqrt_bc2, grad_scale;
    int nesterov;
};

__device__ __forceinline__ void unpack8f(const uint4& u, float (&f)[8]) {
    const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const float2 t = unpack_bf16x2(w[j]);
        f[2 * j] = t.x;
        f[2 * j + 1] = t.y;
    }
}

// Update 8 consecutive elements of (master, s1, s2) at index i with gradient g (already
// summed over ranks, not yet scaled); returns the new weights packed as 8 bf16.
__device__ __forceinline__ uint4 optim_update8(const OptHyper& h, float* __restrict__ master,
                                               float* __restrict__ s1, float* __restrict__ s2,
                                               size_t i, float (&g)[8], float wd) {
    float4 p0 = *reinterpret_cast<float4*>(master + i), p1 = *reinterpret_cast<float4*>(master + i + 4);
    float4 m0 = *reinterpret_cast<float4*>(s1 + i), m1 = *reinterpret_cast<float4*>(s1 + i + 4);
    float p[8] = {p0.x, p0.y, p0.z, p0.w, p1.x, p1.y, p1.z, p1.w};
    float m[8] = {m0.x, m0.y, m0.z, m0.w, m1.x, m1.y, m1.z, m1.w};
    if (h.optimizer == TONY_OPT_ADAMW) {
        float4 v0 = *reinterpret_cast<float4*>(s2 + i), v1 = *reinterpret_cast<float4*>(s2 + i + 4);
        float v[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float gj = g[j] * h.grad_scale;
            m[j] = h.beta1 * m[j] + (1.0f - h.beta1) * gj;
            v[j] = h.beta2 * v[j] + (1.0f - h.beta2) * gj * gj;
            const float denom = sqrtf(v[j]) * h.inv_sqrt_bc2 + h.eps;
            p[j] = p[j] * (1.0f - h.lr * wd) - h.lr * (m[j] * h.inv_bc1) / denom;
        }
        *reinterpret_cast<float4*>(s2 + i) = make_float4(v[0], v[1], v[2], v[3]);
        *reinterpret_cast<float4*>(s2 + i + 4) = make_float4(v[4], v[5], v[6], v[7]);
    } else {  // SGD with momentum (beta1 = momentum)
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float gj = g[j] * h.grad_scale + wd * p[j];
            m[j] = h.beta1 * m[j] + gj;
            const float step = h.nesterov ? gj + h.beta1 * m[j] : m[j];
            p[j] -= h.lr * step;
        }
    }
    *reinterpret_cast<float4*>(master + i) = make_float4(p[0], p[1], p[2], p[3]);
    *reinterpret_cast<float4*>(master + i + 4) = make_float4(p[4], p[5], p[6], p[7]);
    *reinterpret_cast<float4*>(s1 + i) = make_float4(m[0], m[1], m[2], m[3]);
    *reinterpret_cast<float4*>(s1 + i + 4) = make_float4(m[4], m[5], m[6], m[7]);
    uint4 o;
    o.x = pack_bf16x2(p[0], p[1]);
    o.y = pack_bf16x2(p[2], p[3]);
    o.z = pack_bf16x2(p[4], p[5]);
    o.w = pack_bf16x2(p[6], p[7]);
    return o;
}

__global__ void __launch_bounds__(512)
flat_optim_kernel(OptHyper h, float* __restrict__ master, float* __restrict__ s1,
                  float* __restrict__ s2, const __nv_bfloat16* __restrict__ grad,
                  __nv_bfloat16* __restrict__ param, size_t n, const unsigned char* __restrict__ wd_mask,
                  size_t mask_offset) {
    const size_t nvec = n / 8;
    for (size_t v = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; v < nvec;
         v += static_cast<size_t>(gridDim.x) * blockDim.x) {
        const size_t i = v * 8;
        float g[8];
        unpack8f(ld_global_nc_v4(grad + i), g);
        const float wd = (wd_mask == nullptr || wd_mask[(mask_offset + i) >> 8]) ? h.weight_decay : 0.f;
        const uint4 w = optim_update8(h, master, s1, s2, i, g, wd);
        st_global_v4(param + i, w);
    }
}

struct FusedArgs {
    tony_fused_comm comm;
    size_t begin, count, state_offset;
    float* master;
    float* s1;
    float* s2;
    const unsigned char* wd_mask;
    int broadcast;
    uint32_t epoch;
    unsigned long long timeout_ns;
    float ar_scale;  // plain all-reduce: out = sum * ar_scale
    // Dedicated parameter servers: the bucket is cut into num_owners shards (not world) and only rank
    // owner_index >= 0 reduces / updates its shard; every rank still joins both flag phases and every
    // rank's gradient buffer is summed (servers that compute no gradient keep theirs zero).
    // num_owners == 0: the symmetric default (every rank owns 1/world).
    int num_owners, owner_index;
};

// Phase A/C helper: tell every peer we reached `epoch` in flag slot `slot`, then wait for all.
__device__ __forceinline__ bool arrive_and_wait_all(const FusedArgs& a, int slot, int tid) {
    const int world = a.comm.world, rank = a.comm.rank;
    bool ok = true;
    if (tid < world && tid != rank) {
        st_release_sys(a.comm.flags[tid] + slot * TONY_MAX_RANKS + rank, a.epoch);
        ok = wait_flag_sys(a.comm.flags[rank] + slot * TONY_MAX_RANKS + tid, a.epoch, a.timeout_ns);
        if (!ok) *a.comm.status = TONY_ERR_TIMEOUT;
    }
    return ok;
}

// U = 16-byte vectors per thread per loop trip.  The peer loads of one vector take microseconds over
// NVLink; with one vector in flight per thread the 32 CTAs this kernel is allowed next to the backward
// GEMMs moved ~80 GB/s per direction (measured, 2 ranks).  U vectors x `world` peers of loads are issued
// before the first is consumed (U = 4 / 2 / 1 for world <= 2 / 4 / 8 keeps it within the register file).
template <int U>
__global__ void __launch_bounds__(512)
fused_reduce_optim_kernel(OptHyper h, FusedArgs a) {
    __shared__ int sh_flag;
    const int world = a.comm.world, rank = a.comm.rank;
    const int tid = threadIdx.x;

    // ---- A: every rank's bucket is complete (stream order covers our own) -------------
    if (world > 1) {
        if (tid == 0) sh_flag = 1;
        __syncthreads();
        if (blockIdx.x == 0) {
            __threadfence_system();
            if (!arrive_and_wait_all(a, 0, tid)) sh_flag = 0;
        } else if (tid < world && tid != rank) {
            if (!wait_flag_sys(a.comm.flags[rank] + tid, a.epoch, a.timeout_ns)) {
                *a.comm.status = TONY_ERR_TIMEOUT;
                sh_flag = 0;
            }
        }
        __syncthreads();
        if (!sh_flag) return;  // a peer never arrived: leave state untouched, host sees status
    }

    // ---- B: reduce my shard from all peers, update, broadcast ---------------------------
    const int nsh = a.num_owners > 0 ? a.num_owners : world;
    const int my = a.num_owners > 0 ? a.owner_index : rank;
    const size_t chunk = a.count / nsh;
    const size_t my_begin = a.begin + static_cast<size_t>(my < 0 ? 0 : my) * chunk;
    const size_t nvec = my < 0 ? 0 : chunk / 8;      // a pure worker owns nothing
    constexpr int MAXW = U == 4 ? 2 : (U == 2 ? 4 : TONY_MAX_RANKS);   // peers this instantiation can hold
    const size_t tstride = static_cast<size_t>(gridDim.x) * blockDim.x;
    for (size_t v0 = blockIdx.x * static_cast<size_t>(blockDim.x) + tid; v0 < nvec; v0 += tstride * U) {
        uint4 gv[U][MAXW];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const size_t v = v0 + u * tstride;
            if (v < nvec) {
                const size_t idx = my_begin + v * 8;
#pragma unroll
                for (int j = 0; j < MAXW; ++j) {
                    if (j < world) {
                        int r = rank + j;
                        if (r >= world) r -= world;
                        gv[u][j] = ld_global_v4(static_cast<const __nv_bfloat16*>(a.comm.grad[r]) + idx);
                    }
                }
            }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const size_t v = v0 + u * tstride;
            if (v >= nvec) break;
            const size_t idx = my_begin + v * 8;
            float g[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int j = 0; j < MAXW; ++j) {
                if (j < world) {
                    float t[8];
                    unpack8f(gv[u][j], t);
#pragma unroll
                    for (int k = 0; k < 8; ++k) g[k] += t[k];
                }
            }
            uint4 w;
            if (h.optimizer >= 0) {
                const size_t si = a.state_offset + v * 8;
                const float wd = (a.wd_mask == nullptr || a.wd_mask[idx >> 8]) ? h.weight_decay : 0.f;
                w = optim_update8(h, a.master, a.s1, a.s2, si, g, wd);
            } else {
                w.x = pack_bf16x2(g[0] * a.ar_scale, g[1] * a.ar_scale);
                w.y = pack_bf16x2(g[2] * a.ar_scale, g[3] * a.ar_scale);
                w.z = pack_bf16x2(g[4] * a.ar_scale, g[5] * a.ar_scale);
                w.w = pack_bf16x2(g[6] * a.ar_scale, g[7] * a.ar_scale);
            }
            if (a.broadcast) {
#pragma unroll
                for (int j = 0; j < MAXW; ++j) {
                    if (j < world) {
                        int r = rank + j;
                        if (r >= world) r -= world;
                        st_global_v4(static_cast<__nv_bfloat16*>(a.comm.param[r]) + idx, w);
                    }
                }
            } else {
                st_global_v4(static_cast<__nv_bfloat16*>(a.comm.param[rank]) + idx, w);
            }
        }
    }

    // ---- C: last CTA out signals "done" and waits for every peer's "done" ------------
    if (world > 1) {
        __threadfence_system();
        __syncthreads();
        if (tid == 0) {
            const unsigned int prev = atomicAdd(a.comm.counter, 1u);
            sh_flag = (prev == gridDim.x - 1) ? 2 : 1;
        }
        __syncthreads();
        if (sh_flag == 2) {
            __threadfence_system();
            if (tid == 0) *a.comm.counter = 0;  // ready for the next launch
            arrive_and_wait_all(a, 1, tid);
        }
    }
}

// ---- NVLS variant ---------------------------------------------------------------------------
__device__ __forceinline__ uint4 multimem_ld_reduce_bf16x8(const void* mc_addr) {
    uint4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0, %1, %2, %3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                 : "l"(mc_addr)
                 : "memory");
    return v;
}
__device__ __forceinline__ void multimem_st_bf16x8(void* mc_addr, const uint4& v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.bf16x2 [%0], {%1, %2, %3, %4};" ::"l"(mc_addr), "r"(v.x),
                 "r"(v.y), "r"(v.z), "r"(v.w)
                 : "memory");
}
__device__ __forceinline__ void multimem_red_add_release(unsigned int* mc_addr, unsigned int v) {
    asm volatile("multimem.red.release.sys.global.add.u32 [%0], %1;" ::"l"(mc_addr), "r"(v) : "memory");
}

struct NvlsArgs {
    tony_nvls_comm comm;
    size_t begin, count, state_offset;
    float* master;
    float* s1;
    float* s2;
    const unsigned char* wd_mask;
    int broadcast;
    uint32_t target;  // world * call_index: every rank added 1 per call to the multicast flag
    unsigned long long timeout_ns;
    int num_owners, owner_index;   // see FusedArgs
};

constexpr int NVLS_FLAG_A = 40, NVLS_FLAG_B = 41;  // words of the flag block used by this path

__global__ void __launch_bounds__(512)
fused_reduce_optim_nvls_kernel(OptHyper h, NvlsArgs a) {
    __shared__ int sh_flag;
    const int world = a.comm.world, rank = a.comm.rank;
    const int tid = threadIdx.x;
    // ---- A: all ranks' buckets complete: one multicast add, everyone polls its own copy ----
    if (tid == 0) {
        sh_flag = 1;
        if (blockIdx.x == 0) {
            __threadfence_system();
            multimem_red_add_release(a.comm.flags_mc + NVLS_FLAG_A, 1u);
        }
        if (!wait_flag_sys(a.comm.flags_local + NVLS_FLAG_A, a.target, a.timeout_ns)) {
            *a.comm.status = TONY_ERR_TIMEOUT;
            sh_flag = 0;
        }
    }
    __syncthreads();
    if (!sh_flag) return;

    // ---- B: in-switch reduction of my shard, optimizer, multicast store of the new weights ---
    const int nsh = a.num_owners > 0 ? a.num_owners : world;
    const int my = a.num_owners > 0 ? a.owner_index : rank;
    const size_t chunk = a.count / nsh;
    const size_t my_begin = a.begin + static_cast<size_t>(my < 0 ? 0 : my) * chunk;
    const size_t nvec = my < 0 ? 0 : chunk / 8;
    const __nv_bfloat16* gmc = static_cast<const __nv_bfloat16*>(a.comm.grad_mc);
    // four in-switch reductions in flight per thread (each is a round trip through the NVSwitch)
    constexpr int U = 4;
    const size_t tstride = static_cast<size_t>(gridDim.x) * blockDim.x;
    for (size_t v0 = blockIdx.x * static_cast<size_t>(blockDim.x) + tid; v0 < nvec; v0 += tstride * U) {
        uint4 sum[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const size_t v = v0 + u * tstride;
            if (v < nvec) sum[u] = multimem_ld_reduce_bf16x8(gmc + my_begin + v * 8);
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const size_t v = v0 + u * tstride;
            if (v >= nvec) break;
            const size_t idx = my_begin + v * 8;
            float g[8];
            unpack8f(sum[u], g);
            uint4 w;
            if (h.optimizer >= 0) {
                const float wd = (a.wd_mask == nullptr || a.wd_mask[idx >> 8]) ? h.weight_decay : 0.f;
                w = optim_update8(h, a.master, a.s1, a.s2, a.state_offset + v * 8, g, wd);
            } else {
                w.x = pack_bf16x2(g[0] * h.grad_scale, g[1] * h.grad_scale);
                w.y = pack_bf16x2(g[2] * h.grad_scale, g[3] * h.grad_scale);
                w.z = pack_bf16x2(g[4] * h.grad_scale, g[5] * h.grad_scale);
                w.w = pack_bf16x2(g[6] * h.grad_scale, g[7] * h.grad_scale);
            }
            if (a.broadcast)
                multimem_st_bf16x8(static_cast<__nv_bfloat16*>(a.comm.param_mc) + idx, w);
            else
                st_global_v4(static_cast<__nv_bfloat16*>(a.comm.param_local) + idx, w);
        }
    }

    // ---- C: last CTA out publishes "done" and waits for every rank's "done" ------------------
    __threadfence_system();
    __syncthreads();
    if (tid == 0) {
        const unsigned int prev = atomicAdd(a.comm.counter, 1u);
        if (prev == gridDim.x - 1) {
            __threadfence_system();
            *a.comm.counter = 0;
            multimem_red_add_release(a.comm.flags_mc + NVLS_FLAG_B, 1u);
            if (!wait_flag_sys(a.comm.flags_local + NVLS_FLAG_B, a.target, a.timeout_ns))
                *a.comm.status = TONY_ERR_TIMEOUT;
        }
    }
}

// ---- stand-alone collectives over peer memory (K0c): all-reduce / reduce-scatter in bf16 or fp32,
// all-gather and broadcast of raw bytes.  Input is every rank's `grad` buffer, output its `param` buffer;
// `begin` / `bytes` delimit the same byte range on every rank (bytes % (world * 16) == 0).
//   all-reduce     out[range]            = scale * sum_r in_r[range]           (every rank)
//   reduce-scatter out[my 1/world shard] = scale * sum_r in_r[my shard]
//   all-gather     out[shard r]          = in_r[shard r]                       for every r
//   broadcast      out[range]            = in_root[range]
// Pull design: a rank only ever writes its own output; flag barrier before (inputs ready) and after (nobody
// still reads my input), as in the fused kernel above.
struct CollArgs {
    tony_fused_comm comm;
    size_t begin, bytes;
    int op, fp32, root;
    float scale;
    uint32_t epoch;
    unsigned long long timeout_ns;
};

__device__ __forceinline__ bool coll_arrive_and_wait(const CollArgs& a, int slot, int tid) {
    const int world = a.comm.world, rank = a.comm.rank;
    bool ok = true;
    if (tid < world && tid != rank) {
        st_release_sys(a.comm.flags[tid] + slot * TONY_MAX_RANKS + rank, a.epoch);
        ok = wait_flag_sys(a.comm.flags[rank] + slot * TONY_MAX_RANKS + tid, a.epoch, a.timeout_ns);
        if (!ok) *a.comm.status = TONY_ERR_TIMEOUT;
    }
    return ok;
}

__global__ void __launch_bounds__(512)
symm_collective_kernel(CollArgs a) {
    __shared__ int sh_flag;
    const int world = a.comm.world, rank = a.comm.rank;
    const int tid = threadIdx.x;
    if (world > 1) {
        if (tid == 0) sh_flag = 1;
        __syncthreads();
        if (blockIdx.x == 0) {
            __threadfence_system();
            if (!coll_arrive_and_wait(a, 0, tid)) sh_flag = 0;
        } else if (tid < world && tid != rank) {
            if (!wait_flag_sys(a.comm.flags[rank] + tid, a.epoch, a.timeout_ns)) {
                *a.comm.status = TONY_ERR_TIMEOUT;
                sh_flag = 0;
            }
        }
        __syncthreads();
        if (!sh_flag) return;
    }
    const size_t chunk = a.bytes / world;
    const size_t gtid = blockIdx.x * static_cast<size_t>(blockDim.x) + tid;
    const size_t gstride = static_cast<size_t>(gridDim.x) * blockDim.x;
    if (a.op == TONY_COLL_ALLREDUCE || a.op == TONY_COLL_REDUCE_SCATTER) {
        const size_t my = a.begin + static_cast<size_t>(rank) * chunk;
        for (size_t v = gtid; v < chunk / 16; v += gstride) {
            const size_t off = my + v * 16;
            uint4 gv[TONY_MAX_RANKS];
#pragma unroll
            for (int j = 0; j < TONY_MAX_RANKS; ++j) {
                if (j < world) {
                    int r = rank + j;
                    if (r >= world) r -= world;
                    gv[j] = ld_global_v4(static_cast<const char*>(a.comm.grad[r]) + off);
                }
            }
            uint4 w;
            if (a.fp32) {
                float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
                for (int j = 0; j < TONY_MAX_RANKS; ++j) {
                    if (j < world) {
                        acc.x += __uint_as_float(gv[j].x);
                        acc.y += __uint_as_float(gv[j].y);
                        acc.z += __uint_as_float(gv[j].z);
                        acc.w += __uint_as_float(gv[j].w);
                    }
                }
                w = make_uint4(__float_as_uint(acc.x * a.scale), __float_as_uint(acc.y * a.scale),
                               __float_as_uint(acc.z * a.scale), __float_as_uint(acc.w * a.scale));
            } else {
                float g[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
                for (int j = 0; j < TONY_MAX_RANKS; ++j) {
                    if (j < world) {
                        float t[8];
                        unpack8f(gv[j], t);
#pragma unroll
                        for (int k = 0; k < 8; ++k) g[k] += t[k];
                    }
                }
                w.x = pack_bf16x2(g[0] * a.scale, g[1] * a.scale);
                w.y = pack_bf16x2(g[2] * a.scale, g[3] * a.scale);
                w.z = pack_bf16x2(g[4] * a.scale, g[5] * a.scale);
                w.w = pack_bf16x2(g[6] * a.scale, g[7] * a.scale);
            }
            if (a.op == TONY_COLL_ALLREDUCE) {
#pragma unroll
                for (int j = 0; j < TONY_MAX_RANKS; ++j) {
                    if (j < world) {
                        int r = rank + j;
                        if (r >= world) r -= world;
                        st_global_v4(static_cast<char*>(a.comm.param[r]) + off, w);
                    }
                }
            } else {
                st_global_v4(static_cast<char*>(a.comm.param[rank]) + off, w);
            }
        }
    } else {
        // all-gather: segment j comes from rank j's input; broadcast: every segment from the root's
        char* out = static_cast<char*>(a.comm.param[rank]);
        for (size_t v = gtid; v < a.bytes / 16; v += gstride) {
            const size_t off = a.begin + v * 16;
            int src = a.root;
            if (a.op == TONY_COLL_ALLGATHER) {
                src = static_cast<int>((v * 16) / chunk);
                src = (src + rank) % world;      // start with the local shard, spread the peers
                const size_t o2 = a.begin + static_cast<size_t>(src) * chunk + (v * 16) % chunk;
                st_global_v4(out + o2, ld_global_v4(static_cast<const char*>(a.comm.grad[src]) + o2));
                continue;
            }
            st_global_v4(out + off, ld_global_v4(static_cast<const char*>(a.comm.grad[src]) + off));
        }
    }
    if (world > 1) {
        __threadfence_system();
        __syncthreads();
        if (tid == 0) {
            const unsigned int prev = atomicAdd(a.comm.counter, 1u);
            sh_flag = (prev == gridDim.x - 1) ? 2 : 1;
        }
        __syncthreads();
        if (sh_flag == 2) {
            __threadfence_system();
            if (tid == 0) *a.comm.counter = 0;
            coll_arrive_and_wait(a, 1, tid);
        }
    }
}

// NVLS broadcast: the root reads its own parameter buffer and publishes it to every rank with multimem.st
// (one store per 16 bytes leaves the root, the switch replicates it); everyone joins the two flag phases.
__global__ void __launch_bounds__(512)
nvls_broadcast_kernel(NvlsArgs a, int root) {
    __shared__ int sh_flag;
    const int tid = threadIdx.x;
    if (tid == 0) {
        sh_flag = 1;
        if (blockIdx.x == 0) {
            __threadfence_system();
            multimem_red_add_release(a.comm.flags_mc + NVLS_FLAG_A, 1u);
        }
        if (!wait_flag_sys(a.comm.flags_local + NVLS_FLAG_A, a.target, a.timeout_ns)) {
            *a.comm.status = TONY_ERR_TIMEOUT;
            sh_flag = 0;
        }
    }
    __syncthreads();
    if (!sh_flag) return;
    if (a.comm.rank == root) {
        const __nv_bfloat16* src = static_cast<const __nv_bfloat16*>(a.comm.param_local);
        for (size_t v = blockIdx.x * static_cast<size_t>(blockDim.x) + tid; v < a.count / 8;
             v += static_cast<size_t>(gridDim.x) * blockDim.x) {
            const size_t idx = a.begin + v * 8;
            multimem_st_bf16x8(static_cast<__nv_bfloat16*>(a.comm.param_mc) + idx, ld_global_v4(src + idx));
        }
    }
    __threadfence_system();
    __syncthreads();
    if (tid == 0) {
        const unsigned int prev = atomicAdd(a.comm.counter, 1u);
        if (prev == gridDim.x - 1) {
            __threadfence_system();
            *a.comm.counter = 0;
            multimem_red_add_release(a.comm.flags_mc + NVLS_FLAG_B, 1u);
            if (!wait_flag_sys(a.comm.flags_local + NVLS_FLAG_B, a.target, a.timeout_ns))
                *a.comm.status = TONY_ERR_TIMEOUT;
        }
    }
}

__global__ void symm_barrier_kernel(FusedArgs a) {
    __threadfence_system();
    arrive_and_wait_all(a, 0, threadIdx.x);
}

static unsigned long long comm_timeout_ns() {
    static unsigned long long cached = 0;
    if (cached == 0) {
        const char* e = getenv("TONY_B200_COMM_TIMEOUT_MS");
        long long ms = e ? atoll(e) : 60000;
        if (ms <= 0) ms = 60000;
        cached = static_cast<unsigned long long>(ms) * 1000000ull;
    }
    return cached;
}

static int check_comm(const tony_fused_comm* c, const char* who) {
    if (!c || c->world < 1 || c->world > TONY_MAX_RANKS || c->rank < 0 || c->rank >= c->world) {
        set_last_error(who, "bad comm descriptor (world must be 1..8)");
        return TONY_ERR_INVALID;
    }
    if (c->world > 1 && (!c->counter || !c->status)) {
        set_last_error(who, "comm.counter / comm.status must be set when world > 1");
        return TONY_ERR_INVALID;
    }
    for (int r = 0; r < c->world; ++r)
        if (!c->grad[r] || !c->param[r] || (c->world > 1 && !c->flags[r])) {
            set_last_error(who, "comm has a null peer pointer");
            return TONY_ERR_INVALID;
        }
    return TONY_OK;
}

}  // namespace tony

using namespace tony;

static OptHyper make_hyper(int optimizer, float lr, float b1, float b2, float eps, float wd, float bc1,
                           float bc2, float grad_scale, int nesterov) {
    OptHyper h;
    h.optimizer = optimizer;
    h.lr = lr;
    h.beta1 = b1;
    h.beta2 = b2;
    h.eps = eps;
    h.weight_decay = wd;
    h.inv_bc1 = bc1 > 0.f ? 1.0f / bc1 : 1.0f;
    h.inv_sqrt_bc2 = bc2 > 0.f ? 1.0f / sqrtf(bc2) : 1.0f;
    h.grad_scale = grad_scale;
    h.nesterov = nesterov;
    return h;
}

static int launch_flat(const OptHyper& h, float* master, float* s1, float* s2, const void* grad,
                       void* param, size_t n, const unsigned char* wd_mask, size_t mask_offset,
                       void* stream) {
    if (!master || !s1 || !grad || !param || n == 0 || (n % 8) != 0) {
        set_last_error("flat optimizer", "null pointer, or n not a multiple of 8");
        return TONY_ERR_INVALID;
    }
    const int threads = 512;
    const size_t nvec = n / 8;
    const int grid = static_cast<int>(
        std::min<size_t>((nvec + threads - 1) / threads, static_cast<size_t>(device_sm_count()) * 4));
    flat_optim_kernel<<<grid, threads, 0, static_cast<cudaStream_t>(stream)>>>(
        h, master, s1, s2, static_cast<const __nv_bfloat16*>(grad), static_cast<__nv_bfloat16*>(param),
        n, wd_mask, mask_offset);
    count_launch();
    TONY_CUDA_CHECK(cudaGetLastError());
    return TONY_OK;
}

extern "C" int tony_adamw_flat(float* master, float* exp_avg, float* exp_avg_sq,
                               const void* grad_bf16, void* param_bf16, size_t n, float lr,
                               float beta1, float beta2, float eps, float weight_decay, float bc1,
                               float bc2, float grad_scale, const unsigned char* wd_mask,
                               size_t mask_offset_elems, void* stream) {
    if (!exp_avg_sq) return TONY_ERR_INVALID;
    return launch_flat(make_hyper(TONY_OPT_ADAMW, lr, beta1, beta2, eps, weight_decay, bc1, bc2,
                                  grad_scale, 0),
                       master, exp_avg, exp_avg_sq, grad_bf16, param_bf16, n, wd_mask,
                       mask_offset_elems, stream);
}

extern "C" int tony_sgd_flat(float* master, float* momentum_buf, const void* grad_bf16,
                             void* param_bf16, size_t n, float lr, float momentum,
                             float weight_decay, int nesterov, float grad_scale,
                             const unsigned char* wd_mask, size_t mask_offset_elems, void* stream) {
    return launch_flat(make_hyper(TONY_OPT_SGD, lr, momentum, 0.f, 0.f, weight_decay, 1.f, 1.f,
                                  grad_scale, nesterov),
                       master, momentum_buf, nullptr, grad_bf16, param_bf16, n, wd_mask,
                       mask_offset_elems, stream);
}

static int launch_fused(const OptHyper& h, FusedArgs& a, int num_ctas, void* stream) {
    a.timeout_ns = comm_timeout_ns();
    const int threads = 512;
    const size_t nvec = a.count / (a.num_owners > 0 ? a.num_owners : a.comm.world) / 8;
    // world == 1 is pure HBM streaming: fill the machine.  With peers the kernel is NVLink
    // bound and shares the GPU with backward, so it stays on the caller's CTA budget.
    int grid = a.comm.world == 1 ? device_sm_count() * 4 : (num_ctas > 0 ? num_ctas : 32);
    grid = static_cast<int>(std::min<size_t>(grid, std::max<size_t>(1, (nvec + threads - 1) / threads)));
    if (a.comm.world > 1) grid = std::min(grid, device_sm_count());
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (a.comm.world <= 2)
        fused_reduce_optim_kernel<4><<<grid, threads, 0, st>>>(h, a);
    else if (a.comm.world <= 4)
        fused_reduce_optim_kernel<2><<<grid, threads, 0, st>>>(h, a);
    else
        fused_reduce_optim_kernel<1><<<grid, threads, 0, st>>>(h, a);
    count_launch();
    TONY_CUDA_CHECK(cudaGetLastError());
    return TONY_OK;
}

static int fused_reduce_optim_impl(const tony_fused_comm* comm, size_t begin, size_t count, float* master,
                                   float* state1, float* state2, size_t state_offset, int optimizer, float lr,
                                   float beta1_or_momentum, float beta2, float eps, float weight_decay, float bc1,
                                   float bc2, float grad_scale, const unsigned char* wd_mask, int broadcast,
                                   unsigned int epoch, int num_ctas, void* stream, int num_owners,
                                   int owner_index) {
    int rc = check_comm(comm, "tony_fused_reduce_optim");
    if (rc) return rc;
    const int nsh = num_owners > 0 ? num_owners : comm->world;
    if (count == 0 || (count % (static_cast<size_t>(nsh) * 8)) != 0 || (begin % 8) != 0 ||
        (state_offset % 8) != 0 || num_owners < 0 || num_owners > comm->world ||
        (num_owners > 0 && (owner_index < -1 || owner_index >= num_owners || broadcast))) {
        set_last_error("tony_fused_reduce_optim",
                       "count must be a multiple of shards*8; begin/state_offset of 8; owner shards imply broadcast = 0");
        return TONY_ERR_INVALID;
    }
    const bool owns = num_owners == 0 || owner_index >= 0;
    if ((owns && (!master || !state1 || (optimizer == TONY_OPT_ADAMW && !state2))) ||
        (optimizer != TONY_OPT_ADAMW && optimizer != TONY_OPT_SGD))
        return TONY_ERR_INVALID;
    FusedArgs a;
    a.num_owners = num_owners;
    a.owner_index = owner_index;
    a.comm = *comm;
    a.begin = begin;
    a.count = count;
    a.state_offset = state_offset;
    a.master = master;
    a.s1 = state1;
    a.s2 = state2;
    a.wd_mask = wd_mask;
    a.broadcast = broadcast;
    a.epoch = epoch;
    a.ar_scale = 1.f;
    return launch_fused(make_hyper(optimizer, lr, beta1_or_momentum, beta2, eps, weight_decay, bc1, bc2,
                                   grad_scale, 0),
                        a, num_ctas, stream);
}

extern "C" int tony_fused_reduce_optim(const tony_fused_comm* comm, size_t begin, size_t count,
                                       float* master, float* state1, float* state2,
                                       size_t state_offset, int optimizer, float lr,
                                       float beta1_or_momentum, float beta2, float eps,
                                       float weight_decay, float bc1, float bc2, float grad_scale,
                                       const unsigned char* wd_mask, int broadcast,
                                       unsigned int epoch, int num_ctas, void* stream) {
    return fused_reduce_optim_impl(comm, begin, count, master, state1, state2, state_offset, optimizer, lr,
                                   beta1_or_momentum, beta2, eps, weight_decay, bc1, bc2, grad_scale, wd_mask,
                                   broadcast, epoch, num_ctas, stream, 0, 0);
}

extern "C" int tony_fused_reduce_optim_owners(const tony_fused_comm* comm, size_t begin, size_t count,
                                              float* master, float* state1, float* state2, size_t state_offset,
                                              int optimizer, float lr, float beta1_or_momentum, float beta2,
                                              float eps, float weight_decay, float bc1, float bc2,
                                              float grad_scale, const unsigned char* wd_mask, int num_owners,
                                              int owner_index, unsigned int epoch, int num_ctas, void* stream) {
    if (num_owners <= 0) return TONY_ERR_INVALID;
    return fused_reduce_optim_impl(comm, begin, count, master, state1, state2, state_offset, optimizer, lr,
                                   beta1_or_momentum, beta2, eps, weight_decay, bc1, bc2, grad_scale, wd_mask, 0,
                                   epoch, num_ctas, stream, num_owners, owner_index);
}

extern "C" int tony_symm_allreduce_bf16(const tony_fused_comm* comm, size_t begin, size_t count,
                                        float scale, unsigned int epoch, int num_ctas,
                                        void* stream) {
    int rc = check_comm(comm, "tony_symm_allreduce_bf16");
    if (rc) return rc;
    if (count == 0 || (count % (static_cast<size_t>(comm->world) * 8)) != 0 || (begin % 8) != 0) {
        set_last_error("tony_symm_allreduce_bf16", "count must be a multiple of world*8; begin of 8");
        return TONY_ERR_INVALID;
    }
    FusedArgs a;
    a.num_owners = 0;
    a.owner_index = 0;
    a.comm = *comm;
    a.begin = begin;
    a.count = count;
    a.state_offset = 0;
    a.master = a.s1 = a.s2 = nullptr;
    a.wd_mask = nullptr;
    a.broadcast = 1;
    a.epoch = epoch;
    a.ar_scale = scale;
    return launch_fused(make_hyper(-1, 0, 0, 0, 0, 0, 1, 1, 1, 0), a, num_ctas, stream);
}

extern "C" int tony_symm_barrier(const tony_fused_comm* comm, unsigned int epoch, void* stream) {
    if (!comm || comm->world < 1 || comm->world > TONY_MAX_RANKS) return TONY_ERR_INVALID;
    if (comm->world == 1) return TONY_OK;
    for (int r = 0; r < comm->world; ++r)
        if (!comm->flags[r]) return TONY_ERR_INVALID;
    if (!comm->status) return TONY_ERR_INVALID;
    FusedArgs a;
    a.num_owners = 0;
    a.owner_index = 0;
    a.comm = *comm;
    a.begin = a.count = a.state_offset = 0;
    a.master = a.s1 = a.s2 = nullptr;
    a.wd_mask = nullptr;
    a.broadcast = 0;
    a.epoch = epoch;
    a.ar_scale = 1.f;
    a.timeout_ns = comm_timeout_ns();
    symm_barrier_kernel<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(a);
    count_launch();
    TONY_CUDA_CHECK(cudaGetLastError());
    return TONY_OK;
}


extern "C" int tony_fused_reduce_optim_nvls(const tony_nvls_comm* comm, size_t begin, size_t count,
                                            float* master, float* state1, float* state2,
                                            size_t state_offset, int optimizer, float lr,
                                            float beta1_or_momentum, float beta2, float eps,
                                            float weight_decay, float bc1, float bc2, float grad_scale,
                                            const unsigned char* wd_mask, int broadcast,
                                            unsigned int call_index, int num_ctas, void* stream) {
    if (!comm || comm->world < 2 || comm->world > TONY_MAX_RANKS || comm->rank < 0 || comm->rank >= comm->world ||
        !comm->grad_mc || !comm->param_mc || !comm->param_local || !comm->flags_mc || !comm->flags_local ||
        !comm->counter || !comm->status || call_index == 0) {
        set_last_error("tony_fused_reduce_optim_nvls", "bad comm descriptor (needs world >= 2 and multicast mappings)");
        return TONY_ERR_INVALID;
    }
    if (count == 0 || (count % (static_cast<size_t>(comm->world) * 8)) != 0 || (begin % 8) != 0 ||
        (state_offset % 8) != 0) {
        set_last_error("tony_fused_reduce_optim_nvls", "count must be a multiple of world*8; begin/state_offset of 8");
        return TONY_ERR_INVALID;
    }
    if (optimizer >= 0 && (!master || !state1 || (optimizer == TONY_OPT_ADAMW && !state2))) return TONY_ERR_INVALID;
    NvlsArgs a;
    a.num_owners = 0;
    a.owner_index = 0;
    a.comm = *comm;
    a.begin = begin;
    a.count = count;
    a.state_offset = state_offset;
    a.master = master;
    a.s1 = state1;
    a.s2 = state2;
    a.wd_mask = wd_mask;
    a.broadcast = broadcast;
    a.target = static_cast<uint32_t>(comm->world) * call_index;
    a.timeout_ns = comm_timeout_ns();
    const int threads = 512;
    const size_t nvec = count / comm->world / 8;
    int grid = num_ctas > 0 ? num_ctas : 32;
    grid = static_cast<int>(std::min<size_t>(grid, std::max<size_t>(1, (nvec + threads - 1) / threads)));
    grid = std::min(grid, device_sm_count());
    fused_reduce_optim_nvls_kernel<<<grid, threads, 0, static_cast<cudaStream_t>(stream)>>>(
        make_hyper(optimizer, lr, beta1_or_momentum, beta2, eps, weight_decay, bc1, bc2, grad_scale, 0), a);
    count_launch();
    TONY_CUDA_CHECK(cudaGetLastError());
    return TONY_OK;
}

extern "C" int tony_symm_collective(const tony_fused_comm* comm, int op, int fp32, size_t begin_bytes, size_t bytes,
                                    float scale, int root, unsigned int epoch, int num_ctas, void* stream) {
    int rc = check_comm(comm, "tony_symm_collective");
    if (rc) return rc;
    if (op < TONY_COLL_ALLREDUCE || op > TONY_COLL_BROADCAST || bytes == 0 ||
        (bytes % (static_cast<size_t>(comm->world) * 16)) != 0 || (begin_bytes % 16) != 0 || root < 0 ||
        root >= comm->world) {
        set_last_error("tony_symm_collective", "bytes must be a multiple of world*16, begin of 16, root < world");
        return TONY_ERR_INVALID;
    }
    CollArgs a;
    a.comm = *comm;
    a.begin = begin_bytes;
    a.bytes = bytes;
    a.op = op;
    a.fp32 = fp32;
    a.root = root;
    a.scale = scale;
    a.epoch = epoch;
    a.timeout_ns = comm_timeout_ns();
    const int threads = 512;
    const size_t nvec = bytes / 16;
    int grid = comm->world == 1 ? device_sm_count() * 2 : (num_ctas > 0 ? num_ctas : 32);
    grid = static_cast<int>(std::min<size_t>(grid, std::max<size_t>(1, (nvec + threads - 1) / threads)));
    symm_collective_kernel<<<grid, threads, 0, static_cast<cudaStream_t>(stream)>>>(a);
    count_launch();
    TONY_CUDA_CHECK(cudaGetLastError());
    return TONY_OK;
}

extern "C" int tony_nvls_broadcast(const tony_nvls_comm* comm, size_t begin, size_t count, int root,
                                   unsigned int call_index, int num_ctas, void* stream) {
    if (!comm || comm->world < 2 || comm->world > TONY_MAX_RANKS || comm->rank < 0 || comm->rank >= comm->world ||
        !comm->param_mc || !comm->param_local || !comm->flags_mc || !comm->flags_local || !comm->counter ||
        !comm->status || call_index == 0 || root < 0 || root >= comm->world || count == 0 || (count % 8) ||
        (begin % 8)) {
        set_last_error("tony_nvls_broadcast", "bad comm descriptor or range (count, begin multiples of 8)");
        return TONY_ERR_INVALID;
    }
    NvlsArgs a;
    a.num_owners = 0;
    a.owner_index = 0;
    a.comm = *comm;
    a.begin = begin;
    a.count = count;
    a.state_offset = 0;
    a.master = a.s1 = a.s2 = nullptr;
    a.wd_mask = nullptr;
    a.broadcast = 1;
    a.target = static_cast<uint32_t>(comm->world) * call_index;
    a.timeout_ns = comm_timeout_ns();
    const int threads = 512;
    int grid = num_ctas > 0 ? num_ctas : 32;
    grid = static_cast<int>(std::min<size_t>(grid, std::max<size_t>(1, (count / 8 + threads - 1) / threads)));
    nvls_broadcast_kernel<<<grid, threads, 0, static_cast<cudaStream_t>(stream)>>>(a, root);
    count_launch();
    TONY_CUDA_CHECK(cudaGetLastError());
    return TONY_OK;
}
