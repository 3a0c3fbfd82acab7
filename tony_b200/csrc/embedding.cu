// tony_b200 — small data-movement kernels that sit between the library attention call and
// our GEMMs.
//
// tony_pack3_rows: interleave three [rows, row_elems] bf16 matrices (dq, dk, dv as the
// flash-attention backward returns them) into one packed [rows, 3, row_elems] gradient for
// the fused QKV projection, in ONE launch of 16-byte copies.  Stock autograd does this with
// three zero-filled select_backward tensors plus adds; a plain strided copy_ goes through
// ATen's non-vectorised elementwise kernel (measured 36 us each x 72 per BERT-large step).
#include "common.cuh"
#include "tony_b200_kernels.h"

#include <algorithm>

namespace tony {

__global__ void __launch_bounds__(256)
pack3_rows_kernel(const uint4* __restrict__ a, const uint4* __restrict__ b, const uint4* __restrict__ c,
                  long long lda, long long ldb, long long ldc, uint4* __restrict__ out, int rows,
                  int vec_per_row) {
    const long long total = static_cast<long long>(rows) * 3 * vec_per_row;
    for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        const int v = static_cast<int>(i % vec_per_row);
        const long long rw = i / vec_per_row;
        const int which = static_cast<int>(rw % 3);
        const long long row = rw / 3;
        const uint4* src = which == 0 ? a + row * lda : (which == 1 ? b + row * ldb : c + row * ldc);
        out[i] = ld_global_nc_v4(src + v);
    }
}

}  // namespace tony

using namespace tony;

extern "C" int tony_pack3_rows(const void* a, const void* b, const void* c, long long lda, long long ldb,
                               long long ldc, void* out, int rows, int row_elems, void* stream) {
    if (!a || !b || !c || !out || rows <= 0 || row_elems <= 0) return TONY_ERR_INVALID;
    if ((row_elems % 8) || (lda % 8) || (ldb % 8) || (ldc % 8) ||
        ((reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(b) | reinterpret_cast<uintptr_t>(c) |
          reinterpret_cast<uintptr_t>(out)) & 15)) {
        set_last_error("tony_pack3_rows", "row_elems / strides must be multiples of 8, pointers 16B aligned");
        return TONY_ERR_INVALID;
    }
    const int vec = row_elems / 8;
    const long long total = static_cast<long long>(rows) * 3 * vec;
    const int grid = static_cast<int>(std::min<long long>((total + 255) / 256, device_sm_count() * 16LL));
    pack3_rows_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        static_cast<const uint4*>(a), static_cast<const uint4*>(b), static_cast<const uint4*>(c), lda / 8,
        ldb / 8, ldc / 8, static_cast<uint4*>(out), rows, vec);
    count_launch();
    TONY_CUDA_CHECK(cudaGetLastError());
    return TONY_OK;
}

// ---------------------------------------------------------------------------------------
// Fused input embedding (K9 of the kernel worklist).
//   forward : out[t,:] = word[ids[t],:] (+ type[tt[t],:]) (+ pos[t mod S,:])      one pass
//   backward: one pass over dy produces all three gradients:
//               dword[ids[t],:] += dy[t,:]     bf16x2 reductions (red.add.noftz.bf16x2)
//               dpos[s,:]        = sum_b dy[b*S+s,:]          register accumulation, one store
//               dtype[k,:]       = sum_{t: tt[t]=k} dy[t,:]   per-position partial rows in scratch,
//                                                              folded by a second small kernel
// Replaces three ATen gathers + two adds (forward) and three index_add_ scatters whose
// type-embedding form serialises 16 K tokens onto 2 rows with atomics (backward).
namespace tony {

constexpr int EMB_MAX_TYPES = 4;

__global__ void __launch_bounds__(256)
embedding_sum_fwd_kernel(const __nv_bfloat16* __restrict__ word, const long long* __restrict__ ids,
                         const __nv_bfloat16* __restrict__ type, const long long* __restrict__ tt,
                         const __nv_bfloat16* __restrict__ pos, __nv_bfloat16* __restrict__ out,
                         long long tokens, int seq, int H) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (long long t = static_cast<long long>(blockIdx.x) * 8 + warp; t < tokens;
         t += static_cast<long long>(gridDim.x) * 8) {
        const __nv_bfloat16* w = word + ids[t] * H;
        const __nv_bfloat16* ty = type ? type + tt[t] * H : nullptr;
        const __nv_bfloat16* ps = pos ? pos + (t % seq) * H : nullptr;
        for (int c = lane * 8; c < H; c += 256) {
            const uint4 a = ld_global_nc_v4(w + c);
            const uint32_t aw[4] = {a.x, a.y, a.z, a.w};
            float v[8];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float2 f = unpack_bf16x2(aw[j]);
                v[2 * j] = f.x;
                v[2 * j + 1] = f.y;
            }
            if (ty) {
                const uint4 b = ld_global_nc_v4(ty + c);
                const uint32_t bw[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const float2 f = unpack_bf16x2(bw[j]);
                    v[2 * j] += f.x;
                    v[2 * j + 1] += f.y;
                }
            }
            if (ps) {
                const uint4 b = ld_global_nc_v4(ps + c);
                const uint32_t bw[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const float2 f = unpack_bf16x2(bw[j]);
                    v[2 * j] += f.x;
                    v[2 * j + 1] += f.y;
                }
            }
            uint4 o;
            o.x = pack_bf16x2(v[0], v[1]);
            o.y = pack_bf16x2(v[2], v[3]);
            o.z = pack_bf16x2(v[4], v[5]);
            o.w = pack_bf16x2(v[6], v[7]);
            st_global_v4(out + t * H + c, o);
        }
    }
}

__device__ __forceinline__ void red_add_bf16x2(__nv_bfloat16* addr, uint32_t packed) {
    asm volatile("red.global.add.noftz.bf16x2 [%0], %1;" ::"l"(addr), "r"(packed) : "memory");
}

// One CTA per (sequence position, 256-column tile); 32 lanes x 8 columns, 8 warps split the batch.
__global__ void __launch_bounds__(256)
embedding_sum_bwd_kernel(const __nv_bfloat16* __restrict__ dy, const long long* __restrict__ ids,
                         const long long* __restrict__ tt, __nv_bfloat16* __restrict__ dword,
                         __nv_bfloat16* __restrict__ dpos, int pos_accumulate,
                         float* __restrict__ type_partials, int n_types, int batch, int seq, int H) {
    __shared__ float red[8][256];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int s = blockIdx.x;
    const int col = blockIdx.y * 256 + lane * 8;
    float pacc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    float tacc[EMB_MAX_TYPES][8];
#pragma unroll
    for (int k = 0; k < EMB_MAX_TYPES; ++k)
#pragma unroll
        for (int j = 0; j < 8; ++j) tacc[k][j] = 0.f;
    if (col < H) {
        for (int b = warp; b < batch; b += 8) {
            const long long t = static_cast<long long>(b) * seq + s;
            const uint4 u = ld_global_nc_v4(dy + t * H + col);
            const uint32_t w[4] = {u.x, u.y, u.z, u.w};
            if (dword) {
                __nv_bfloat16* dst = dword + ids[t] * H + col;
#pragma unroll
                for (int j = 0; j < 4; ++j) red_add_bf16x2(dst + 2 * j, w[j]);
            }
            const int k = type_partials ? static_cast<int>(tt[t]) : 0;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float2 f = unpack_bf16x2(w[j]);
                pacc[2 * j] += f.x;
                pacc[2 * j + 1] += f.y;
#pragma unroll
                for (int q = 0; q < EMB_MAX_TYPES; ++q) {
                    if (q == k) {
                        tacc[q][2 * j] += f.x;
                        tacc[q][2 * j + 1] += f.y;
                    }
                }
            }
        }
    }
    // fold the 8 warps (batch slices) of this CTA: position gradient, then each type row
    auto cta_sum = [&](const float (&v)[8]) -> float {
#pragma unroll
        for (int j = 0; j < 8; ++j) red[warp][lane * 8 + j] = v[j];
        __syncthreads();
        float t = 0.f;
#pragma unroll
        for (int w = 0; w < 8; ++w) t += red[w][threadIdx.x];
        __syncthreads();
        return t;
    };
    const int gc = blockIdx.y * 256 + threadIdx.x;
    if (dpos) {
        float t = cta_sum(pacc);
        if (gc < H) {
            __nv_bfloat16* d = dpos + static_cast<size_t>(s) * H + gc;
            if (pos_accumulate) t += __bfloat162float(*d);
            *d = __float2bfloat16(t);
        }
    }
    if (type_partials) {
        for (int k = 0; k < n_types; ++k) {
            const float t = cta_sum(tacc[k]);
            if (gc < H) type_partials[(static_cast<size_t>(s) * n_types + k) * H + gc] = t;
        }
    }
}

__global__ void __launch_bounds__(256)
embedding_type_fold_kernel(const float* __restrict__ partials, int seq, int n_types, int H,
                           __nv_bfloat16* __restrict__ dtype, int accumulate) {
    __shared__ float red[8][256];
    const int k = blockIdx.y;
    float t = fold_partials_256(partials + static_cast<size_t>(k) * H, static_cast<size_t>(n_types) * H, seq,
                                blockIdx.x * 256, H, red);
    const int gc = blockIdx.x * 256 + threadIdx.x;
    if (gc < H) {
        __nv_bfloat16* d = dtype + static_cast<size_t>(k) * H + gc;
        if (accumulate) t += __bfloat162float(*d);
        *d = __float2bfloat16(t);
    }
}

}  // namespace tony

extern "C" int tony_embedding_sum_fwd(const void* word, const long long* ids, const void* type,
                                      const long long* tt, const void* pos, void* out, long long tokens,
                                      int seq, int H, void* stream) {
    if (!word || !ids || !out || tokens <= 0 || seq <= 0 || H <= 0 || (type && !tt)) return TONY_ERR_INVALID;
    if (H % 8) {
        set_last_error("tony_embedding_sum_fwd", "H must be a multiple of 8");
        return TONY_ERR_INVALID;
    }
    const int grid = static_cast<int>(std::min<long long>((tokens + 7) / 8, device_sm_count() * 16LL));
    embedding_sum_fwd_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        static_cast<const __nv_bfloat16*>(word), ids, static_cast<const __nv_bfloat16*>(type), tt,
        static_cast<const __nv_bfloat16*>(pos), static_cast<__nv_bfloat16*>(out), tokens, seq, H);
    count_launch();
    TONY_CUDA_CHECK(cudaGetLastError());
    return TONY_OK;
}

extern "C" size_t tony_embedding_bwd_scratch_bytes(int seq, int n_types, int H) {
    return static_cast<size_t>(seq) * static_cast<size_t>(n_types > 0 ? n_types : 1) * H * sizeof(float);
}

extern "C" int tony_embedding_sum_bwd(const void* dy, const long long* ids, const long long* tt, void* dword,
                                      long long vocab_rows, int word_accumulate, void* dtype, int n_types,
                                      int type_accumulate, void* dpos, int pos_accumulate, float* scratch,
                                      size_t scratch_bytes, int batch, int seq, int H, void* stream) {
    if (!dy || batch <= 0 || seq <= 0 || H <= 0 || (dword && !ids) || (dtype && !tt)) return TONY_ERR_INVALID;
    if (H % 8 || n_types > EMB_MAX_TYPES) {
        set_last_error("tony_embedding_sum_bwd", "H must be a multiple of 8 and at most 4 token types");
        return TONY_ERR_UNSUPPORTED;
    }
    if (dtype && (!scratch || scratch_bytes < tony_embedding_bwd_scratch_bytes(seq, n_types, H))) {
        set_last_error("tony_embedding_sum_bwd", "scratch too small (see tony_embedding_bwd_scratch_bytes)");
        return TONY_ERR_INVALID;
    }
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (dword && !word_accumulate)
        TONY_CUDA_CHECK(cudaMemsetAsync(dword, 0, static_cast<size_t>(vocab_rows) * H * 2, st));
    const int col_tiles = (H + 255) / 256;
    embedding_sum_bwd_kernel<<<dim3(seq, col_tiles), 256, 0, st>>>(
        static_cast<const __nv_bfloat16*>(dy), ids, tt, static_cast<__nv_bfloat16*>(dword),
        static_cast<__nv_bfloat16*>(dpos), pos_accumulate, dtype ? scratch : nullptr, n_types, batch, seq, H);
    TONY_CUDA_CHECK(cudaGetLastError());
    count_launch();
    if (dtype) {
        embedding_type_fold_kernel<<<dim3(col_tiles, n_types), 256, 0, st>>>(
            scratch, seq, n_types, H, static_cast<__nv_bfloat16*>(dtype), type_accumulate);
        TONY_CUDA_CHECK(cudaGetLastError());
        count_launch();
    }
    return TONY_OK;
}
