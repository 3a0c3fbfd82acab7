// tony_b200 — column sum of a bf16 matrix (bias gradient: db[c] = sum_r dY[r,c]).
//
// Two-stage: a 2-D grid of CTAs (column tile x row slab) writes fp32 partial rows into
// caller scratch, a finalize kernel folds the slabs and casts to bf16 (optionally
// accumulating).  Reads dY exactly once with 16-byte loads.
// Replaces the sum(dim=0) ATen reduction inside addmm backward in launched jobs.
#include "common.cuh"
#include "tony_b200_kernels.h"

#include <algorithm>
#include <mutex>

namespace tony {

constexpr int CS_MAX_SLABS = 64;         // static-scratch variant (fixed 2 MiB array)
constexpr int CS_MAX_SLABS_FUSED = 160;  // single-launch variant: enough for 4 CTAs per SM on one column tile
constexpr int CS_THREADS = 256;  // 8 warps: 32 lanes x 8 columns = 256 columns per CTA

__global__ void __launch_bounds__(CS_THREADS)
colsum_partial_kernel(const __nv_bfloat16* __restrict__ x, int rows, int cols, long long ld,
                      int rows_per_slab, float* __restrict__ partials) {
    __shared__ float red[CS_THREADS / 32][256];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int col = blockIdx.x * 256 + lane * 8;
    const int r0 = blockIdx.y * rows_per_slab;
    const int r1 = min(rows, r0 + rows_per_slab);
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    if (col < cols) {
        for (int r = r0 + warp; r < r1; r += CS_THREADS / 32) {
            const uint4 u = ld_global_nc_v4(x + static_cast<size_t>(r) * ld + col);
            const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float2 f = unpack_bf16x2(w[j]);
                acc[2 * j] += f.x;
                acc[2 * j + 1] += f.y;
            }
        }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) red[warp][lane * 8 + j] = acc[j];
    __syncthreads();
    const int c = threadIdx.x;  // 256 threads <-> 256 columns
    float s = 0.f;
#pragma unroll
    for (int w = 0; w < CS_THREADS / 32; ++w) s += red[w][c];
    const int gc = blockIdx.x * 256 + c;
    if (gc < cols) partials[static_cast<size_t>(blockIdx.y) * cols + gc] = s;
}

__global__ void colsum_finalize_kernel(const float* __restrict__ partials, int slabs, int cols,
                                       __nv_bfloat16* __restrict__ out, int accumulate) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= cols) return;
    float s = 0.f;
    for (int k = 0; k < slabs; ++k) s += partials[static_cast<size_t>(k) * cols + c];
    if (accumulate) s += __bfloat162float(out[c]);
    out[c] = __float2bfloat16(s);
}

// Single-launch form used by tony_colsum_bf16: every CTA publishes its partial row and the slabs are folded
// in two levels by the CTAs that arrive last (fold_two_level_256 in common.cuh).  Saves one ~6 us launch per bias
// gradient (121 per BERT-large step).  Counters live in the first 4 KB of the caller's scratch; the host clears
// them on the stream before each launch.
//
// ACT != 0 turns the pass into the backward of a fused Linear + activation whose incoming gradient arrives from
// outside the library: g = x * act'(pre) is written to `gout` and its column sums (the bias gradient) come out of
// the same read - one pass instead of ATen's three fp32 passes plus a separate column sum.
// ACT: 1 ReLU (pre may be the ReLU output), 2 exact GELU, 3 tanh GELU.
template <int ACT>
__global__ void __launch_bounds__(CS_THREADS, 4)   // with no minimum ptxas aims lower and serialises the row loads
colsum_fused_kernel(const __nv_bfloat16* __restrict__ x, int rows, int cols, long long ld,
                    int rows_per_slab, float* __restrict__ partials, float* __restrict__ gparts,
                    unsigned int* __restrict__ counters, __nv_bfloat16* __restrict__ out, int accumulate,
                    const __nv_bfloat16* __restrict__ pre, __nv_bfloat16* __restrict__ gout) {
    pdl_wait();                // launched with programmatic stream serialization: see launch_pdl_kernel
    pdl_launch_dependents();
    __shared__ float red[CS_THREADS / 32][256];
    __shared__ unsigned int is_last;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int col = blockIdx.x * 256 + lane * 8;
    const int r0 = blockIdx.y * rows_per_slab;
    const int r1 = min(rows, r0 + rows_per_slab);
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    if (col < cols) {
        // rows in flight per warp: eight (four with the second operand); the last batch is predicated instead of
        // falling back to a one-row-at-a-time loop (each of those iterations was a full memory round trip)
        constexpr int W = CS_THREADS / 32;
        if constexpr (ACT == 0) {
            for (int r = r0 + warp; r < r1; r += 8 * W) {
                uint4 u[8];
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    const int rr = r + k * W;
                    u[k] = rr < r1 ? ld_global_nc_v4(x + static_cast<size_t>(rr) * ld + col) : make_uint4(0, 0, 0, 0);
                }
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    const uint32_t w[4] = {u[k].x, u[k].y, u[k].z, u[k].w};
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const float2 f = unpack_bf16x2(w[j]);
                        acc[2 * j] += f.x;
                        acc[2 * j + 1] += f.y;
                    }
                }
            }
        } else {
            for (int r = r0 + warp; r < r1; r += 4 * W) {
                uint4 u[4], v[4];
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const int rr = r + k * W;
                    const size_t o = static_cast<size_t>(rr) * ld + col;
                    u[k] = rr < r1 ? ld_global_nc_v4(x + o) : make_uint4(0, 0, 0, 0);
                    v[k] = rr < r1 ? ld_global_nc_v4(pre + o) : make_uint4(0, 0, 0, 0);
                }
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const uint32_t w[4] = {u[k].x, u[k].y, u[k].z, u[k].w};
                    const uint32_t pw[4] = {v[k].x, v[k].y, v[k].z, v[k].w};
                    uint32_t ow[4];
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        float2 f = unpack_bf16x2(w[j]);
                        const float2 q = unpack_bf16x2(pw[j]);
                        if (ACT == 1) {
                            f.x = q.x > 0.f ? f.x : 0.f;
                            f.y = q.y > 0.f ? f.y : 0.f;
                        } else if (ACT == 2) {
                            f.x *= dgelu_erf(q.x);
                            f.y *= dgelu_erf(q.y);
                        } else {
                            f.x *= dgelu_tanh(q.x);
                            f.y *= dgelu_tanh(q.y);
                        }
                        ow[j] = pack_bf16x2(f.x, f.y);
                        const float2 g = unpack_bf16x2(ow[j]);     // sum what was stored
                        acc[2 * j] += g.x;
                        acc[2 * j + 1] += g.y;
                    }
                    const int rr = r + k * W;
                    if (rr < r1) st_global_v4(gout + static_cast<size_t>(rr) * ld + col, make_uint4(ow[0], ow[1], ow[2], ow[3]));
                }
            }
        }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) red[warp][lane * 8 + j] = acc[j];
    __syncthreads();
    const int c = threadIdx.x;
    float s = 0.f;
#pragma unroll
    for (int w = 0; w < CS_THREADS / 32; ++w) s += red[w][c];
    __syncthreads();
    const int gc = blockIdx.x * 256 + c;
    const float pp[1] = {s};
    float tt[1];
    if (!fold_two_level_256<1>(pp, partials, gparts, counters, cols, red, &is_last, tt)) return;
    if (gc < cols && out != nullptr) {
        float t = tt[0];
        if (accumulate) t += __bfloat162float(out[gc]);
        out[gc] = __float2bfloat16(t);
    }
}

// Interleave three [rows, H] matrices (the separate dq / dk / dv a library attention backward returns) into the
// packed [rows, 3H] gradient of the fused QKV projection AND produce that projection's bias gradient - the column
// sums of what is written - in the same pass.  Same structure as colsum_fused_kernel; a 256-column tile never
// straddles two sources because H is a multiple of 256 here.
__global__ void __launch_bounds__(CS_THREADS, 4)
pack3_colsum_kernel(const __nv_bfloat16* __restrict__ a, const __nv_bfloat16* __restrict__ b,
                    const __nv_bfloat16* __restrict__ c, long long lda, long long ldb, long long ldc, int H, int rows,
                    int rows_per_slab, __nv_bfloat16* __restrict__ packed, float* __restrict__ partials,
                    float* __restrict__ gparts, unsigned int* __restrict__ counters, __nv_bfloat16* __restrict__ out,
                    int accumulate) {
    pdl_wait();
    pdl_launch_dependents();
    __shared__ float red[CS_THREADS / 32][256];
    __shared__ unsigned int is_last;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int cols = 3 * H;
    const int col = blockIdx.x * 256 + lane * 8;
    const int part = (blockIdx.x * 256) / H;
    const int pc = col - part * H;
    const __nv_bfloat16* src = part == 0 ? a : (part == 1 ? b : c);
    const long long ld = part == 0 ? lda : (part == 1 ? ldb : ldc);
    const int r0 = blockIdx.y * rows_per_slab;
    const int r1 = min(rows, r0 + rows_per_slab);
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    constexpr int W = CS_THREADS / 32;
    for (int r = r0 + warp; r < r1; r += 8 * W) {
        uint4 u[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const int rr = r + k * W;
            u[k] = rr < r1 ? ld_global_nc_v4(src + static_cast<size_t>(rr) * ld + pc) : make_uint4(0, 0, 0, 0);
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const int rr = r + k * W;
            if (rr < r1) st_global_v4(packed + static_cast<size_t>(rr) * cols + col, u[k]);
            const uint32_t w[4] = {u[k].x, u[k].y, u[k].z, u[k].w};
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float2 f = unpack_bf16x2(w[j]);
                acc[2 * j] += f.x;
                acc[2 * j + 1] += f.y;
            }
        }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) red[warp][lane * 8 + j] = acc[j];
    __syncthreads();
    float s = 0.f;
#pragma unroll
    for (int w = 0; w < W; ++w) s += red[w][threadIdx.x];
    __syncthreads();
    const float pp[1] = {s};
    float tt[1];
    if (!fold_two_level_256<1>(pp, partials, gparts, counters, cols, red, &is_last, tt)) return;
    const int gc = blockIdx.x * 256 + threadIdx.x;
    float t = tt[0];
    if (accumulate) t += __bfloat162float(out[gc]);
    out[gc] = __float2bfloat16(t);
}

}  // namespace tony

using namespace tony;

constexpr int CS_MAX_GROUPS = 16;   // second fold level: 160 slabs -> groups of 13
static size_t colsum_partials_bytes(int cols) {
    return static_cast<size_t>(CS_MAX_SLABS_FUSED + CS_MAX_GROUPS) * static_cast<size_t>(cols) * sizeof(float);
}

extern "C" size_t tony_colsum_scratch_bytes(int cols) {
    // 4 KB of per-column-tile counters, then the partial rows
    return 4096 + colsum_partials_bytes(cols);
}

static int colsum_launch(const char* what, const void* x, const void* pre, void* gout, int act, int rows, int cols,
                         long long ld, void* out, float* scratch, size_t scratch_bytes, int accumulate, void* stream) {
    if (!x || !scratch || rows <= 0 || cols <= 0) return TONY_ERR_INVALID;
    if ((cols % 8) != 0 || (ld % 8) != 0 || (reinterpret_cast<uintptr_t>(x) & 15) ||
        (reinterpret_cast<uintptr_t>(pre) & 15) || (reinterpret_cast<uintptr_t>(gout) & 15)) {
        set_last_error(what, "cols and ld must be multiples of 8; matrices 16B aligned");
        return TONY_ERR_INVALID;
    }
    if (scratch_bytes < tony_colsum_scratch_bytes(cols)) {
        set_last_error(what, "scratch too small (see tony_colsum_scratch_bytes)");
        return TONY_ERR_INVALID;
    }
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const int col_tiles = (cols + 255) / 256;
    int slabs = std::max(1, std::min(CS_MAX_SLABS_FUSED, (device_sm_count() * 4) / col_tiles));
    slabs = std::min(slabs, std::max(1, rows / 64));
    const int rows_per_slab = (rows + slabs - 1) / slabs;
    slabs = (rows + rows_per_slab - 1) / rows_per_slab;
    // scratch layout: [4 KB of counters][partial rows][group rows].  The counters are cleared on the stream
    // before every call, so nothing depends on what the caller's scratch held before (the closing CTAs also
    // leave them at zero).
    if (fold_counter_words(col_tiles, slabs) > 1024) {
        set_last_error(what, "too many columns for the 4 KB of counters");
        return TONY_ERR_TOO_LARGE;
    }
    unsigned int* counters = reinterpret_cast<unsigned int*>(scratch);
    float* partials = scratch + 1024;
    float* gparts = partials + static_cast<size_t>(CS_MAX_SLABS_FUSED) * cols;
    if (clear_counters_each_call()) TONY_CUDA_CHECK(cudaMemsetAsync(counters, 0, 4096, s));
    const dim3 grid(col_tiles, slabs);
    const __nv_bfloat16* xx = static_cast<const __nv_bfloat16*>(x);
    const __nv_bfloat16* pp = static_cast<const __nv_bfloat16*>(pre);
    __nv_bfloat16* oo = static_cast<__nv_bfloat16*>(out);
    __nv_bfloat16* gg = static_cast<__nv_bfloat16*>(gout);
    switch (act) {
        case 0: TONY_CUDA_CHECK(launch_pdl_kernel(colsum_fused_kernel<0>, grid, dim3(CS_THREADS), 0, s, xx, rows, cols, ld, rows_per_slab, partials, gparts, counters, oo, accumulate, pp, gg)); break;
        case 1: TONY_CUDA_CHECK(launch_pdl_kernel(colsum_fused_kernel<1>, grid, dim3(CS_THREADS), 0, s, xx, rows, cols, ld, rows_per_slab, partials, gparts, counters, oo, accumulate, pp, gg)); break;
        case 2: TONY_CUDA_CHECK(launch_pdl_kernel(colsum_fused_kernel<2>, grid, dim3(CS_THREADS), 0, s, xx, rows, cols, ld, rows_per_slab, partials, gparts, counters, oo, accumulate, pp, gg)); break;
        default: TONY_CUDA_CHECK(launch_pdl_kernel(colsum_fused_kernel<3>, grid, dim3(CS_THREADS), 0, s, xx, rows, cols, ld, rows_per_slab, partials, gparts, counters, oo, accumulate, pp, gg)); break;
    }
    count_launch();
    TONY_CUDA_CHECK(cudaGetLastError());
    return TONY_OK;
}

extern "C" int tony_colsum_bf16(const void* x, int rows, int cols, long long ld, void* out,
                                float* scratch, size_t scratch_bytes, int accumulate,
                                void* stream) {
    if (!out) return TONY_ERR_INVALID;
    return colsum_launch("tony_colsum_bf16", x, nullptr, nullptr, 0, rows, cols, ld, out, scratch, scratch_bytes,
                         accumulate, stream);
}

extern "C" int tony_act_bwd_colsum_bf16(const void* dy, const void* pre, void* dpre, int act, int rows, int cols,
                                        long long ld, void* dbias, float* scratch, size_t scratch_bytes,
                                        int accumulate, void* stream) {
    if (!pre || !dpre || act < 1 || act > 3) return TONY_ERR_INVALID;
    return colsum_launch("tony_act_bwd_colsum_bf16", dy, pre, dpre, act, rows, cols, ld, dbias, scratch, scratch_bytes,
                         accumulate, stream);
}

extern "C" int tony_pack3_rows_colsum(const void* a, const void* b, const void* c, long long lda, long long ldb,
                                      long long ldc, void* packed, int rows, int H, void* dbias, float* scratch,
                                      size_t scratch_bytes, int accumulate, void* stream) {
    if (!a || !b || !c || !packed || !dbias || !scratch || rows <= 0 || H <= 0) return TONY_ERR_INVALID;
    if ((H % 256) || (lda % 8) || (ldb % 8) || (ldc % 8) ||
        ((reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(b) | reinterpret_cast<uintptr_t>(c) |
          reinterpret_cast<uintptr_t>(packed)) & 15)) {
        set_last_error("tony_pack3_rows_colsum", "H must be a multiple of 256, strides multiples of 8, pointers 16B aligned");
        return TONY_ERR_UNSUPPORTED;
    }
    const int cols = 3 * H;
    if (scratch_bytes < tony_colsum_scratch_bytes(cols)) {
        set_last_error("tony_pack3_rows_colsum", "scratch too small (see tony_colsum_scratch_bytes(3 * H))");
        return TONY_ERR_INVALID;
    }
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const int col_tiles = cols / 256;
    int slabs = std::max(1, std::min(CS_MAX_SLABS_FUSED, (device_sm_count() * 4) / col_tiles));
    slabs = std::min(slabs, std::max(1, rows / 64));
    const int rows_per_slab = (rows + slabs - 1) / slabs;
    slabs = (rows + rows_per_slab - 1) / rows_per_slab;
    if (fold_counter_words(col_tiles, slabs) > 1024) return TONY_ERR_TOO_LARGE;
    unsigned int* counters = reinterpret_cast<unsigned int*>(scratch);
    float* partials = scratch + 1024;
    float* gparts = partials + static_cast<size_t>(CS_MAX_SLABS_FUSED) * cols;
    if (clear_counters_each_call()) TONY_CUDA_CHECK(cudaMemsetAsync(counters, 0, 4096, s));
    TONY_CUDA_CHECK(launch_pdl_kernel(pack3_colsum_kernel, dim3(col_tiles, slabs), dim3(CS_THREADS), 0, s,
                                      static_cast<const __nv_bfloat16*>(a), static_cast<const __nv_bfloat16*>(b),
                                      static_cast<const __nv_bfloat16*>(c), lda, ldb, ldc, H, rows, rows_per_slab,
                                      static_cast<__nv_bfloat16*>(packed), partials, gparts, counters,
                                      static_cast<__nv_bfloat16*>(dbias), accumulate));
    count_launch();
    return TONY_OK;
}

// ---------------------------------------------------------------------------------------
// Allocation-free variant.
//
// Who needs it: a host program whose GPU memory is already spoken for.  TonY launches
// TensorFlow, MXNet and Horovod jobs as well as PyTorch ones, and several of those
// runtimes reserve almost the whole device for their own pool the moment they start.
// A helper library loaded into such a process (ctypes / dlopen) that then calls
// cudaMalloc for its scratch gets cudaErrorMemoryAllocation — there is nothing left
// outside the pool — and it has no way to ask the framework's pool for memory through a
// C ABI.  The same holds for callers that are forbidden to allocate on their hot path
// because cudaMalloc synchronises the device.  For them the partial-sum scratch lives in
// the library image itself: a file-scope device array that every GPU which loads this
// library carries until the process exits.
//
// Limit: TONY_COLSUM_STATIC_MAX_COLS columns (scratch = 64 slabs x 8192 cols x 4 B = 2 MiB
// per GPU).  Wider matrices are refused with TONY_ERR_TOO_LARGE; use tony_colsum_bf16 with
// caller-provided scratch for those.
//
// Sharing: there is ONE array per GPU and every call uses it.  It still holds whatever
// the previous call left, but nothing here depends on a zeroed start: the first kernel
// overwrites every word the second kernel goes on to read, so no clearing is needed.
// Two calls that overlap on the same GPU (different streams, or different host threads
// on different streams) would write the same words and race; keep calls to this entry
// point ordered on one stream per GPU.
//
// Set-up: in host code the array's name is only a stand-in, not an address a kernel can
// use, so the first call on each device asks the runtime for the real device address
// (cudaGetSymbolAddress, with that device current) and caches it.  That first call does
// a little more work than later ones; make it before you begin capturing a stream.
namespace tony {

static __device__ __align__(256) float g_colsum_static_scratch[CS_MAX_SLABS * TONY_COLSUM_STATIC_MAX_COLS];

static std::mutex g_colsum_static_mu;
static float* g_colsum_static_ptr[64] = {nullptr};

// Device address of the static scratch on the current device (resolved once per device).
static int colsum_static_scratch(float** out) {
    int dev = 0;
    TONY_CUDA_CHECK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) {
        set_last_error("tony_colsum_bf16_static", "device ordinal out of range");
        return TONY_ERR_UNSUPPORTED;
    }
    std::lock_guard<std::mutex> lock(g_colsum_static_mu);
    if (g_colsum_static_ptr[dev] == nullptr) {
        float* p = nullptr;
        TONY_CUDA_CHECK(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_colsum_static_scratch));
        g_colsum_static_ptr[dev] = p;
    }
    *out = g_colsum_static_ptr[dev];
    return TONY_OK;
}

}  // namespace tony

extern "C" int tony_colsum_bf16_static(const void* x, int rows, int cols, long long ld, void* out,
                                       int accumulate, void* stream) {
    if (!x || !out || rows <= 0 || cols <= 0) return TONY_ERR_INVALID;
    if ((cols % 8) != 0 || (ld % 8) != 0 || (reinterpret_cast<uintptr_t>(x) & 15)) {
        set_last_error("tony_colsum_bf16_static", "cols and ld must be multiples of 8; x 16B aligned");
        return TONY_ERR_INVALID;
    }
    if (cols > TONY_COLSUM_STATIC_MAX_COLS) {
        set_last_error("tony_colsum_bf16_static",
                       "cols exceeds TONY_COLSUM_STATIC_MAX_COLS (8192): the built-in scratch is 2 MiB; "
                       "call tony_colsum_bf16 with your own scratch instead");
        return TONY_ERR_TOO_LARGE;
    }
    float* scratch = nullptr;
    int rc = colsum_static_scratch(&scratch);
    if (rc) return rc;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const int col_tiles = (cols + 255) / 256;
    int slabs = std::max(1, std::min(CS_MAX_SLABS, (device_sm_count() * 4) / col_tiles));
    slabs = std::min(slabs, std::max(1, rows / 32));
    const int rows_per_slab = (rows + slabs - 1) / slabs;
    slabs = (rows + rows_per_slab - 1) / rows_per_slab;
    colsum_partial_kernel<<<dim3(col_tiles, slabs), CS_THREADS, 0, s>>>(
        static_cast<const __nv_bfloat16*>(x), rows, cols, ld, rows_per_slab, scratch);
    TONY_CUDA_CHECK(cudaGetLastError());
    colsum_finalize_kernel<<<(cols + 255) / 256, 256, 0, s>>>(
        scratch, slabs, cols, static_cast<__nv_bfloat16*>(out), accumulate);
    count_launch(2);
    TONY_CUDA_CHECK(cudaGetLastError());
    return TONY_OK;
}
