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

// Single-launch form used by tony_colsum_bf16: every CTA publishes its partial row, the last
// CTA of each column tile (found with a counter that it resets to zero again) folds the
// slabs and writes the bf16 result.  Saves one ~6 us launch per bias gradient (121 per
// BERT-large step).  Counters live in the first 4 KB of the caller's scratch; the host clears
// them on the stream before each launch.
__global__ void __launch_bounds__(CS_THREADS)
colsum_fused_kernel(const __nv_bfloat16* __restrict__ x, int rows, int cols, long long ld,
                    int rows_per_slab, float* __restrict__ partials, unsigned int* __restrict__ counters,
                    __nv_bfloat16* __restrict__ out, int accumulate) {
    __shared__ float red[CS_THREADS / 32][256];
    __shared__ unsigned int is_last;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int col = blockIdx.x * 256 + lane * 8;
    const int r0 = blockIdx.y * rows_per_slab;
    const int r1 = min(rows, r0 + rows_per_slab);
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    if (col < cols) {
        // eight rows in flight per warp; the last batch is predicated instead of falling back to a
        // one-row-at-a-time loop (each of those iterations was a full memory round trip)
        constexpr int W = CS_THREADS / 32;
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
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) red[warp][lane * 8 + j] = acc[j];
    __syncthreads();
    const int c = threadIdx.x;
    float s = 0.f;
#pragma unroll
    for (int w = 0; w < CS_THREADS / 32; ++w) s += red[w][c];
    const int gc = blockIdx.x * 256 + c;
    if (gc < cols) partials[static_cast<size_t>(blockIdx.y) * cols + gc] = s;
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned int prev = atomicAdd(&counters[blockIdx.x], 1u);
        is_last = (prev == gridDim.y - 1) ? 1u : 0u;
    }
    __syncthreads();
    if (!is_last) return;
    __threadfence();
    float t = fold_partials_256(partials, static_cast<size_t>(cols), static_cast<int>(gridDim.y),
                                blockIdx.x * 256, cols, red);
    if (gc < cols) {
        if (accumulate) t += __bfloat162float(out[gc]);
        out[gc] = __float2bfloat16(t);
    }
    if (threadIdx.x == 0) counters[blockIdx.x] = 0;
}

}  // namespace tony

using namespace tony;

static size_t colsum_partials_bytes(int cols) {
    return static_cast<size_t>(CS_MAX_SLABS_FUSED) * static_cast<size_t>(cols) * sizeof(float);
}

extern "C" size_t tony_colsum_scratch_bytes(int cols) {
    // 4 KB of per-column-tile counters, then the partial rows
    return 4096 + colsum_partials_bytes(cols);
}

extern "C" int tony_colsum_bf16(const void* x, int rows, int cols, long long ld, void* out,
                                float* scratch, size_t scratch_bytes, int accumulate,
                                void* stream) {
    if (!x || !out || !scratch || rows <= 0 || cols <= 0) return TONY_ERR_INVALID;
    if ((cols % 8) != 0 || (ld % 8) != 0 || (reinterpret_cast<uintptr_t>(x) & 15)) {
        set_last_error("tony_colsum_bf16", "cols and ld must be multiples of 8; x 16B aligned");
        return TONY_ERR_INVALID;
    }
    if (scratch_bytes < tony_colsum_scratch_bytes(cols)) {
        set_last_error("tony_colsum_bf16", "scratch too small (see tony_colsum_scratch_bytes)");
        return TONY_ERR_INVALID;
    }
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const int col_tiles = (cols + 255) / 256;
    int slabs = std::max(1, std::min(CS_MAX_SLABS_FUSED, (device_sm_count() * 4) / col_tiles));
    slabs = std::min(slabs, std::max(1, rows / 64));
    const int rows_per_slab = (rows + slabs - 1) / slabs;
    slabs = (rows + rows_per_slab - 1) / rows_per_slab;
    // scratch layout: [4 KB of per-column-tile counters][partial rows].  The counters are
    // cleared on the stream before every call, so nothing depends on what the caller's
    // scratch held before (the last CTA also leaves them at zero).
    if (col_tiles > 1024) {
        set_last_error("tony_colsum_bf16", "more than 262144 columns");
        return TONY_ERR_TOO_LARGE;
    }
    unsigned int* counters = reinterpret_cast<unsigned int*>(scratch);
    float* partials = scratch + 1024;
    TONY_CUDA_CHECK(cudaMemsetAsync(counters, 0, static_cast<size_t>(col_tiles) * 4, s));
    colsum_fused_kernel<<<dim3(col_tiles, slabs), CS_THREADS, 0, s>>>(
        static_cast<const __nv_bfloat16*>(x), rows, cols, ld, rows_per_slab, partials, counters,
        static_cast<__nv_bfloat16*>(out), accumulate);
    count_launch();
    TONY_CUDA_CHECK(cudaGetLastError());
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
