// tony_b200 — implicit-GEMM 2-D convolution on tcgen05 tensor cores (K8a of SURVEY.md §8), sm_100a only.
//
// Activations are NHWC bf16 (a channels-last tensor), filters KRSC bf16 ([Cout][R][S][Cin]).  No im2col
// matrix is ever materialised: the TMA unit's im2col mode gathers, for one filter tap, the 128 (or 64)
// input pixels that a tile of consecutive output pixels needs — padding, stride and image borders are
// resolved by the tensor map (out-of-bounds pixels arrive as zeros) — straight into the 128-byte
// swizzled shared-memory layout tcgen05.mma reads.  One persistent warp-specialised kernel serves
//
//   fprop  Y[pix, co]  = sum_{tap, ci}  X[pix (+) tap, ci] * W[co, tap, ci]
//          A = im2col(X) K-major, B = W K-major, K = taps x Cin
//   dgrad  dX[pix, ci] = sum_{tap, co} dY[pix (-) tap, co] * W[co, tap, ci]
//          A = im2col(dY) K-major, B = W MN-major (ci contiguous, no transposed copy of the filter);
//          a stride-2 layer runs one launch per output parity class (each class is a small
//          stride-1 convolution over dY), writing its rows at stride 2
//   wgrad  dW[co, tap, ci] = sum_{pix} dY[pix, co] * X[pix (+) tap, ci]
//          A = dY MN-major (plain 2-D TMA), B = im2col(X) MN-major, K = output pixels, split-K
//
// Roles per CTA (576 threads): warp 0 TMA producer, warp 1 TMEM allocator + MMA issuer, warps 2-17
// epilogue (TMEM -> registers -> swizzled smem transpose -> 16-byte global stores).  The fp32
// accumulator is double-buffered in TMEM so the epilogue of tile i overlaps the MMAs of tile i+1.
// The forward epilogue can also accumulate the per-channel sum and sum of squares of the bf16 output
// (BatchNorm batch statistics) so the following BatchNorm needs no statistics pass over Y.
//
// The reference (zuston/TonY) has no device code; this replaces the cuDNN convolutions of the
// CNN jobs it launches (tony-examples/horovod-on-tony/tensorflow2_mnist.py:48-49,
// tony-examples/mnist-tensorflow/mnist_distributed.py:87-97).
#include "common.cuh"
#include "tony_b200_kernels.h"

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <mutex>

namespace tony {

constexpr int CV_BLOCK_M = 128;
constexpr int CV_BLOCK_K = 64;
constexpr int CV_UMMA_K = 16;
constexpr int CV_THREADS = 576;
constexpr int CV_EPI_WARPS = 16;
constexpr int CV_STAGING_BYTES = CV_EPI_WARPS * 2048;
constexpr int CV_CHUNK_BYTES = 64 * 128;   // [64 rows][128 B]
constexpr int CV_MAX_TAPS = 16;

// FastDiv / fdiv (division by launch-time constants as multiply + shift) live in host_math.h

struct ConvTap {
    unsigned short off_w, off_h;   // im2col offsets from the tile's base pixel
    int wsel;                      // tap slot (r * S + s) inside the filter
};

struct ConvParams {
    int M, N;
    int num_m_blocks, num_n_blocks, num_k_blocks, splits, kb_per_split;
    int b_mn_major;                // MODE 0: filter operand K-major (fprop) or MN-major (dgrad)
    int ntaps, cblocks;            // 64-channel blocks per tap of the im2col'd tensor
    int w_tap_stride;              // elements between two taps along the filter's inner dimension
    int P, Q;                      // positions the im2col traversal visits per image (rows, columns)
    int lower_w, lower_h, stride_w, stride_h;
    ConvTap taps[CV_MAX_TAPS];
    int out_strided, OH, OW, osh, osw, oph, opw;   // MODE 0 output row of pixel (n, p, q)
    __nv_bfloat16* C;
    long long ldc;
    uint32_t flags;
    float* workspace;              // split-K partials [splits][M][N]
    float* stat_sum;               // optional [N]: += column sums of the bf16 output
    float* stat_sqsum;             // optional [N]: += column sums of squares
    int dbg;                       // timing experiments only (TONY_B200_CONV_DBG): 1 = skip A loads, 2 = skip B loads
    // MODE 1, tiled form: a k-block is a bw x bh patch of output pixels of one image (bw * bh = 64) instead of
    // 64 consecutive pixels, and both operands come through plain 4-D tiled maps - the filter tap is a
    // coordinate shift and the padding is the map's out-of-range zero fill.  An experiment kept as an option:
    // it is correct but not faster than the im2col form (see tony_conv2d_wgrad_bf16).
    int tiled, bw, bh, nbw, nbh, pad_w, pad_h;
    // halo form (conv3x3_halo_tcgen05_kernel): a tile is a 16 x 8 patch of one image's output pixels
    int halo, tiles_x, tiles_y, b_resident;
    FastDiv fd_splits, fd_nblk, fd_tx, fd_ty, fd_pq, fd_q;   // filled by fill_fastdiv() right before the launch
};
static inline void fill_fastdiv(ConvParams& p) {
    p.fd_splits = make_fastdiv(p.splits);
    p.fd_nblk = make_fastdiv(p.num_n_blocks);
    p.fd_tx = make_fastdiv(p.tiles_x);
    p.fd_ty = make_fastdiv(p.tiles_y);
    p.fd_pq = make_fastdiv(p.P * p.Q);
    p.fd_q = make_fastdiv(p.Q);
}

__device__ __forceinline__ void tma_load_im2col_4d(const CUtensorMap* map, uint64_t* bar, void* smem_dst,
                                                   int32_t c, int32_t w, int32_t h, int32_t n,
                                                   uint16_t off_w, uint16_t off_h) {
    // the two 16-bit offsets enter as 32-bit operands and are narrowed inside the asm block: as "h" operands
    // fed from the uniform datapath they came out wrong
    asm volatile(
        "{\n\t"
        ".reg .b16 ow, oh;\n\t"
        "cvt.u16.u32 ow, %7;\n\t"
        "cvt.u16.u32 oh, %8;\n\t"
        "cp.async.bulk.tensor.4d.shared::cluster.global.im2col.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%3, %4, %5, %6}], [%2], {ow, oh};\n\t"
        "}\n"
        :
        : "r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c), "r"(w),
          "r"(h), "r"(n), "r"(static_cast<uint32_t>(off_w)), "r"(static_cast<uint32_t>(off_h))
        : "memory");
}

__device__ __forceinline__ uint32_t cv_stg_addr(uint32_t base, int row, int unit) {
    return base + row * 64 + ((unit ^ ((row >> 1) & 3)) << 4);
}

// pixel index -> row index of the output tensor (rows of `ldc` elements)
__device__ __forceinline__ long long cv_out_row(const ConvParams& p, int m) {
    if (p.halo) {      // row (y, x) = (loc / 8, loc % 8) of tile (img, ty, tx); rows past the image are dropped
        const int mt = m >> 7, loc = m & 127;
        const int t2 = fdiv(mt, p.fd_tx), tx = mt - t2 * p.tiles_x;
        const int img = fdiv(t2, p.fd_ty), ty = t2 - img * p.tiles_y;
        const int y = ty * 16 + (loc >> 3), x = tx * 8 + (loc & 7);
        if (y >= p.P || x >= p.Q) return -1;
        return (static_cast<long long>(img) * p.P + y) * p.Q + x;
    }
    if (!p.out_strided) return m;
    const int pq = p.P * p.Q;
    const int n = fdiv(m, p.fd_pq);
    const int rem = m - n * pq;
    const int pp = fdiv(rem, p.fd_q);
    const int qq = rem - pp * p.Q;
    return (static_cast<long long>(n) * p.OH + pp * p.osh + p.oph) * p.OW + qq * p.osw + p.opw;
}

// Epilogue role (warps 2..17 of both convolution kernels): TMEM -> registers -> per-warp transposed staging tile
// -> 16-byte global stores of whole rows, with the optional accumulate, split-K partial and BatchNorm-statistics
// paths.  `staging` is CV_STAGING_BYTES of shared memory, `stat_acc` 2 * BLOCK_N floats (zeroed by the caller).
template <int BLOCK_N, int CS>
__device__ __forceinline__ void cv_epilogue_role(const ConvParams& p, uint8_t* staging, float* stat_acc,
                                                 uint32_t tmem_base, uint64_t* tmem_full_bar,
                                                 uint64_t* tmem_empty_bar, int worker, int num_workers, int num_units,
                                                 uint32_t crank, int warp_idx, int lane) {
    const int quarter = warp_idx & 3;
    const int group = (warp_idx - 2) >> 2;
    const uint32_t stg = smem_u32(staging + (warp_idx - 2) * 2048);
    const int unit16 = lane & 3;
    const int etid = threadIdx.x - 64;     // 0..511 among the epilogue warps
    int stat_nblk = -1;
    // one column of BatchNorm partials per lane and 32-column chunk this warp owns (see the epilogue loop)
    constexpr int STAT_SLOTS = (BLOCK_N / 32 + CV_EPI_WARPS / 4 - 1) / (CV_EPI_WARPS / 4);
    static_assert(STAT_SLOTS <= 2, "the slot select below assumes at most two chunks per warp");
    float acc1[STAT_SLOTS], acc2[STAT_SLOTS];
#pragma unroll
    for (int i = 0; i < STAT_SLOTS; ++i) acc1[i] = acc2[i] = 0.f;
    // the four quarter warps of a group own the same columns: meet in shared memory, once per column block
    auto stat_fold = [&]() {
#pragma unroll
        for (int i = 0; i < STAT_SLOTS; ++i) {
            const int lc = (group + i * (CV_EPI_WARPS / 4)) * 32 + unit16 * 8 + (lane >> 2);
            if (lc < BLOCK_N) {
                atomicAdd(stat_acc + lc, acc1[i]);
                atomicAdd(stat_acc + BLOCK_N + lc, acc2[i]);
            }
            acc1[i] = acc2[i] = 0.f;
        }
    };
    int iter = 0;
    for (int unit = worker; unit < num_units; unit += num_workers, ++iter) {
        const int tile = fdiv(unit, p.fd_splits);
        const int split = unit - tile * p.splits;
        const int m_grp = fdiv(tile, p.fd_nblk);
        const int m_blk = m_grp * CS + static_cast<int>(crank);
        const int n_blk = tile - m_grp * p.num_n_blocks;
        if (p.stat_sum != nullptr && n_blk != stat_nblk) {
            if (stat_nblk >= 0) {
                // this CTA moves on to another column block: hand the finished partials over
                stat_fold();
                asm volatile("bar.sync 1, 512;" ::: "memory");
                for (int i = etid; i < 2 * BLOCK_N; i += 512) {
                    const int col = stat_nblk * BLOCK_N + (i % BLOCK_N);
                    if (col < p.N) atomicAdd((i < BLOCK_N ? p.stat_sum : p.stat_sqsum) + col, stat_acc[i]);
                    stat_acc[i] = 0.f;
                }
                asm volatile("bar.sync 1, 512;" ::: "memory");
            }
            stat_nblk = n_blk;
        }
        const int as = iter & 1;
        const uint32_t aphase = (iter >> 1) & 1;
        const int row0 = m_blk * CV_BLOCK_M + quarter * 32;
        // the four output rows this lane stores (and, for ACCUMULATE, loads) in every 32-column chunk
        long long orow[4];
        if (p.halo) {
            // a tile is one 16 x 8 patch: its image / patch position once per tile (warp-uniform divisions), the
            // row inside it by shifts.  (cv_out_row per row - four divisions each - made the epilogue the limit.)
            const int t2 = fdiv(m_blk, p.fd_tx), tx = m_blk - t2 * p.tiles_x;
            const int img = fdiv(t2, p.fd_ty), ty = t2 - img * p.tiles_y;
            const long long base = static_cast<long long>(img) * p.P * p.Q;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int loc = quarter * 32 + i * 8 + (lane >> 2);
                const int y = ty * 16 + (loc >> 3), x = tx * 8 + (loc & 7);
                orow[i] = (y < p.P && x < p.Q) ? base + static_cast<long long>(y) * p.Q + x : -1;
            }
        } else {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int m = row0 + i * 8 + (lane >> 2);
                orow[i] = m < p.M ? cv_out_row(p, m) : -1;
            }
        }
        mbar_wait(&tmem_full_bar[as], aphase);
        tcgen05_fence_after();
        const uint32_t taddr0 = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + as * BLOCK_N;
#pragma unroll 1
        for (int c = group; c < BLOCK_N / 32; c += CV_EPI_WARPS / 4) {
            uint32_t r[32];
            tmem_ld_32x32(taddr0 + c * 32, r);
            tmem_ld_wait();
            const int col0 = n_blk * BLOCK_N + c * 32;
            if (col0 >= p.N || row0 >= p.M) continue;   // warp-uniform
            if (p.splits > 1) {
                const int row = row0 + lane;
                if (row < p.M) {
                    float* dst = p.workspace + (static_cast<size_t>(split) * p.M + row) * p.N + col0;
#pragma unroll
                    for (int g = 0; g < 8; ++g) {
                        if (col0 + g * 4 < p.N)
                            *reinterpret_cast<float4*>(dst + g * 4) =
                                make_float4(__uint_as_float(r[g * 4]), __uint_as_float(r[g * 4 + 1]),
                                            __uint_as_float(r[g * 4 + 2]), __uint_as_float(r[g * 4 + 3]));
                    }
                }
                continue;
            }
            const int gc = col0 + unit16 * 8;
            if (p.flags & TONY_EPI_ACCUMULATE) {
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    uint4 v = make_uint4(0, 0, 0, 0);
                    if (orow[i] >= 0 && gc < p.N) v = ld_global_v4(p.C + orow[i] * p.ldc + gc);
                    st_shared_v4(cv_stg_addr(stg, i * 8 + (lane >> 2), unit16), v);
                }
                __syncwarp();
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const uint4 v = ld_shared_v4(cv_stg_addr(stg, lane, u));
                    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const float2 f = unpack_bf16x2(w[j]);
                        r[u * 8 + 2 * j] = __float_as_uint(__uint_as_float(r[u * 8 + 2 * j]) + f.x);
                        r[u * 8 + 2 * j + 1] = __float_as_uint(__uint_as_float(r[u * 8 + 2 * j + 1]) + f.y);
                    }
                }
                __syncwarp();
            }
            uint32_t o[16];
#pragma unroll
            for (int j = 0; j < 16; ++j)
                o[j] = pack_bf16x2(__uint_as_float(r[2 * j]), __uint_as_float(r[2 * j + 1]));
#pragma unroll
            for (int u = 0; u < 4; ++u)
                st_shared_v4(cv_stg_addr(stg, lane, u), make_uint4(o[4 * u], o[4 * u + 1], o[4 * u + 2], o[4 * u + 3]));
            __syncwarp();
            // stores first, then - under ONE warp-uniform branch - the statistics of what was stored: with the
            // test inside the store loop the compiler predicated the 48 unpack / add / fma instructions per chunk
            // and every launch without statistics (all data gradients) issued them for nothing
            uint4 sv[4];
            bool ok[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                sv[i] = ld_shared_v4(cv_stg_addr(stg, i * 8 + (lane >> 2), unit16));
                ok[i] = orow[i] >= 0 && gc < p.N;
                if (ok[i]) st_global_v4(p.C + orow[i] * p.ldc + gc, sv[i]);
            }
            float s1[8], s2[8];
            if (p.stat_sum != nullptr) {
#pragma unroll
                for (int j = 0; j < 8; ++j) s1[j] = s2[j] = 0.f;
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    if (ok[i]) {
                        const uint32_t w[4] = {sv[i].x, sv[i].y, sv[i].z, sv[i].w};
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            const float2 f = unpack_bf16x2(w[j]);
                            s1[2 * j] += f.x;
                            s1[2 * j + 1] += f.y;
                            s2[2 * j] = fmaf(f.x, f.x, s2[2 * j]);
                            s2[2 * j + 1] = fmaf(f.y, f.y, s2[2 * j + 1]);
                        }
                    }
                }
            }
            __syncwarp();
            if (p.stat_sum != nullptr) {
                // lanes with equal (lane & 3) hold the same 8 columns for different rows.  Fold the 8 row groups
                // with a halving exchange (7 shuffles per array, not 24): lane (rg, unit16) ends with the warp's
                // sum for column unit16 * 8 + rg and keeps it in a register until this CTA leaves the column block
                const int rg = lane >> 2;
                float k1[4], k2[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const float snd1 = (rg & 4) ? s1[j] : s1[j + 4], snd2 = (rg & 4) ? s2[j] : s2[j + 4];
                    k1[j] = ((rg & 4) ? s1[j + 4] : s1[j]) + __shfl_xor_sync(0xffffffffu, snd1, 16);
                    k2[j] = ((rg & 4) ? s2[j + 4] : s2[j]) + __shfl_xor_sync(0xffffffffu, snd2, 16);
                }
#pragma unroll
                for (int j = 0; j < 2; ++j) {
                    const float snd1 = (rg & 2) ? k1[j] : k1[j + 2], snd2 = (rg & 2) ? k2[j] : k2[j + 2];
                    k1[j] = ((rg & 2) ? k1[j + 2] : k1[j]) + __shfl_xor_sync(0xffffffffu, snd1, 8);
                    k2[j] = ((rg & 2) ? k2[j + 2] : k2[j]) + __shfl_xor_sync(0xffffffffu, snd2, 8);
                }
                const float snd1 = (rg & 1) ? k1[0] : k1[1], snd2 = (rg & 1) ? k2[0] : k2[1];
                const float f1 = ((rg & 1) ? k1[1] : k1[0]) + __shfl_xor_sync(0xffffffffu, snd1, 4);
                const float f2 = ((rg & 1) ? k2[1] : k2[0]) + __shfl_xor_sync(0xffffffffu, snd2, 4);
                if (STAT_SLOTS == 1 || c == group) {
                    acc1[0] += f1;
                    acc2[0] += f2;
                } else {
                    acc1[STAT_SLOTS - 1] += f1;
                    acc2[STAT_SLOTS - 1] += f2;
                }
            }
        }
        tcgen05_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&tmem_empty_bar[as]);
    }
    if (p.stat_sum != nullptr && stat_nblk >= 0) {
        stat_fold();
        asm volatile("bar.sync 1, 512;" ::: "memory");
        for (int i = etid; i < 2 * BLOCK_N; i += 512) {
            const int col = stat_nblk * BLOCK_N + (i % BLOCK_N);
            if (col < p.N) atomicAdd((i < BLOCK_N ? p.stat_sum : p.stat_sqsum) + col, stat_acc[i]);
        }
    }
}

// CS > 1 (MODE 0 only): CS CTAs of a cluster work on CS consecutive pixel tiles of the same column block in
// lockstep and share the filter operand: every CTA fetches 1/CS of each B tile and the TMA unit multicasts
// it into all CS shared memories, so the filter crosses L2 -> SM once per cluster instead of once per CTA
// (at 256 output channels the filter is 2/3 of a CTA's operand traffic, and that traffic - not the MMAs -
// bounds the kernel).  A slot is refilled only after every CTA of the cluster has consumed it: each MMA
// thread's tcgen05.commit arrives on the `empty` barrier of all CS CTAs.
template <int BLOCK_N, int STAGES, int MODE, int CS>
__global__ void __launch_bounds__(CV_THREADS, 1)
conv_bf16_tcgen05_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                         const ConvParams p) {
    static_assert(CS == 1 || MODE == 0, "operand multicast is implemented for the fprop / dgrad form");
    constexpr uint32_t A_BYTES = CV_BLOCK_M * CV_BLOCK_K * 2;
    constexpr uint32_t B_BYTES = BLOCK_N * CV_BLOCK_K * 2;
    constexpr uint32_t STAGE_BYTES = A_BYTES + B_BYTES;
    constexpr uint32_t TMEM_COLS = (2 * BLOCK_N < 32) ? 32 : 2 * BLOCK_N;
    constexpr int N_CHUNKS = BLOCK_N / 64;

    // 128-byte swizzled operand tiles need a 1024-byte aligned base.  The dynamic shared-memory window of a
    // kernel without static shared memory starts 1024-aligned; there is no room left to round up by hand
    // (the 256-column variant uses 226.25 of the 227 KB), so a misaligned base traps instead.
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    if ((smem_u32(smem_raw) & 1023u) != 0) __trap();
    uint8_t* smem = smem_raw;
    uint8_t* staging = smem + STAGES * STAGE_BYTES;
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(staging + CV_STAGING_BYTES);
    uint64_t* empty_bar = full_bar + STAGES;
    uint64_t* tmem_full_bar = empty_bar + STAGES;
    uint64_t* tmem_empty_bar = tmem_full_bar + 2;
    uint32_t* tmem_base_slot = reinterpret_cast<uint32_t*>(tmem_empty_bar + 2);
    // per-CTA partial BatchNorm statistics [2][BLOCK_N]: folded over every tile this CTA produces for
    // one column block and added to global memory once (one atomic per column per CTA, not per warp)
    float* stat_acc = reinterpret_cast<float*>(tmem_base_slot + 4);

    const int warp_idx = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);   // provably warp-uniform: role branches and their loops stay on the uniform datapath
    const int lane = threadIdx.x & 31;

    if (p.stat_sum != nullptr)
        for (int i = threadIdx.x; i < 2 * BLOCK_N; i += CV_THREADS) stat_acc[i] = 0.f;
    if (warp_idx == 0 && lane == 0) {
        tma_prefetch_desc(&tmap_a);
        tma_prefetch_desc(&tmap_b);
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], CS);      // one tcgen05.commit arrival per CTA of the cluster
        }
        for (int s = 0; s < 2; ++s) {
            mbar_init(&tmem_full_bar[s], 1);
            mbar_init(&tmem_empty_bar[s], CV_EPI_WARPS);
        }
        fence_barrier_init();
    }
    if (warp_idx == 1) tmem_alloc<TMEM_COLS>(tmem_base_slot);
    tcgen05_fence_before();
    if (CS > 1) cluster_sync_all(); else __syncthreads();   // peers' barriers exist before any multicast
    tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_base_slot;
    pdl_launch_dependents();

    // Work units.  CS == 1: unit = (tile, split), tile = m_blk * num_n + n_blk, CTA-strided.  CS > 1: a unit
    // is a GROUP of CS consecutive m-blocks of one column block; the cluster's CTAs take one m-block each
    // (a CTA whose m-block lies past the end still loads its share of B and runs the barriers).
    const uint32_t crank = CS > 1 ? cluster_ctarank() : 0u;
    const int num_tiles = CS > 1 ? ((p.num_m_blocks + CS - 1) / CS) * p.num_n_blocks : p.num_m_blocks * p.num_n_blocks;
    const int num_units = num_tiles * p.splits;
    const int worker = static_cast<int>(blockIdx.x) / CS;
    const int num_workers = static_cast<int>(gridDim.x) / CS;
    const int pq = p.P * p.Q;
    constexpr uint16_t MC_MASK = static_cast<uint16_t>((1u << CS) - 1u);

    if (warp_idx == 0) {
        // ---------------------------------------------------------------- TMA producer
        // (whole warp in the loop, one elected lane issues: see the note at the MMA issuer)
        {
            int stage = 0;
            uint32_t phase = 0;
            pdl_wait();
            for (int unit = worker; unit < num_units; unit += num_workers) {
                const int tile = fdiv(unit, p.fd_splits);
                const int split = unit - tile * p.splits;
                const int m_grp = fdiv(tile, p.fd_nblk);
                const int m_blk = m_grp * CS + static_cast<int>(crank);
                const int n_blk = tile - m_grp * p.num_n_blocks;
                const int kb0 = split * p.kb_per_split;
                const int kb1 = min(p.num_k_blocks, kb0 + p.kb_per_split);
                if (MODE == 0) {
                    const int m0 = m_blk * CV_BLOCK_M;
                    const int img = fdiv(m0, p.fd_pq);
                    const int rem = m0 - img * pq;
                    const int p0 = fdiv(rem, p.fd_q);
                    const int q0 = rem - p0 * p.Q;
                    const int cw = p.lower_w + q0 * p.stride_w;
                    const int ch = p.lower_h + p0 * p.stride_h;
                    int t = kb0 / p.cblocks;
                    int cc = kb0 - t * p.cblocks;
                    for (int kb = kb0; kb < kb1; ++kb) {
                        mbar_wait(&empty_bar[stage], phase ^ 1);
                        uint8_t* sa = smem + stage * STAGE_BYTES;
                        uint8_t* sb = sa + A_BYTES;
                        const ConvTap tap = p.taps[t];
                        const int winner = tap.wsel * p.w_tap_stride;
                        if (elect_one()) {
                        if (p.dbg) {   // wrong results on purpose: which operand's loads bound the kernel?
                            const bool la = !(p.dbg & 1), lb = !(p.dbg & 2);
                            mbar_expect_tx(&full_bar[stage], (la ? A_BYTES : 0) + (lb ? B_BYTES : 0));
                            if (la) tma_load_im2col_4d(&tmap_a, &full_bar[stage], sa, cc * 64, cw, ch, img, tap.off_w, tap.off_h);
                            if (lb) tma_load_2d(&tmap_b, &full_bar[stage], sb, winner + cc * 64, n_blk * BLOCK_N);
                        } else {
                        mbar_expect_tx(&full_bar[stage], STAGE_BYTES);
                        tma_load_im2col_4d(&tmap_a, &full_bar[stage], sa, cc * 64, cw, ch, img, tap.off_w, tap.off_h);
                        if (CS == 1) {
                            if (!p.b_mn_major) {
                                tma_load_2d(&tmap_b, &full_bar[stage], sb, winner + cc * 64, n_blk * BLOCK_N);
                            } else {
#pragma unroll
                                for (int c = 0; c < N_CHUNKS; ++c)
                                    tma_load_2d(&tmap_b, &full_bar[stage], sb + c * CV_CHUNK_BYTES,
                                                winner + n_blk * BLOCK_N + c * 64, cc * 64);
                            }
                        } else if (!p.b_mn_major) {
                            // K-major filter tile [BLOCK_N rows][64 k]: this CTA's BLOCK_N / CS rows, to everyone
                            constexpr int ROWS = BLOCK_N / CS;
                            tma_load_2d_multicast(&tmap_b, &full_bar[stage], sb + crank * ROWS * 128, winner + cc * 64,
                                                  n_blk * BLOCK_N + static_cast<int>(crank) * ROWS, MC_MASK);
                        } else if (N_CHUNKS >= CS) {
                            // MN-major filter tile: whole [64 k][64 n] chunks dealt round-robin to the CTAs
#pragma unroll
                            for (int c = 0; c < N_CHUNKS / CS; ++c) {
                                const int ch = c * CS + static_cast<int>(crank);
                                tma_load_2d_multicast(&tmap_b, &full_bar[stage], sb + ch * CV_CHUNK_BYTES,
                                                      winner + n_blk * BLOCK_N + ch * 64, cc * 64, MC_MASK);
                            }
                        } else {
                            // fewer chunks than CTAs: each chunk's 64 k-rows are cut into CS / N_CHUNKS parts
                            constexpr int PPC = (CS / (N_CHUNKS > 0 ? N_CHUNKS : 1)) > 0 ? CS / (N_CHUNKS > 0 ? N_CHUNKS : 1) : 1;
                            constexpr int ROWS = 64 / PPC;
                            const int ch = static_cast<int>(crank) / PPC;
                            const int part = static_cast<int>(crank) % PPC;
                            tma_load_2d_multicast(&tmap_b, &full_bar[stage],
                                                  sb + ch * CV_CHUNK_BYTES + part * ROWS * 128,
                                                  winner + n_blk * BLOCK_N + ch * 64, cc * 64 + part * ROWS, MC_MASK);
                        }
                        }
                        }
                        __syncwarp();
                        if (++cc == p.cblocks) {
                            cc = 0;
                            ++t;
                        }
                        if (++stage == STAGES) {
                            stage = 0;
                            phase ^= 1;
                        }
                    }
                } else {
                    const int total_chunks = p.ntaps * p.cblocks;
                    const int g0 = n_blk * N_CHUNKS;
                    const int valid = min(N_CHUNKS, total_chunks - g0);
                    for (int kb = kb0; kb < kb1; ++kb) {
                        if (p.tiled) {
                            const int per_img = p.nbw * p.nbh;
                            const int img = kb / per_img;
                            const int rem = kb - img * per_img;
                            const int by = rem / p.nbw;
                            const int bx = rem - by * p.nbw;
                            const int ow = bx * p.bw, oh = by * p.bh;        // first output pixel of the patch
                            mbar_wait(&empty_bar[stage], phase ^ 1);
                            uint8_t* sa = smem + stage * STAGE_BYTES;
                            uint8_t* sb = sa + A_BYTES;
                            if (elect_one()) {
                                mbar_expect_tx(&full_bar[stage], A_BYTES + valid * CV_CHUNK_BYTES);
#pragma unroll
                                for (int c = 0; c < CV_BLOCK_M / 64; ++c)
                                    tma_load_4d(&tmap_a, &full_bar[stage], sa + c * CV_CHUNK_BYTES,
                                                m_blk * CV_BLOCK_M + c * 64, ow, oh, img);
#pragma unroll
                                for (int c = 0; c < N_CHUNKS; ++c) {
                                    if (c < valid) {
                                        const int g = g0 + c;
                                        const int t = g / p.cblocks;
                                        const int cc = g - t * p.cblocks;
                                        const ConvTap tap = p.taps[t];
                                        tma_load_4d(&tmap_b, &full_bar[stage], sb + c * CV_CHUNK_BYTES, cc * 64,
                                                    ow * p.stride_w + tap.off_w - p.pad_w,
                                                    oh * p.stride_h + tap.off_h - p.pad_h, img);
                                    }
                                }
                            }
                            __syncwarp();
                            if (++stage == STAGES) {
                                stage = 0;
                                phase ^= 1;
                            }
                            continue;
                        }
                        const int k0 = kb * CV_BLOCK_K;
                        const int img = fdiv(k0, p.fd_pq);
                        const int rem = k0 - img * pq;
                        const int p0 = fdiv(rem, p.fd_q);
                        const int q0 = rem - p0 * p.Q;
                        const int cw = p.lower_w + q0 * p.stride_w;
                        const int ch = p.lower_h + p0 * p.stride_h;
                        mbar_wait(&empty_bar[stage], phase ^ 1);
                        uint8_t* sa = smem + stage * STAGE_BYTES;
                        uint8_t* sb = sa + A_BYTES;
                        if (elect_one()) {
                        const bool la = !(p.dbg & 1), lb = !(p.dbg & 2);    // timing experiments only
                        mbar_expect_tx(&full_bar[stage], (la ? A_BYTES : 0) + (lb ? valid * CV_CHUNK_BYTES : 0));
#pragma unroll
                        for (int c = 0; c < CV_BLOCK_M / 64; ++c)
                            if (la) tma_load_2d(&tmap_a, &full_bar[stage], sa + c * CV_CHUNK_BYTES, m_blk * CV_BLOCK_M + c * 64, k0);
#pragma unroll
                        for (int c = 0; c < N_CHUNKS; ++c) {
                            if (c < valid && lb) {
                                const int g = g0 + c;
                                const int t = g / p.cblocks;
                                const int cc = g - t * p.cblocks;
                                const ConvTap tap = p.taps[t];
                                tma_load_im2col_4d(&tmap_b, &full_bar[stage], sb + c * CV_CHUNK_BYTES, cc * 64, cw, ch,
                                                   img, tap.off_w, tap.off_h);
                            }
                        }
                        }
                        __syncwarp();
                        if (++stage == STAGES) {
                            stage = 0;
                            phase ^= 1;
                        }
                    }
                }
            }
        }
    } else if (warp_idx == 1) {
        // ---------------------------------------------------------------- MMA issuer
        // The WHOLE warp runs this loop and one elected lane issues: every operand of tcgen05.mma / commit is
        // then provably warp-uniform and lives in uniform registers.  Inside an `if (lane == 0)` region the
        // compiler wraps each UTCHMMA in an ELECT / R2UR.BROADCAST / BRA.U.ANY loop (about 25 instructions per
        // MMA), which made this one thread - not the tensor core - the limit: ~700 clk per k-block measured
        // with every load disabled, against 128-512 clk of MMA work.
        {
            const uint32_t a_mn = MODE == 1 ? 1u : 0u;
            const uint32_t b_mn = MODE == 1 ? 1u : static_cast<uint32_t>(p.b_mn_major);
            const uint32_t idesc = make_idesc_bf16(CV_BLOCK_M, BLOCK_N, a_mn, b_mn);
            const uint32_t a_lbo = a_mn ? CV_CHUNK_BYTES : 16;
            const uint32_t b_lbo = b_mn ? CV_CHUNK_BYTES : 16;
            const uint32_t a_kstep = a_mn ? (CV_UMMA_K * 128) >> 4 : (CV_UMMA_K * 2) >> 4;
            const uint32_t b_kstep = b_mn ? (CV_UMMA_K * 128) >> 4 : (CV_UMMA_K * 2) >> 4;
            const uint64_t da0 = make_umma_desc_sw128(smem_u32(smem), a_lbo, 1024);
            const uint64_t db0 = make_umma_desc_sw128(smem_u32(smem) + A_BYTES, b_lbo, 1024);
            int stage = 0;
            uint32_t phase = 0;
            int iter = 0;
            for (int unit = worker; unit < num_units; unit += num_workers, ++iter) {
                const int tile = fdiv(unit, p.fd_splits);
                const int split = unit - tile * p.splits;
                const int as = iter & 1;
                const uint32_t aphase = (iter >> 1) & 1;
                mbar_wait(&tmem_empty_bar[as], aphase ^ 1);
                tcgen05_fence_after();
                const uint32_t d_tmem = tmem_base + as * BLOCK_N;
                const int kb0 = split * p.kb_per_split;
                const int kb1 = min(p.num_k_blocks, kb0 + p.kb_per_split);
                for (int kb = kb0; kb < kb1; ++kb) {
                    mbar_wait(&full_bar[stage], phase);
                    tcgen05_fence_after();
                    // descriptors of stage 0 are built once; a stage is a fixed number of 16-byte units further
                    // (the start-address field cannot carry: shared memory ends below 2^18 bytes)
                    const uint64_t da = da0 + static_cast<uint64_t>(stage) * (STAGE_BYTES >> 4);
                    const uint64_t db = db0 + static_cast<uint64_t>(stage) * (STAGE_BYTES >> 4);
                    if (elect_one()) {
#pragma unroll
                        for (int k = 0; k < CV_BLOCK_K / CV_UMMA_K; ++k)
                            umma_bf16(d_tmem, da + k * a_kstep, db + k * b_kstep, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
                        if (CS > 1) umma_commit_multicast(&empty_bar[stage], MC_MASK);   // frees the slot cluster-wide
                        else umma_commit(&empty_bar[stage]);
                    }
                    __syncwarp();
                    if (++stage == STAGES) {
                        stage = 0;
                        phase ^= 1;
                    }
                }
                if (elect_one()) umma_commit(&tmem_full_bar[as]);
                __syncwarp();
            }
        }
    } else {
        // ---------------------------------------------------------------- epilogue warps (2..17)
        cv_epilogue_role<BLOCK_N, CS>(p, staging, stat_acc, tmem_base, tmem_full_bar, tmem_empty_bar, worker, num_workers,
                                      num_units, crank, warp_idx, lane);
    }

    __syncwarp();
    tcgen05_fence_before();
    if (CS > 1) cluster_sync_all(); else __syncthreads();   // no CTA leaves while peers may still write to it
    if (warp_idx == 1) {
        tcgen05_fence_after();
        tmem_dealloc<TMEM_COLS>(tmem_base);
    }
}

// ---------------------------------------------------------------------------------------------------------
// 3x3 / stride 1 convolution (forward, and the data gradient as the same traversal with mirrored taps) with the
// activation read ONCE per tile.
//
// The im2col kernel above fetches the 128 x 64 activation tile of every filter tap separately, so a 3x3 layer
// pulls each input pixel through L2 -> shared memory nine times; at 64 / 128 channels that traffic (24-32 KB per
// k-block against ~43 B/clk/SM) is what bounds it.  Here a tile is a 16 x 8 patch of one image's output pixels
// and ONE tiled TMA load brings its 18 x 10 halo (64 channels, 128 B per pixel, SWIZZLE_128B, out-of-range pixels
// zero = the padding) into shared memory.  Tap (r, s) is then only a different START ADDRESS for the same bytes:
// row m = (y, x) of the A operand is halo pixel (y + r) * 10 + (x + s), so the descriptor starts (r * 10 + s) * 128
// bytes into the patch and its 8-row groups are one halo row = 1280 bytes apart.  The hardware applies the
// 128-byte swizzle to the absolute shared-memory address, so a start that is off the 1024-byte period and a group
// stride that is not a multiple of it read exactly what TMA wrote (tools/probes/umma_halo_probe.cu: all nine taps
// exact with the descriptor's base-offset field left 0).
// Filter tiles travel through their own ring; when the whole filter of the column block fits (64 -> 64 channels:
// 72 KB) it is loaded once per CTA and stays.
constexpr int HALO_TH = 16, HALO_TW = 8, HALO_PH = HALO_TH + 2, HALO_PW = HALO_TW + 2;
constexpr uint32_t HALO_PATCH_BYTES = HALO_PH * HALO_PW * 128;          // 23040
constexpr uint32_t HALO_A_STAGE = (HALO_PATCH_BYTES + 1023) / 1024 * 1024;   // 23552 -> keep stages 1024-aligned
constexpr uint32_t HALO_BAR_BYTES = 512;
constexpr int HALO_TAPS = 9;

template <int BLOCK_N, int A_STAGES, int B_STAGES>
__global__ void __launch_bounds__(CV_THREADS, 1)
conv3x3_halo_tcgen05_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                            const ConvParams p) {
    constexpr uint32_t B_BYTES = BLOCK_N * CV_BLOCK_K * 2;
    constexpr uint32_t TMEM_COLS = (2 * BLOCK_N < 32) ? 32 : 2 * BLOCK_N;
    constexpr int N_CHUNKS = BLOCK_N / 64;
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    if ((smem_u32(smem_raw) & 1023u) != 0) __trap();
    uint8_t* smem_a = smem_raw;
    uint8_t* smem_b = smem_a + A_STAGES * HALO_A_STAGE;
    uint8_t* staging = smem_b + B_STAGES * B_BYTES;
    uint64_t* full_a = reinterpret_cast<uint64_t*>(staging + CV_STAGING_BYTES);
    uint64_t* empty_a = full_a + A_STAGES;
    uint64_t* full_b = empty_a + A_STAGES;
    uint64_t* empty_b = full_b + B_STAGES;
    uint64_t* tmem_full_bar = empty_b + B_STAGES;
    uint64_t* tmem_empty_bar = tmem_full_bar + 2;
    uint32_t* tmem_base_slot = reinterpret_cast<uint32_t*>(tmem_empty_bar + 2);
    static_assert((2 * A_STAGES + 2 * B_STAGES + 4) * 8 + 16 <= HALO_BAR_BYTES, "barrier block too small");
    float* stat_acc = reinterpret_cast<float*>(staging + CV_STAGING_BYTES + HALO_BAR_BYTES);

    const int warp_idx = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);   // provably warp-uniform: role branches and their loops stay on the uniform datapath
    const int lane = threadIdx.x & 31;
    if (p.stat_sum != nullptr)
        for (int i = threadIdx.x; i < 2 * BLOCK_N; i += CV_THREADS) stat_acc[i] = 0.f;
    if (warp_idx == 0 && lane == 0) {
        tma_prefetch_desc(&tmap_a);
        tma_prefetch_desc(&tmap_b);
        for (int s = 0; s < A_STAGES; ++s) {
            mbar_init(&full_a[s], 1);
            mbar_init(&empty_a[s], 1);
        }
        for (int s = 0; s < B_STAGES; ++s) {
            mbar_init(&full_b[s], 1);
            mbar_init(&empty_b[s], 1);
        }
        for (int s = 0; s < 2; ++s) {
            mbar_init(&tmem_full_bar[s], 1);
            mbar_init(&tmem_empty_bar[s], CV_EPI_WARPS);
        }
        fence_barrier_init();
    }
    if (warp_idx == 1) tmem_alloc<TMEM_COLS>(tmem_base_slot);
    tcgen05_fence_before();
    __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_base_slot;
    pdl_launch_dependents();

    const int num_units = p.num_m_blocks * p.num_n_blocks;
    const int worker = static_cast<int>(blockIdx.x);
    const int num_workers = static_cast<int>(gridDim.x);
    const int resident = p.b_resident;          // B stage of (channel block, tap) = cb * ntaps + tap, loaded once

    if (warp_idx == 0) {
        // ---------------------------------------------------------------- TMA producer
        int sa = 0, sb = 0;
        uint32_t pa = 0, pb = 0;
        bool first = true;
        pdl_wait();
        for (int unit = worker; unit < num_units; unit += num_workers) {
            const int m_tile = fdiv(unit, p.fd_nblk);
            const int n_blk = unit - m_tile * p.num_n_blocks;
            const int t2 = fdiv(m_tile, p.fd_tx), tx = m_tile - t2 * p.tiles_x;
            const int img = fdiv(t2, p.fd_ty), ty = t2 - img * p.tiles_y;
            for (int cb = 0; cb < p.cblocks; ++cb) {
                mbar_wait(&empty_a[sa], pa ^ 1);
                if (elect_one()) {
                    if (p.dbg & 1) {      // timing experiments only: no activation loads
                        mbar_arrive(&full_a[sa]);
                    } else {
                        mbar_expect_tx(&full_a[sa], HALO_PATCH_BYTES);
                        tma_load_4d(&tmap_a, &full_a[sa], smem_a + sa * HALO_A_STAGE, cb * 64, tx * HALO_TW + p.lower_w,
                                    ty * HALO_TH + p.lower_h, img);
                    }
                }
                __syncwarp();
                if (++sa == A_STAGES) {
                    sa = 0;
                    pa ^= 1;
                }
                if (resident && !first) continue;
                for (int t = 0; t < HALO_TAPS; ++t) {
                    const int slot = resident ? cb * HALO_TAPS + t : sb;
                    if (!resident) mbar_wait(&empty_b[slot], pb ^ 1);
                    const int winner = p.taps[t].wsel * p.w_tap_stride;
                    uint8_t* dst = smem_b + slot * B_BYTES;
                    if (elect_one()) {
                        mbar_expect_tx(&full_b[slot], B_BYTES);
                        if (!p.b_mn_major) {
                            tma_load_2d(&tmap_b, &full_b[slot], dst, winner + cb * 64, n_blk * BLOCK_N);
                        } else {
#pragma unroll
                            for (int c = 0; c < N_CHUNKS; ++c)
                                tma_load_2d(&tmap_b, &full_b[slot], dst + c * CV_CHUNK_BYTES,
                                            winner + n_blk * BLOCK_N + c * 64, cb * 64);
                        }
                    }
                    __syncwarp();
                    if (!resident && ++sb == B_STAGES) {
                        sb = 0;
                        pb ^= 1;
                    }
                }
            }
            first = false;
        }
    } else if (warp_idx == 1) {
        // ---------------------------------------------------------------- MMA issuer (whole warp, elected lane)
        // Everything an MMA needs beyond two 64-bit adds is prepared outside the tile loop: the nine tap offsets,
        // the descriptors of A stage 0 and B slot 0.  With N <= 128 a k-block is only 192-256 clk of tensor work,
        // so the instructions between two UTCHMMA decide the speed of this kernel.
        const uint32_t b_mn = static_cast<uint32_t>(p.b_mn_major);
        const uint32_t idesc = make_idesc_bf16(CV_BLOCK_M, BLOCK_N, 0u, b_mn);
        const uint32_t b_lbo = b_mn ? CV_CHUNK_BYTES : 16;
        constexpr uint32_t A_KSTEP = (CV_UMMA_K * 2) >> 4;
        const uint32_t b_kstep = b_mn ? (CV_UMMA_K * 128) >> 4 : (CV_UMMA_K * 2) >> 4;
        uint32_t toff[HALO_TAPS];
#pragma unroll
        for (int t = 0; t < HALO_TAPS; ++t) toff[t] = ((p.taps[t].off_h * HALO_PW + p.taps[t].off_w) * 128) >> 4;
        const uint64_t da00 = make_umma_desc_sw128(smem_u32(smem_a), 16, HALO_PW * 128);
        const uint64_t db00 = make_umma_desc_sw128(smem_u32(smem_b), b_lbo, 1024);
        int sa = 0, sb = 0;
        uint32_t pa = 0, pb = 0;
        bool first = true;
        int iter = 0;
        for (int unit = worker; unit < num_units; unit += num_workers, ++iter) {
            const int as = iter & 1;
            const uint32_t aphase = (iter >> 1) & 1;
            mbar_wait(&tmem_empty_bar[as], aphase ^ 1);
            tcgen05_fence_after();
            const uint32_t d_tmem = tmem_base + as * BLOCK_N;
            for (int cb = 0; cb < p.cblocks; ++cb) {
                mbar_wait(&full_a[sa], pa);
                tcgen05_fence_after();
                const uint64_t da0 = da00 + static_cast<uint64_t>(sa) * (HALO_A_STAGE >> 4);
                if (resident) {
                    if (first) {
#pragma unroll 1
                        for (int t = 0; t < HALO_TAPS; ++t) mbar_wait(&full_b[cb * HALO_TAPS + t], 0u);
                        tcgen05_fence_after();
                    }
                    const uint64_t dbc = db00 + static_cast<uint64_t>(cb * HALO_TAPS) * (B_BYTES >> 4);
                    if (elect_one()) {
                        if (!(p.dbg & 4)) {   // timing experiments only: dbg 4 issues no MMAs
#pragma unroll
                            for (int t = 0; t < HALO_TAPS; ++t) {
                                const uint64_t da = da0 + toff[t];
                                const uint64_t db = dbc + static_cast<uint64_t>(t) * (B_BYTES >> 4);
#pragma unroll
                                for (int k = 0; k < CV_BLOCK_K / CV_UMMA_K; ++k)
                                    umma_bf16(d_tmem, da + k * A_KSTEP, db + k * b_kstep, idesc,
                                              (cb > 0 || t > 0 || k > 0) ? 1u : 0u);
                            }
                        }
                        umma_commit(&empty_a[sa]);
                    }
                    __syncwarp();
                } else {
#pragma unroll
                    for (int t = 0; t < HALO_TAPS; ++t) {
                        mbar_wait(&full_b[sb], pb);
                        tcgen05_fence_after();
                        const uint64_t da = da0 + toff[t];
                        const uint64_t db = db00 + static_cast<uint64_t>(sb) * (B_BYTES >> 4);
                        if (elect_one()) {
                            if (!(p.dbg & 4)) {
#pragma unroll
                                for (int k = 0; k < CV_BLOCK_K / CV_UMMA_K; ++k)
                                    umma_bf16(d_tmem, da + k * A_KSTEP, db + k * b_kstep, idesc,
                                              (cb > 0 || t > 0 || k > 0) ? 1u : 0u);
                            }
                            umma_commit(&empty_b[sb]);
                            if (t == HALO_TAPS - 1) umma_commit(&empty_a[sa]);
                        }
                        __syncwarp();
                        if (++sb == B_STAGES) {
                            sb = 0;
                            pb ^= 1;
                        }
                    }
                }
                if (++sa == A_STAGES) {
                    sa = 0;
                    pa ^= 1;
                }
            }
            if (elect_one()) umma_commit(&tmem_full_bar[as]);
            __syncwarp();
            first = false;
        }
    } else {
        cv_epilogue_role<BLOCK_N, 1>(p, staging, stat_acc, tmem_base, tmem_full_bar, tmem_empty_bar, worker, num_workers,
                                     num_units, 0u, warp_idx, lane);
    }

    __syncwarp();
    tcgen05_fence_before();
    __syncthreads();
    if (warp_idx == 1) {
        tcgen05_fence_after();
        tmem_dealloc<TMEM_COLS>(tmem_base);
    }
}

// 2x2 space-to-depth of a 3-channel NCHW image batch for the 7x7 / stride-2 stem convolution:
//   xs[n, i, 2 + j, (dy*2 + dx)*4 + c] = x[n, c, 2i + dy, 2j + dx]   (c < 3; channel 3 of each group is 0)
// with two zero pixels before and one after every row.  On this tensor the stem is a stride-1 convolution
// with 4 x 4 taps over 16 channels, and because NHWC keeps a row's pixels adjacent, the 4 horizontal taps
// of an output pixel are 64 CONTIGUOUS elements: the convolution kernel reads them as one 64-"channel"
// pixel through a tensor map whose pixel stride (16 elements) is smaller than its channel extent (64).
__global__ void __launch_bounds__(256)
stem_space_to_depth_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ xs, int N, int H, int W) {
    const int H2 = H / 2, W2 = W / 2, WP = W2 + 3;
    const long long total = static_cast<long long>(N) * H2 * WP;
    const size_t plane = static_cast<size_t>(H) * W;
    for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        const int jp = static_cast<int>(i % WP);
        const long long t = i / WP;
        const int ii = static_cast<int>(t % H2);
        const int n = static_cast<int>(t / H2);
        uint32_t o[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        const int j = jp - 2;
        if (j >= 0 && j < W2) {
            const __nv_bfloat16* base = x + static_cast<size_t>(n) * 3 * plane + static_cast<size_t>(2 * ii) * W + 2 * j;
            uint32_t v[3][2];   // [c][dy] = pair (dx = 0, 1)
#pragma unroll
            for (int c = 0; c < 3; ++c)
#pragma unroll
                for (int dy = 0; dy < 2; ++dy)
                    v[c][dy] = __ldg(reinterpret_cast<const uint32_t*>(base + c * plane + dy * W));
#pragma unroll
            for (int dy = 0; dy < 2; ++dy)
#pragma unroll
                for (int dx = 0; dx < 2; ++dx) {
                    const int sh = dx * 16;
                    const uint32_t c0 = (v[0][dy] >> sh) & 0xffffu, c1 = (v[1][dy] >> sh) & 0xffffu,
                                   c2 = (v[2][dy] >> sh) & 0xffffu;
                    o[(dy * 2 + dx) * 2] = c0 | (c1 << 16);
                    o[(dy * 2 + dx) * 2 + 1] = c2;
                }
        }
        uint4* dst = reinterpret_cast<uint4*>(xs + static_cast<size_t>(i) * 16);
        dst[0] = make_uint4(o[0], o[1], o[2], o[3]);
        dst[1] = make_uint4(o[4], o[5], o[6], o[7]);
    }
}

// Same transform with the horizontal window written out: xw[n, i, q, b*16 + ch] = xs[n, i, q + b, ch]
// (b = 0..3), a dense [N, H/2, W/2, 64] tensor (4x the bytes of xs).  Every thread produces one
// 32-byte group (n, i, q, b).
__global__ void __launch_bounds__(256)
stem_space_to_depth_window_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ xw, int N, int H,
                                  int W) {
    const int H2 = H / 2, W2 = W / 2;
    const long long total = static_cast<long long>(N) * H2 * W2 * 4;
    const size_t plane = static_cast<size_t>(H) * W;
    for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        const int b = static_cast<int>(i & 3);
        long long t = i >> 2;
        const int q = static_cast<int>(t % W2);
        t /= W2;
        const int ii = static_cast<int>(t % H2);
        const int n = static_cast<int>(t / H2);
        uint32_t o[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        const int j = q + b - 2;
        if (j >= 0 && j < W2) {
            const __nv_bfloat16* base = x + static_cast<size_t>(n) * 3 * plane + static_cast<size_t>(2 * ii) * W + 2 * j;
            uint32_t v[3][2];
#pragma unroll
            for (int c = 0; c < 3; ++c)
#pragma unroll
                for (int dy = 0; dy < 2; ++dy)
                    v[c][dy] = __ldg(reinterpret_cast<const uint32_t*>(base + c * plane + dy * W));
#pragma unroll
            for (int dy = 0; dy < 2; ++dy)
#pragma unroll
                for (int dx = 0; dx < 2; ++dx) {
                    const int sh = dx * 16;
                    const uint32_t c0 = (v[0][dy] >> sh) & 0xffffu, c1 = (v[1][dy] >> sh) & 0xffffu,
                                   c2 = (v[2][dy] >> sh) & 0xffffu;
                    o[(dy * 2 + dx) * 2] = c0 | (c1 << 16);
                    o[(dy * 2 + dx) * 2 + 1] = c2;
                }
        }
        uint4* dst = reinterpret_cast<uint4*>(xw + static_cast<size_t>(i) * 16);
        dst[0] = make_uint4(o[0], o[1], o[2], o[3]);
        dst[1] = make_uint4(o[4], o[5], o[6], o[7]);
    }
}

// ----------------------------------------------------------------------------------------------
// Host side
// ----------------------------------------------------------------------------------------------
typedef CUresult (*PFN_encodeIm2col)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                     const cuuint64_t*, const int*, const int*, cuuint32_t, cuuint32_t,
                                     const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                     CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static PFN_encodeIm2col get_im2col_fn() {
    static PFN_encodeIm2col fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* ptr = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeIm2col", &ptr, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<PFN_encodeIm2col>(ptr);
    });
    return fn;
}

// im2col map over an NHWC bf16 tensor: 64 channels x `pixels` pixels per load
static int make_im2col_tmap(CUtensorMap* map, const void* ptr, int N, int H, int W, int C, int lower_w, int lower_h,
                            int upper_w, int upper_h, int stride_w, int stride_h, int pixels, long long es_w = 0,
                            long long es_h = 0, long long es_n = 0) {
    static thread_local bool tls_ctx_bound = false;
    if (!tls_ctx_bound) {
        cudaFree(nullptr);
        tls_ctx_bound = true;
    }
    PFN_encodeIm2col enc = get_im2col_fn();
    if (!enc) {
        set_last_error("cuTensorMapEncodeIm2col", "driver entry point not found");
        return TONY_ERR_UNSUPPORTED;
    }
    cuuint64_t gdim[4] = {static_cast<cuuint64_t>(C), static_cast<cuuint64_t>(W), static_cast<cuuint64_t>(H),
                          static_cast<cuuint64_t>(N)};
    if (es_w == 0) es_w = C;
    if (es_h == 0) es_h = es_w * W;
    if (es_n == 0) es_n = es_h * H;
    cuuint64_t gstride[3] = {static_cast<cuuint64_t>(es_w) * 2, static_cast<cuuint64_t>(es_h) * 2,
                             static_cast<cuuint64_t>(es_n) * 2};
    int lower[2] = {lower_w, lower_h};
    int upper[2] = {upper_w, upper_h};
    cuuint32_t estride[4] = {1, static_cast<cuuint32_t>(stride_w), static_cast<cuuint32_t>(stride_h), 1};
    CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(ptr), gdim, gstride, lower, upper,
                     64, static_cast<cuuint32_t>(pixels), estride, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        char buf[160];
        snprintf(buf, sizeof(buf), "CUresult %d (N %d H %d W %d C %d corners %d %d %d %d stride %d %d)",
                 static_cast<int>(r), N, H, W, C, lower_w, lower_h, upper_w, upper_h, stride_w, stride_h);
        set_last_error("cuTensorMapEncodeIm2col", buf);
        return TONY_ERR_CUDA;
    }
    return TONY_OK;
}

typedef CUresult (*PFN_encodeTiled4)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                     const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                     CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// 4-D tiled map over a dense NHWC bf16 tensor: box = 64 channels x (bw x bh) pixels of one image, visited with
// element strides (sw, sh) - a patch of the output grid seen through the convolution's stride
static int make_patch_tmap(CUtensorMap* map, const void* ptr, int N, int H, int W, int C, int bw, int bh, int sw,
                           int sh) {
    static thread_local bool tls_ctx_bound = false;
    if (!tls_ctx_bound) {
        cudaFree(nullptr);
        tls_ctx_bound = true;
    }
    static PFN_encodeTiled4 enc = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            enc = reinterpret_cast<PFN_encodeTiled4>(fn);
    });
    if (!enc) {
        set_last_error("cuTensorMapEncodeTiled", "driver entry point not found");
        return TONY_ERR_UNSUPPORTED;
    }
    cuuint64_t gdim[4] = {static_cast<cuuint64_t>(C), static_cast<cuuint64_t>(W), static_cast<cuuint64_t>(H),
                          static_cast<cuuint64_t>(N)};
    cuuint64_t gstride[3] = {static_cast<cuuint64_t>(C) * 2, static_cast<cuuint64_t>(W) * C * 2,
                             static_cast<cuuint64_t>(H) * W * C * 2};
    cuuint32_t box[4] = {64, static_cast<cuuint32_t>(bw * sw), static_cast<cuuint32_t>(bh * sh), 1};
    cuuint32_t estride[4] = {1, static_cast<cuuint32_t>(sw), static_cast<cuuint32_t>(sh), 1};
    CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(ptr), gdim, gstride, box, estride,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        char buf[128];
        snprintf(buf, sizeof(buf), "CUresult %d (4-D patch map, box %d x %d, stride %d)", static_cast<int>(r), bw, bh, sw);
        set_last_error("cuTensorMapEncodeTiled", buf);
        return TONY_ERR_CUDA;
    }
    return TONY_OK;
}

template <int BLOCK_N, int STAGES, int MODE, int CS>
static int launch_conv(const CUtensorMap& ta, const CUtensorMap& tb, const ConvParams& p, int grid, cudaStream_t stream) {
    constexpr size_t SMEM =
        static_cast<size_t>(STAGES) * (CV_BLOCK_M + BLOCK_N) * CV_BLOCK_K * 2 + CV_STAGING_BYTES + 256 +
        2 * BLOCK_N * 4;
    static bool configured[64] = {false};
    int dev = 0;
    TONY_CUDA_CHECK(cudaGetDevice(&dev));
    if (dev >= 0 && dev < 64 && !configured[dev]) {
        TONY_CUDA_CHECK(cudaFuncSetAttribute(conv_bf16_tcgen05_kernel<BLOCK_N, STAGES, MODE, CS>,
                                             cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(SMEM)));
        configured[dev] = true;
    }
    if (CS > 1) {
        // persistent clusters: no more of them than can be resident at once (cluster placement may leave
        // a few SMs unused), each takes whole groups of CS tiles
        static int max_clusters[64] = {0};
        if (dev >= 0 && dev < 64 && max_clusters[dev] == 0) {
            cudaLaunchConfig_t q = {};
            q.gridDim = dim3(device_sm_count() / CS * CS);
            q.blockDim = dim3(CV_THREADS);
            q.dynamicSmemBytes = SMEM;
            cudaLaunchAttribute qa[1];
            qa[0].id = cudaLaunchAttributeClusterDimension;
            qa[0].val.clusterDim.x = CS;
            qa[0].val.clusterDim.y = 1;
            qa[0].val.clusterDim.z = 1;
            q.attrs = qa;
            q.numAttrs = 1;
            int n = 0;
            if (cudaOccupancyMaxActiveClusters(&n, conv_bf16_tcgen05_kernel<BLOCK_N, STAGES, MODE, CS>, &q) !=
                    cudaSuccess || n <= 0)
                n = device_sm_count() / CS;
            max_clusters[dev] = n;
        }
        const int groups = ((p.num_m_blocks + CS - 1) / CS) * p.num_n_blocks;
        grid = std::min(groups, max_clusters[dev >= 0 && dev < 64 ? dev : 0]) * CS;
    }
    static int pdl = -1;
    if (pdl < 0) {
        const char* e = getenv("TONY_B200_PDL");
        pdl = e ? atoi(e) : 1;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(CV_THREADS);
    cfg.dynamicSmemBytes = SMEM;
    cfg.stream = stream;
    cudaLaunchAttribute attr[2];
    int na = 0;
    if (CS > 1) {
        attr[na].id = cudaLaunchAttributeClusterDimension;
        attr[na].val.clusterDim.x = CS;
        attr[na].val.clusterDim.y = 1;
        attr[na].val.clusterDim.z = 1;
        ++na;
    }
    if (pdl) {
        attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[na].val.programmaticStreamSerializationAllowed = 1;
        ++na;
    }
    cfg.attrs = attr;
    cfg.numAttrs = na;
    ConvParams pf = p;
    fill_fastdiv(pf);
    TONY_CUDA_CHECK(cudaLaunchKernelEx(&cfg, conv_bf16_tcgen05_kernel<BLOCK_N, STAGES, MODE, CS>, ta, tb, pf));
    count_launch();
    return TONY_OK;
}

// cluster size of the fprop / dgrad kernels (filter multicast); TONY_B200_CONV_CLUSTER = 1 | 2 | 4
static int conv_cluster_size() {
    static int cs = -1;
    if (cs < 0) {
        const char* e = getenv("TONY_B200_CONV_CLUSTER");
        // default 1: with the issue loops fixed the multicast forms measure equal (2) or slower (4) - the kernel
        // sits at its MMA + issue floor, not on L2 -> SM bytes; kept as validated options
        cs = e ? atoi(e) : 1;
        if (cs != 1 && cs != 2 && cs != 4) cs = 1;
    }
    return cs;
}

template <int MODE, int CS>
static int dispatch_conv_cs(int block_n, const CUtensorMap& ta, const CUtensorMap& tb, const ConvParams& p, int grid,
                            cudaStream_t stream) {
    if (block_n == 64) return launch_conv<64, 8, MODE, CS>(ta, tb, p, grid, stream);
    if (block_n == 128) return launch_conv<128, 6, MODE, CS>(ta, tb, p, grid, stream);
    return launch_conv<256, 4, MODE, CS>(ta, tb, p, grid, stream);
}

template <int MODE>
static int dispatch_conv(int block_n, const CUtensorMap& ta, const CUtensorMap& tb, const ConvParams& p, int grid,
                         cudaStream_t stream, int cs = 1) {
    if (MODE == 0 && cs == 2) return dispatch_conv_cs<0, 2>(block_n, ta, tb, p, grid, stream);
    if (MODE == 0 && cs == 4) return dispatch_conv_cs<0, 4>(block_n, ta, tb, p, grid, stream);
    return dispatch_conv_cs<MODE, 1>(block_n, ta, tb, p, grid, stream);
}

// rows of one filter-operand TMA box when `cs` CTAs share the tile (see the kernel's B loads)
static int filter_box_rows(int block_n, int b_mn_major, int cs) {
    if (!b_mn_major) return block_n / cs;
    const int chunks = block_n / 64;
    return chunks >= cs ? 64 : 64 / (cs / chunks);
}

template <int BLOCK_N, int A_STAGES, int B_STAGES>
static int launch_halo(const CUtensorMap& ta, const CUtensorMap& tb, const ConvParams& p, int grid, cudaStream_t stream) {
    constexpr size_t SMEM = static_cast<size_t>(A_STAGES) * HALO_A_STAGE + static_cast<size_t>(B_STAGES) * BLOCK_N * 128 +
                            CV_STAGING_BYTES + HALO_BAR_BYTES + 2 * BLOCK_N * 4;
    static_assert(SMEM <= 232448, "over the 227 KB of shared memory a CTA can have");
    static bool configured[64] = {false};
    int dev = 0;
    TONY_CUDA_CHECK(cudaGetDevice(&dev));
    if (dev >= 0 && dev < 64 && !configured[dev]) {
        TONY_CUDA_CHECK(cudaFuncSetAttribute(conv3x3_halo_tcgen05_kernel<BLOCK_N, A_STAGES, B_STAGES>,
                                             cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(SMEM)));
        configured[dev] = true;
    }
    ConvParams pf = p;
    fill_fastdiv(pf);
    conv3x3_halo_tcgen05_kernel<BLOCK_N, A_STAGES, B_STAGES><<<grid, CV_THREADS, SMEM, stream>>>(ta, tb, pf);
    count_launch();
    TONY_CUDA_CHECK(cudaGetLastError());
    return TONY_OK;
}

// Which 3x3 / stride-1 layers with <= 128 output channels take the halo kernel.  Unset: only those whose whole filter
// stays resident in shared memory (64 -> 64 channels: 74.7 us against 87.0 us for the im2col kernel at 256 x 56 x 56);
// with the filter streaming through its ring the 12-23 % of MMA rows its 16 x 8 tiles waste on 28 x 28 images
// outweigh the saved activation traffic (62.5 vs 58.4 us).  TONY_B200_CONV_HALO=1: every eligible layer, 0: none.
// Read on every call (tests switch it); a getenv is ~100 ns.
static int conv_halo_mode() {
    const char* e = getenv("TONY_B200_CONV_HALO");
    return e ? (atoi(e) != 0 ? 2 : 0) : 1;
}

// Fills the halo-form fields of `p` (taps, M, tile grid) and launches; `mirror`: data-gradient traversal.
// x: the tensor the patch is cut from ([N, H, W, C] dense), output [N, H, W, p.N].
constexpr int HALO_DECLINED = -12345;   // run_halo: this layer stays on the im2col kernel
static int run_halo(ConvParams p, const void* x, int N, int H, int W, int C, const CUtensorMap& tb, int block_n, bool mirror,
                    cudaStream_t stream) {
    p.halo = 1;
    p.tiles_x = (W + HALO_TW - 1) / HALO_TW;
    p.tiles_y = (H + HALO_TH - 1) / HALO_TH;
    p.num_m_blocks = N * p.tiles_x * p.tiles_y;
    p.M = p.num_m_blocks * CV_BLOCK_M;
    p.P = H;
    p.Q = W;
    p.lower_w = p.lower_h = -1;
    p.out_strided = 0;
    p.ntaps = 9;
    p.cblocks = (C + 63) / 64;
    p.splits = 1;
    for (int r = 0; r < 3; ++r)
        for (int s = 0; s < 3; ++s) {
            ConvTap& t = p.taps[r * 3 + s];
            t.off_h = static_cast<unsigned short>(mirror ? 2 - r : r);
            t.off_w = static_cast<unsigned short>(mirror ? 2 - s : s);
            t.wsel = r * 3 + s;
        }
    const int b_stages = block_n == 64 ? 9 : 6;
    p.b_resident = (p.ntaps * p.cblocks <= b_stages && p.num_n_blocks == 1) ? 1 : 0;
    if (!p.b_resident && conv_halo_mode() < 2) return HALO_DECLINED;
    CUtensorMap ta;
    int rc = make_patch_tmap(&ta, x, N, H, W, C, HALO_PW, HALO_PH, 1, 1);
    if (rc) return rc;
    const long long units = static_cast<long long>(p.num_m_blocks) * p.num_n_blocks;
    const int grid = static_cast<int>(std::min<long long>(units, device_sm_count()));
    if (block_n == 64) return launch_halo<64, 4, 9>(ta, tb, p, grid, stream);
    return launch_halo<128, 3, 6>(ta, tb, p, grid, stream);
}

static int pick_block_n(int n) { return n <= 64 ? 64 : (n <= 128 ? 128 : 256); }

static int check_desc(const char* what, const tony_conv_desc* d) {
    if (!d || d->N <= 0 || d->H <= 0 || d->W <= 0 || d->C <= 0 || d->K <= 0 || d->R <= 0 || d->S <= 0 ||
        d->stride <= 0 || d->pad < 0) {
        set_last_error(what, "bad convolution descriptor");
        return TONY_ERR_INVALID;
    }
    if ((d->C % 8) || (d->K % 8) || d->R * d->S > CV_MAX_TAPS || d->stride > 2) {
        set_last_error(what, "needs C % 8 == 0, K % 8 == 0, R*S <= 16, stride 1 or 2");
        return TONY_ERR_UNSUPPORTED;
    }
    return TONY_OK;
}

}  // namespace tony

using namespace tony;

extern "C" int tony_conv2d_out_size(int in, int filter, int stride, int pad) {
    return (in + 2 * pad - filter) / stride + 1;
}

// Y[N, P, Q, K] = conv(X[N, H, W, C], Wt[K, R, S, C])
extern "C" int tony_conv2d_fprop_bf16(const tony_conv_desc* d, const void* x, const void* w, void* y,
                                      float* stat_sum, float* stat_sqsum, void* stream_v) {
    int rc = check_desc("tony_conv2d_fprop_bf16", d);
    if (rc) return rc;
    if (!x || !w || !y || ((stat_sum == nullptr) != (stat_sqsum == nullptr))) return TONY_ERR_INVALID;
    cudaStream_t stream = static_cast<cudaStream_t>(stream_v);
    const int pad_w = d->pad_h_only ? 0 : d->pad;
    const int pad_hi_h = d->pad_hi < 0 ? d->pad : d->pad_hi;
    const int pad_hi_w = d->pad_h_only ? 0 : pad_hi_h;
    const int P = (d->H + d->pad + pad_hi_h - d->R) / d->stride + 1;
    const int Q = (d->W + pad_w + pad_hi_w - d->S) / d->stride + 1;
    if (P <= 0 || Q <= 0) return TONY_ERR_INVALID;
    ConvParams p = {};
    p.M = d->N * P * Q;
    p.N = d->K;
    const int block_n = pick_block_n(p.N);
    p.num_m_blocks = (p.M + CV_BLOCK_M - 1) / CV_BLOCK_M;
    p.num_n_blocks = (p.N + block_n - 1) / block_n;
    p.ntaps = d->R * d->S;
    p.cblocks = (d->C + 63) / 64;
    p.num_k_blocks = p.ntaps * p.cblocks;
    p.splits = 1;
    p.kb_per_split = p.num_k_blocks;
    p.b_mn_major = 0;
    p.w_tap_stride = d->C;
    p.P = P;
    p.Q = Q;
    p.lower_w = -pad_w;
    p.lower_h = -d->pad;
    p.stride_w = p.stride_h = d->stride;
    for (int r = 0; r < d->R; ++r)
        for (int s = 0; s < d->S; ++s) {
            ConvTap& t = p.taps[r * d->S + s];
            t.off_w = static_cast<unsigned short>(s);
            t.off_h = static_cast<unsigned short>(r);
            t.wsel = r * d->S + s;
        }
    p.out_strided = 0;
    p.C = static_cast<__nv_bfloat16*>(y);
    p.ldc = d->K;
    p.flags = 0;
    p.stat_sum = stat_sum;
    p.stat_sqsum = stat_sqsum;
    {
        const char* e = getenv("TONY_B200_CONV_DBG");
        p.dbg = e ? atoi(e) : 0;
    }
    if (conv_halo_mode() && d->R == 3 && d->S == 3 && d->stride == 1 && d->pad == 1 && !d->pad_h_only &&
        (d->pad_hi < 0 || d->pad_hi == 1) && !d->x_stride_w && !d->x_stride_h && !d->x_stride_n && d->C % 64 == 0 &&
        block_n <= 128) {
        CUtensorMap tbh;
        rc = make_tmap_2d_bf16(&tbh, w, static_cast<long long>(p.ntaps) * d->C, d->K, static_cast<long long>(p.ntaps) * d->C,
                               64, filter_box_rows(block_n, 0, 1));
        if (rc) return rc;
        rc = run_halo(p, x, d->N, d->H, d->W, d->C, tbh, block_n, false, stream);
        if (rc != HALO_DECLINED) return rc;
    }
    CUtensorMap ta, tb;
    rc = make_im2col_tmap(&ta, x, d->N, d->H, d->W, d->C, -pad_w, -d->pad, pad_hi_w - (d->S - 1),
                          pad_hi_h - (d->R - 1), d->stride, d->stride, CV_BLOCK_M, d->x_stride_w, d->x_stride_h,
                          d->x_stride_n);
    if (rc) return rc;
    const int cs = (p.num_m_blocks >= 8 && !p.dbg) ? conv_cluster_size() : 1;
    rc = make_tmap_2d_bf16(&tb, w, static_cast<long long>(p.ntaps) * d->C, d->K, static_cast<long long>(p.ntaps) * d->C,
                           64, filter_box_rows(block_n, 0, cs));
    if (rc) return rc;
    const long long units = static_cast<long long>(p.num_m_blocks) * p.num_n_blocks;
    const int grid = static_cast<int>(std::min<long long>(units, device_sm_count()));
    return dispatch_conv<0>(block_n, ta, tb, p, grid, stream, cs);
}

// dX[N, H, W, C] (+)= conv_transpose(dY[N, P, Q, K], Wt[K, R, S, C]);  flags: TONY_EPI_ACCUMULATE
extern "C" int tony_conv2d_dgrad_bf16(const tony_conv_desc* d, const void* dy, const void* w, void* dx,
                                      unsigned int flags, void* stream_v) {
    int rc = check_desc("tony_conv2d_dgrad_bf16", d);
    if (rc) return rc;
    if (!dy || !w || !dx || (flags & ~TONY_EPI_ACCUMULATE)) return TONY_ERR_INVALID;
    if (d->pad_h_only || (d->pad_hi >= 0 && d->pad_hi != d->pad) || d->x_stride_w || d->x_stride_h || d->x_stride_n) {
        set_last_error("tony_conv2d_dgrad_bf16", "dense NHWC input with symmetric padding only");
        return TONY_ERR_UNSUPPORTED;
    }
    cudaStream_t stream = static_cast<cudaStream_t>(stream_v);
    const int P = tony_conv2d_out_size(d->H, d->R, d->stride, d->pad);
    const int Q = tony_conv2d_out_size(d->W, d->S, d->stride, d->pad);
    if (P <= 0 || Q <= 0) return TONY_ERR_INVALID;
    const int st = d->stride;
    const int block_n = pick_block_n(d->C);
    CUtensorMap tb;
    // the filter seen as [K rows][R*S*C] : B(n = ci, k = co) of tap t sits at inner offset t*C + ci
    const long long min_tiles = static_cast<long long>(d->N) * (d->H / st) * (d->W / st) / CV_BLOCK_M;
    const int cs = min_tiles >= 8 ? conv_cluster_size() : 1;
    rc = make_tmap_2d_bf16(&tb, w, static_cast<long long>(d->R) * d->S * d->C, d->K,
                           static_cast<long long>(d->R) * d->S * d->C, 64, filter_box_rows(block_n, 1, cs));
    if (rc) return rc;
    if (conv_halo_mode() && d->R == 3 && d->S == 3 && st == 1 && d->pad == 1 && d->K % 64 == 0 && block_n <= 128) {
        // dX = 3x3 traversal of dY with mirrored taps and the filter read MN-major
        ConvParams p = {};
        p.N = d->C;
        p.num_n_blocks = (p.N + block_n - 1) / block_n;
        p.b_mn_major = 1;
        p.w_tap_stride = d->C;
        p.C = static_cast<__nv_bfloat16*>(dx);
        p.ldc = d->C;
        p.flags = flags;
        CUtensorMap tbh;
        rc = make_tmap_2d_bf16(&tbh, w, static_cast<long long>(d->R) * d->S * d->C, d->K,
                               static_cast<long long>(d->R) * d->S * d->C, 64, filter_box_rows(block_n, 1, 1));
        if (rc) return rc;
        rc = run_halo(p, dy, d->N, P, Q, d->K, tbh, block_n, true, stream);
        if (rc != HALO_DECLINED) return rc;
    }
    // a parity class with no filter tap (1x1 stride 2: every odd row / column) receives no gradient
    bool need_zero = false;
    for (int ph = 0; ph < st; ++ph) {
        int nr = 0, ns = 0;
        for (int r = 0; r < d->R; ++r) nr += ((ph + d->pad - r) % st == 0);
        for (int s = 0; s < d->S; ++s) ns += ((ph + d->pad - s) % st == 0);
        if (nr == 0 || ns == 0) need_zero = true;
    }
    if (need_zero && !(flags & TONY_EPI_ACCUMULATE))
        TONY_CUDA_CHECK(cudaMemsetAsync(dx, 0, static_cast<size_t>(d->N) * d->H * d->W * d->C * 2, stream));
    for (int ph = 0; ph < st; ++ph)
        for (int pw = 0; pw < st; ++pw) {
            // positions h = st*i + ph; contributing filter rows r with (h + pad - r) % st == 0 read
            // dY row i + (ph + pad - r) / st
            int dh[CV_MAX_TAPS], rr[CV_MAX_TAPS], nh = 0, dw[CV_MAX_TAPS], ss[CV_MAX_TAPS], nw = 0;
            for (int r = 0; r < d->R; ++r) {
                const int num = ph + d->pad - r;
                if (((num % st) + st) % st == 0) {
                    dh[nh] = (num >= 0 ? num / st : -((-num) / st));
                    rr[nh++] = r;
                }
            }
            for (int s = 0; s < d->S; ++s) {
                const int num = pw + d->pad - s;
                if (((num % st) + st) % st == 0) {
                    dw[nw] = (num >= 0 ? num / st : -((-num) / st));
                    ss[nw++] = s;
                }
            }
            const int HP = (d->H - ph + st - 1) / st;   // rows of dX in this class
            const int WP = (d->W - pw + st - 1) / st;
            if (nh == 0 || nw == 0 || HP <= 0 || WP <= 0) continue;
            if (nh * nw > CV_MAX_TAPS) return TONY_ERR_UNSUPPORTED;
            int min_h = dh[0], min_w = dw[0];
            for (int i = 1; i < nh; ++i) min_h = std::min(min_h, dh[i]);
            for (int i = 1; i < nw; ++i) min_w = std::min(min_w, dw[i]);
            ConvParams p = {};
            p.M = d->N * HP * WP;
            p.N = d->C;
            p.num_m_blocks = (p.M + CV_BLOCK_M - 1) / CV_BLOCK_M;
            p.num_n_blocks = (p.N + block_n - 1) / block_n;
            p.ntaps = nh * nw;
            p.cblocks = (d->K + 63) / 64;
            p.num_k_blocks = p.ntaps * p.cblocks;
            p.splits = 1;
            p.kb_per_split = p.num_k_blocks;
            p.b_mn_major = 1;
            p.w_tap_stride = d->C;
            p.P = HP;
            p.Q = WP;
            p.lower_w = min_w;
            p.lower_h = min_h;
            p.stride_w = p.stride_h = 1;
            for (int i = 0; i < nh; ++i)
                for (int j = 0; j < nw; ++j) {
                    ConvTap& t = p.taps[i * nw + j];
                    t.off_h = static_cast<unsigned short>(dh[i] - min_h);
                    t.off_w = static_cast<unsigned short>(dw[j] - min_w);
                    t.wsel = rr[i] * d->S + ss[j];
                }
            p.out_strided = (st > 1) ? 1 : 0;
            p.OH = d->H;
            p.OW = d->W;
            p.osh = p.osw = st;
            p.oph = ph;
            p.opw = pw;
            p.C = static_cast<__nv_bfloat16*>(dx);
            p.ldc = d->C;
            p.flags = flags;
            CUtensorMap ta;
            // traversal visits exactly HP x WP base pixels: count = dim + upper - lower  (stride 1)
            rc = make_im2col_tmap(&ta, dy, d->N, P, Q, d->K, min_w, min_h, WP - Q + min_w, HP - P + min_h, 1, 1,
                                  CV_BLOCK_M);
            if (rc) return rc;
            const long long units = static_cast<long long>(p.num_m_blocks) * p.num_n_blocks;
            const int grid = static_cast<int>(std::min<long long>(units, device_sm_count()));
            rc = dispatch_conv<0>(block_n, ta, tb, p, grid, stream, cs);
            if (rc) return rc;
        }
    return TONY_OK;
}

extern "C" size_t tony_conv2d_wgrad_workspace_bytes(const tony_conv_desc* d) {
    if (!d) return 0;
    // worst case: one split per SM over a single tile row
    return static_cast<size_t>(160) * std::max(d->K, 128) * (static_cast<size_t>(d->R) * d->S * d->C + 256) * 4;
}

// dW[K, R, S, C] (+)= sum over output pixels of dY[pix, K]^T . im2col(X)[pix, (r, s, c)]
extern "C" int tony_conv2d_wgrad_bf16(const tony_conv_desc* d, const void* dy, const void* x, void* dw,
                                      unsigned int flags, void* workspace, size_t workspace_bytes, void* stream_v) {
    int rc = check_desc("tony_conv2d_wgrad_bf16", d);
    if (rc) return rc;
    if (!dy || !x || !dw || (flags & ~TONY_EPI_ACCUMULATE)) return TONY_ERR_INVALID;
    if (d->C % 64) {
        set_last_error("tony_conv2d_wgrad_bf16", "needs C % 64 == 0");
        return TONY_ERR_UNSUPPORTED;
    }
    cudaStream_t stream = static_cast<cudaStream_t>(stream_v);
    const int pad_w = d->pad_h_only ? 0 : d->pad;
    const int pad_hi_h = d->pad_hi < 0 ? d->pad : d->pad_hi;
    const int pad_hi_w = d->pad_h_only ? 0 : pad_hi_h;
    const int P = (d->H + d->pad + pad_hi_h - d->R) / d->stride + 1;
    const int Q = (d->W + pad_w + pad_hi_w - d->S) / d->stride + 1;
    if (P <= 0 || Q <= 0) return TONY_ERR_INVALID;
    const long long pixels = static_cast<long long>(d->N) * P * Q;
    ConvParams p = {};
    p.M = d->K;
    p.N = d->R * d->S * d->C;
    const int block_n = pick_block_n(p.N);
    p.num_m_blocks = (p.M + CV_BLOCK_M - 1) / CV_BLOCK_M;
    p.num_n_blocks = (p.N + block_n - 1) / block_n;
    p.ntaps = d->R * d->S;
    p.cblocks = d->C / 64;
    p.num_k_blocks = static_cast<int>((pixels + CV_BLOCK_K - 1) / CV_BLOCK_K);
    // tiled form (dense input; TONY_B200_CONV_WGRAD=tiled enables): pick the 64-pixel patch shape that wastes
    // the fewest out-of-range pixels on this output grid.  Validated, but measured SLOWER than the im2col form
    // (128 vs 99 us at 256 -> 256 channels, 14 x 14): the limit is not the im2col mode but the 48 KB of
    // operands per 512 clk of MMAs - either way 384 rows of 128 bytes per k-block through the L2 -> SM path.
    static int wgrad_tiled = -1;
    if (wgrad_tiled < 0) {
        const char* e = getenv("TONY_B200_CONV_WGRAD");
        wgrad_tiled = (e && e[0] == 't') ? 1 : 0;
    }
    p.tiled = 0;
    if (wgrad_tiled && !d->x_stride_w && !d->x_stride_h && !d->x_stride_n) {
        double best = 0.0;
        for (int bw = 1; bw <= 64; bw *= 2) {
            const int bh = 64 / bw;
            if (bw * d->stride > 256 || bh * d->stride > 256) continue;
            const int nbw = (Q + bw - 1) / bw, nbh = (P + bh - 1) / bh;
            const double eff = (static_cast<double>(Q) / (nbw * bw)) * (static_cast<double>(P) / (nbh * bh));
            if (eff > best + 1e-9) {
                best = eff;
                p.bw = bw;
                p.bh = bh;
                p.nbw = nbw;
                p.nbh = nbh;
            }
        }
        if (best > 0.0) {
            p.tiled = 1;
            p.pad_w = pad_w;
            p.pad_h = d->pad;
            p.num_k_blocks = d->N * p.nbw * p.nbh;
        }
    }
    const int sms = device_sm_count();
    const long long tiles = static_cast<long long>(p.num_m_blocks) * p.num_n_blocks;
    int splits = 1;
    if (workspace != nullptr) {
        splits = static_cast<int>(std::max<long long>(1, sms / tiles));
        splits = std::min(splits, std::max(1, p.num_k_blocks / 8));
        while (splits > 1 && static_cast<size_t>(splits) * p.M * static_cast<size_t>(p.N) * 4 > workspace_bytes) --splits;
    }
    p.kb_per_split = (p.num_k_blocks + splits - 1) / splits;
    p.splits = (p.num_k_blocks + p.kb_per_split - 1) / p.kb_per_split;
    p.P = P;
    p.Q = Q;
    p.lower_w = -pad_w;
    p.lower_h = -d->pad;
    p.stride_w = p.stride_h = d->stride;
    for (int r = 0; r < d->R; ++r)
        for (int s = 0; s < d->S; ++s) {
            ConvTap& t = p.taps[r * d->S + s];
            t.off_w = static_cast<unsigned short>(s);
            t.off_h = static_cast<unsigned short>(r);
            t.wsel = r * d->S + s;
        }
    p.out_strided = 0;
    p.C = static_cast<__nv_bfloat16*>(dw);
    p.ldc = p.N;
    p.flags = flags;
    p.workspace = static_cast<float*>(workspace);
    {
        const char* e = getenv("TONY_B200_CONV_DBG");
        p.dbg = e ? atoi(e) : 0;
    }
    CUtensorMap ta, tb;
    if (p.tiled) {
        // both operands as 64-pixel patches of one image; the tap is a coordinate shift, padding the zero fill
        rc = make_patch_tmap(&ta, dy, d->N, P, Q, d->K, p.bw, p.bh, 1, 1);
        if (rc) return rc;
        rc = make_patch_tmap(&tb, x, d->N, d->H, d->W, d->C, p.bw, p.bh, d->stride, d->stride);
        if (rc) return rc;
    } else {
        // dY as [pixels][K]: A(m = co, k = pixel) is MN-major, 64 x 64 boxes
        rc = make_tmap_2d_bf16(&ta, dy, d->K, pixels, d->K, 64, 64);
        if (rc) return rc;
        rc = make_im2col_tmap(&tb, x, d->N, d->H, d->W, d->C, -pad_w, -d->pad, pad_hi_w - (d->S - 1),
                              pad_hi_h - (d->R - 1), d->stride, d->stride, CV_BLOCK_K, d->x_stride_w, d->x_stride_h,
                              d->x_stride_n);
        if (rc) return rc;
    }
    const long long units = tiles * p.splits;
    const int grid = static_cast<int>(std::min<long long>(units, sms));
    rc = dispatch_conv<1>(block_n, ta, tb, p, grid, stream);
    if (rc) return rc;
    if (p.splits > 1)
        return launch_splitk_reduce(static_cast<const float*>(workspace), p.splits, p.M, p.N,
                                    static_cast<__nv_bfloat16*>(dw), p.ldc, 1.0f, flags, stream);
    return TONY_OK;
}

// x: [N, 3, H, W] bf16 NCHW (H, W even) -> xs: [N, H/2, W/2 + 3, 16] bf16 (see stem_space_to_depth_kernel)
extern "C" int tony_stem_space_to_depth(const void* x, void* xs, int N, int H, int W, void* stream) {
    if (!x || !xs || N <= 0 || H <= 0 || W <= 0 || (H & 1) || (W & 1)) return TONY_ERR_INVALID;
    const long long total = static_cast<long long>(N) * (H / 2) * (W / 2 + 3);
    const int grid = static_cast<int>(std::min<long long>((total + 255) / 256, device_sm_count() * 32LL));
    stem_space_to_depth_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        static_cast<const __nv_bfloat16*>(x), static_cast<__nv_bfloat16*>(xs), N, H, W);
    count_launch();
    TONY_CUDA_CHECK(cudaGetLastError());
    return TONY_OK;
}

// x: [N, 3, H, W] bf16 NCHW -> xw: [N, H/2, W/2, 64] bf16, the 4-pixel horizontal window of the space-to-depth
// image written out per output column (dense input for tony_conv2d_fprop_bf16 with R = 4, S = 1, C = 64)
extern "C" int tony_stem_space_to_depth_window(const void* x, void* xw, int N, int H, int W, void* stream) {
    if (!x || !xw || N <= 0 || H <= 0 || W <= 0 || (H & 1) || (W & 1)) return TONY_ERR_INVALID;
    const long long total = static_cast<long long>(N) * (H / 2) * (W / 2) * 4;
    const int grid = static_cast<int>(std::min<long long>((total + 255) / 256, device_sm_count() * 32LL));
    stem_space_to_depth_window_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        static_cast<const __nv_bfloat16*>(x), static_cast<__nv_bfloat16*>(xw), N, H, W);
    count_launch();
    TONY_CUDA_CHECK(cudaGetLastError());
    return TONY_OK;
}
