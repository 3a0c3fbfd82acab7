// tony_b200 — bf16 GEMM on 5th-gen tensor cores (tcgen05 + TMEM + TMA), sm_100a only.
//
//   C[m,n] = epilogue( alpha * sum_k A(m,k) * B(n,k) )
//
// * persistent kernel, one CTA per SM, 576 threads = 18 warps, warp-specialised:
//     warp 0      TMA producer (one elected lane): global -> 128B-swizzled smem ring
//     warp 1      TMEM allocator + tcgen05.mma issuer (one lane), commits to mbarriers
//     warps 2..17 epilogue: tcgen05.ld (TMEM -> registers) -> bias/activation/residual
//                 -> smem transpose -> fully coalesced 16-byte global stores; runs
//                 concurrently with the next tile's MMAs because the accumulator is
//                 double-buffered in TMEM (2 x BLOCK_N columns)
// * both operands may be K-major (reduction dim contiguous) or MN-major (the other
//   dim contiguous), selected at run time; this is what lets forward (X.W^T), dgrad
//   (dY.W) and wgrad (dY^T.X) of a Linear layer run without any transpose copy.
// * optional split-K (fp32 partials in caller-provided workspace + a reduce kernel)
//   for the small-output / long-reduction wgrad shapes that cannot fill 148 SMs.
// * optional "B readiness" flags: the producer acquires a per-n-block flag before
//   loading B tiles; the all-gather-fused GEMM (parallel/ps path, K3 in SURVEY.md §8)
//   uses this to consume weight shards as copy CTAs pull them over NVLink.
//
// The reference (zuston/TonY) contains no GEMM or any device code;
// this kernel replaces cuBLASLt + separate elementwise kernels in the jobs it launches.
#include "common.cuh"
#include "tony_b200_kernels.h"

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <mutex>

namespace tony {

constexpr int BLOCK_M = 128;
constexpr int BLOCK_K = 64;  // 64 bf16 = one 128-byte swizzle row
constexpr int UMMA_K = 16;
constexpr int GEMM_THREADS = 576;   // TMA warp + MMA warp + 16 epilogue warps
constexpr int EPI_WARPS = 16;
constexpr int EPI_STAGING_BYTES = EPI_WARPS * 2048;
constexpr int CHUNK_BYTES = BLOCK_K * 128;  // one [64 x 128B] MN-major chunk = 8 KB

struct GemmParams {
    int M, N, K;
    int a_mn_major, b_mn_major;
    int num_m_blocks, num_n_blocks, num_k_blocks;
    int splits, kb_per_split;
    uint32_t flags;
    float alpha;
    __nv_bfloat16* C;
    long long ldc;
    const __nv_bfloat16* bias;
    const __nv_bfloat16* aux;
    long long ld_aux;
    __nv_bfloat16* aux_out;
    float* workspace;              // split-K partials [splits][M][N]
    const uint32_t* b_ready;       // optional per-n-block readiness flags (gpu scope)
    uint32_t b_ready_epoch;
    int copy_ctas;                 // CTAs [0, copy_ctas) are not GEMM workers (fused AG)
    // fused all-gather: copy CTAs pull peer-owned rows of B into ag_dst ([N,K] row-major)
    const __nv_bfloat16* ag_src[TONY_MAX_RANKS];
    __nv_bfloat16* ag_dst;
    uint32_t* ag_flags;
    int ag_rows_per_rank;
    int n_rot;                     // n-block sweep starts here (the local shard)
    // CTA-pair kernel only: rowsum[m] = sum_k A(m,k) (the bias gradient when A = dY^T in a wgrad
    // GEMM), produced by one extra N=16 MMA per k-step against a tile of ones
    __nv_bfloat16* rowsum_out;     // nullptr = off
    int rowsum_accumulate;
    float* rowsum_ws;              // split-K: partial row sums [splits][M]
    int tma_store;                 // epilogue writes C / aux_out with TMA tensor stores (UTMASTG)
    int group_m;                   // m-blocks per n-sweep of the tile order
};

__device__ __forceinline__ void tile_coords(const GemmParams& p, int tile, int& m_blk, int& n_blk) {
    // grouped ordering: group_m m-blocks share each sweep over n so A stays L2 resident (16 by default;
    // the all-gather GEMM uses 32 so that a few waves of tiles run on the local weight shard before the
    // first remote slab is needed - with 16 the sweep reached the remote n-blocks within the first two
    // waves and the gather was exposed, not hidden)
    const int GROUP_M = p.group_m;
    const int per_group = GROUP_M * p.num_n_blocks;
    const int group = tile / per_group;
    const int within = tile - group * per_group;
    const int first_m = group * GROUP_M;
    const int gm = min(GROUP_M, p.num_m_blocks - first_m);
    m_blk = first_m + within % gm;
    n_blk = within / gm + p.n_rot;
    if (n_blk >= p.num_n_blocks) n_blk -= p.num_n_blocks;
}

// Copy-CTA role of the all-gather-fused GEMM: walk the n-blocks (slabs of BLOCK_N weight rows) in the same
// rotated order the GEMM workers consume them and bring every row that lives on a peer into the local
// full matrix, then publish the slab with a release increment.  Rows are contiguous on both sides, so a
// slab is one (or two) contiguous byte ranges per owner: they move as 16 KB bulk copies - peer memory ->
// shared memory (cp.async.bulk + mbarrier), shared -> local global (cp.async.bulk store) - through a ring
// in the CTA's otherwise unused operand buffers, RING / 2 loads in flight per CTA.  (The first version
// copied with register loads, 8 x 16 B per thread per trip: one NVLink round trip per 74 KB per CTA.)
template <int RING>
__device__ __forceinline__ void allgather_copy_role(const GemmParams& p, int block_n, uint8_t* ring, uint64_t* bars) {
    constexpr uint32_t CHUNK = 16384;
    constexpr int AHEAD = RING / 2;
    constexpr int QN = 32;                      // event queue (loads in flight + slab markers between them)
    if (threadIdx.x != 0) return;
    const size_t row_bytes = static_cast<size_t>(p.K) * 2;
    const int me = static_cast<int>(blockIdx.x), ncopy = p.copy_ctas;
    uint8_t* dst_base = reinterpret_cast<uint8_t*>(p.ag_dst);

    // ---- walker: this CTA's chunks in sweep order, one division per slab / owner range, none per chunk ----
    int wj = -1;                 // slab position in the rotated order (-1: not started)
    int w_nb = 0;                // its n-block index
    int w_row = 0, w_row_end = 0;   // rows of the slab not yet turned into a range
    size_t r_off = 0, r_len = 0; // current owner range: byte offset in the full matrix, length
    const uint8_t* r_src = nullptr;
    size_t r_k = 0, r_chunks = 0;   // next chunk of the range that is mine, number of chunks in the range
    int rr = 0;                  // round-robin phase so ranges shorter than ncopy chunks still spread evenly
    bool w_done = false;

    // events
    size_t q_off[QN];
    uint32_t q_bytes[QN];        // 0 = slab marker (q_off holds the n-block)
    int qh = 0, qt = 0;
    long long n_loaded = 0, n_stored = 0;
    int in_flight = 0;

    auto next_event = [&](size_t& off, uint32_t& bytes, const uint8_t*& src) -> bool {
        // true: produced an event (bytes == 0: slab marker with off = n-block); false: sweep finished
        for (;;) {
            if (r_k < r_chunks) {
                const size_t o = r_k * CHUNK;
                off = r_off + o;
                bytes = static_cast<uint32_t>(r_len - o < CHUNK ? r_len - o : CHUNK);
                src = r_src + o;
                r_k += ncopy;
                return true;
            }
            if (w_row < w_row_end) {
                // next owner range of the slab
                const int owner = w_row / p.ag_rows_per_rank;
                const int stop = min(w_row_end, (owner + 1) * p.ag_rows_per_rank);
                r_off = static_cast<size_t>(w_row) * row_bytes;
                r_len = static_cast<size_t>(stop - w_row) * row_bytes;
                r_src = reinterpret_cast<const uint8_t*>(p.ag_src[owner]) +
                        static_cast<size_t>(w_row - owner * p.ag_rows_per_rank) * row_bytes;
                w_row = stop;
                if (r_src == dst_base + r_off) {          // rows this rank owns are already in place
                    r_k = r_chunks = 0;
                    continue;
                }
                r_chunks = (r_len + CHUNK - 1) / CHUNK;
                r_k = static_cast<size_t>((me - rr + ncopy) % ncopy);
                rr = static_cast<int>((rr + r_chunks) % ncopy);
                continue;
            }
            if (wj >= 0 && w_row_end > 0) {               // slab finished: marker, once
                off = static_cast<size_t>(w_nb);
                bytes = 0;
                src = nullptr;
                w_row_end = 0;
                return true;
            }
            if (++wj >= p.num_n_blocks) {
                w_done = true;
                return false;
            }
            w_nb = p.n_rot + wj;
            if (w_nb >= p.num_n_blocks) w_nb -= p.num_n_blocks;
            w_row = w_nb * block_n;
            w_row_end = min(p.N, w_row + block_n);
            r_k = r_chunks = 0;
        }
    };

    for (;;) {
        // fill: keep AHEAD loads in flight (markers ride along in the queue)
        while (!w_done && in_flight < AHEAD && ((qt + 1) % QN) != qh) {
            size_t off;
            uint32_t bytes;
            const uint8_t* src;
            if (!next_event(off, bytes, src)) break;
            q_off[qt] = off;
            q_bytes[qt] = bytes;
            qt = (qt + 1) % QN;
            if (bytes) {
                const int slot = static_cast<int>(n_loaded % RING);
                mbar_expect_tx(&bars[slot], bytes);
                bulk_load_1d(smem_u32(ring + slot * CHUNK), src, bytes, &bars[slot]);
                ++n_loaded;
                ++in_flight;
            }
        }
        if (qh == qt) break;
        const size_t off = q_off[qh];
        const uint32_t bytes = q_bytes[qh];
        qh = (qh + 1) % QN;
        if (bytes) {
            const int slot = static_cast<int>(n_stored % RING);
            mbar_wait(&bars[slot], static_cast<uint32_t>((n_stored / RING) & 1));
            bulk_store_1d(dst_base + off, smem_u32(ring + slot * CHUNK), bytes);
            tma_store_commit();
            ++n_stored;
            --in_flight;
            bulk_wait_read<AHEAD - 1>();      // the slot the next load reuses has been read out
        } else {
            tma_store_wait_all();             // this CTA's part of the slab is in local memory
            fence_proxy_async();
            __threadfence();
            asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(p.ag_flags + off) : "memory");
        }
    }
}

// ---- epilogue --------------------------------------------------------------------------
// Each epilogue warp owns 32 accumulator rows (its TMEM lane quarter) and walks 32-column
// chunks.  TMEM hands every lane one ROW; global memory wants every store instruction to
// cover whole sectors of consecutive rows.  A per-warp 2 KB staging tile ([32 rows][64 B],
// 16-byte units XOR-swizzled by row pair so both access patterns are bank-conflict free)
// transposes between the two: every global load/store below moves 8 rows x 64 B per
// instruction.  Sixteen such warps (four per SM sub-partition) hide each other's latency.
__device__ __forceinline__ uint32_t stg_addr(uint32_t base, int row, int unit) {
    return base + row * 64 + ((unit ^ ((row >> 1) & 3)) << 4);
}

// global [32 x 32] bf16 tile -> this lane's row as 16 packed bf16x2 words
__device__ __forceinline__ void load_tile_rows(const __nv_bfloat16* g, long long ld, int row0, int col0,
                                               int M, int N, uint32_t stg, int lane, uint32_t (&out)[16]) {
    const int unit = lane & 3;
    const int gc = col0 + unit * 8;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int row = i * 8 + (lane >> 2);
        const int gr = row0 + row;
        uint4 v = make_uint4(0, 0, 0, 0);
        if (gr < M && gc < N) v = ld_global_v4(g + static_cast<size_t>(gr) * ld + gc);
        st_shared_v4(stg_addr(stg, row, unit), v);
    }
    __syncwarp();
#pragma unroll
    for (int u = 0; u < 4; ++u) {
        const uint4 v = ld_shared_v4(stg_addr(stg, lane, u));
        out[4 * u] = v.x;
        out[4 * u + 1] = v.y;
        out[4 * u + 2] = v.z;
        out[4 * u + 3] = v.w;
    }
    __syncwarp();
}

// this lane's row (16 packed words) -> global [32 x 32] bf16 tile, coalesced
__device__ __forceinline__ void store_tile_rows(__nv_bfloat16* g, long long ld, int row0, int col0, int M,
                                                int N, uint32_t stg, int lane, const uint32_t (&in)[16]) {
#pragma unroll
    for (int u = 0; u < 4; ++u)
        st_shared_v4(stg_addr(stg, lane, u), make_uint4(in[4 * u], in[4 * u + 1], in[4 * u + 2], in[4 * u + 3]));
    __syncwarp();
    const int unit = lane & 3;
    const int gc = col0 + unit * 8;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int row = i * 8 + (lane >> 2);
        const int gr = row0 + row;
        const uint4 v = ld_shared_v4(stg_addr(stg, row, unit));
        if (gr < M && gc < N) st_global_v4(g + static_cast<size_t>(gr) * ld + gc, v);
    }
    __syncwarp();
}

// this lane's row (16 packed words) -> the warp's staging tile -> ONE TMA tensor store of the 32 x 32 tile
// (64-byte swizzle = the staging tile's XOR pattern; rows / columns past M / N are clipped by the map)
__device__ __forceinline__ void store_tile_tma(const CUtensorMap* map, int row0, int col0, uint32_t stg, int lane,
                                               const uint32_t (&in)[16]) {
#pragma unroll
    for (int u = 0; u < 4; ++u)
        st_shared_v4(stg_addr(stg, lane, u), make_uint4(in[4 * u], in[4 * u + 1], in[4 * u + 2], in[4 * u + 3]));
    fence_proxy_async_smem();      // generic-proxy writes -> async-proxy (TMA) read
    __syncwarp();
    if (elect_one()) {     // converged warp, uniform operands: no per-lane broadcast loop around UTMASTG
        tma_store_2d(map, stg, col0, row0);
        tma_store_commit();
    }
}

// before the staging tile is written again: the previous TMA store must have finished reading it
__device__ __forceinline__ void staging_acquire(int lane) {
    if (elect_one()) tma_store_wait_read();
    __syncwarp();
}

__device__ __forceinline__ void epilogue_chunk32(const GemmParams& p, const CUtensorMap* tmap_c,
                                                 const CUtensorMap* tmap_aux, const uint32_t (&r)[32], int row0,
                                                 int col0, uint32_t stg, int lane) {
    const uint32_t flags = p.flags;
    if (p.tma_store) staging_acquire(lane);
    float v[32];
    if (flags & TONY_EPI_BIAS) {
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            uint4 b = make_uint4(0, 0, 0, 0);
            if (col0 + u * 8 < p.N) b = __ldg(reinterpret_cast<const uint4*>(p.bias + col0 + u * 8));
            const uint32_t bw[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float2 f = unpack_bf16x2(bw[j]);
                v[u * 8 + 2 * j] = fmaf(__uint_as_float(r[u * 8 + 2 * j]), p.alpha, f.x);
                v[u * 8 + 2 * j + 1] = fmaf(__uint_as_float(r[u * 8 + 2 * j + 1]), p.alpha, f.y);
            }
        }
    } else {
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]) * p.alpha;
    }
    if (flags & (TONY_EPI_RESIDUAL | TONY_EPI_DGELU_ERF | TONY_EPI_DGELU_TANH | TONY_EPI_DRELU)) {
        uint32_t a[16];
        load_tile_rows(p.aux, p.ld_aux, row0, col0, p.M, p.N, stg, lane, a);
#pragma unroll
        for (int j = 0; j < 16; ++j) {
            const float2 f = unpack_bf16x2(a[j]);
            if (flags & TONY_EPI_RESIDUAL) {
                v[2 * j] += f.x;
                v[2 * j + 1] += f.y;
            } else if (flags & TONY_EPI_DGELU_ERF) {
                v[2 * j] *= dgelu_erf(f.x);
                v[2 * j + 1] *= dgelu_erf(f.y);
            } else if (flags & TONY_EPI_DGELU_TANH) {
                v[2 * j] *= dgelu_tanh(f.x);
                v[2 * j + 1] *= dgelu_tanh(f.y);
            } else {
                v[2 * j] = f.x > 0.f ? v[2 * j] : 0.f;
                v[2 * j + 1] = f.y > 0.f ? v[2 * j + 1] : 0.f;
            }
        }
    }
    uint32_t o[16];
    if (flags & TONY_EPI_SAVE_PREACT) {
#pragma unroll
        for (int j = 0; j < 16; ++j) o[j] = pack_bf16x2(v[2 * j], v[2 * j + 1]);
        if (p.tma_store) {
            store_tile_tma(tmap_aux, row0, col0, stg, lane, o);
            staging_acquire(lane);
        } else {
            store_tile_rows(p.aux_out, p.ldc, row0, col0, p.M, p.N, stg, lane, o);
        }
    }
    if (flags & TONY_EPI_GELU_ERF) {
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = gelu_erf(v[j]);
    } else if (flags & TONY_EPI_GELU_TANH) {
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = gelu_tanh(v[j]);
    } else if (flags & TONY_EPI_RELU) {
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = fmaxf(v[j], 0.f);
    }
    if (flags & TONY_EPI_ACCUMULATE) {
        uint32_t c[16];
        load_tile_rows(p.C, p.ldc, row0, col0, p.M, p.N, stg, lane, c);
#pragma unroll
        for (int j = 0; j < 16; ++j) {
            const float2 f = unpack_bf16x2(c[j]);
            v[2 * j] += f.x;
            v[2 * j + 1] += f.y;
        }
    }
#pragma unroll
    for (int j = 0; j < 16; ++j) o[j] = pack_bf16x2(v[2 * j], v[2 * j + 1]);
    if (p.tma_store)
        store_tile_tma(tmap_c, row0, col0, stg, lane, o);
    else
        store_tile_rows(p.C, p.ldc, row0, col0, p.M, p.N, stg, lane, o);
}

template <int BLOCK_N, int STAGES>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_bf16_tcgen05_kernel(const __grid_constant__ CUtensorMap tmap_a,
                         const __grid_constant__ CUtensorMap tmap_b, const __grid_constant__ CUtensorMap tmap_c,
                         const __grid_constant__ CUtensorMap tmap_aux, const GemmParams p) {
    constexpr uint32_t A_BYTES = BLOCK_M * BLOCK_K * 2;
    constexpr uint32_t B_BYTES = BLOCK_N * BLOCK_K * 2;
    constexpr uint32_t STAGE_BYTES = A_BYTES + B_BYTES;
    constexpr uint32_t TMEM_COLS = 2 * BLOCK_N;  // double-buffered fp32 accumulator

    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>(
        (reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
    uint8_t* staging = smem + STAGES * STAGE_BYTES;
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(staging + EPI_STAGING_BYTES);
    uint64_t* empty_bar = full_bar + STAGES;
    uint64_t* tmem_full_bar = empty_bar + STAGES;
    uint64_t* tmem_empty_bar = tmem_full_bar + 2;
    uint32_t* tmem_base_slot = reinterpret_cast<uint32_t*>(tmem_empty_bar + 2);

    const int warp_idx = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);   // provably warp-uniform: role branches and their loops stay on the uniform datapath
    const int lane = threadIdx.x & 31;

    if (warp_idx == 0 && lane == 0) {
        tma_prefetch_desc(&tmap_a);
        tma_prefetch_desc(&tmap_b);
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        for (int s = 0; s < 2; ++s) {
            mbar_init(&tmem_full_bar[s], 1);
            mbar_init(&tmem_empty_bar[s], EPI_WARPS);  // one arrive per epilogue warp
        }
        fence_barrier_init();
    }
    if (warp_idx == 1) tmem_alloc<TMEM_COLS>(tmem_base_slot);
    tcgen05_fence_before();
    __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_base_slot;
    pdl_launch_dependents();  // the next kernel may start its own prologue on idle SMs

    const int num_tiles = p.num_m_blocks * p.num_n_blocks;
    const int num_units = num_tiles * p.splits;
    const int worker = static_cast<int>(blockIdx.x) - p.copy_ctas;
    const int num_workers = static_cast<int>(gridDim.x) - p.copy_ctas;

    if (worker >= 0) {
        if (warp_idx == 0) {
            // ------------------------------------------------ TMA producer
            // The whole warp runs the loop and ONE elected lane issues (here and in the MMA warp): the
            // operands of UTMALDG / UTCHMMA / UTCBAR are then warp-uniform values in uniform registers.
            // Inside an `if (lane == 0)` region the compiler cannot prove that and wraps every such
            // instruction in an ELECT / R2UR.BROADCAST / BRA.U.ANY loop - about 25 instructions per MMA,
            // which made the issuing thread, not the tensor core, the limit of the main loop.
            {
                int stage = 0;
                uint32_t phase = 0;
                pdl_wait();  // everything above overlapped the previous kernel's tail
                for (int unit = worker; unit < num_units; unit += num_workers) {
                    const int tile = unit / p.splits;
                    const int split = unit - tile * p.splits;
                    int m_blk, n_blk;
                    tile_coords(p, tile, m_blk, n_blk);
                    if (p.b_ready != nullptr) {
                        if (elect_one()) {
                            while (static_cast<int32_t>(ld_acquire_gpu(p.b_ready + n_blk) -
                                                        p.b_ready_epoch) < 0) {
                                __nanosleep(64);
                            }
                            fence_proxy_async();  // generic-proxy copies -> async-proxy TMA reads
                        }
                        __syncwarp();
                    }
                    const int kb0 = split * p.kb_per_split;
                    const int kb1 = min(p.num_k_blocks, kb0 + p.kb_per_split);
                    for (int kb = kb0; kb < kb1; ++kb) {
                        mbar_wait(&empty_bar[stage], phase ^ 1);
                        uint8_t* sa = smem + stage * STAGE_BYTES;
                        uint8_t* sb = sa + A_BYTES;
                        if (elect_one()) {
                        mbar_expect_tx(&full_bar[stage], STAGE_BYTES);
                        if (!p.a_mn_major) {
                            tma_load_2d(&tmap_a, &full_bar[stage], sa, kb * BLOCK_K, m_blk * BLOCK_M);
                        } else {
#pragma unroll
                            for (int c = 0; c < BLOCK_M / 64; ++c)
                                tma_load_2d(&tmap_a, &full_bar[stage], sa + c * CHUNK_BYTES,
                                            m_blk * BLOCK_M + c * 64, kb * BLOCK_K);
                        }
                        if (!p.b_mn_major) {
                            tma_load_2d(&tmap_b, &full_bar[stage], sb, kb * BLOCK_K, n_blk * BLOCK_N);
                        } else {
#pragma unroll
                            for (int c = 0; c < BLOCK_N / 64; ++c)
                                tma_load_2d(&tmap_b, &full_bar[stage], sb + c * CHUNK_BYTES,
                                            n_blk * BLOCK_N + c * 64, kb * BLOCK_K);
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
        } else if (warp_idx == 1) {
            // ------------------------------------------------ MMA issuer (whole warp, elected lane issues)
            {
                const uint32_t idesc = make_idesc_bf16(BLOCK_M, BLOCK_N, p.a_mn_major, p.b_mn_major);
                // K-major: 8-row groups 1024 B apart, k-step = 32 B inside the swizzle row.
                // MN-major: 64-wide chunks CHUNK_BYTES apart (LBO), 8 k-rows = 1024 B (SBO),
                //           k-step = 16 rows = 2048 B.
                const uint32_t a_lbo = p.a_mn_major ? CHUNK_BYTES : 16;
                const uint32_t b_lbo = p.b_mn_major ? CHUNK_BYTES : 16;
                const uint32_t a_kstep = p.a_mn_major ? (UMMA_K * 128) >> 4 : (UMMA_K * 2) >> 4;
                const uint32_t b_kstep = p.b_mn_major ? (UMMA_K * 128) >> 4 : (UMMA_K * 2) >> 4;
                int stage = 0;
                uint32_t phase = 0;
                int iter = 0;
                for (int unit = worker; unit < num_units; unit += num_workers, ++iter) {
                    const int tile = unit / p.splits;
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
                        const uint32_t sa = smem_u32(smem + stage * STAGE_BYTES);
                        const uint32_t sb = sa + A_BYTES;
                        const uint64_t da = make_umma_desc_sw128(sa, a_lbo, 1024);
                        const uint64_t db = make_umma_desc_sw128(sb, b_lbo, 1024);
                        if (elect_one()) {
#pragma unroll
                            for (int k = 0; k < BLOCK_K / UMMA_K; ++k) {
                                umma_bf16(d_tmem, da + k * a_kstep, db + k * b_kstep, idesc,
                                          (kb > kb0 || k > 0) ? 1u : 0u);
                            }
                            umma_commit(&empty_bar[stage]);  // frees the smem slot when MMAs retire
                        }
                        __syncwarp();
                        if (++stage == STAGES) {
                            stage = 0;
                            phase ^= 1;
                        }
                    }
                    if (elect_one()) umma_commit(&tmem_full_bar[as]);  // accumulator complete -> epilogue
                    __syncwarp();
                }
            }
        } else {
            // ------------------------------------------------ epilogue warps (2..17)
            // TMEM lane quarter = warp_idx % 4 (hardware rule); four warps share a quarter and
            // deal its 32-column chunks round-robin.
            const int quarter = warp_idx & 3;
            const int group = (warp_idx - 2) >> 2;
            const uint32_t stg = smem_u32(staging + (warp_idx - 2) * 2048);
            int iter = 0;
            for (int unit = worker; unit < num_units; unit += num_workers, ++iter) {
                const int tile = unit / p.splits;
                const int split = unit - tile * p.splits;
                int m_blk, n_blk;
                tile_coords(p, tile, m_blk, n_blk);
                const int as = iter & 1;
                const uint32_t aphase = (iter >> 1) & 1;
                mbar_wait(&tmem_full_bar[as], aphase);
                tcgen05_fence_after();
                const int row0 = m_blk * BLOCK_M + quarter * 32;
                const uint32_t taddr0 =
                    tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + as * BLOCK_N;
#pragma unroll 1
                for (int c = group; c < BLOCK_N / 32; c += EPI_WARPS / 4) {
                    uint32_t r[32];
                    tmem_ld_32x32(taddr0 + c * 32, r);
                    tmem_ld_wait();
                    const int col0 = n_blk * BLOCK_N + c * 32;
                    if (col0 >= p.N || row0 >= p.M) continue;  // warp-uniform
                    if (p.splits > 1) {
                        const int row = row0 + lane;
                        if (row < p.M) {
                            float* dst = p.workspace + (static_cast<size_t>(split) * p.M + row) * p.N + col0;
#pragma unroll
                            for (int g = 0; g < 8; ++g) {
                                if (col0 + g * 4 < p.N)
                                    *reinterpret_cast<float4*>(dst + g * 4) = make_float4(
                                        __uint_as_float(r[g * 4]), __uint_as_float(r[g * 4 + 1]),
                                        __uint_as_float(r[g * 4 + 2]), __uint_as_float(r[g * 4 + 3]));
                            }
                        }
                        continue;
                    }
                    epilogue_chunk32(p, &tmap_c, &tmap_aux, r, row0, col0, stg, lane);
                }
                tcgen05_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&tmem_empty_bar[as]);
            }
            if (p.tma_store && elect_one()) tma_store_wait_all();   // shared memory stays valid until read
        }
    }
    else {
        // ring slots: the 2 * STAGES single-arrival barriers (full + empty) and the operand buffers, unused here
        pdl_wait();
        allgather_copy_role<2 * STAGES>(p, BLOCK_N, smem, full_bar);
    }

    __syncwarp();
    tcgen05_fence_before();
    __syncthreads();
    if (warp_idx == 1) {
        tcgen05_fence_after();
        tmem_dealloc<TMEM_COLS>(tmem_base);
    }
}

// ----------------------------------------------------------------------------------------
// CTA-pair variant: two SMs (one cluster of 2) compute one 256 x 256 tile.
//   * each CTA TMA-loads only ITS half of both operands (128 rows of A, 128 rows of B), so
//     a k-block costs 32 KB of L2->smem traffic per SM instead of 48 KB, and the smem ring is
//     6 deep instead of 4;
//   * the leader CTA issues tcgen05.mma.cta_group::2 (M = 256, N = 256): the tensor cores of
//     both SMs read A/B halves from both shared memories; each CTA's TMEM receives its 128
//     rows x 256 columns of the accumulator;
//   * TMA completions of both CTAs are credited to the leader's `full` barrier; the leader's
//     tcgen05.commit multicasts the `empty` / `tmem_full` arrivals to both CTAs; the epilogue
//     warps of both CTAs arrive remotely on the leader's `tmem_empty` barrier.
// Per-CTA shared-memory layout and descriptors are exactly those of the 128 x 128 1-CTA kernel.
template <int STAGES, bool ROWSUM>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(GEMM_THREADS, 1)
gemm_bf16_tcgen05_2cta_kernel(const __grid_constant__ CUtensorMap tmap_a,
                              const __grid_constant__ CUtensorMap tmap_b, const __grid_constant__ CUtensorMap tmap_c,
                              const __grid_constant__ CUtensorMap tmap_aux, const GemmParams p) {
    constexpr int HALF_N = 128;                       // B rows held by each CTA
    constexpr int TILE_N = 256;
    constexpr int TILE_M = 256;
    constexpr uint32_t A_BYTES = BLOCK_M * BLOCK_K * 2;
    constexpr uint32_t B_BYTES = HALF_N * BLOCK_K * 2;
    constexpr uint32_t STAGE_BYTES = A_BYTES + B_BYTES;
    constexpr uint32_t TMEM_COLS = 2 * TILE_N;

    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>(
        (reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
    uint8_t* staging = smem + STAGES * STAGE_BYTES;
    // ROWSUM only: 8 rows x 128 B of bf16 1.0 (this CTA's half of a 16 x 64 operand of ones)
    uint8_t* ones_tile = staging + EPI_STAGING_BYTES;
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(ones_tile + (ROWSUM ? 1024 : 0));
    uint64_t* empty_bar = full_bar + STAGES;
    uint64_t* tmem_full_bar = empty_bar + STAGES;
    uint64_t* tmem_empty_bar = tmem_full_bar + 2;
    uint32_t* tmem_base_slot = reinterpret_cast<uint32_t*>(tmem_empty_bar + 2);

    const int warp_idx = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);   // provably warp-uniform: role branches and their loops stay on the uniform datapath
    const int lane = threadIdx.x & 31;
    const uint32_t cta_rank = cluster_ctarank();
    const bool leader = cta_rank == 0;
    // Row-sum mode keeps ONE accumulator stage (columns [0,256)) and puts the 16-column row-sum
    // accumulator at column 256; the reduction is so long in the GEMMs that use it (wgrad, K = tokens)
    // that the lost epilogue overlap is noise.
    constexpr bool rowsum = ROWSUM;      // a separate instantiation: the plain kernel's code is untouched
    constexpr int nstage = ROWSUM ? 1 : 2;
    if (ROWSUM && warp_idx == 2) {
        reinterpret_cast<uint4*>(ones_tile)[lane] = make_uint4(0x3F803F80u, 0x3F803F80u, 0x3F803F80u, 0x3F803F80u);
        reinterpret_cast<uint4*>(ones_tile)[lane + 32] = make_uint4(0x3F803F80u, 0x3F803F80u, 0x3F803F80u, 0x3F803F80u);
        fence_proxy_async();   // generic-proxy stores -> tensor-core (async proxy) reads
    }

    if (warp_idx == 0 && lane == 0) {
        tma_prefetch_desc(&tmap_a);
        tma_prefetch_desc(&tmap_b);
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(&full_bar[s], 1);    // leader's producer arrive.expect_tx (both CTAs' bytes)
            mbar_init(&empty_bar[s], 1);   // multicast commit from the leader's MMA thread
        }
        for (int s = 0; s < 2; ++s) {
            mbar_init(&tmem_full_bar[s], 1);
            mbar_init(&tmem_empty_bar[s], 2 * EPI_WARPS);  // epilogue warps of BOTH CTAs
        }
        fence_barrier_init();
    }
    if (warp_idx == 1) tmem_alloc_2cta<TMEM_COLS>(tmem_base_slot);
    tcgen05_fence_before();
    cluster_sync_all();   // peer barriers initialised before any remote arrive / TMA credit
    tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_base_slot;
    pdl_launch_dependents();

    const int num_tiles = p.num_m_blocks * p.num_n_blocks;   // blocks of 256 x 256 here
    const int num_units = num_tiles * p.splits;
    // fused all-gather: the first copy_ctas CTAs (whole clusters) pull the remote weight shards, the rest work
    const int copy_clusters = p.copy_ctas >> 1;
    const int cluster_id = static_cast<int>(blockIdx.x >> 1) - copy_clusters;
    const int num_clusters = static_cast<int>(gridDim.x >> 1) - copy_clusters;

    if (cluster_id < 0) {
        pdl_wait();
        allgather_copy_role<2 * STAGES>(p, TILE_N, smem, full_bar);
    } else if (warp_idx == 0) {
        // ---------------------------------------------------- TMA producer (both CTAs; whole warp, one lane issues)
        {
            int stage = 0;
            uint32_t phase = 0;
            pdl_wait();
            for (int unit = cluster_id; unit < num_units; unit += num_clusters) {
                const int tile = unit / p.splits;
                const int split = unit - tile * p.splits;
                int m_blk, n_blk;
                tile_coords(p, tile, m_blk, n_blk);
                if (p.b_ready != nullptr) {
                    if (elect_one()) {
                        while (static_cast<int32_t>(ld_acquire_gpu(p.b_ready + n_blk) - p.b_ready_epoch) < 0)
                            __nanosleep(64);
                        fence_proxy_async();  // generic / bulk-copy writes -> this thread's TMA reads
                    }
                    __syncwarp();
                }
                const int m0 = m_blk * TILE_M + static_cast<int>(cta_rank) * BLOCK_M;
                const int n0 = n_blk * TILE_N + static_cast<int>(cta_rank) * HALF_N;
                const int kb0 = split * p.kb_per_split;
                const int kb1 = min(p.num_k_blocks, kb0 + p.kb_per_split);
                for (int kb = kb0; kb < kb1; ++kb) {
                    mbar_wait(&empty_bar[stage], phase ^ 1);
                    uint8_t* sa = smem + stage * STAGE_BYTES;
                    uint8_t* sb = sa + A_BYTES;
                    const uint32_t bar = mapa_u32(smem_u32(&full_bar[stage]), 0);  // leader's barrier
                    if (elect_one()) {
                    if (leader) mbar_expect_tx(&full_bar[stage], 2 * STAGE_BYTES);
                    if (!p.a_mn_major) {
                        tma_load_2d_2cta(&tmap_a, bar, sa, kb * BLOCK_K, m0);
                    } else {
#pragma unroll
                        for (int c = 0; c < BLOCK_M / 64; ++c)
                            tma_load_2d_2cta(&tmap_a, bar, sa + c * CHUNK_BYTES, m0 + c * 64, kb * BLOCK_K);
                    }
                    if (!p.b_mn_major) {
                        tma_load_2d_2cta(&tmap_b, bar, sb, kb * BLOCK_K, n0);
                    } else {
#pragma unroll
                        for (int c = 0; c < HALF_N / 64; ++c)
                            tma_load_2d_2cta(&tmap_b, bar, sb + c * CHUNK_BYTES, n0 + c * 64, kb * BLOCK_K);
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
    } else if (warp_idx == 1) {
        // ---------------------------------------------------- MMA issuer (leader CTA only; whole warp, one lane issues)
        if (leader) {
            const uint32_t idesc = make_idesc_bf16(TILE_M, TILE_N, p.a_mn_major, p.b_mn_major);
            const uint32_t a_lbo = p.a_mn_major ? CHUNK_BYTES : 16;
            const uint32_t b_lbo = p.b_mn_major ? CHUNK_BYTES : 16;
            const uint32_t a_kstep = p.a_mn_major ? (UMMA_K * 128) >> 4 : (UMMA_K * 2) >> 4;
            const uint32_t b_kstep = p.b_mn_major ? (UMMA_K * 128) >> 4 : (UMMA_K * 2) >> 4;
            const uint32_t idesc_ones = make_idesc_bf16(TILE_M, 16, p.a_mn_major, 0);
            const uint64_t d_ones = make_umma_desc_sw128(smem_u32(ones_tile), 16, 1024);
            int stage = 0;
            uint32_t phase = 0;
            int iter = 0;
            for (int unit = cluster_id; unit < num_units; unit += num_clusters, ++iter) {
                const int tile = unit / p.splits;
                const int split = unit - tile * p.splits;
                const int as = iter % nstage;
                const uint32_t aphase = (iter / nstage) & 1;
                mbar_wait(&tmem_empty_bar[as], aphase ^ 1);
                tcgen05_fence_after();
                const uint32_t d_tmem = tmem_base + as * TILE_N;
                int m_blk_i, n_blk_i;
                tile_coords(p, tile, m_blk_i, n_blk_i);
                const bool do_rowsum = rowsum && n_blk_i == 0;   // one column of tiles owns the row sums
                const int kb0 = split * p.kb_per_split;
                const int kb1 = min(p.num_k_blocks, kb0 + p.kb_per_split);
                for (int kb = kb0; kb < kb1; ++kb) {
                    mbar_wait(&full_bar[stage], phase);
                    tcgen05_fence_after();
                    const uint32_t sa = smem_u32(smem + stage * STAGE_BYTES);
                    const uint32_t sb = sa + A_BYTES;
                    const uint64_t da = make_umma_desc_sw128(sa, a_lbo, 1024);
                    const uint64_t db = make_umma_desc_sw128(sb, b_lbo, 1024);
                    if (elect_one()) {
#pragma unroll
                        for (int k = 0; k < BLOCK_K / UMMA_K; ++k) {
                            umma_bf16_2cta(d_tmem, da + k * a_kstep, db + k * b_kstep, idesc,
                                           (kb > kb0 || k > 0) ? 1u : 0u);
                        }
                        if (do_rowsum) {
                            // D2[256 x 16] += A_tile[256 x 64] . ones[64 x 16]: every column is the row sum
#pragma unroll
                            for (int k = 0; k < BLOCK_K / UMMA_K; ++k)
                                umma_bf16_2cta(tmem_base + TILE_N, da + k * a_kstep, d_ones, idesc_ones,
                                               (kb > kb0 || k > 0) ? 1u : 0u);
                        }
                        umma_commit_2cta(&empty_bar[stage], 3);   // frees the slot in both CTAs
                    }
                    __syncwarp();
                    if (++stage == STAGES) {
                        stage = 0;
                        phase ^= 1;
                    }
                }
                if (elect_one()) umma_commit_2cta(&tmem_full_bar[as], 3);      // accumulators complete in both CTAs
                __syncwarp();
            }
        }
    } else {
        // ---------------------------------------------------- epilogue warps (both CTAs)
        const int quarter = warp_idx & 3;
        const int group = (warp_idx - 2) >> 2;
        const uint32_t stg = smem_u32(staging + (warp_idx - 2) * 2048);
        int iter = 0;
        for (int unit = cluster_id; unit < num_units; unit += num_clusters, ++iter) {
            const int tile = unit / p.splits;
            const int split = unit - tile * p.splits;
            int m_blk, n_blk;
            tile_coords(p, tile, m_blk, n_blk);
            const int as = iter % nstage;
            const uint32_t aphase = (iter / nstage) & 1;
            mbar_wait(&tmem_full_bar[as], aphase);
            tcgen05_fence_after();
            const int row0 = m_blk * TILE_M + static_cast<int>(cta_rank) * BLOCK_M + quarter * 32;
            const uint32_t taddr0 = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + as * TILE_N;
#pragma unroll 1
            for (int c = group; c < TILE_N / 32; c += EPI_WARPS / 4) {
                uint32_t r[32];
                tmem_ld_32x32(taddr0 + c * 32, r);
                tmem_ld_wait();
                const int col0 = n_blk * TILE_N + c * 32;
                if (col0 >= p.N || row0 >= p.M) continue;  // warp-uniform
                if (p.splits > 1) {
                    const int row = row0 + lane;
                    if (row < p.M) {
                        float* dst = p.workspace + (static_cast<size_t>(split) * p.M + row) * p.N + col0;
#pragma unroll
                        for (int g = 0; g < 8; ++g) {
                            if (col0 + g * 4 < p.N)
                                *reinterpret_cast<float4*>(dst + g * 4) = make_float4(
                                    __uint_as_float(r[g * 4]), __uint_as_float(r[g * 4 + 1]),
                                    __uint_as_float(r[g * 4 + 2]), __uint_as_float(r[g * 4 + 3]));
                        }
                    }
                    continue;
                }
                epilogue_chunk32(p, &tmap_c, &tmap_aux, r, row0, col0, stg, lane);
            }
            if (rowsum && n_blk == 0 && group == 0) {
                // one warp per TMEM lane quarter: column 0 of the row-sum accumulator is this lane's row
                uint32_t r[32];
                tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + TILE_N, r);
                tmem_ld_wait();
                const int row = row0 + lane;
                if (row < p.M) {
                    const float v = __uint_as_float(r[0]);
                    if (p.splits > 1) {
                        p.rowsum_ws[static_cast<size_t>(split) * p.M + row] = v;
                    } else {
                        float t = v;
                        if (p.rowsum_accumulate) t += __bfloat162float(p.rowsum_out[row]);
                        p.rowsum_out[row] = __float2bfloat16(t);
                    }
                }
            }
            tcgen05_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_cluster(mapa_u32(smem_u32(&tmem_empty_bar[as]), 0));
        }
        if (p.tma_store && elect_one()) tma_store_wait_all();
    }

    __syncwarp();
    tcgen05_fence_before();
    cluster_sync_all();   // both CTAs done with TMEM and with each other's barriers
    if (warp_idx == 1) {
        tcgen05_fence_after();
        tmem_dealloc_2cta<TMEM_COLS>(tmem_base);
    }
}

// out = alpha * sum_s ws[s] (+ old out) (+ bias), fp32 partials -> bf16
__global__ void splitk_reduce_kernel(const float* __restrict__ ws, int splits, int M, int N,
                                     __nv_bfloat16* __restrict__ C, long long ldc, float alpha,
                                     uint32_t flags, const __nv_bfloat16* __restrict__ bias,
                                     const float* __restrict__ rowsum_ws, __nv_bfloat16* __restrict__ rowsum_out,
                                     int rowsum_accumulate) {
    pdl_wait();                // launched with programmatic stream serialization: see launch_pdl_kernel
    pdl_launch_dependents();
    if (rowsum_out) {   // partial row sums of the split-K units (see GemmParams::rowsum_out)
        for (size_t r = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; r < static_cast<size_t>(M);
             r += static_cast<size_t>(gridDim.x) * blockDim.x) {
            float t = 0.f;
            for (int sp = 0; sp < splits; ++sp) t += rowsum_ws[static_cast<size_t>(sp) * M + r];
            if (rowsum_accumulate) t += __bfloat162float(rowsum_out[r]);
            rowsum_out[r] = __float2bfloat16(t);
        }
    }
    const size_t total4 = static_cast<size_t>(M) * N / 4;
    const size_t plane = static_cast<size_t>(M) * N;
    for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < total4;
         i += static_cast<size_t>(gridDim.x) * blockDim.x) {
        const size_t e = i * 4;
        const int row = static_cast<int>(e / N);
        const int col = static_cast<int>(e - static_cast<size_t>(row) * N);
        float4 acc = *reinterpret_cast<const float4*>(ws + e);
        for (int s = 1; s < splits; ++s) {
            const float4 v = *reinterpret_cast<const float4*>(ws + s * plane + e);
            acc.x += v.x;
            acc.y += v.y;
            acc.z += v.z;
            acc.w += v.w;
        }
        acc.x *= alpha;
        acc.y *= alpha;
        acc.z *= alpha;
        acc.w *= alpha;
        if (flags & TONY_EPI_BIAS) {
            acc.x += __bfloat162float(bias[col]);
            acc.y += __bfloat162float(bias[col + 1]);
            acc.z += __bfloat162float(bias[col + 2]);
            acc.w += __bfloat162float(bias[col + 3]);
        }
        __nv_bfloat16* cptr = C + static_cast<size_t>(row) * ldc + col;
        if (flags & TONY_EPI_ACCUMULATE) {
            const uint2 o = *reinterpret_cast<const uint2*>(cptr);
            const float2 a = unpack_bf16x2(o.x), b = unpack_bf16x2(o.y);
            acc.x += a.x;
            acc.y += a.y;
            acc.z += b.x;
            acc.w += b.y;
        }
        uint2 o;
        o.x = pack_bf16x2(acc.x, acc.y);
        o.y = pack_bf16x2(acc.z, acc.w);
        *reinterpret_cast<uint2*>(cptr) = o;
    }
}

// ----------------------------------------------------------------------------
// Host side
// ----------------------------------------------------------------------------
typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                    const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                    const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static PFN_encodeTiled get_encode_fn() {
    static PFN_encodeTiled fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* ptr = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) ==
                cudaSuccess &&
            qres == cudaDriverEntryPointSuccess) {
            fn = reinterpret_cast<PFN_encodeTiled>(ptr);
        }
    });
    return fn;
}

// Operand of logical shape [mn_extent, k_extent]; `mn_major` says which dim is contiguous.
int make_operand_tmap(CUtensorMap* map, const void* ptr, int mn_extent, int k_extent, long long ld,
                      int mn_major, int block_mn) {
    // cuTensorMapEncodeTiled is a raw driver call: it needs a context bound to THIS thread.
    // Autograd worker threads for device 0 may not have one yet (torch binds lazily), so
    // force the runtime to bind the primary context once per thread.
    static thread_local bool tls_ctx_bound = false;
    if (!tls_ctx_bound) {
        cudaFree(nullptr);
        tls_ctx_bound = true;
    }
    PFN_encodeTiled enc = get_encode_fn();
    if (!enc) {
        set_last_error("cuTensorMapEncodeTiled", "driver entry point not found");
        return TONY_ERR_UNSUPPORTED;
    }
    cuuint64_t gdim[2];
    cuuint32_t box[2];
    if (!mn_major) {
        gdim[0] = static_cast<cuuint64_t>(k_extent);
        gdim[1] = static_cast<cuuint64_t>(mn_extent);
        box[0] = BLOCK_K;
        box[1] = static_cast<cuuint32_t>(block_mn);
    } else {
        gdim[0] = static_cast<cuuint64_t>(mn_extent);
        gdim[1] = static_cast<cuuint64_t>(k_extent);
        box[0] = 64;
        box[1] = BLOCK_K;
    }
    cuuint64_t gstride[1] = {static_cast<cuuint64_t>(ld) * 2};
    cuuint32_t estride[2] = {1, 1};
    CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), gdim, gstride,
                     box, estride, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                     CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        char buf[64];
        snprintf(buf, sizeof(buf), "CUresult %d", static_cast<int>(r));
        set_last_error("cuTensorMapEncodeTiled", buf);
        return TONY_ERR_CUDA;
    }
    return TONY_OK;
}

static int make_tmap_2d_bf16_mode(CUtensorMap* map, const void* ptr, long long inner_extent, long long outer_extent,
                                  long long ld, int box_inner, int box_outer, CUtensorMapSwizzle swz);

// Generic 2-D bf16 map with 128-byte swizzle: `inner` is the contiguous dimension.
int make_tmap_2d_bf16(CUtensorMap* map, const void* ptr, long long inner_extent, long long outer_extent,
                      long long ld, int box_inner, int box_outer) {
    return make_tmap_2d_bf16_mode(map, ptr, inner_extent, outer_extent, ld, box_inner, box_outer,
                                  CU_TENSOR_MAP_SWIZZLE_128B);
}

int make_tmap_2d_bf16_sw64(CUtensorMap* map, const void* ptr, long long inner_extent, long long outer_extent,
                           long long ld, int box_inner, int box_outer) {
    return make_tmap_2d_bf16_mode(map, ptr, inner_extent, outer_extent, ld, box_inner, box_outer,
                                  CU_TENSOR_MAP_SWIZZLE_64B);
}

static int make_tmap_2d_bf16_mode(CUtensorMap* map, const void* ptr, long long inner_extent, long long outer_extent,
                                  long long ld, int box_inner, int box_outer, CUtensorMapSwizzle swz) {
    static thread_local bool tls_ctx_bound = false;
    if (!tls_ctx_bound) {
        cudaFree(nullptr);
        tls_ctx_bound = true;
    }
    PFN_encodeTiled enc = get_encode_fn();
    if (!enc) {
        set_last_error("cuTensorMapEncodeTiled", "driver entry point not found");
        return TONY_ERR_UNSUPPORTED;
    }
    cuuint64_t gdim[2] = {static_cast<cuuint64_t>(inner_extent), static_cast<cuuint64_t>(outer_extent)};
    cuuint32_t box[2] = {static_cast<cuuint32_t>(box_inner), static_cast<cuuint32_t>(box_outer)};
    cuuint64_t gstride[1] = {static_cast<cuuint64_t>(ld) * 2};
    cuuint32_t estride[2] = {1, 1};
    CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), gdim, gstride,
                     box, estride, CU_TENSOR_MAP_INTERLEAVE_NONE, swz,
                     CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        char buf[64];
        snprintf(buf, sizeof(buf), "CUresult %d", static_cast<int>(r));
        set_last_error("cuTensorMapEncodeTiled", buf);
        return TONY_ERR_CUDA;
    }
    return TONY_OK;
}

bool pdl_enabled() {
    static int v = -1;
    if (v < 0) {
        const char* e = getenv("TONY_B200_PDL");
        v = (e && atoi(e) == 0) ? 0 : 1;
    }
    return v == 1;
}

bool clear_counters_each_call() {
    static int v = -1;
    if (v < 0) {
        const char* e = getenv("TONY_B200_CLEAR_COUNTERS");
        v = (e && atoi(e) != 0) ? 1 : 0;
    }
    return v == 1;
}

int device_sm_count() {
    static int cached[64] = {0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    if (cached[dev] == 0) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
            n = 148;
        cached[dev] = n;
    }
    return cached[dev];
}

// Launch with programmatic stream serialization: the kernel may begin (prologue only, see
// pdl_wait) before the previous kernel in the stream has fully drained.  TONY_B200_PDL=0 disables.
template <typename Kern>
static cudaError_t launch_pdl(Kern kern, int grid, size_t smem, cudaStream_t stream, const CUtensorMap& ta,
                              const CUtensorMap& tb, const CUtensorMap& tc, const CUtensorMap& tx,
                              const GemmParams& p) {
    static int pdl = -1;
    if (pdl < 0) {
        const char* e = getenv("TONY_B200_PDL");
        pdl = e ? atoi(e) : 1;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(GEMM_THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kern, ta, tb, tc, tx, p);
}

template <int BLOCK_N, int STAGES>
static int launch_gemm(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& tc, const CUtensorMap& tx,
                       const GemmParams& p, int grid, cudaStream_t stream) {
    constexpr size_t SMEM =
        static_cast<size_t>(STAGES) * (BLOCK_M + BLOCK_N) * BLOCK_K * 2 + EPI_STAGING_BYTES + 1024 + 256;
    static bool configured[64] = {false};
    int dev = 0;
    TONY_CUDA_CHECK(cudaGetDevice(&dev));
    if (dev >= 0 && dev < 64 && !configured[dev]) {
        TONY_CUDA_CHECK(cudaFuncSetAttribute(gemm_bf16_tcgen05_kernel<BLOCK_N, STAGES>,
                                             cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             static_cast<int>(SMEM)));
        configured[dev] = true;
    }
    TONY_CUDA_CHECK(launch_pdl(gemm_bf16_tcgen05_kernel<BLOCK_N, STAGES>, grid, SMEM, stream, ta, tb, tc, tx, p));
    count_launch();
    return TONY_OK;
}

template <int STAGES, bool ROWSUM>
static int launch_gemm_2cta(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& tc,
                            const CUtensorMap& tx, const GemmParams& p, int grid, cudaStream_t stream) {
    constexpr size_t SMEM = static_cast<size_t>(STAGES) * (BLOCK_M + 128) * BLOCK_K * 2 + EPI_STAGING_BYTES +
                            (ROWSUM ? 1024 : 0) + 1024 + 256;
    static bool configured[64] = {false};
    int dev = 0;
    TONY_CUDA_CHECK(cudaGetDevice(&dev));
    if (dev >= 0 && dev < 64 && !configured[dev]) {
        TONY_CUDA_CHECK(cudaFuncSetAttribute(gemm_bf16_tcgen05_2cta_kernel<STAGES, ROWSUM>,
                                             cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             static_cast<int>(SMEM)));
        configured[dev] = true;
    }
    TONY_CUDA_CHECK(launch_pdl(gemm_bf16_tcgen05_2cta_kernel<STAGES, ROWSUM>, grid, SMEM, stream, ta, tb, tc, tx, p));
    count_launch();
    return TONY_OK;
}

// TONY_B200_GEMM_2CTA=0 disables the CTA-pair kernel (A/B testing); default on
static int two_cta_mode() {
    static int mode = -1;
    if (mode < 0) {
        const char* e = getenv("TONY_B200_GEMM_2CTA");
        mode = e ? atoi(e) : 1;
    }
    return mode;
}

// wave_efficiency: host_math.h

}  // namespace tony

using namespace tony;

struct AgHost {  // fused all-gather description (host side)
    const void* const* src;
    void* dst;
    unsigned int* flags;
    int world, rank, rows_per_rank;
};

static int gemm_host(const void* A, const void* B, void* C, int M, int N, int K, int a_mn_major,
                     int b_mn_major, long long lda, long long ldb, long long ldc,
                     unsigned int flags, float alpha, const void* bias, const void* aux,
                     long long ld_aux, void* aux_out, void* workspace, size_t workspace_bytes,
                     int force_block_n, int force_splits, const unsigned int* b_ready,
                     unsigned int b_ready_epoch, int copy_ctas, int max_ctas, const AgHost* ag,
                     void* stream_v, void* rowsum_out = nullptr, int rowsum_accumulate = 0) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_v);
    if (!A || !B || !C || M <= 0 || N <= 0 || K <= 0) {
        set_last_error("tony_gemm_bf16", "null pointer or non-positive dimension");
        return TONY_ERR_INVALID;
    }
    if ((N % 8) != 0 || (lda % 8) != 0 || (ldb % 8) != 0 || (ldc % 8) != 0 ||
        (reinterpret_cast<uintptr_t>(A) & 15) || (reinterpret_cast<uintptr_t>(B) & 15) ||
        (reinterpret_cast<uintptr_t>(C) & 15)) {
        set_last_error("tony_gemm_bf16", "N and leading dims must be multiples of 8; pointers 16B aligned");
        return TONY_ERR_INVALID;
    }
    if ((flags & TONY_EPI_BIAS) && !bias) return TONY_ERR_INVALID;
    if ((flags & (TONY_EPI_RESIDUAL | TONY_EPI_DGELU_ERF | TONY_EPI_DGELU_TANH | TONY_EPI_DRELU)) &&
        (!aux || (ld_aux % 8) != 0))
        return TONY_ERR_INVALID;
    if ((flags & TONY_EPI_SAVE_PREACT) && !aux_out) return TONY_ERR_INVALID;

    const int sms = device_sm_count();
    int workers = sms - copy_ctas;
    if (max_ctas > 0) workers = std::min(workers, max_ctas - copy_ctas);
    if (workers <= 0) return TONY_ERR_INVALID;

    int num_m = (M + BLOCK_M - 1) / BLOCK_M;
    const int num_k = (K + BLOCK_K - 1) / BLOCK_K;
    int block_n = force_block_n;
    // CTA-pair kernel (256 x 256 tiles over 2 SMs): force_block_n == 512 forces it, 0 lets the
    // heuristic decide, 128 / 256 force the 1-CTA kernels.
    bool use_2cta = false;
    if (two_cta_mode() > 0 && (workers % 2) == 0 && (copy_ctas % 2) == 0 &&
        (ag != nullptr || (copy_ctas == 0 && b_ready == nullptr))) {
        if (force_block_n == 512) {
            use_2cta = true;
        } else if ((force_block_n == 0 || (ag != nullptr && force_block_n == 256)) && N >= 256 && M >= 256) {
            // measured on BERT-large (profiles/): the pair kernel wins on every such shape with many tiles,
            // even where its wave quantisation is worse, because each SM moves a third less
            // operand traffic and the smem ring is 6 deep
            use_2cta = true;
            if (ag == nullptr && force_block_n == 0 && rowsum_out == nullptr) {
                // few tiles (ResNet's 7x7 / 14x14 pointwise gradients, N not a multiple of 256): compare SM-time per
                // wave x waves of the three tilings instead of assuming the pair kernel
                const int pick = choose_gemm_tiling(M, N, workers);     // host_math.h
                if (pick != 512) {
                    use_2cta = false;
                    block_n = pick;
                }
            }
        }
    }
    if (rowsum_out && !use_2cta) {
        set_last_error("tony_gemm_bf16_rowsum", "row sums ride only in the CTA-pair kernel (M, N >= 256)");
        return TONY_ERR_UNSUPPORTED;
    }
    if (use_2cta) {
        block_n = 256;
        num_m = (M + 255) / 256;
        workers /= 2;  // clusters
    } else if (block_n != 128 && block_n != 256) {
        const long long t256 = static_cast<long long>(num_m) * ((N + 255) / 256);
        const long long t128 = static_cast<long long>(num_m) * ((N + 127) / 128);
        block_n = (N <= 128 || wave_efficiency(t128, workers) > 1.12 * wave_efficiency(t256, workers))
                      ? 128
                      : 256;
    }
    const int num_n = (N + block_n - 1) / block_n;
    const long long tiles = static_cast<long long>(num_m) * num_n;

    // split-K only for plain / accumulate / bias epilogues, and only if it fills the machine better
    int splits = 1;
    const bool splittable =
        (flags & ~(TONY_EPI_ACCUMULATE | TONY_EPI_BIAS)) == 0 && workspace != nullptr && (N % 4) == 0;
    if (splittable) {
        if (force_splits > 1) {
            splits = force_splits;
        } else if (force_splits == 0 &&
                   (tiles * 2 <= workers ||
                    (wave_efficiency(tiles, workers) < 0.75 && static_cast<long long>(M) * N <= (4ll << 20)))) {
            // (a big output pays for its fp32 partials: 12544 x 512 split three ways wrote and re-read 77 MB)
            // up to 16 splits normally; very long reductions over one or two output tiles (pointwise-convolution
            // weight gradients: K = every pixel of the batch) may use one split per SM
            int max_s = 16;
            if (tiles * 32 <= workers) max_s = static_cast<int>(workers / tiles);
            // k-block time of this tiling relative to the pair kernel's 256 x 256 x 64 block
            const double tile_scale = use_2cta ? 1.0 : (block_n == 256 ? 1.12 : 0.6);
            splits = choose_splitk(tiles, num_k, workers, static_cast<long long>(M) * N, max_s, tile_scale);   // host_math.h
        }
    }
    if (splittable) {
        splits = std::min(splits, num_k);
        while (splits > 1 &&
               static_cast<size_t>(splits) * M * (static_cast<size_t>(N) + (rowsum_out ? 1 : 0)) * 4 > workspace_bytes)
            --splits;
    }
    int kb_per_split = (num_k + splits - 1) / splits;
    splits = (num_k + kb_per_split - 1) / kb_per_split;  // no empty splits

    CUtensorMap ta, tb;
    int rc = make_operand_tmap(&ta, A, M, K, lda, a_mn_major, BLOCK_M);
    if (rc) return rc;
    rc = make_operand_tmap(&tb, B, N, K, ldb, b_mn_major, use_2cta ? 128 : block_n);
    if (rc) return rc;

    // epilogue output maps: 32 x 32 boxes, 64-byte swizzle (the staging tile's own XOR pattern)
    static int tma_store_mode = -1;
    if (tma_store_mode < 0) {
        const char* e = getenv("TONY_B200_GEMM_TMA_STORE");
        tma_store_mode = e ? atoi(e) : 1;
    }
    CUtensorMap tc = ta, tx = ta;   // placeholders when unused
    const int use_tma_store = tma_store_mode && splits == 1;
    if (use_tma_store) {
        rc = make_tmap_2d_bf16_sw64(&tc, C, N, M, ldc, 32, 32);
        if (rc) return rc;
        if (flags & TONY_EPI_SAVE_PREACT) {
            rc = make_tmap_2d_bf16_sw64(&tx, aux_out, N, M, ldc, 32, 32);
            if (rc) return rc;
        }
    }

    GemmParams p;
    p.tma_store = use_tma_store;
    // all-gather GEMM: 32 m-blocks per sweep keep A's group + B inside the L2 (64 did not: +270 us of HBM
    // re-reads at 16384 x 3344) while still giving the copy CTAs a few waves of local-shard tiles as head start
    p.group_m = ag ? 32 : 16;
    p.M = M;
    p.N = N;
    p.K = K;
    p.a_mn_major = a_mn_major ? 1 : 0;
    p.b_mn_major = b_mn_major ? 1 : 0;
    p.num_m_blocks = num_m;
    p.num_n_blocks = num_n;
    p.num_k_blocks = num_k;
    p.splits = splits;
    p.kb_per_split = kb_per_split;
    p.flags = flags;
    p.alpha = alpha;
    p.C = static_cast<__nv_bfloat16*>(C);
    p.ldc = ldc;
    p.bias = static_cast<const __nv_bfloat16*>(bias);
    p.aux = static_cast<const __nv_bfloat16*>(aux);
    p.ld_aux = ld_aux;
    p.aux_out = static_cast<__nv_bfloat16*>(aux_out);
    p.workspace = static_cast<float*>(workspace);
    p.b_ready = b_ready;
    p.b_ready_epoch = b_ready_epoch;
    p.copy_ctas = copy_ctas;
    for (int r = 0; r < TONY_MAX_RANKS; ++r) p.ag_src[r] = nullptr;
    p.ag_dst = nullptr;
    p.ag_flags = nullptr;
    p.ag_rows_per_rank = 1;
    p.n_rot = 0;
    p.rowsum_out = static_cast<__nv_bfloat16*>(rowsum_out);
    p.rowsum_accumulate = rowsum_accumulate;
    // split-K partial row sums live behind the [splits][M][N] partial products
    p.rowsum_ws = (rowsum_out && splits > 1)
                      ? static_cast<float*>(workspace) + static_cast<size_t>(splits) * M * static_cast<size_t>(N)
                      : nullptr;
    if (ag) {
        for (int r = 0; r < ag->world; ++r) p.ag_src[r] = static_cast<const __nv_bfloat16*>(ag->src[r]);
        p.ag_dst = static_cast<__nv_bfloat16*>(ag->dst);
        p.ag_flags = ag->flags;
        p.ag_rows_per_rank = ag->rows_per_rank;
        p.n_rot = (ag->rank * ag->rows_per_rank) / block_n;
    }

    const long long units = tiles * splits;
    const int grid = static_cast<int>(std::min<long long>(units, workers)) + copy_ctas;
    const int grid_pairs = static_cast<int>(std::min<long long>(units, workers)) * 2 + copy_ctas;   // CTAs
    if (use_2cta)
        rc = rowsum_out ? launch_gemm_2cta<6, true>(ta, tb, tc, tx, p, grid_pairs, stream)
                        : launch_gemm_2cta<6, false>(ta, tb, tc, tx, p, grid_pairs, stream);
    else if (block_n == 256)
        rc = launch_gemm<256, 4>(ta, tb, tc, tx, p, grid, stream);
    else
        rc = launch_gemm<128, 6>(ta, tb, tc, tx, p, grid, stream);
    if (rc) return rc;

    if (splits > 1) {
        const size_t total4 = static_cast<size_t>(M) * N / 4;
        const int threads = 256;
        const int blocks = static_cast<int>(std::min<size_t>((total4 + threads - 1) / threads,
                                                             static_cast<size_t>(sms) * 8));
        TONY_CUDA_CHECK(launch_pdl_kernel(
            splitk_reduce_kernel, dim3(blocks), dim3(threads), 0, stream, static_cast<const float*>(workspace), splits, M,
            N, static_cast<__nv_bfloat16*>(C), static_cast<long long>(ldc), alpha, static_cast<uint32_t>(flags),
            static_cast<const __nv_bfloat16*>(bias), static_cast<const float*>(p.rowsum_ws),
            static_cast<__nv_bfloat16*>(rowsum_out), rowsum_accumulate));
        count_launch();
    }
    return TONY_OK;
}

// Fold split-K partials (shared with the convolution weight-gradient kernel).
int tony::launch_splitk_reduce(const float* ws, int splits, int M, int N, __nv_bfloat16* C, long long ldc, float alpha,
                               unsigned int flags, cudaStream_t stream) {
    const size_t total4 = static_cast<size_t>(M) * N / 4;
    const int threads = 256;
    const int blocks = static_cast<int>(std::min<size_t>((total4 + threads - 1) / threads,
                                                         static_cast<size_t>(device_sm_count()) * 8));
    TONY_CUDA_CHECK(launch_pdl_kernel(splitk_reduce_kernel, dim3(blocks), dim3(threads), 0, stream, ws, splits, M, N, C, ldc,
                                      alpha, static_cast<uint32_t>(flags), static_cast<const __nv_bfloat16*>(nullptr),
                                      static_cast<const float*>(nullptr), static_cast<__nv_bfloat16*>(nullptr), 0));
    count_launch();
    return TONY_OK;
}

extern "C" int tony_gemm_bf16_ex(const void* A, const void* B, void* C, int M, int N, int K,
                                 int a_mn_major, int b_mn_major, long long lda, long long ldb,
                                 long long ldc, unsigned int flags, float alpha, const void* bias,
                                 const void* aux, long long ld_aux, void* aux_out, void* workspace,
                                 size_t workspace_bytes, int force_block_n, int force_splits,
                                 const unsigned int* b_ready, unsigned int b_ready_epoch,
                                 int copy_ctas, int max_ctas, void* stream) {
    if (copy_ctas != 0) {
        set_last_error("tony_gemm_bf16_ex", "copy_ctas is reserved for tony_allgather_gemm_bf16");
        return TONY_ERR_INVALID;
    }
    return gemm_host(A, B, C, M, N, K, a_mn_major, b_mn_major, lda, ldb, ldc, flags, alpha, bias, aux,
                     ld_aux, aux_out, workspace, workspace_bytes, force_block_n, force_splits, b_ready,
                     b_ready_epoch, 0, max_ctas, nullptr, stream);
}

extern "C" int tony_allgather_gemm_bf16(const void* X, const void* const* w_shard, void* w_full,
                                        void* Y, int M, int N, int K, int rank, int world,
                                        unsigned int flags, const void* bias, void* aux_out,
                                        unsigned int* ready_flags, unsigned int epoch,
                                        int copy_ctas, void* stream) {
    if (!w_shard || !w_full || !ready_flags || world < 1 || world > TONY_MAX_RANKS || rank < 0 ||
        rank >= world || (N % world) != 0 || (K % 8) != 0 || epoch == 0) {
        set_last_error("tony_allgather_gemm_bf16", "bad arguments (N % world, K % 8, epoch >= 1)");
        return TONY_ERR_INVALID;
    }
    if (flags & ~(TONY_EPI_BIAS | TONY_EPI_GELU_ERF | TONY_EPI_GELU_TANH | TONY_EPI_RELU |
                  TONY_EPI_SAVE_PREACT)) {
        set_last_error("tony_allgather_gemm_bf16", "unsupported epilogue flag");
        return TONY_ERR_INVALID;
    }
    // bulk copies keep 96 KB in flight per copy CTA: a handful saturate what the GEMM can consume
    if (copy_ctas <= 0) copy_ctas = world <= 2 ? 4 : 8;   // measured: 364 us at 2 ranks, 374 us at 8 ranks
    copy_ctas = std::min(copy_ctas, device_sm_count() / 2) & ~1;
    if (copy_ctas < 2) copy_ctas = 2;
    AgHost ag;
    ag.src = w_shard;
    ag.dst = w_full;
    ag.flags = ready_flags;
    ag.world = world;
    ag.rank = rank;
    ag.rows_per_rank = N / world;
    // every copy CTA bumps a slab's flag once per call, so call #epoch completes at epoch*copy_ctas
    // the tile width must be a function of N alone: the per-slab flags count calls, so every
    // call on the same weight has to walk the same set of slabs
    const int block_n = N >= 256 ? 256 : 128;
    return gemm_host(X, w_full, Y, M, N, K, 0, 0, K, K, N, flags, 1.0f, bias, nullptr, 0, aux_out,
                     nullptr, 0, block_n, 1, ready_flags, epoch * static_cast<unsigned int>(copy_ctas),
                     copy_ctas, 0, &ag, stream);
}

extern "C" int tony_gemm_bf16(const void* A, const void* B, void* C, int M, int N, int K,
                              int a_mn_major, int b_mn_major, long long lda, long long ldb,
                              long long ldc, unsigned int flags, float alpha, const void* bias,
                              const void* aux, long long ld_aux, void* aux_out, void* workspace,
                              size_t workspace_bytes, void* stream) {
    return gemm_host(A, B, C, M, N, K, a_mn_major, b_mn_major, lda, ldb, ldc, flags, alpha, bias, aux,
                     ld_aux, aux_out, workspace, workspace_bytes, 0, 0, nullptr, 0, 0, 0, nullptr,
                     stream);
}

extern "C" int tony_gemm_bf16_rowsum(const void* A, const void* B, void* C, int M, int N, int K, int a_mn_major,
                                     int b_mn_major, long long lda, long long ldb, long long ldc,
                                     unsigned int flags, float alpha, const void* bias, const void* aux,
                                     long long ld_aux, void* aux_out, void* workspace, size_t workspace_bytes,
                                     void* rowsum_out, int rowsum_accumulate, void* stream) {
    if (!rowsum_out) return TONY_ERR_INVALID;
    return gemm_host(A, B, C, M, N, K, a_mn_major, b_mn_major, lda, ldb, ldc, flags, alpha, bias, aux, ld_aux,
                     aux_out, workspace, workspace_bytes, 0, 0, nullptr, 0, 0, 0, nullptr, stream, rowsum_out,
                     rowsum_accumulate);
}
