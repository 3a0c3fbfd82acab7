// Fused multi-head attention for sm_100a on a PACKED qkv activation.
//
// Input is the output of the fused QKV projection, [tokens, 3, heads, 64] bf16, read in place
// through tensor maps (no q/k/v split or transposes); the context comes back as
// [tokens, heads*64], the layout the output projection's GEMM consumes.
//
// Forward: a persistent kernel, two CTAs resident per SM, each looping over work items
// (batch, head, 128 query rows); all pipelines run across item boundaries, so the loads and
// the first score MMAs of the next item overlap the epilogue of the current one.
//   warp 0     TMA producer: Q tile per item, K_j / V_j tiles (64 keys x 64) through two rings of
//              three 8 KB slots
//   warp 1     tcgen05.mma issuer, an event loop over two cursors: S = Q K_j^T (M128 N64 K64)
//              into one of TWO 64-column TMEM buffers, up to two blocks ahead of the softmax;
//              O += P_j V_j (M128 N64 K64) with P read FROM TENSOR MEMORY as the A operand and
//              V consumed as an MN-major B operand straight from its row-major tile.  Whichever
//              MMA has its inputs ready is issued, so neither waits behind the other's barrier.
//   warps 2-5  softmax: one thread per query row (= one TMEM lane).  A block's 64 scores are
//              pulled into registers with two tcgen05.ld and the TMEM buffer is handed back at
//              once.  P goes back to TMEM as bf16 pairs (tcgen05.st, 32 columns, double
//              buffered): the probabilities never touch shared memory.  The running max only
//              advances when it grows by more than 2^8 (exact: the stale offset cancels in
//              O / l), so the TMEM-resident O is rescaled (tcgen05.ld -> mul -> tcgen05.st)
//              only then.
// TMEM columns per CTA: S0 [0,64) S1 [64,128) | O [128,192) | P0 [192,224) P1 [224,256).
// The causal mask is applied only in the (at most two) diagonal blocks, which run a separate
// instantiation of the block body; blocks above the diagonal are never loaded, and causal work
// items are handed out heaviest first.
// Measured (profiles/r1_attention_tcgen05_vs_library.md): correct against fp32, 430 TFLOP/s
// forward on B200 vs 575 for the cuDNN sm_100 kernel, so ops/functional.py keeps the library
// core as the default and this path is opt-in (TONY_B200_ATTENTION=tcgen05).
//
// The reference (zuston/TonY) has no device code; this replaces the attention
// core that stock PyTorch jobs launched by the reference run through a library kernel.
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdlib.h>

#include <type_traits>

#include "common.cuh"

namespace tony {

constexpr int ATT_BM = 128;        // rows per CTA (queries in fwd / dq, keys in dkdv)
constexpr int ATT_BK = 64;         // streamed block (keys in fwd / dq, queries in dkdv)
constexpr int ATT_D = 64;          // head dim
constexpr int ATT_THREADS = 192;   // backward: TMA warp, MMA warp, 4 elementwise warps
constexpr uint32_t ATT_TILE_BYTES = ATT_BM * ATT_D * 2;   // 16 KB
constexpr uint32_t ATT_HALF_BYTES = ATT_BK * ATT_D * 2;   // 8 KB
constexpr int ATT_FWD_SLOTS = 3;   // depth of the K ring and of the V ring
constexpr uint32_t ATT_FWD_SMEM =
    ATT_TILE_BYTES + 2 * ATT_FWD_SLOTS * ATT_HALF_BYTES + ATT_TILE_BYTES + 256 + 1024;
constexpr float ATT_RESCALE_THRESHOLD = 8.0f;

struct AttnFwdParams {
    int batch, seq, heads, hidden;
    int causal;
    int num_items;                  // batch * heads * (seq / 128)
    float scale_log2;               // softmax scale * log2(e)
    __nv_bfloat16* out;             // [tokens, hidden]
    float* lse;                     // [batch*heads, seq], log2 domain
};

// 16-byte unit `u` (0..7) of row `row` in a [128 x 64] bf16 K-major swizzle-128B panel
__device__ __forceinline__ uint32_t panel_unit_offset(int row, int u) {
    return static_cast<uint32_t>(row * 128 + ((u ^ (row & 7)) << 4));
}

struct AttnItem {
    int bh, qt, nblk;
};

__device__ __forceinline__ AttnItem attn_item(const AttnFwdParams& p, int w) {
    const int nq = p.seq / ATT_BM;
    AttnItem it;
    if (p.causal) {   // heaviest query tiles first, heads interleaved
        const int nbh = p.batch * p.heads;
        it.qt = nq - 1 - w / nbh;
        it.bh = w - (w / nbh) * nbh;
        it.nblk = (it.qt + 1) * (ATT_BM / ATT_BK);
    } else {          // tiles of one head back to back: its K/V stay hot in L2
        it.bh = w / nq;
        it.qt = w - it.bh * nq;
        it.nblk = p.seq / ATT_BK;
    }
    return it;
}

__global__ void __launch_bounds__(ATT_THREADS, 2)
attention_fwd_kernel(const __grid_constant__ CUtensorMap tmap_qkv128,
                     const __grid_constant__ CUtensorMap tmap_qkv64, const AttnFwdParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>(
        (reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
    uint8_t* sQ = smem;
    uint8_t* sK = sQ + ATT_TILE_BYTES;                               // ring of 8 KB key tiles
    uint8_t* sV = sK + ATT_FWD_SLOTS * ATT_HALF_BYTES;               // ring of 8 KB value tiles
    uint8_t* sStage = sV + ATT_FWD_SLOTS * ATT_HALF_BYTES;           // 16 KB epilogue staging
    uint64_t* q_full = reinterpret_cast<uint64_t*>(sStage + ATT_TILE_BYTES);
    uint64_t* q_empty = q_full + 1;
    uint64_t* k_full = q_empty + 1;
    uint64_t* k_empty = k_full + ATT_FWD_SLOTS;
    uint64_t* v_full = k_empty + ATT_FWD_SLOTS;
    uint64_t* v_empty = v_full + ATT_FWD_SLOTS;
    uint64_t* s_full = v_empty + ATT_FWD_SLOTS;      // [2]
    uint64_t* s_free = s_full + 2;                   // [2]
    uint64_t* p_full = s_free + 2;                   // [2]
    uint64_t* pv_done = p_full + 2;                  // [2]
    uint64_t* o_free = pv_done + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(o_free + 1);

    const int warp_idx = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);   // provably warp-uniform: role branches and their loops stay on the uniform datapath
    const int lane = threadIdx.x & 31;

    if (warp_idx == 0 && lane == 0) {
        tma_prefetch_desc(&tmap_qkv128);
        tma_prefetch_desc(&tmap_qkv64);
        mbar_init(q_full, 1);
        mbar_init(q_empty, 1);
        for (int s = 0; s < ATT_FWD_SLOTS; ++s) {
            mbar_init(&k_full[s], 1);
            mbar_init(&k_empty[s], 1);
            mbar_init(&v_full[s], 1);
            mbar_init(&v_empty[s], 1);
        }
        for (int s = 0; s < 2; ++s) {
            mbar_init(&s_full[s], 1);
            mbar_init(&s_free[s], 4);
            mbar_init(&p_full[s], 4);
            mbar_init(&pv_done[s], 1);
        }
        mbar_init(o_free, 4);
        fence_barrier_init();
    }
    if (warp_idx == 1) tmem_alloc<256>(tmem_slot);
    tcgen05_fence_before();
    __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    // TMEM columns: S0 [0,64) S1 [64,128) fp32 scores | O [128,192) fp32 | P0 [192,224) P1 [224,256) bf16x2
    const uint32_t tO = tmem_base + 2 * ATT_BK;
    const uint32_t tP = tmem_base + 2 * ATT_BK + ATT_D;
    const int stride = static_cast<int>(gridDim.x);

    // Producer and MMA warps: the WHOLE warp runs the loop and one elected lane issues, so the operands of
    // UTMALDG / UTCHMMA / UTCBAR are warp-uniform values in uniform registers (inside `if (lane == 0)` the
    // compiler wraps each of them in an ELECT / R2UR.BROADCAST / BRA.U.ANY loop, ~25 instructions per MMA).
    if (warp_idx == 0) {
        {
            uint32_t g = 0;   // running block index across work items
            uint32_t n = 0;
            for (int w = blockIdx.x; w < p.num_items; w += stride, ++n) {
                const AttnItem a = attn_item(p, w);
                const int b = a.bh / p.heads, h = a.bh - b * p.heads;
                const int row_base = b * p.seq;
                mbar_wait_relaxed(q_empty, (n & 1) ^ 1);   // every S of the previous item has retired
                if (elect_one()) {
                    mbar_expect_tx(q_full, ATT_TILE_BYTES);
                    tma_load_2d(&tmap_qkv128, q_full, sQ, h * ATT_D, row_base + a.qt * ATT_BM);
                }
                __syncwarp();
                for (int j = 0; j < a.nblk; ++j, ++g) {
                    const uint32_t slot = g % ATT_FWD_SLOTS;
                    const uint32_t par = ((g / ATT_FWD_SLOTS) & 1) ^ 1;
                    mbar_wait_relaxed(&k_empty[slot], par);
                    if (elect_one()) {
                        mbar_expect_tx(&k_full[slot], ATT_HALF_BYTES);
                        tma_load_2d(&tmap_qkv64, &k_full[slot], sK + slot * ATT_HALF_BYTES,
                                    p.hidden + h * ATT_D, row_base + j * ATT_BK);
                    }
                    __syncwarp();
                    mbar_wait_relaxed(&v_empty[slot], par);
                    if (elect_one()) {
                        mbar_expect_tx(&v_full[slot], ATT_HALF_BYTES);
                        tma_load_2d(&tmap_qkv64, &v_full[slot], sV + slot * ATT_HALF_BYTES,
                                    2 * p.hidden + h * ATT_D, row_base + j * ATT_BK);
                    }
                    __syncwarp();
                }
            }
        }
    } else if (warp_idx == 1) {
        {
            // Event loop over two cursors on the flattened (item, block) sequence: the score
            // MMAs run up to two blocks ahead of the P V MMAs; whichever has all its inputs
            // ready is issued, so neither waits behind the other's barrier.
            const uint32_t idesc_s = make_idesc_bf16(ATT_BM, ATT_BK, 0, 0);
            const uint32_t idesc_o = make_idesc_bf16(ATT_BM, ATT_D, 0, 1);   // B = V, MN-major
            const uint64_t dq = make_umma_desc_sw128(smem_u32(sQ), 16, 1024);
            int ws = blockIdx.x, js = 0;
            uint32_t ns = 0, gs = 0;
            int nblk_s = ws < p.num_items ? attn_item(p, ws).nblk : 0;
            int wp = blockIdx.x, jp = 0;
            uint32_t np = 0, gp = 0;
            int nblk_p = nblk_s;
            while (nblk_p) {
                bool progressed = false;
                if (nblk_s && gs < gp + 2) {
                    const uint32_t slot = gs % ATT_FWD_SLOTS;
                    bool ok = gs < 2 || mbar_test_wait(&s_free[gs & 1], ((gs - 2) >> 1) & 1);
                    if (ok && js == 0) ok = mbar_test_wait(q_full, ns & 1);
                    if (ok) ok = mbar_test_wait(&k_full[slot], (gs / ATT_FWD_SLOTS) & 1);
                    ok = __all_sync(0xffffffffu, ok);     // one warp-uniform decision
                    if (ok) {
                        tcgen05_fence_after();
                        const uint64_t dk = make_umma_desc_sw128(smem_u32(sK + slot * ATT_HALF_BYTES), 16, 1024);
                        const uint32_t tS = tmem_base + (gs & 1) * ATT_BK;
                        const bool last = js + 1 == nblk_s;
                        if (elect_one()) {
#pragma unroll
                            for (int k = 0; k < ATT_D / 16; ++k) umma_bf16(tS, dq + k * 2, dk + k * 2, idesc_s, k > 0);
                            umma_commit(&k_empty[slot]);
                            umma_commit(&s_full[gs & 1]);
                            if (last) umma_commit(q_empty);   // Q may be replaced once these MMAs retire
                        }
                        __syncwarp();
                        progressed = true;
                        ++gs;
                        if (++js == nblk_s) {
                            js = 0;
                            ++ns;
                            ws += stride;
                            nblk_s = ws < p.num_items ? attn_item(p, ws).nblk : 0;
                        }
                    }
                }
                {
                    const uint32_t bsel = gp & 1;
                    const uint32_t slot = gp % ATT_FWD_SLOTS;
                    bool ok = mbar_test_wait(&p_full[bsel], (gp >> 1) & 1);
                    if (ok) ok = mbar_test_wait(&v_full[slot], (gp / ATT_FWD_SLOTS) & 1);
                    if (ok && jp == 0 && np > 0) ok = mbar_test_wait(o_free, (np - 1) & 1);   // previous O read out
                    ok = __all_sync(0xffffffffu, ok);
                    if (ok) {
                        tcgen05_fence_after();
                        const uint64_t dv = make_umma_desc_sw128(smem_u32(sV + slot * ATT_HALF_BYTES), 8192, 1024);
                        if (elect_one()) {
#pragma unroll
                            for (int k = 0; k < ATT_BK / 16; ++k)   // A: P from TMEM, 8 columns per k-step; B: 16 key rows = 2048 B
                                umma_bf16_ts(tO, tP + bsel * (ATT_BK / 2) + k * 8, dv + k * 128, idesc_o,
                                             (jp > 0 || k > 0) ? 1u : 0u);
                            umma_commit(&v_empty[slot]);
                            umma_commit(&pv_done[bsel]);
                        }
                        __syncwarp();
                        progressed = true;
                        ++gp;
                        if (++jp == nblk_p) {
                            jp = 0;
                            ++np;
                            wp += stride;
                            nblk_p = wp < p.num_items ? attn_item(p, wp).nblk : 0;
                        }
                    }
                }
                if (!progressed) __nanosleep(40);   // keep the poll off the softmax warps' issue slots
            }
        }
    } else {
        const int quarter = warp_idx & 3;
        const int row = quarter * 32 + lane;
        const uint32_t lane_addr = static_cast<uint32_t>(quarter * 32) << 16;
        const uint32_t stage = smem_u32(sStage);
        uint32_t g = 0;   // running block index across work items
        uint32_t n = 0;
        float m_used = 0.f, l = 0.f;
        int q0 = 0;
#ifdef ATT_PROFILE
        long long tacc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        long long tprev = clock64();
#define ATT_T(k) { long long _t = clock64(); tacc[k] += _t - tprev; tprev = _t; }
#else
#define ATT_T(k)
#endif

        auto block = [&](int j, auto masked_tag) {
            constexpr bool MASKED = decltype(masked_tag)::value;
            const uint32_t bsel = g & 1;
            const uint32_t ph = (g >> 1) & 1;
            const int rel = q0 + row - j * ATT_BK;   // keys with column index > rel are masked
            ATT_T(0);
            mbar_wait(&s_full[bsel], ph);
            ATT_T(1);
            tcgen05_fence_after();
            uint32_t r[64];
            tmem_ld_32x32(tmem_base + lane_addr + bsel * ATT_BK, r);
            tmem_ld_32x32(tmem_base + lane_addr + bsel * ATT_BK + 32, r + 32);
            tmem_ld_wait();
            tcgen05_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&s_free[bsel]);
            ATT_T(2);
            float mx[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
            for (int i = 0; i < 64; ++i) {
                if (MASKED && i > rel) r[i] = __float_as_uint(-INFINITY);
                mx[i & 3] = fmaxf(mx[i & 3], __uint_as_float(r[i]));
            }
            const float m_new = fmaxf(m_used, fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3])) * p.scale_log2);
            float alpha = 1.f;
            bool need = false;
            if (j == 0) {
                m_used = m_new;
            } else if (m_new > m_used + ATT_RESCALE_THRESHOLD) {
                need = true;
                alpha = ex2_approx(m_used - m_new);
                m_used = m_new;
            }
            if (__any_sync(0xffffffffu, need)) {
                mbar_wait(&pv_done[bsel ^ 1], ((g - 1) >> 1) & 1);   // previous P V retired
                tcgen05_fence_after();
#pragma unroll 1
                for (int c = 0; c < 2; ++c) {
                    uint32_t o[32];
                    tmem_ld_32x32(tO + lane_addr + c * 32, o);
                    tmem_ld_wait();
#pragma unroll
                    for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
                    tmem_st_32x32(tO + lane_addr + c * 32, o);
                }
                tmem_st_wait();
                l *= alpha;
            }
            const float neg_m = -m_used;
            float sum[4] = {0.f, 0.f, 0.f, 0.f};
            ATT_T(3);
#pragma unroll
            for (int i = 0; i < 32; ++i) {   // packed P overwrites the scores it was made from
                const float a = ex2_approx(fmaf(__uint_as_float(r[2 * i]), p.scale_log2, neg_m));
                const float c = ex2_approx(fmaf(__uint_as_float(r[2 * i + 1]), p.scale_log2, neg_m));
                sum[i & 1] += a;
                sum[2 + (i & 1)] += c;
                r[i] = pack_bf16x2(a, c);
            }
            l += (sum[0] + sum[1]) + (sum[2] + sum[3]);
            ATT_T(4);
            if (g >= 2) mbar_wait(&pv_done[bsel], ph ^ 1);
            ATT_T(5);   // P V two blocks back retired: its P columns are free
            tcgen05_fence_after();
            // P stays in tensor memory (32 columns of bf16 pairs): the second MMA reads its A
            // operand from there, so the probabilities never touch shared memory
            tmem_st_32x32(tP + lane_addr + bsel * (ATT_BK / 2), r);
            tmem_st_wait();
            tcgen05_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&p_full[bsel]);
            ATT_T(6);
        };

        for (int w = blockIdx.x; w < p.num_items; w += stride, ++n) {
            const AttnItem a = attn_item(p, w);
            const int b = a.bh / p.heads, h = a.bh - b * p.heads;
            const int row_base = b * p.seq;
            q0 = a.qt * ATT_BM;
            m_used = -INFINITY;
            l = 0.f;
            for (int j = 0; j < a.nblk; ++j, ++g) {
                if (p.causal && (j * ATT_BK + ATT_BK - 1 > q0))
                    block(j, std::true_type{});
                else
                    block(j, std::false_type{});
            }
            ATT_T(7);
            // ---- item epilogue: O / l -> bf16 context rows (staged through this warp's 32 rows of a
            // swizzled panel so every global store covers whole 128-byte rows), log2-sum-exp
            const uint32_t gl = g - 1;
            mbar_wait(&pv_done[gl & 1], (gl >> 1) & 1);
            tcgen05_fence_after();
            uint32_t o[64];
            tmem_ld_32x32(tO + lane_addr, o);
            tmem_ld_32x32(tO + lane_addr + 32, o + 32);
            tmem_ld_wait();
            tcgen05_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(o_free);   // the next item's first P V may overwrite O
            const float inv = 1.f / l;
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                uint4 q;
                q.x = pack_bf16x2(__uint_as_float(o[8 * u]) * inv, __uint_as_float(o[8 * u + 1]) * inv);
                q.y = pack_bf16x2(__uint_as_float(o[8 * u + 2]) * inv, __uint_as_float(o[8 * u + 3]) * inv);
                q.z = pack_bf16x2(__uint_as_float(o[8 * u + 4]) * inv, __uint_as_float(o[8 * u + 5]) * inv);
                q.w = pack_bf16x2(__uint_as_float(o[8 * u + 6]) * inv, __uint_as_float(o[8 * u + 7]) * inv);
                st_shared_v4(stage + panel_unit_offset(row, u), q);
            }
            __syncwarp();
            __nv_bfloat16* obase =
                p.out + static_cast<size_t>(row_base + q0 + quarter * 32) * p.hidden + h * ATT_D;
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int rr = i * 4 + (lane >> 3);
                const uint4 v = ld_shared_v4(stage + panel_unit_offset(quarter * 32 + rr, lane & 7));
                st_global_v4(obase + static_cast<size_t>(rr) * p.hidden + (lane & 7) * 8, v);
            }
            __syncwarp();   // the staging rows are rewritten by this warp's next epilogue
            p.lse[static_cast<size_t>(a.bh) * p.seq + q0 + row] = m_used + lg2_approx(l);
            ATT_T(0);
        }
#ifdef ATT_PROFILE
        if (lane == 0 && blockIdx.x < 4) {
            for (int k = 0; k < 8; ++k)
                reinterpret_cast<long long*>(p.lse)[(blockIdx.x * 4 + (warp_idx - 2)) * 8 + k] = tacc[k];
        }
#endif
    }

    __syncwarp();
    tcgen05_fence_before();
    __syncthreads();
    if (warp_idx == 1) {
        tcgen05_fence_after();
        tmem_dealloc<256>(tmem_base);
    }
}

// ----------------------------------------------------------------------------------------
// Backward.  Three launches, no atomics, dq/dk/dv written straight into the PACKED gradient
// [tokens, 3, heads, 64] of the fused QKV projection:
//   delta   D[q] = sum_d dO[q,d] O[q,d]
//   dq      one CTA per 128 query rows, loops over 64-key blocks:
//             S = Q K_j^T, dP = dO V_j^T (TMEM) -> dS = P o (dP - D) (bf16, smem) -> dQ += dS K_j
//   dkdv    one CTA per 128 keys, loops over 64-query blocks, everything transposed so the
//           key is the TMEM lane:  S^T = K Q_i^T, dP^T = V dO_i^T -> P^T, dS^T (smem) ->
//             dV += P^T dO_i, dK += dS^T Q_i
// P is recomputed from the forward's log2-sum-exp.  Every "second" MMA takes the row-major
// K_j / Q_i / dO_i tile as an MN-major B operand, so no tile is ever transposed in memory.
// TMEM: 192 (dq) / 256 (dkdv) columns per CTA, two CTAs per SM.
constexpr int ATT_BWD_SLOTS = 6;    // dq kernel: three key blocks of K_j / V_j in flight
constexpr int ATT_KV_SLOTS = 5;     // dkdv kernel (two CTAs per SM leave room for five)
constexpr uint32_t ATT_DQ_SMEM = 2 * ATT_TILE_BYTES + ATT_BWD_SLOTS * ATT_HALF_BYTES + ATT_TILE_BYTES + 128 + 1024;
constexpr uint32_t ATT_DKDV_SMEM =
    2 * ATT_TILE_BYTES + ATT_KV_SLOTS * ATT_HALF_BYTES + 2 * ATT_TILE_BYTES + 1024 + 128 + 1024;

struct AttnBwdParams {
    int batch, seq, heads, hidden;
    int causal;
    float scale_log2, scale;
    const float* lse;      // [batch*heads, seq] log2 domain
    const float* delta;    // [batch*heads, seq]
    __nv_bfloat16* dqkv;   // [tokens, 3*hidden]
};

__global__ void attention_bwd_delta_kernel(const __nv_bfloat16* __restrict__ out,
                                           const __nv_bfloat16* __restrict__ dout, float* __restrict__ delta,
                                           int batch, int seq, int heads) {
    // eight lanes per (token, head): 16 bytes of O and dO each
    const long long gid = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
    const long long pair = gid >> 3;
    const int part = static_cast<int>(gid & 7);
    const long long total = static_cast<long long>(batch) * seq * heads;
    float acc = 0.f;
    if (pair < total) {
        const uint4 a = ld_global_nc_v4(out + pair * ATT_D + part * 8);
        const uint4 g = ld_global_nc_v4(dout + pair * ATT_D + part * 8);
        const uint32_t aw[4] = {a.x, a.y, a.z, a.w}, gw[4] = {g.x, g.y, g.z, g.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const float2 x = unpack_bf16x2(aw[i]), y = unpack_bf16x2(gw[i]);
            acc = fmaf(x.x, y.x, fmaf(x.y, y.y, acc));
        }
    }
    acc += __shfl_xor_sync(0xffffffffu, acc, 1);
    acc += __shfl_xor_sync(0xffffffffu, acc, 2);
    acc += __shfl_xor_sync(0xffffffffu, acc, 4);
    if (pair < total && part == 0) {
        const long long token = pair / heads;
        const int h = static_cast<int>(pair - token * heads);
        const long long b = token / seq;
        const int s = static_cast<int>(token - b * seq);
        delta[(b * heads + h) * seq + s] = acc;
    }
}

__global__ void __launch_bounds__(ATT_THREADS, 2)
attention_bwd_dq_kernel(const __grid_constant__ CUtensorMap tmap_qkv128,
                        const __grid_constant__ CUtensorMap tmap_qkv64,
                        const __grid_constant__ CUtensorMap tmap_do128, const AttnBwdParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>(
        (reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
    uint8_t* sQ = smem;
    uint8_t* sdO = sQ + ATT_TILE_BYTES;
    uint8_t* sRing = sdO + ATT_TILE_BYTES;
    uint8_t* sdS = sRing + ATT_BWD_SLOTS * ATT_HALF_BYTES;
    uint64_t* qdo_full = reinterpret_cast<uint64_t*>(sdS + ATT_TILE_BYTES);
    uint64_t* ring_full = qdo_full + 1;
    uint64_t* ring_empty = ring_full + ATT_BWD_SLOTS;
    uint64_t* sc_full = ring_empty + ATT_BWD_SLOTS;
    uint64_t* sc_free = sc_full + 1;
    uint64_t* ds_full = sc_free + 1;
    uint64_t* dq_done = ds_full + 1;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(dq_done + 1);

    const int warp_idx = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);   // provably warp-uniform: role branches and their loops stay on the uniform datapath
    const int lane = threadIdx.x & 31;
    const int nq = p.seq / ATT_BM;
    const int bh = static_cast<int>(blockIdx.x) / nq;
    const int qt = nq - 1 - (static_cast<int>(blockIdx.x) - bh * nq);
    const int b = bh / p.heads;
    const int h = bh - b * p.heads;
    const int q0 = qt * ATT_BM;
    const int row_base = b * p.seq;
    const int nblk = p.causal ? (q0 + ATT_BM) / ATT_BK : p.seq / ATT_BK;

    if (warp_idx == 0 && lane == 0) {
        tma_prefetch_desc(&tmap_qkv128);
        tma_prefetch_desc(&tmap_qkv64);
        tma_prefetch_desc(&tmap_do128);
        mbar_init(qdo_full, 1);
        for (int s = 0; s < ATT_BWD_SLOTS; ++s) {
            mbar_init(&ring_full[s], 1);
            mbar_init(&ring_empty[s], 1);
        }
        mbar_init(sc_full, 1);
        mbar_init(sc_free, 4);
        mbar_init(ds_full, 4);
        mbar_init(dq_done, 1);
        fence_barrier_init();
    }
    if (warp_idx == 1) tmem_alloc<256>(tmem_slot);
    tcgen05_fence_before();
    __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const uint32_t tS = tmem_base;
    const uint32_t tdP = tmem_base + ATT_BK;
    const uint32_t tdQ = tmem_base + 2 * ATT_BK;

    if (warp_idx == 0) {   // whole warp in the loop, one elected lane issues (see attention_fwd_kernel)
        {
            if (elect_one()) {
                mbar_expect_tx(qdo_full, 2 * ATT_TILE_BYTES);
                tma_load_2d(&tmap_qkv128, qdo_full, sQ, h * ATT_D, row_base + q0);
                tma_load_2d(&tmap_do128, qdo_full, sdO, h * ATT_D, row_base + q0);
            }
            __syncwarp();
            for (int i = 0; i < 2 * nblk; ++i) {
                const int slot = i % ATT_BWD_SLOTS;
                const uint32_t use = static_cast<uint32_t>(i / ATT_BWD_SLOTS);
                mbar_wait(&ring_empty[slot], (use & 1) ^ 1);
                const int col = ((i & 1) ? 2 * p.hidden : p.hidden) + h * ATT_D;   // K_j then V_j
                if (elect_one()) {
                    mbar_expect_tx(&ring_full[slot], ATT_HALF_BYTES);
                    tma_load_2d(&tmap_qkv64, &ring_full[slot], sRing + slot * ATT_HALF_BYTES, col,
                                row_base + (i >> 1) * ATT_BK);
                }
                __syncwarp();
            }
        }
    } else if (warp_idx == 1) {
        {
            const uint32_t idesc_sc = make_idesc_bf16(ATT_BM, ATT_BK, 0, 0);
            const uint32_t idesc_dq = make_idesc_bf16(ATT_BM, ATT_D, 0, 1);
            const uint64_t dq = make_umma_desc_sw128(smem_u32(sQ), 16, 1024);
            const uint64_t ddo = make_umma_desc_sw128(smem_u32(sdO), 16, 1024);
            const uint64_t dds = make_umma_desc_sw128(smem_u32(sdS), 16, 1024);
            auto issue_sc = [&](int j) {
                const int ik = 2 * j, iv = 2 * j + 1;
                const int sk = ik % ATT_BWD_SLOTS, sv = iv % ATT_BWD_SLOTS;
                mbar_wait(&ring_full[sk], static_cast<uint32_t>(ik / ATT_BWD_SLOTS) & 1);
                mbar_wait(&ring_full[sv], static_cast<uint32_t>(iv / ATT_BWD_SLOTS) & 1);
                tcgen05_fence_after();
                const uint64_t dk = make_umma_desc_sw128(smem_u32(sRing + sk * ATT_HALF_BYTES), 16, 1024);
                const uint64_t dv = make_umma_desc_sw128(smem_u32(sRing + sv * ATT_HALF_BYTES), 16, 1024);
                if (elect_one()) {
#pragma unroll
                    for (int k = 0; k < ATT_D / 16; ++k) umma_bf16(tS, dq + k * 2, dk + k * 2, idesc_sc, k > 0);
#pragma unroll
                    for (int k = 0; k < ATT_D / 16; ++k) umma_bf16(tdP, ddo + k * 2, dv + k * 2, idesc_sc, k > 0);
                    umma_commit(&ring_empty[sv]);
                    umma_commit(sc_full);
                }
                __syncwarp();
            };
            mbar_wait(qdo_full, 0);
            issue_sc(0);
            for (int j = 0; j < nblk; ++j) {
                if (j + 1 < nblk) {
                    mbar_wait(sc_free, j & 1);
                    tcgen05_fence_after();
                    issue_sc(j + 1);
                }
                mbar_wait(ds_full, j & 1);
                tcgen05_fence_after();
                const int sk = (2 * j) % ATT_BWD_SLOTS;
                const uint64_t dkmn =
                    make_umma_desc_sw128(smem_u32(sRing + sk * ATT_HALF_BYTES), 8192, 1024);
                if (elect_one()) {
#pragma unroll
                    for (int k = 0; k < ATT_BK / 16; ++k)
                        umma_bf16(tdQ, dds + k * 2, dkmn + k * 128, idesc_dq, (j > 0 || k > 0) ? 1u : 0u);
                    umma_commit(&ring_empty[sk]);
                    umma_commit(dq_done);
                }
                __syncwarp();
            }
        }
    } else {
        const int quarter = warp_idx & 3;
        const int row = quarter * 32 + lane;
        const uint32_t lane_addr = static_cast<uint32_t>(quarter * 32) << 16;
        const uint32_t sdS_u32 = smem_u32(sdS);
        const size_t stat = static_cast<size_t>(bh) * p.seq + q0 + row;
        const float neg_l = -p.lse[stat];
        const float drow = p.delta[stat];
        for (int j = 0; j < nblk; ++j) {
            const int rel = q0 + row - j * ATT_BK;   // keys with column index > rel are masked
            const bool diag = p.causal && (j * ATT_BK + ATT_BK - 1 > q0);
            mbar_wait(sc_full, j & 1);
            tcgen05_fence_after();
            // pull the whole block (64 scores + 64 dP values) into registers and hand the TMEM columns
            // back at once: the next block's score MMAs run while this one's exponentials do
            uint32_t rs[64], rp[64];
            tmem_ld_32x32(tS + lane_addr, rs);
            tmem_ld_32x32(tS + lane_addr + 32, rs + 32);
            tmem_ld_32x32(tdP + lane_addr, rp);
            tmem_ld_32x32(tdP + lane_addr + 32, rp + 32);
            tmem_ld_wait();
            tcgen05_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(sc_free);
            if (j > 0) mbar_wait(dq_done, (j - 1) & 1);   // previous dS tile consumed
#pragma unroll
            for (int c = 0; c < 2; ++c) {
                uint32_t w[16];
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    const int e = c * 32 + 2 * i;
                    float p0 = ex2_approx(fmaf(__uint_as_float(rs[e]), p.scale_log2, neg_l));
                    float p1 = ex2_approx(fmaf(__uint_as_float(rs[e + 1]), p.scale_log2, neg_l));
                    if (diag) {
                        if (e > rel) p0 = 0.f;
                        if (e + 1 > rel) p1 = 0.f;
                    }
                    w[i] = pack_bf16x2(p0 * (__uint_as_float(rp[e]) - drow),
                                       p1 * (__uint_as_float(rp[e + 1]) - drow));
                }
#pragma unroll
                for (int v = 0; v < 4; ++v)
                    st_shared_v4(sdS_u32 + panel_unit_offset(row, c * 4 + v),
                                 make_uint4(w[4 * v], w[4 * v + 1], w[4 * v + 2], w[4 * v + 3]));
            }
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) mbar_arrive(ds_full);
        }
        mbar_wait(dq_done, (nblk - 1) & 1);
        tcgen05_fence_after();
        const size_t token = static_cast<size_t>(row_base + q0 + row);
        __nv_bfloat16* grow = p.dqkv + token * (3 * p.hidden) + h * ATT_D;
#pragma unroll 1
        for (int c = 0; c < 2; ++c) {
            uint32_t r[32];
            tmem_ld_32x32(tdQ + lane_addr + c * 32, r);
            tmem_ld_wait();
#pragma unroll
            for (int v = 0; v < 4; ++v) {
                uint4 o;
                o.x = pack_bf16x2(__uint_as_float(r[8 * v]) * p.scale, __uint_as_float(r[8 * v + 1]) * p.scale);
                o.y = pack_bf16x2(__uint_as_float(r[8 * v + 2]) * p.scale, __uint_as_float(r[8 * v + 3]) * p.scale);
                o.z = pack_bf16x2(__uint_as_float(r[8 * v + 4]) * p.scale, __uint_as_float(r[8 * v + 5]) * p.scale);
                o.w = pack_bf16x2(__uint_as_float(r[8 * v + 6]) * p.scale, __uint_as_float(r[8 * v + 7]) * p.scale);
                st_global_v4(grow + c * 32 + v * 8, o);
            }
        }
    }

    __syncwarp();
    tcgen05_fence_before();
    __syncthreads();
    if (warp_idx == 1) {
        tcgen05_fence_after();
        tmem_dealloc<256>(tmem_base);
    }
}

__global__ void __launch_bounds__(ATT_THREADS, 2)
attention_bwd_dkdv_kernel(const __grid_constant__ CUtensorMap tmap_qkv128,
                          const __grid_constant__ CUtensorMap tmap_qkv64,
                          const __grid_constant__ CUtensorMap tmap_do64, const AttnBwdParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>(
        (reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
    uint8_t* sK = smem;
    uint8_t* sV = sK + ATT_TILE_BYTES;
    uint8_t* sRing = sV + ATT_TILE_BYTES;
    uint8_t* sPt = sRing + ATT_KV_SLOTS * ATT_HALF_BYTES;
    uint8_t* sdSt = sPt + ATT_TILE_BYTES;
    float* sStat = reinterpret_cast<float*>(sdSt + ATT_TILE_BYTES);   // [2 buffers][lse 64 | delta 64]
    uint64_t* kv_full = reinterpret_cast<uint64_t*>(sStat + 256);
    uint64_t* ring_full = kv_full + 1;
    uint64_t* ring_empty = ring_full + ATT_KV_SLOTS;
    uint64_t* sc_full = ring_empty + ATT_KV_SLOTS;
    uint64_t* sc_free = sc_full + 1;
    uint64_t* pd_full = sc_free + 1;
    uint64_t* acc_done = pd_full + 1;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_done + 1);

    const int warp_idx = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);   // provably warp-uniform: role branches and their loops stay on the uniform datapath
    const int lane = threadIdx.x & 31;
    const int nk = p.seq / ATT_BM;
    const int bh = static_cast<int>(blockIdx.x) / nk;
    const int kt = static_cast<int>(blockIdx.x) - bh * nk;   // causal: early key tiles are the heavy ones
    const int b = bh / p.heads;
    const int h = bh - b * p.heads;
    const int k0 = kt * ATT_BM;
    const int row_base = b * p.seq;
    const int i0 = p.causal ? k0 / ATT_BK : 0;          // first 64-query block that sees these keys
    const int nblk = p.seq / ATT_BK - i0;

    if (warp_idx == 0 && lane == 0) {
        tma_prefetch_desc(&tmap_qkv128);
        tma_prefetch_desc(&tmap_qkv64);
        tma_prefetch_desc(&tmap_do64);
        mbar_init(kv_full, 1);
        for (int s = 0; s < ATT_KV_SLOTS; ++s) {
            mbar_init(&ring_full[s], 1);
            mbar_init(&ring_empty[s], 1);
        }
        mbar_init(sc_full, 1);
        mbar_init(sc_free, 4);
        mbar_init(pd_full, 4);
        mbar_init(acc_done, 1);
        fence_barrier_init();
    }
    if (warp_idx == 1) tmem_alloc<256>(tmem_slot);
    tcgen05_fence_before();
    __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const uint32_t tSt = tmem_base;
    const uint32_t tdPt = tmem_base + ATT_BK;
    const uint32_t tdV = tmem_base + 2 * ATT_BK;
    const uint32_t tdK = tmem_base + 3 * ATT_BK;

    if (warp_idx == 0) {   // whole warp in the loop, one elected lane issues (see attention_fwd_kernel)
        {
            if (elect_one()) {
                mbar_expect_tx(kv_full, 2 * ATT_TILE_BYTES);
                tma_load_2d(&tmap_qkv128, kv_full, sK, p.hidden + h * ATT_D, row_base + k0);
                tma_load_2d(&tmap_qkv128, kv_full, sV, 2 * p.hidden + h * ATT_D, row_base + k0);
            }
            __syncwarp();
            for (int n = 0; n < 2 * nblk; ++n) {
                const int slot = n % ATT_KV_SLOTS;
                const uint32_t use = static_cast<uint32_t>(n / ATT_KV_SLOTS);
                mbar_wait(&ring_empty[slot], (use & 1) ^ 1);
                const int qrow = row_base + (i0 + (n >> 1)) * ATT_BK;
                if (elect_one()) {
                    mbar_expect_tx(&ring_full[slot], ATT_HALF_BYTES);
                    if (n & 1)
                        tma_load_2d(&tmap_do64, &ring_full[slot], sRing + slot * ATT_HALF_BYTES, h * ATT_D, qrow);
                    else
                        tma_load_2d(&tmap_qkv64, &ring_full[slot], sRing + slot * ATT_HALF_BYTES, h * ATT_D, qrow);
                }
                __syncwarp();
            }
        }
    } else if (warp_idx == 1) {
        {
            const uint32_t idesc_sc = make_idesc_bf16(ATT_BM, ATT_BK, 0, 0);
            const uint32_t idesc_acc = make_idesc_bf16(ATT_BM, ATT_D, 0, 1);
            const uint64_t dk = make_umma_desc_sw128(smem_u32(sK), 16, 1024);
            const uint64_t dv = make_umma_desc_sw128(smem_u32(sV), 16, 1024);
            const uint64_t dpt = make_umma_desc_sw128(smem_u32(sPt), 16, 1024);
            const uint64_t ddst = make_umma_desc_sw128(smem_u32(sdSt), 16, 1024);
            auto issue_sc = [&](int n) {
                const int iq = 2 * n, ido = 2 * n + 1;
                const int sq = iq % ATT_KV_SLOTS, sdo = ido % ATT_KV_SLOTS;
                mbar_wait(&ring_full[sq], static_cast<uint32_t>(iq / ATT_KV_SLOTS) & 1);
                mbar_wait(&ring_full[sdo], static_cast<uint32_t>(ido / ATT_KV_SLOTS) & 1);
                tcgen05_fence_after();
                const uint64_t dq = make_umma_desc_sw128(smem_u32(sRing + sq * ATT_HALF_BYTES), 16, 1024);
                const uint64_t ddo = make_umma_desc_sw128(smem_u32(sRing + sdo * ATT_HALF_BYTES), 16, 1024);
                if (elect_one()) {
#pragma unroll
                    for (int k = 0; k < ATT_D / 16; ++k) umma_bf16(tSt, dk + k * 2, dq + k * 2, idesc_sc, k > 0);
#pragma unroll
                    for (int k = 0; k < ATT_D / 16; ++k) umma_bf16(tdPt, dv + k * 2, ddo + k * 2, idesc_sc, k > 0);
                    umma_commit(sc_full);
                }
                __syncwarp();
            };
            mbar_wait(kv_full, 0);
            issue_sc(0);
            for (int n = 0; n < nblk; ++n) {
                if (n + 1 < nblk) {
                    mbar_wait(sc_free, n & 1);
                    tcgen05_fence_after();
                    issue_sc(n + 1);
                }
                mbar_wait(pd_full, n & 1);
                tcgen05_fence_after();
                const int sq = (2 * n) % ATT_KV_SLOTS, sdo = (2 * n + 1) % ATT_KV_SLOTS;
                const uint64_t dqmn = make_umma_desc_sw128(smem_u32(sRing + sq * ATT_HALF_BYTES), 8192, 1024);
                const uint64_t ddomn = make_umma_desc_sw128(smem_u32(sRing + sdo * ATT_HALF_BYTES), 8192, 1024);
                if (elect_one()) {
#pragma unroll
                    for (int k = 0; k < ATT_BK / 16; ++k)
                        umma_bf16(tdV, dpt + k * 2, ddomn + k * 128, idesc_acc, (n > 0 || k > 0) ? 1u : 0u);
#pragma unroll
                    for (int k = 0; k < ATT_BK / 16; ++k)
                        umma_bf16(tdK, ddst + k * 2, dqmn + k * 128, idesc_acc, (n > 0 || k > 0) ? 1u : 0u);
                    umma_commit(&ring_empty[sq]);
                    umma_commit(&ring_empty[sdo]);
                    umma_commit(acc_done);
                }
                __syncwarp();
            }
        }
    } else {
        const int quarter = warp_idx & 3;
        const int row = quarter * 32 + lane;             // key within the tile
        const int tid128 = (warp_idx - 2) * 32 + lane;
        const uint32_t lane_addr = static_cast<uint32_t>(quarter * 32) << 16;
        const uint32_t sPt_u32 = smem_u32(sPt), sdSt_u32 = smem_u32(sdSt);
        const float* stat_base = (tid128 < 64 ? p.lse : p.delta) + static_cast<size_t>(bh) * p.seq + (tid128 & 63);
        for (int n = 0; n < nblk; ++n) {
            const int qb = (i0 + n) * ATT_BK;            // first query of this block
            const int buf = n & 1;
            sStat[buf * 128 + tid128] = __ldg(stat_base + qb);
            asm volatile("bar.sync 1, 128;" ::: "memory");
            const bool diag = p.causal && (qb < k0 + ATT_BM - 1);
            const int rel = k0 + row - qb;               // query columns < rel cannot see this key
            mbar_wait(sc_full, n & 1);
            tcgen05_fence_after();
            if (n > 0) mbar_wait(acc_done, (n - 1) & 1);   // P^T / dS^T tiles consumed
#pragma unroll 1
            for (int c = 0; c < 2; ++c) {
                uint32_t rs[32], rp[32];
                tmem_ld_32x32(tSt + lane_addr + c * 32, rs);
                tmem_ld_32x32(tdPt + lane_addr + c * 32, rp);
                tmem_ld_wait();
                uint32_t wp[16], wd[16];
                const float4* l4 = reinterpret_cast<const float4*>(sStat + buf * 128 + c * 32);
                const float4* d4 = reinterpret_cast<const float4*>(sStat + buf * 128 + 64 + c * 32);
#pragma unroll
                for (int g = 0; g < 8; ++g) {
                    const float4 lq = l4[g];
                    const float4 dq4 = d4[g];
                    const float lv[4] = {lq.x, lq.y, lq.z, lq.w};
                    const float dvv[4] = {dq4.x, dq4.y, dq4.z, dq4.w};
                    float pr[4], ds[4];
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        const int col = g * 4 + e;
                        float pv = ex2_approx(fmaf(__uint_as_float(rs[col]), p.scale_log2, -lv[e]));
                        if (diag && (c * 32 + col < rel)) pv = 0.f;
                        pr[e] = pv;
                        ds[e] = pv * (__uint_as_float(rp[col]) - dvv[e]);
                    }
                    wp[2 * g] = pack_bf16x2(pr[0], pr[1]);
                    wp[2 * g + 1] = pack_bf16x2(pr[2], pr[3]);
                    wd[2 * g] = pack_bf16x2(ds[0], ds[1]);
                    wd[2 * g + 1] = pack_bf16x2(ds[2], ds[3]);
                }
#pragma unroll
                for (int v = 0; v < 4; ++v) {
                    const uint32_t off = panel_unit_offset(row, c * 4 + v);
                    st_shared_v4(sPt_u32 + off, make_uint4(wp[4 * v], wp[4 * v + 1], wp[4 * v + 2], wp[4 * v + 3]));
                    st_shared_v4(sdSt_u32 + off, make_uint4(wd[4 * v], wd[4 * v + 1], wd[4 * v + 2], wd[4 * v + 3]));
                }
            }
            tcgen05_fence_before();
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) {
                mbar_arrive(sc_free);
                mbar_arrive(pd_full);
            }
        }
        mbar_wait(acc_done, (nblk - 1) & 1);
        tcgen05_fence_after();
        const size_t token = static_cast<size_t>(row_base + k0 + row);
        __nv_bfloat16* grow = p.dqkv + token * (3 * p.hidden) + h * ATT_D;
#pragma unroll 1
        for (int c = 0; c < 4; ++c) {
            // c = 0,1: dV columns; c = 2,3: dK columns (scaled)
            uint32_t r[32];
            tmem_ld_32x32(tdV + lane_addr + c * 32, r);
            tmem_ld_wait();
            const float f = c < 2 ? 1.f : p.scale;
            __nv_bfloat16* dst = grow + (c < 2 ? 2 * p.hidden + c * 32 : p.hidden + (c - 2) * 32);
#pragma unroll
            for (int v = 0; v < 4; ++v) {
                uint4 o;
                o.x = pack_bf16x2(__uint_as_float(r[8 * v]) * f, __uint_as_float(r[8 * v + 1]) * f);
                o.y = pack_bf16x2(__uint_as_float(r[8 * v + 2]) * f, __uint_as_float(r[8 * v + 3]) * f);
                o.z = pack_bf16x2(__uint_as_float(r[8 * v + 4]) * f, __uint_as_float(r[8 * v + 5]) * f);
                o.w = pack_bf16x2(__uint_as_float(r[8 * v + 6]) * f, __uint_as_float(r[8 * v + 7]) * f);
                st_global_v4(dst + v * 8, o);
            }
        }
    }

    __syncwarp();
    tcgen05_fence_before();
    __syncthreads();
    if (warp_idx == 1) {
        tcgen05_fence_after();
        tmem_dealloc<256>(tmem_base);
    }
}

static int set_smem_once(const void* kernel, uint32_t bytes, int which) {
    static bool done[8][64] = {};
    int dev = 0;
    TONY_CUDA_CHECK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return TONY_ERR_UNSUPPORTED;
    if (!done[which][dev]) {
        TONY_CUDA_CHECK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             static_cast<int>(bytes)));
        done[which][dev] = true;
    }
    return TONY_OK;
}

static int check_attn_shape(const char* what, int batch, int seq, int heads, int head_dim) {
    if (head_dim != ATT_D || seq <= 0 || seq % ATT_BM != 0 || batch <= 0 || heads <= 0) {
        set_last_error(what, "needs head_dim == 64 and seq a multiple of 128");
        return TONY_ERR_UNSUPPORTED;
    }
    return TONY_OK;
}

}  // namespace tony

using namespace tony;

extern "C" int tony_attention_fwd_bf16(const void* qkv, void* out, float* lse, int batch, int seq,
                                       int heads, int head_dim, int causal, float scale, void* stream) {
    int rc = check_attn_shape("tony_attention_fwd_bf16", batch, seq, heads, head_dim);
    if (rc != TONY_OK) return rc;
    if ((reinterpret_cast<uintptr_t>(qkv) | reinterpret_cast<uintptr_t>(out)) & 15) {
        set_last_error("tony_attention_fwd_bf16", "qkv/out must be 16-byte aligned");
        return TONY_ERR_INVALID;
    }
    const int hidden = heads * ATT_D;
    const long long tokens = static_cast<long long>(batch) * seq;
    CUtensorMap tmap128, tmap64;
    rc = make_tmap_2d_bf16(&tmap128, qkv, 3LL * hidden, tokens, 3LL * hidden, ATT_D, ATT_BM);
    if (rc != TONY_OK) return rc;
    rc = make_tmap_2d_bf16(&tmap64, qkv, 3LL * hidden, tokens, 3LL * hidden, ATT_D, ATT_BK);
    if (rc != TONY_OK) return rc;
    rc = set_smem_once(reinterpret_cast<const void*>(attention_fwd_kernel), ATT_FWD_SMEM, 0);
    if (rc != TONY_OK) return rc;
    AttnFwdParams p;
    p.batch = batch;
    p.seq = seq;
    p.heads = heads;
    p.hidden = hidden;
    p.causal = causal;
    p.scale_log2 = scale * 1.4426950408889634f;
    p.out = static_cast<__nv_bfloat16*>(out);
    p.lse = lse;
    p.num_items = batch * heads * (seq / ATT_BM);
    const int grid = p.num_items < 2 * device_sm_count() ? p.num_items : 2 * device_sm_count();
    attention_fwd_kernel<<<grid, ATT_THREADS, ATT_FWD_SMEM, static_cast<cudaStream_t>(stream)>>>(tmap128, tmap64, p);
    TONY_CUDA_CHECK(cudaGetLastError());
    count_launch();
    return TONY_OK;
}

extern "C" int tony_attention_bwd_bf16(const void* qkv, const void* out, const void* dout, const float* lse,
                                       float* delta, void* dqkv, int batch, int seq, int heads,
                                       int head_dim, int causal, float scale, void* stream) {
    int rc = check_attn_shape("tony_attention_bwd_bf16", batch, seq, heads, head_dim);
    if (rc != TONY_OK) return rc;
    if ((reinterpret_cast<uintptr_t>(qkv) | reinterpret_cast<uintptr_t>(out) |
         reinterpret_cast<uintptr_t>(dout) | reinterpret_cast<uintptr_t>(dqkv)) & 15) {
        set_last_error("tony_attention_bwd_bf16", "tensors must be 16-byte aligned");
        return TONY_ERR_INVALID;
    }
    const int hidden = heads * ATT_D;
    const long long tokens = static_cast<long long>(batch) * seq;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    CUtensorMap qkv128, qkv64, do128, do64;
    if ((rc = make_tmap_2d_bf16(&qkv128, qkv, 3LL * hidden, tokens, 3LL * hidden, ATT_D, ATT_BM)) != TONY_OK) return rc;
    if ((rc = make_tmap_2d_bf16(&qkv64, qkv, 3LL * hidden, tokens, 3LL * hidden, ATT_D, ATT_BK)) != TONY_OK) return rc;
    if ((rc = make_tmap_2d_bf16(&do128, dout, hidden, tokens, hidden, ATT_D, ATT_BM)) != TONY_OK) return rc;
    if ((rc = make_tmap_2d_bf16(&do64, dout, hidden, tokens, hidden, ATT_D, ATT_BK)) != TONY_OK) return rc;
    if ((rc = set_smem_once(reinterpret_cast<const void*>(attention_bwd_dq_kernel), ATT_DQ_SMEM, 1)) != TONY_OK) return rc;
    if ((rc = set_smem_once(reinterpret_cast<const void*>(attention_bwd_dkdv_kernel), ATT_DKDV_SMEM, 2)) != TONY_OK) return rc;

    const long long pairs = tokens * heads;
    const int dthreads = 256;
    const long long dblocks = (pairs * 8 + dthreads - 1) / dthreads;
    attention_bwd_delta_kernel<<<static_cast<unsigned>(dblocks), dthreads, 0, st>>>(
        static_cast<const __nv_bfloat16*>(out), static_cast<const __nv_bfloat16*>(dout), delta, batch, seq, heads);
    TONY_CUDA_CHECK(cudaGetLastError());

    AttnBwdParams p;
    p.batch = batch;
    p.seq = seq;
    p.heads = heads;
    p.hidden = hidden;
    p.causal = causal;
    p.scale = scale;
    p.scale_log2 = scale * 1.4426950408889634f;
    p.lse = lse;
    p.delta = delta;
    p.dqkv = static_cast<__nv_bfloat16*>(dqkv);
    const int grid = batch * heads * (seq / ATT_BM);
    attention_bwd_dq_kernel<<<grid, ATT_THREADS, ATT_DQ_SMEM, st>>>(qkv128, qkv64, do128, p);
    TONY_CUDA_CHECK(cudaGetLastError());
    attention_bwd_dkdv_kernel<<<grid, ATT_THREADS, ATT_DKDV_SMEM, st>>>(qkv128, qkv64, do64, p);
    TONY_CUDA_CHECK(cudaGetLastError());
    count_launch(3);
    return TONY_OK;
}
