// tony_b200 kernel library — common device helpers for sm_100a (B200).
//
// Everything here is raw PTX wrapped in small inline functions: mbarrier, TMA
// (cp.async.bulk.tensor), tcgen05 (alloc / mma / commit / ld), proxy fences,
// system-scope acquire/release for the cross-GPU flag protocol, and the few
// numeric helpers the epilogues share.  The reference (zuston/TonY) has no
// device code at all, so nothing here has a counterpart there.
#pragma once

#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "host_math.h"
#include "tony_b200_kernels.h"

namespace tony {

#define TONY_CUDA_CHECK(expr)                                  \
    do {                                                       \
        cudaError_t _e = (expr);                               \
        if (_e != cudaSuccess) {                               \
            tony::set_last_error(#expr, cudaGetErrorString(_e)); \
            return TONY_ERR_CUDA;                        \
        }                                                      \
    } while (0)

void set_last_error(const char* what, const char* detail);
void count_launch(int n = 1);  // bench accounting: kernels launched by this library
int device_sm_count();         // SM count of the current device (cached)
// The column reductions keep their arrival counters in the first 4 KB of the caller's scratch and leave them at
// zero, so the caller zeroes that head ONCE; TONY_B200_CLEAR_COUNTERS=1 clears it on the stream before every call
// again (one memset node per launch: ~250 per BERT-large step).
bool clear_counters_each_call();
bool pdl_enabled();            // TONY_B200_PDL != 0 (default on)

// Launch with programmatic stream serialization: the kernel's CTAs may be placed while the previous kernel of
// the stream is still draining.  Every kernel launched this way starts with pdl_wait() (blocks until the previous
// kernel has completed and its writes are visible) followed by pdl_launch_dependents(), so a chain of them
// overlaps each launch latency with the predecessor's tail and nothing else.
template <typename Kern, typename... Args>
static inline cudaError_t launch_pdl_kernel(Kern kern, dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                                            Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl_enabled() ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kern, args...);
}
// 2-D bf16 tensor map, 128-byte swizzle, `inner` contiguous (defined in gemm_tcgen05.cu)
int make_tmap_2d_bf16(CUtensorMap* map, const void* ptr, long long inner_extent, long long outer_extent,
                      long long ld, int box_inner, int box_outer);
// same with a 64-byte swizzle (32 bf16 per row): the layout of the epilogue's per-warp staging tiles
int make_tmap_2d_bf16_sw64(CUtensorMap* map, const void* ptr, long long inner_extent, long long outer_extent,
                           long long ld, int box_inner, int box_outer);
// C (+)= alpha * sum over `splits` fp32 partial planes [M][N] (defined in gemm_tcgen05.cu)
int launch_splitk_reduce(const float* ws, int splits, int M, int N, __nv_bfloat16* C, long long ldc, float alpha,
                         unsigned int flags, cudaStream_t stream);

// ----------------------------------------------------------------------------
// Small device utilities
// ----------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

__device__ __forceinline__ uint32_t lane_id() {
    uint32_t l;
    asm volatile("mov.u32 %0, %%laneid;" : "=r"(l));
    return l;
}

__device__ __forceinline__ bool elect_one() {
    uint32_t pred = 0;
    asm volatile(
        "{\n\t"
        ".reg .pred P;\n\t"
        "elect.sync _|P, 0xffffffff;\n\t"
        "selp.u32 %0, 1, 0, P;\n\t"
        "}\n"
        : "=r"(pred));
    return pred != 0;
}

// ----------------------------------------------------------------------------
// mbarrier
// ----------------------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}

__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
                 "r"(bytes)
                 : "memory");
}

__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t"
        ".reg .pred P;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 P, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, P;\n\t"
        "}\n"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}

// Non-blocking probe (no hardware suspend): for issuers that poll several barriers.
__device__ __forceinline__ bool mbar_test_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t"
        ".reg .pred P;\n\t"
        "mbarrier.test_wait.parity.shared::cta.b64 P, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, P;\n\t"
        "}\n"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}

__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    while (!mbar_try_wait(bar, parity)) {
    }
}

// For producer / issuer threads whose wake-up latency is not critical: sleep between probes so
// the spin does not compete with the math warps for issue slots and the MIO queue.
__device__ __forceinline__ void mbar_wait_relaxed(uint64_t* bar, uint32_t parity, uint32_t ns = 64) {
    while (!mbar_try_wait(bar, parity)) __nanosleep(ns);
}

// Make mbarrier.init visible to the async proxy (TMA / tcgen05.commit).
__device__ __forceinline__ void fence_barrier_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}

// Order generic-proxy writes (st.global / st.shared) before later async-proxy
// reads (TMA loads) issued by this thread.
__device__ __forceinline__ void fence_proxy_async() {
    asm volatile("fence.proxy.async;" ::: "memory");
}

// Same, restricted to this CTA's shared memory (st.shared operand tiles -> tcgen05.mma reads);
// does not wait for outstanding global stores.
__device__ __forceinline__ void fence_proxy_async_smem() {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

// ----------------------------------------------------------------------------
// TMA: 2-D tiled bulk tensor load global -> shared, completion on an mbarrier
// ----------------------------------------------------------------------------
__device__ __forceinline__ void tma_load_2d(const CUtensorMap* map, uint64_t* bar, void* smem_dst,
                                            int32_t c0, int32_t c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%3, %4}], [%2];"
        :
        : "r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)),
          "r"(c0), "r"(c1)
        : "memory");
}

// TMA store: shared (this CTA) -> global tile, tracked by the thread's bulk async-group.  The map clips
// rows / columns beyond the tensor, so ragged tiles need no predicates.
__device__ __forceinline__ void tma_store_2d(const CUtensorMap* map, uint32_t smem_src, int32_t c0, int32_t c1) {
    asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
                 :
                 : "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_src), "r"(c0), "r"(c1)
                 : "memory");
}
__device__ __forceinline__ void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
// all of this thread's committed stores have finished READING shared memory (the buffer may be rewritten)
__device__ __forceinline__ void tma_store_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
// ... and have completed (writes performed)
__device__ __forceinline__ void tma_store_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

// Tiled 2-D load delivered to the same shared-memory offset of every CTA in `cta_mask` of this cluster;
// each destination's mbarrier (same offset) is credited with the bytes.  Loads of this kind read only
// this GPU's memory (weights / activations of the local rank).
__device__ __forceinline__ void tma_load_2d_multicast(const CUtensorMap* map, uint64_t* bar, void* smem_dst,
                                                      int32_t c0, int32_t c1, uint16_t cta_mask) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster"
        " [%0], [%1, {%3, %4}], [%2], %5;"
        :
        : "r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1),
          "h"(cta_mask)
        : "memory");
}

// tcgen05.commit whose arrival goes to the barrier at this offset in every CTA of `mask` (1-CTA MMAs)
__device__ __forceinline__ void umma_commit_multicast(uint64_t* bar, uint16_t mask) {
    asm volatile(
        "tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
            smem_u32(bar)),
        "h"(mask)
        : "memory");
}

// 1-D bulk copies (UBLKCP): global -> shared with mbarrier completion, shared -> global in a bulk group.
// `src` may be peer-mapped memory (a plain, non-multicast load); sizes / addresses multiples of 16 bytes.
__device__ __forceinline__ void bulk_load_1d(uint32_t smem_dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 :
                 : "r"(smem_dst), "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void bulk_store_1d(void* dst, uint32_t smem_src, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
                 :
                 : "l"(dst), "r"(smem_src), "r"(bytes)
                 : "memory");
}
template <int N>
__device__ __forceinline__ void bulk_wait_read() {
    asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}

// 4-D tiled load (channels, width, height, image); out-of-range coordinates are zero-filled
__device__ __forceinline__ void tma_load_4d(const CUtensorMap* map, uint64_t* bar, void* smem_dst, int32_t c0,
                                            int32_t c1, int32_t c2, int32_t c3) {
    asm volatile(
        "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%3, %4, %5, %6}], [%2];"
        :
        : "r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1),
          "r"(c2), "r"(c3)
        : "memory");
}

__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* map) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(map)) : "memory");
}

// ----------------------------------------------------------------------------
// tcgen05 / TMEM
// ----------------------------------------------------------------------------
template <uint32_t kCols>
__device__ __forceinline__ void tmem_alloc(uint32_t* smem_result) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                     smem_u32(smem_result)),
                 "n"(kCols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}

template <uint32_t kCols>
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "n"(kCols)
                 : "memory");
}

__device__ __forceinline__ void tcgen05_fence_before() {
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tcgen05_fence_after() {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}

// D[tmem] (+)= A[smem desc] * B[smem desc], bf16 inputs, fp32 accumulate.
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b,
                                          uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
        "}\n"
        :
        : "r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
}

// Same with the A operand read from tensor memory (M = 128 lanes, 16 bf16 = 8 columns per
// instruction): D[tmem] (+)= A[tmem] * B[smem desc].
__device__ __forceinline__ void umma_bf16_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t desc_b,
                                             uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t"
        "}\n"
        :
        : "r"(tmem_d), "r"(tmem_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
}

// Arrive on an mbarrier once every previously issued tcgen05.mma has retired.
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(
                     smem_u32(bar))
                 : "memory");
}

// 32 lanes x 32 consecutive fp32 columns -> 32 registers per thread (r must be a fully
// unrolled register array: every index below is a compile-time constant).
__device__ __forceinline__ void tmem_ld_32x32(uint32_t taddr, uint32_t* r) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]),
          "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]),
          "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]),
          "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]),
          "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
}

__device__ __forceinline__ void tmem_ld_wait() {
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// 32 registers per thread -> 32 lanes x 32 consecutive fp32 columns (inverse of tmem_ld_32x32).
__device__ __forceinline__ void tmem_st_32x32(uint32_t taddr, const uint32_t* r) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
        "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
        "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};"
        :
        : "r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]),
          "r"(r[7]), "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]),
          "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]), "r"(r[18]), "r"(r[19]), "r"(r[20]),
          "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]),
          "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31])
        : "memory");
}

__device__ __forceinline__ void tmem_st_wait() {
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
}

__device__ __forceinline__ float ex2_approx(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

__device__ __forceinline__ float lg2_approx(float x) {
    float y;
    asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// ---- programmatic dependent launch -------------------------------------------------------
// A kernel launched with the programmatic-stream-serialization attribute may start while the
// previous kernel in the stream is still draining; it must not touch global memory before
// pdl_wait() (returns at once when the kernel was launched normally).  pdl_launch_dependents()
// lets the NEXT kernel's CTAs be scheduled as soon as every CTA of this grid has called it.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() {
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}

// ---- CTA-pair (cta_group::2) variants ----------------------------------------------------
__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// shared::cluster address of `local_smem_addr` as seen in CTA `rank` of this cluster
__device__ __forceinline__ uint32_t mapa_u32(uint32_t local_smem_addr, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local_smem_addr), "r"(rank));
    return r;
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
template <uint32_t kCols>
__device__ __forceinline__ void tmem_alloc_2cta(uint32_t* smem_result) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                     smem_u32(smem_result)),
                 "n"(kCols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
template <uint32_t kCols>
__device__ __forceinline__ void tmem_dealloc_2cta(uint32_t taddr) {
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "n"(kCols) : "memory");
}
// TMA load issued by either CTA of a pair; completion bytes are credited to the mbarrier at
// shared::cluster address `bar_cluster_addr` (the leader CTA's barrier).
__device__ __forceinline__ void tma_load_2d_2cta(const CUtensorMap* map, uint32_t bar_cluster_addr,
                                                 void* smem_dst, int32_t c0, int32_t c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%3, %4}], [%2];"
        :
        : "r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(map)), "r"(bar_cluster_addr),
          "r"(c0), "r"(c1)
        : "memory");
}
// D[tmem of both CTAs] (+)= A * B over the CTA pair: M = 256 (128 rows per CTA), each CTA
// holds half of B's N rows.  Issued by one thread of the leader CTA only.
__device__ __forceinline__ void umma_bf16_2cta(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b,
                                               uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t"
        "}\n"
        :
        : "r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
}
// Arrive (once all earlier MMAs retired) on the barrier at this smem offset in every CTA of `mask`.
__device__ __forceinline__ void umma_commit_2cta(uint64_t* bar, uint16_t mask) {
    asm volatile(
        "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
            smem_u32(bar)),
        "h"(mask)
        : "memory");
}

// Shared-memory matrix descriptor for tcgen05.mma, 128-byte swizzle.
//   bits [0,14)  start address >> 4        bits [16,30) leading byte offset >> 4
//   bits [32,46) stride byte offset >> 4   bits [46,48) version = 1 (Blackwell)
//   bits [61,64) layout type: 2 = SWIZZLE_128B
__device__ __forceinline__ uint64_t make_umma_desc_sw128(uint32_t smem_addr, uint32_t lbo_bytes,
                                                         uint32_t sbo_bytes) {
    uint64_t d = 0;
    d |= static_cast<uint64_t>((smem_addr >> 4) & 0x3FFF);
    d |= static_cast<uint64_t>((lbo_bytes >> 4) & 0x3FFF) << 16;
    d |= static_cast<uint64_t>((sbo_bytes >> 4) & 0x3FFF) << 32;
    d |= static_cast<uint64_t>(1) << 46;
    d |= static_cast<uint64_t>(2) << 61;
    return d;
}

// Instruction descriptor for kind::f16 with bf16 A/B, fp32 D.
//   [4,6) D fmt = 1 (f32)  [7,10) A fmt = 1 (bf16)  [10,13) B fmt = 1 (bf16)
//   [15] A major (0 = K, 1 = MN)  [16] B major  [17,23) N >> 3  [24,29) M >> 4
__host__ __device__ constexpr uint32_t make_idesc_bf16(uint32_t m, uint32_t n, uint32_t a_mn_major,
                                                       uint32_t b_mn_major) {
    return (1u << 4) | (1u << 7) | (1u << 10) | (a_mn_major << 15) | (b_mn_major << 16) |
           ((n >> 3) << 17) | ((m >> 4) << 24);
}

// ----------------------------------------------------------------------------
// System-scope flag protocol used by the fused compute+collective kernels
// ----------------------------------------------------------------------------
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_gpu(uint32_t* p, uint32_t v) {
    asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_gpu(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// Spin until *p >= target (wrap-safe), giving up after `budget_ns` of wall time.
// Returns false on timeout so a dead peer can never hang the GPU.
__device__ __forceinline__ bool wait_flag_sys(const uint32_t* p, uint32_t target,
                                              unsigned long long budget_ns) {
    unsigned long long t0 = 0;
    uint32_t spins = 0;
    while (static_cast<int32_t>(ld_acquire_sys(p) - target) < 0) {
        if ((++spins & 0x3FF) == 0) {
            unsigned long long now;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
            if (t0 == 0) t0 = now;
            if (now - t0 > budget_ns) return false;
        }
    }
    return true;
}

// 16-byte streaming loads/stores that work on local and peer-mapped memory.
__device__ __forceinline__ uint4 ld_global_v4(const void* p) {
    uint4 v;
    asm volatile("ld.global.v4.u32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                 : "l"(p));
    return v;
}
__device__ __forceinline__ uint4 ld_global_nc_v4(const void* p) {
    uint4 v;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                 : "l"(p));
    return v;
}
__device__ __forceinline__ void st_global_v4(void* p, const uint4& v) {
    asm volatile("st.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z),
                 "r"(v.w)
                 : "memory");
}

// ----------------------------------------------------------------------------
// bf16 pack/unpack + activations
// ----------------------------------------------------------------------------
__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
    __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ float2 unpack_bf16x2(uint32_t u) {
    __nv_bfloat162 v = *reinterpret_cast<__nv_bfloat162*>(&u);
    return __bfloat1622float2(v);
}

__device__ __forceinline__ void st_shared_v4(uint32_t addr, const uint4& v) {
    asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z),
                 "r"(v.w)
                 : "memory");
}
__device__ __forceinline__ uint4 ld_shared_v4(uint32_t addr) {
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                 : "r"(addr)
                 : "memory");
    return v;
}

// Normal CDF as an odd polynomial: Phi(x) = 0.5 + x * R(x^2) on |x| <= 4.808 (x clamped
// beyond, where Phi is 0 or 1 to 7e-7).  R is a degree-9 Chebyshev fit of erf rescaled; max
// abs error 1.7e-5, two orders below bf16 resolution.  No MUFU, no sign handling: about 13
// instructions per element, because at K = 1024 the GELU epilogue has to stay cheaper than
// the tile's 8192 MMA cycles (an exp/rcp based erf was SFU-bound, erff() issue-bound).
__device__ __forceinline__ float normal_cdf_poly(float x) {
    const float xc = fminf(fmaxf(x, -4.808f), 4.808f);
    const float u = xc * xc;
    float p = -1.2970899060e-12f;
    p = fmaf(p, u, 1.6458187301e-10f);
    p = fmaf(p, u, -9.2165835733e-09f);
    p = fmaf(p, u, 3.0256993702e-07f);
    p = fmaf(p, u, -6.5517750708e-06f);
    p = fmaf(p, u, 1.0031848644e-04f);
    p = fmaf(p, u, -1.1397672244e-03f);
    p = fmaf(p, u, 9.8951850180e-03f);
    p = fmaf(p, u, -6.6439227713e-02f);
    p = fmaf(p, u, 3.9893668986e-01f);
    return fmaf(xc, p, 0.5f);
}
__device__ __forceinline__ float gelu_erf(float x) { return x * normal_cdf_poly(x); }
__device__ __forceinline__ float dgelu_erf(float x) {
    const float pdf = 0.39894228040143268f * __expf(-0.5f * x * x);
    return fmaf(x, pdf, normal_cdf_poly(x));
}
__device__ __forceinline__ float gelu_tanh(float x) {
    const float u = 0.7978845608028654f * (x + 0.044715f * x * x * x);
    return 0.5f * x * (1.0f + tanhf(u));
}
__device__ __forceinline__ float dgelu_tanh(float x) {
    const float x2 = x * x;
    const float u = 0.7978845608028654f * (x + 0.044715f * x * x2);
    const float t = tanhf(u);
    const float du = 0.7978845608028654f * (1.0f + 3.0f * 0.044715f * x2);
    return 0.5f * (1.0f + t) + 0.5f * x * (1.0f - t * t) * du;
}

// ---- two-level tail of the single-launch column reductions -----------------------------------------------------
// Grid (column tile, slab).  Every CTA hands in NS column partials per thread (thread t: column tile * 256 + t);
// exactly one CTA per column tile returns true holding the totals.  Slabs are folded in groups of ~sqrt(slabs) by
// the CTA of the group that arrives last, the group sums by the last of those: a fixed summation order, and no CTA
// reads more than ~25 rows (a single CTA walking 592 rows took 25 us behind an 8 us streaming pass).
// partials: [slabs][NS][cols]; gparts: [groups][NS][cols]; counters: gridDim.x tile counters, then gridDim.x * groups
// group counters, all zero on entry and zero again on exit (at most 1024 words: the hosts check it).
template <int NS>
__device__ __forceinline__ void fold_rows_256(const float* __restrict__ partials, int nparts, int col0, int cols,
                                              float (*red)[256], float (&s)[NS]) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int c = col0 + lane * 8;
    const size_t stride = static_cast<size_t>(NS) * cols;
    float t[NS][8];
#pragma unroll
    for (int i = 0; i < NS; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) t[i][j] = 0.f;
    if (c < cols) {
        constexpr int Q = NS == 1 ? 4 : (NS == 2 ? 2 : 1);   // rows in flight per warp: >= 8 float4 loads per thread
        for (int k = warp; k < nparts; k += 8 * Q) {
            float4 a[Q][NS][2];
#pragma unroll
            for (int q = 0; q < Q; ++q) {
                const int kk = k + q * 8;
#pragma unroll
                for (int i = 0; i < NS; ++i) {
                    if (kk < nparts) {
                        const float4* p = reinterpret_cast<const float4*>(partials + static_cast<size_t>(kk) * stride +
                                                                          static_cast<size_t>(i) * cols + c);
                        a[q][i][0] = __ldcg(p);
                        a[q][i][1] = __ldcg(p + 1);
                    } else {
                        a[q][i][0] = a[q][i][1] = make_float4(0.f, 0.f, 0.f, 0.f);
                    }
                }
            }
#pragma unroll
            for (int q = 0; q < Q; ++q)
#pragma unroll
                for (int i = 0; i < NS; ++i) {
                    t[i][0] += a[q][i][0].x; t[i][1] += a[q][i][0].y; t[i][2] += a[q][i][0].z; t[i][3] += a[q][i][0].w;
                    t[i][4] += a[q][i][1].x; t[i][5] += a[q][i][1].y; t[i][6] += a[q][i][1].z; t[i][7] += a[q][i][1].w;
                }
        }
    }
#pragma unroll
    for (int i = 0; i < NS; ++i) {
#pragma unroll
        for (int j = 0; j < 8; ++j) red[warp][lane * 8 + j] = t[i][j];
        __syncthreads();
        float v = 0.f;
#pragma unroll
        for (int w = 0; w < 8; ++w) v += red[w][threadIdx.x];
        __syncthreads();
        s[i] = v;
    }
}

// fold_group_size / fold_counter_words / fold_group_floats: host_math.h

// (tile, ntiles, slab, nslabs): which 256-column tile / slab this call hands in; -1 = from blockIdx / gridDim of a
// (column tile, slab) grid.  A kernel whose CTAs each cover several column tiles calls it once per tile.
template <int NS>
__device__ __forceinline__ bool fold_two_level_256(const float (&p)[NS], float* __restrict__ partials,
                                                   float* __restrict__ gparts, unsigned int* __restrict__ counters,
                                                   int cols, float (*red)[256], unsigned int* flag, float (&s)[NS],
                                                   int tile = -1, int ntiles = -1, int slab = -1, int nslabs = -1) {
    if (tile < 0) {
        tile = static_cast<int>(blockIdx.x);
        ntiles = static_cast<int>(gridDim.x);
        slab = static_cast<int>(blockIdx.y);
        nslabs = static_cast<int>(gridDim.y);
    }
    const int nparts = nslabs;
    const int gs = fold_group_size(nparts);
    const int ngroups = (nparts + gs - 1) / gs;
    const int grp = slab / gs;
    const int gsize = min(gs, nparts - grp * gs);
    const int gc = tile * 256 + threadIdx.x;
    unsigned int* gcount = counters + ntiles + tile * ngroups + grp;
    float* my = partials + static_cast<size_t>(slab) * NS * cols;
    if (gc < cols) {
#pragma unroll
        for (int i = 0; i < NS; ++i) my[static_cast<size_t>(i) * cols + gc] = p[i];
    }
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) *flag = (atomicAdd(gcount, 1u) == static_cast<unsigned int>(gsize - 1)) ? 1u : 0u;
    __syncthreads();
    if (!*flag) return false;
    __threadfence();
    fold_rows_256<NS>(partials + static_cast<size_t>(grp) * gs * NS * cols, gsize, tile * 256, cols, red, s);
    if (ngroups == 1) {
        if (threadIdx.x == 0) *gcount = 0;
        return true;
    }
    float* gp = gparts + static_cast<size_t>(grp) * NS * cols;
    if (gc < cols) {
#pragma unroll
        for (int i = 0; i < NS; ++i) gp[static_cast<size_t>(i) * cols + gc] = s[i];
    }
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        *gcount = 0;
        *flag = (atomicAdd(&counters[tile], 1u) == static_cast<unsigned int>(ngroups - 1)) ? 1u : 0u;
    }
    __syncthreads();
    if (!*flag) return false;
    __threadfence();
    fold_rows_256<NS>(gparts, ngroups, tile * 256, cols, red, s);
    if (threadIdx.x == 0) counters[tile] = 0;
    return true;
}

// Tail of the single-launch column reductions: the last CTA of a 256-column tile folds `nparts`
// fp32 partial rows (`stride` floats apart) with ALL of its 8 warps — warp w takes parts w, w+8,
// ..., lane l the columns [8l, 8l+8) with two 16-byte L2 loads per part, two parts in flight —
// and a shared-memory pass adds the 8 warp sums.  (One thread per column walking every part in
// turn was a chain of ~150 dependent L2 round trips: 10+ us for a 5 us reduction.)
// Returns the total for column `col0 + threadIdx.x`; `red` is [8][256] floats, 256 threads.
__device__ __forceinline__ float fold_partials_256(const float* __restrict__ partials, size_t stride,
                                                   int nparts, int col0, int cols, float (*red)[256]) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int c = col0 + lane * 8;
    float t[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    if (c < cols) {
        int k = warp;
        for (; k + 8 < nparts; k += 16) {
            const float4* p0 = reinterpret_cast<const float4*>(partials + static_cast<size_t>(k) * stride + c);
            const float4* p1 = reinterpret_cast<const float4*>(partials + static_cast<size_t>(k + 8) * stride + c);
            const float4 a0 = __ldcg(p0), a1 = __ldcg(p0 + 1), b0 = __ldcg(p1), b1 = __ldcg(p1 + 1);
            t[0] += a0.x + b0.x; t[1] += a0.y + b0.y; t[2] += a0.z + b0.z; t[3] += a0.w + b0.w;
            t[4] += a1.x + b1.x; t[5] += a1.y + b1.y; t[6] += a1.z + b1.z; t[7] += a1.w + b1.w;
        }
        for (; k < nparts; k += 8) {
            const float4* p0 = reinterpret_cast<const float4*>(partials + static_cast<size_t>(k) * stride + c);
            const float4 a0 = __ldcg(p0), a1 = __ldcg(p0 + 1);
            t[0] += a0.x; t[1] += a0.y; t[2] += a0.z; t[3] += a0.w;
            t[4] += a1.x; t[5] += a1.y; t[6] += a1.z; t[7] += a1.w;
        }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) red[warp][lane * 8 + j] = t[j];
    __syncthreads();
    float s = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) s += red[w][threadIdx.x];
    __syncthreads();
    return s;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

}  // namespace tony
