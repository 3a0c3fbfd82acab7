// Probe: clocks per tcgen05.mma (kind::f16, cta_group::1, M = 128, K = 16, operands in shared memory) as a function
// of N, issued back to back by one thread with one commit at the end.  One CTA per SM on every SM (so the numbers
// include whatever the chip does under full load) or a single CTA (argv[1] = 1).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -I../../include -I../../tony_b200/csrc umma_rate_probe.cu
#include "common.cuh"

#include <cstdio>
#include <cstdlib>

using namespace tony;

template <int N>
__global__ void __launch_bounds__(128) rate_kernel(long long* out, int iters, int same_k, int a_sbo, int a_off) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t* a = smem;                    // room for a 18 x 10 pixel patch (23040 B)
    uint8_t* b = smem + 24576;            // N rows x 128 B
    uint64_t* bar = reinterpret_cast<uint64_t*>(b + N * 128);
    uint32_t* slot = reinterpret_cast<uint32_t*>(bar + 1);
    for (int i = threadIdx.x; i < (24576 + N * 128) / 4; i += 128) reinterpret_cast<uint32_t*>(smem)[i] = 0;
    if (threadIdx.x == 0) mbar_init(bar, 1);
    if (threadIdx.x < 32) tmem_alloc<256>(slot);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    tcgen05_fence_before();
    __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem = *slot;
    if (threadIdx.x < 32) {
        const uint32_t idesc = make_idesc_bf16(128, N, 0, 0);
        const uint64_t da = make_umma_desc_sw128(smem_u32(a) + a_off, 16, a_sbo);
        const uint64_t db = make_umma_desc_sw128(smem_u32(b), 16, 1024);
        long long t0 = 0, t1 = 0;
        if (elect_one()) {
            t0 = clock64();
            for (int it = 0; it < iters; ++it) {
#pragma unroll
                for (int k = 0; k < 4; ++k) umma_bf16(tmem, da + (same_k ? 0 : k * 2), db + (same_k ? 0 : k * 2), idesc, 1u);
            }
            umma_commit(bar);
        }
        __syncwarp();
        mbar_wait(bar, 0);
        if (elect_one()) {
            t1 = clock64();
            if (blockIdx.x == 0) out[0] = t1 - t0;
        }
        __syncwarp();
    } else {
        mbar_wait(bar, 0);
    }
    tcgen05_fence_before();
    __syncthreads();
    if (threadIdx.x < 32) tmem_dealloc<256>(tmem);
}

template <int N>
static void run(int grid, int iters, int a_sbo = 1024, int a_off = 0) {
    long long* d;
    cudaMalloc(&d, 8);
    const size_t smem = 24576 + N * 128 + 64;
    cudaFuncSetAttribute(rate_kernel<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
    for (int rep = 0; rep < 2; ++rep) rate_kernel<N><<<grid, 128, smem>>>(d, iters, 0, a_sbo, a_off);
    cudaError_t e = cudaDeviceSynchronize();
    long long clk = 0;
    cudaMemcpy(&clk, d, 8, cudaMemcpyDeviceToHost);
    printf("N=%3d grid=%3d A group stride %4d start +%4d: %.1f clk per MMA (M128 x N%d x K16)  [%s]\n", N, grid, a_sbo, a_off,
           double(clk) / (4.0 * iters), N, cudaGetErrorString(e));
    cudaFree(d);
}

int main(int argc, char** argv) {
    const int single = argc > 1 ? atoi(argv[1]) : 0;
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    const int grid = single ? 1 : sms;
    run<32>(grid, 2000);
    run<64>(grid, 2000);
    run<128>(grid, 2000);
    run<256>(grid, 2000);
    // the halo form's A operand: 8-row groups 1280 bytes apart, start off the 1024-byte swizzle period
    run<64>(grid, 2000, 1280, 0);
    run<64>(grid, 2000, 1280, 128);
    run<64>(grid, 2000, 1280, 11 * 128);
    run<128>(grid, 2000, 1280, 11 * 128);
    run<64>(grid, 2000, 1024, 128);
    return 0;
}
