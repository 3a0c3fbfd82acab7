// MUFU.EX2 throughput as a function of resident warps per SM sub-partition (independent ops,
// 16 per loop iteration per thread).  One block per SM.
#include <cuda_runtime.h>
#include <stdio.h>

template <int MIX>
__global__ void probe(float* out, int iters) {
    float x[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) x[i] = -1e-3f * (threadIdx.x + i + 1);
    float acc = 0.f;
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 16; ++i) {
            asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x[i]));
            if (MIX) {   // FFMA + FADD per exponential, as in a softmax inner loop
                x[i] = fmaf(x[i], -0.5f, -0.25f);
                acc += x[i];
            }
        }
    }
    long long t1 = clock64();
    float s = acc;
#pragma unroll
    for (int i = 0; i < 16; ++i) s += x[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) out[0] = (float)(t1 - t0);
}

int main() {
    float* d;
    cudaMalloc(&d, 148 * 1024 * 4);
    const int iters = 2048;
    for (int mix = 0; mix < 2; ++mix)
        for (int warps = 1; warps <= 8; warps *= 2) {
            const int threads = warps * 128;
            if (mix) probe<1><<<148, threads>>>(d, iters); else probe<0><<<148, threads>>>(d, iters);
            cudaDeviceSynchronize();
            float clk;
            cudaMemcpy(&clk, d, 4, cudaMemcpyDeviceToHost);
            printf("mix=%d warps/SMSP=%d  %.2f ex2/clk/SM  (%.1f clk per warp-instr per SMSP)\n", mix, warps,
                   (double)iters * 16 * threads / clk, clk / ((double)iters * 16 * warps));
        }
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
