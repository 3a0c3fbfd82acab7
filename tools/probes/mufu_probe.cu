// Throughput probe: MUFU.EX2 (f32 / f16x2 / bf16x2), F2FP pack, and a polynomial exp2 on the FMA
// pipe.  Prints operations per clock per SM.  nvcc -gencode arch=compute_100a,code=sm_100a
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

template <int MODE>
__global__ void probe(float* out, int iters, float seed) {
    float x[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) x[i] = seed * (threadIdx.x + i + 1) * 1e-3f - 1.0f;
    uint32_t h[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) h[i] = 0xbc00bc00u + threadIdx.x + i;
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (MODE == 0) {
                asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x[i]));
            } else if (MODE == 1) {
                asm volatile("ex2.approx.f16x2 %0, %0;" : "+r"(h[i]));
            } else if (MODE == 2) {
                asm volatile("ex2.approx.ftz.bf16x2 %0, %0;" : "+r"(h[i]));
            } else if (MODE == 3) {
                // polynomial exp2 for x <= 0: round-to-int via magic add, cubic on the fraction
                float t = x[i] + 12582912.f;
                float f = x[i] - (t - 12582912.f);
                float pz = fmaf(f, 0.0555041f, 0.2402265f);
                pz = fmaf(pz, f, 0.6931472f);
                pz = fmaf(pz, f, 1.0f);
                x[i] = __int_as_float(__float_as_int(pz) + (__float_as_int(t) << 23)) - 1.5f;
            } else if (MODE == 4) {
                asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(h[i]) : "f"(x[i]), "f"(x[(i + 1) & 7]));
                x[i] += __uint_as_float(h[i] & 0xff);
            }
        }
    }
    long long t1 = clock64();
    float acc = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) acc += x[i] + __uint_as_float(h[i]);
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc + (float)(t1 - t0);
    if (threadIdx.x == 0 && blockIdx.x == 0) out[0] = (float)(t1 - t0);
}

template <int MODE>
void run(const char* name, int per_thread_ops, float* d) {
    const int iters = 4096, threads = 512, blocks = 148 * 2;
    probe<MODE><<<blocks, threads>>>(d, 16, 1.f);
    cudaDeviceSynchronize();
    cudaEvent_t a, b;
    cudaEventCreate(&a);
    cudaEventCreate(&b);
    cudaEventRecord(a);
    probe<MODE><<<blocks, threads>>>(d, iters, 1.f);
    cudaEventRecord(b);
    cudaDeviceSynchronize();
    float ms, clk;
    cudaEventElapsedTime(&ms, a, b);
    cudaMemcpy(&clk, d, 4, cudaMemcpyDeviceToHost);
    double ops_per_sm = (double)iters * 8 * per_thread_ops * threads * 2;
    printf("%-28s %8.2f elem/clk/SM  (%.0f clk, %.3f ms)\n", name, ops_per_sm / clk, clk, ms);
}

int main() {
    float* d;
    cudaMalloc(&d, 148 * 2 * 512 * 4);
    run<0>("ex2.approx.ftz.f32", 1, d);
    run<1>("ex2.approx.f16x2 (2 elem)", 2, d);
    run<2>("ex2.approx.ftz.bf16x2 (2 elem)", 2, d);
    run<3>("poly exp2 on FMA pipe", 1, d);
    run<4>("cvt.rn.bf16x2.f32 + fadd", 2, d);
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
