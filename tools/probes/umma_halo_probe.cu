// Probe: can one tcgen05.mma shared-memory descriptor (K-major, 128-byte swizzle) read a SHIFTED WINDOW of a
// halo patch that a single tiled TMA load would leave in shared memory?
//
// Patch: PH x PW pixels of 64 bf16 channels (128 B per pixel), pixel p at byte offset p * 128 from a 1024-aligned
// base, 16-byte chunk c stored at chunk (c ^ (p & 7)) - the layout TMA's SWIZZLE_128B produces.
// A operand of tap (r, s): row m = (y, x) = (m / 8, m % 8) is pixel (y + r) * PW + (x + s); groups of 8 rows are
// PW * 128 = 1280 bytes apart (the descriptor's stride byte offset), and the start address is off the 1024-byte
// swizzle period by (r * PW + s) * 128 bytes.  Variants: base-offset field 0, or ((start >> 7) & 7).
// Prints, per tap and variant, the largest difference from the fp32 reference.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -I../../include -I../../tony_b200/csrc umma_halo_probe.cu -o /tmp/halo_probe
#include "common.cuh"

#include <cstdio>
#include <cstdlib>
#include <vector>

using namespace tony;

constexpr int PH = 18, PW = 10, NPIX = PH * PW, CH = 64, NN = 64;

__global__ void __launch_bounds__(128)
halo_probe_kernel(const __nv_bfloat16* __restrict__ xin, const __nv_bfloat16* __restrict__ bin, float* __restrict__ out,
                  int r, int s, int bo_mode, int sbo_bytes) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t* patch = smem;                                   // NPIX * 128 bytes (rounded up to 1024)
    uint8_t* btile = smem + ((NPIX * 128 + 1023) / 1024) * 1024;   // 64 rows * 128 bytes
    uint64_t* bar = reinterpret_cast<uint64_t*>(btile + NN * 128);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar + 1);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    for (int i = tid; i < NPIX * 8; i += 128) {
        const int p = i >> 3, c = i & 7;
        const uint4 v = *reinterpret_cast<const uint4*>(xin + p * CH + c * 8);
        *reinterpret_cast<uint4*>(patch + p * 128 + ((c ^ (p & 7)) << 4)) = v;
    }
    for (int i = tid; i < NN * 8; i += 128) {
        const int n = i >> 3, c = i & 7;
        const uint4 v = *reinterpret_cast<const uint4*>(bin + n * CH + c * 8);
        *reinterpret_cast<uint4*>(btile + n * 128 + ((c ^ (n & 7)) << 4)) = v;
    }
    if (tid == 0) mbar_init(bar, 1);
    if (warp == 0) tmem_alloc<64>(tmem_slot);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    tcgen05_fence_before();
    __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    if (warp == 0) {
        if (elect_one()) {
            const uint32_t a0 = smem_u32(patch) + (r * PW + s) * 128;
            const uint32_t b0 = smem_u32(btile);
            const uint32_t idesc = make_idesc_bf16(128, NN, 0, 0);
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                uint64_t da = make_umma_desc_sw128(a0 + k * 32, 16, sbo_bytes);
                if (bo_mode == 1) da |= static_cast<uint64_t>((a0 >> 7) & 7u) << 49;
                const uint64_t db = make_umma_desc_sw128(b0 + k * 32, 16, 1024);
                umma_bf16(tmem_base, da, db, idesc, k > 0);
            }
            umma_commit(bar);
        }
        __syncwarp();
    }
    mbar_wait(bar, 0);
    tcgen05_fence_after();
    uint32_t v[32];
    for (int c = 0; c < NN / 32; ++c) {
        tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(warp * 32) << 16) + c * 32, v);
        tmem_ld_wait();
        for (int j = 0; j < 32; ++j) out[(warp * 32 + lane) * NN + c * 32 + j] = __uint_as_float(v[j]);
    }
    tcgen05_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc<64>(tmem_base);
}

static float bf(const __nv_bfloat16& h) { return __bfloat162float(h); }

int main() {
    std::vector<__nv_bfloat16> x(NPIX * CH), b(NN * CH);
    srand(7);
    for (auto& v : x) v = __float2bfloat16((rand() % 17 - 8) / 8.0f);
    for (auto& v : b) v = __float2bfloat16((rand() % 13 - 6) / 8.0f);
    __nv_bfloat16 *dx, *db;
    float* dout;
    cudaMalloc(&dx, x.size() * 2);
    cudaMalloc(&db, b.size() * 2);
    cudaMalloc(&dout, 128 * NN * 4);
    cudaMemcpy(dx, x.data(), x.size() * 2, cudaMemcpyHostToDevice);
    cudaMemcpy(db, b.data(), b.size() * 2, cudaMemcpyHostToDevice);
    const size_t smem = ((NPIX * 128 + 1023) / 1024) * 1024 + NN * 128 + 64;
    cudaFuncSetAttribute(halo_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
    std::vector<float> got(128 * NN);
    int ok_all[2] = {1, 1};
    for (int bo = 0; bo < 2; ++bo)
        for (int r = 0; r < 3; ++r)
            for (int s = 0; s < 3; ++s) {
                cudaMemset(dout, 0, 128 * NN * 4);
                halo_probe_kernel<<<1, 128, smem>>>(dx, db, dout, r, s, bo, PW * 128);
                cudaError_t e = cudaDeviceSynchronize();
                if (e != cudaSuccess) {
                    printf("tap (%d,%d) base_offset mode %d: CUDA error %s\n", r, s, bo, cudaGetErrorString(e));
                    return 1;
                }
                cudaMemcpy(got.data(), dout, got.size() * 4, cudaMemcpyDeviceToHost);
                double worst = 0;
                int bad_rows = 0;
                for (int m = 0; m < 128; ++m) {
                    const int p = (m / 8 + r) * PW + (m % 8 + s);
                    double row_worst = 0;
                    for (int n = 0; n < NN; ++n) {
                        double ref = 0;
                        for (int c = 0; c < CH; ++c) ref += bf(x[p * CH + c]) * bf(b[n * CH + c]);
                        row_worst = std::max(row_worst, std::abs(ref - got[m * NN + n]));
                    }
                    worst = std::max(worst, row_worst);
                    bad_rows += row_worst > 1e-3;
                }
                printf("tap (%d,%d) base_offset mode %d: max |err| %.4g, rows off %d / 128 -> %s\n", r, s, bo, worst,
                       bad_rows, worst < 1e-3 ? "OK" : "MISMATCH");
                if (worst >= 1e-3) ok_all[bo] = 0;
            }
    printf("summary: base_offset=0 %s, base_offset=(addr>>7)&7 %s\n", ok_all[0] ? "ALL OK" : "fails",
           ok_all[1] ? "ALL OK" : "fails");
    return 0;
}
