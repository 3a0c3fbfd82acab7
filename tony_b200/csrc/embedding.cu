// tony_b200 — small data-movement kernels that sit between the library attention call and
// our GEMMs.
//
// tony_pack3_rows: interleave three [rows, row_elems] bf16 matrices (dq, dk, dv as the
// flash-attention backward returns them) into one packed [rows, 3, row_elems] gradient for
// the fused QKV projection, in ONE launch of 16-byte copies.  Stock autograd does this with
// three zero-filled select_backward tensors plus adds; a plain strided copy_ goes through
// ATen's non-vectorised elementwise kernel (measured 36 us each x 72 per BERT-large step).
#include "common.cuh"
#include "tony_b200_kernels.h"

#include <algorithm>

namespace tony {

__global__ void __launch_bounds__(256)
pack3_rows_kernel(const uint4* __restrict__ a, const uint4* __restrict__ b, const uint4* __restrict__ c,
                  long long lda, long long ldb, long long ldc, uint4* __restrict__ out, int rows,
                  int vec_per_row) {
    const long long total = static_cast<long long>(rows) * 3 * vec_per_row;
    for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        const int v = static_cast<int>(i % vec_per_row);
        const long long rw = i / vec_per_row;
        const int which = static_cast<int>(rw % 3);
        const long long row = rw / 3;
        const uint4* src = which == 0 ? a + row * lda : (which == 1 ? b + row * ldb : c + row * ldc);
        out[i] = ld_global_nc_v4(src + v);
    }
}

}  // namespace tony

using namespace tony;

extern "C" int tony_pack3_rows(const void* a, const void* b, const void* c, long long lda, long long ldb,
                               long long ldc, void* out, int rows, int row_elems, void* stream) {
    if (!a || !b || !c || !out || rows <= 0 || row_elems <= 0) return TONY_ERR_INVALID;
    if ((row_elems % 8) || (lda % 8) || (ldb % 8) || (ldc % 8) ||
        ((reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(b) | reinterpret_cast<uintptr_t>(c) |
          reinterpret_cast<uintptr_t>(out)) & 15)) {
        set_last_error("tony_pack3_rows", "row_elems / strides must be multiples of 8, pointers 16B aligned");
        return TONY_ERR_INVALID;
    }
    const int vec = row_elems / 8;
    const long long total = static_cast<long long>(rows) * 3 * vec;
    const int grid = static_cast<int>(std::min<long long>((total + 255) / 256, device_sm_count() * 16LL));
    pack3_rows_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        static_cast<const uint4*>(a), static_cast<const uint4*>(b), static_cast<const uint4*>(c), lda / 8,
        ldb / 8, ldc / 8, static_cast<uint4*>(out), rows, vec);
    count_launch();
    TONY_CUDA_CHECK(cudaGetLastError());
    return TONY_OK;
}
