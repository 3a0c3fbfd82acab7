// tony_b200 — row-sharded embedding bag over NVLink peer memory (K9 of the kernel worklist).
//
// The recommender's categorical tables are not replicated: row r of every field lives on rank
// r mod world ("ps" ownership), in a symmetric allocation that every rank maps.  Per step:
//   gather  : every rank pulls the rows its batch needs straight from the owners' memory
//             (16-byte loads over NVLink) and writes them into the concatenated activation —
//             the dispatch all-to-all and the lookup are one kernel, no staging buffers;
//   scatter : backward pushes each gradient row into the OWNER's gradient shard with
//             system-scope bf16x2 reductions over NVLink — the combine all-to-all and the
//             sparse accumulation are one kernel;
//   the owner then runs the dense flat optimizer on its 1/world shard only.
// Replaces (per step, per rank) F index_select + cat, F index_add_ into full-size dense gradients,
// a dense all-reduce of every table and a replicated optimizer pass over every table in a stock
// data-parallel recommender; the reference itself (zuston/TonY) only launches such jobs.
#include "common.cuh"
#include "tony_b200_kernels.h"

#include <algorithm>

namespace tony {

struct EmbBagArgs {
    const long long* ids;      // [pairs] = [batch * fields]
    long long pairs;
    int fields;
    long long rows_per_shard;  // rows of ONE field held by each rank
    int units;                 // 16-byte units per row (deep + rest)
    int deep_units;            // leading units that go to the `deep` activation
    int world;
};

__device__ __forceinline__ void red_add_bf16x2_sys(__nv_bfloat16* addr, uint32_t v) {
    asm volatile("red.relaxed.sys.global.add.noftz.bf16x2 [%0], %1;" ::"l"(addr), "r"(v) : "memory");
}

__global__ void __launch_bounds__(256)
embbag_gather_kernel(const tony_embbag_comm comm, const EmbBagArgs a, __nv_bfloat16* __restrict__ deep,
                     __nv_bfloat16* __restrict__ rest) {
    const long long total = a.pairs * a.units;
    const int rest_units = a.units - a.deep_units;
    for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        const long long pair = i / a.units;
        const int u = static_cast<int>(i - pair * a.units);
        const int f = static_cast<int>(pair % a.fields);
        const long long id = a.ids[pair];
        // an id outside [0, rows_per_shard * world) would index another field's rows or memory past the
        // shard on some peer: such a pair reads as a zero row (and scatters nothing in backward)
        uint4 v = make_uint4(0u, 0u, 0u, 0u);
        if (id >= 0 && id < a.rows_per_shard * a.world) {
            const int owner = static_cast<int>(id % a.world);
            const long long lrow = id / a.world;
            const __nv_bfloat16* src = static_cast<const __nv_bfloat16*>(comm.table[owner]) +
                                       ((static_cast<long long>(f) * a.rows_per_shard + lrow) * a.units + u) * 8;
            v = ld_global_v4(src);
        }
        if (u < a.deep_units)
            st_global_v4(deep + (pair * a.deep_units + u) * 8, v);
        else
            st_global_v4(rest + (pair * rest_units + (u - a.deep_units)) * 8, v);
    }
}

__global__ void __launch_bounds__(256)
embbag_scatter_kernel(const tony_embbag_comm comm, const EmbBagArgs a, const __nv_bfloat16* __restrict__ ddeep,
                      const __nv_bfloat16* __restrict__ drest) {
    const long long total = a.pairs * a.units;
    const int rest_units = a.units - a.deep_units;
    for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        const long long pair = i / a.units;
        const int u = static_cast<int>(i - pair * a.units);
        const int f = static_cast<int>(pair % a.fields);
        const long long id = a.ids[pair];
        if (id < 0 || id >= a.rows_per_shard * a.world) continue;   // never reduce into memory we do not own
        const int owner = static_cast<int>(id % a.world);
        const long long lrow = id / a.world;
        const uint4 v = u < a.deep_units
                            ? ld_global_nc_v4(ddeep + (pair * a.deep_units + u) * 8)
                            : ld_global_nc_v4(drest + (pair * rest_units + (u - a.deep_units)) * 8);
        __nv_bfloat16* dst = static_cast<__nv_bfloat16*>(comm.grad[owner]) +
                             ((static_cast<long long>(f) * a.rows_per_shard + lrow) * a.units + u) * 8;
        red_add_bf16x2_sys(dst, v.x);
        red_add_bf16x2_sys(dst + 2, v.y);
        red_add_bf16x2_sys(dst + 4, v.z);
        red_add_bf16x2_sys(dst + 6, v.w);
    }
}

static int check_embbag(const char* what, const tony_embbag_comm* comm, const long long* ids, long long batch,
                        int fields, long long rows_per_shard, int row_elems, int deep_elems) {
    if (!comm || !ids || batch <= 0 || fields <= 0 || rows_per_shard <= 0) return TONY_ERR_INVALID;
    if (comm->world < 1 || comm->world > TONY_MAX_RANKS || row_elems <= 0 || (row_elems % 8) || (deep_elems % 8) ||
        deep_elems < 0 || deep_elems > row_elems) {
        set_last_error(what, "row_elems / deep_elems must be multiples of 8 and world <= TONY_MAX_RANKS");
        return TONY_ERR_INVALID;
    }
    return TONY_OK;
}

}  // namespace tony

using namespace tony;

extern "C" int tony_embbag_gather(const tony_embbag_comm* comm, const long long* ids, long long batch,
                                  int fields, long long rows_per_shard, int row_elems, int deep_elems,
                                  void* deep, void* rest, void* stream) {
    int rc = check_embbag("tony_embbag_gather", comm, ids, batch, fields, rows_per_shard, row_elems, deep_elems);
    if (rc != TONY_OK) return rc;
    EmbBagArgs a{ids, batch * fields, fields, rows_per_shard, row_elems / 8, deep_elems / 8, comm->world};
    const long long total = a.pairs * a.units;
    const int grid = static_cast<int>(std::min<long long>((total + 255) / 256, device_sm_count() * 16LL));
    embbag_gather_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        *comm, a, static_cast<__nv_bfloat16*>(deep), static_cast<__nv_bfloat16*>(rest));
    count_launch();
    TONY_CUDA_CHECK(cudaGetLastError());
    return TONY_OK;
}

extern "C" int tony_embbag_scatter(const tony_embbag_comm* comm, const long long* ids, long long batch,
                                   int fields, long long rows_per_shard, int row_elems, int deep_elems,
                                   const void* ddeep, const void* drest, void* stream) {
    int rc = check_embbag("tony_embbag_scatter", comm, ids, batch, fields, rows_per_shard, row_elems, deep_elems);
    if (rc != TONY_OK) return rc;
    EmbBagArgs a{ids, batch * fields, fields, rows_per_shard, row_elems / 8, deep_elems / 8, comm->world};
    const long long total = a.pairs * a.units;
    const int grid = static_cast<int>(std::min<long long>((total + 255) / 256, device_sm_count() * 16LL));
    embbag_scatter_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        *comm, a, static_cast<const __nv_bfloat16*>(ddeep), static_cast<const __nv_bfloat16*>(drest));
    count_launch();
    TONY_CUDA_CHECK(cudaGetLastError());
    return TONY_OK;
}
