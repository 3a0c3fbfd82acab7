// tony_b200 — library runtime: error reporting, launch accounting, symmetric memory.
//
// Symmetric memory = plain cudaMalloc allocations exported with CUDA IPC so that every
// rank of a job (one process per GPU, as TonY's TaskExecutor launches them —
// reference: tony-core/src/main/java/com/linkedin/tony/TaskExecutor.java:189-237) can map
// every other rank's gradient / parameter / flag buffers and reach them with ordinary
// ld/st over NVLink from inside a kernel.
#include "common.cuh"
#include "tony_b200_kernels.h"

#include <atomic>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <string>

namespace tony {

static std::mutex g_err_mu;
static std::string g_last_error;
std::atomic<unsigned long long> g_launches{0};

void set_last_error(const char* what, const char* detail) {
    std::lock_guard<std::mutex> lock(g_err_mu);
    g_last_error = std::string(what) + ": " + (detail ? detail : "");
}

void count_launch(int n) { g_launches.fetch_add(static_cast<unsigned long long>(n)); }

}  // namespace tony

using namespace tony;

extern "C" const char* tony_last_error(void) {
    static thread_local std::string copy;
    std::lock_guard<std::mutex> lock(g_err_mu);
    copy = g_last_error;
    return copy.c_str();
}

extern "C" int tony_version(void) { return 100; }

extern "C" unsigned long long tony_launch_count(void) { return g_launches.load(); }
extern "C" void tony_launch_count_add(unsigned long long n) { g_launches.fetch_add(n); }

extern "C" int tony_symm_alloc(size_t bytes, void** ptr) {
    if (!ptr || bytes == 0) return TONY_ERR_INVALID;
    void* p = nullptr;
    TONY_CUDA_CHECK(cudaMalloc(&p, bytes));
    TONY_CUDA_CHECK(cudaMemset(p, 0, bytes));
    *ptr = p;
    return TONY_OK;
}

extern "C" int tony_symm_free(void* ptr) {
    if (!ptr) return TONY_OK;
    TONY_CUDA_CHECK(cudaFree(ptr));
    return TONY_OK;
}

extern "C" int tony_ipc_get_handle(void* ptr, unsigned char* handle_out) {
    if (!ptr || !handle_out) return TONY_ERR_INVALID;
    static_assert(sizeof(cudaIpcMemHandle_t) == TONY_IPC_HANDLE_BYTES, "handle size");
    cudaIpcMemHandle_t h;
    TONY_CUDA_CHECK(cudaIpcGetMemHandle(&h, ptr));
    memcpy(handle_out, &h, sizeof(h));
    return TONY_OK;
}

extern "C" int tony_ipc_open(const unsigned char* handle, void** peer_ptr) {
    if (!handle || !peer_ptr) return TONY_ERR_INVALID;
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, sizeof(h));
    void* p = nullptr;
    TONY_CUDA_CHECK(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    *peer_ptr = p;
    return TONY_OK;
}

extern "C" int tony_ipc_close(void* peer_ptr) {
    if (!peer_ptr) return TONY_OK;
    TONY_CUDA_CHECK(cudaIpcCloseMemHandle(peer_ptr));
    return TONY_OK;
}
