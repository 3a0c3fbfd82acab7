// tony_b200 — symmetric heap on CUDA virtual memory management + NVLS multicast.
//
// The CUDA-IPC path (runtime.cu) gives peer pointers; this path additionally gives every
// buffer a MULTICAST address: a store to it lands in all ranks' copies and a
// multimem.ld_reduce from it returns the sum over all ranks' copies, computed inside the
// NVSwitch (NVLS).  Construction, per buffer and per rank:
//   1. tony_vmm_alloc        cuMemCreate (device-pinned, exportable as a POSIX fd) + export
//   2. [fds are exchanged out of band: SCM_RIGHTS over a unix socket, parallel/symm.py]
//   3. tony_vmm_map          import a peer's fd (or use our own handle), reserve VA, map, RW access
//   4. tony_mc_create (rank 0) / tony_mc_import (others)     the multicast object
//   5. tony_mc_add_device    every rank, then a barrier
//   6. tony_mc_bind          bind our physical allocation at offset 0, then a barrier
//   7. tony_mc_map           map the multicast object -> multicast VA
// All driver entry points are resolved at run time (cudaGetDriverEntryPoint), so the library
// still loads on a machine without libcuda (the CPU build box).
// No counterpart in the reference, which has no data plane (SURVEY.md §5.8 sketches this design).
#include "common.cuh"
#include "tony_b200_kernels.h"

#include <cstdio>
#include <cstring>
#include <mutex>

namespace tony {

struct DriverApi {
    CUresult (*MemCreate)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long);
    CUresult (*MemRelease)(CUmemGenericAllocationHandle);
    CUresult (*MemExport)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long);
    CUresult (*MemImport)(CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType);
    CUresult (*MemAddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long);
    CUresult (*MemAddressFree)(CUdeviceptr, size_t);
    CUresult (*MemMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long);
    CUresult (*MemUnmap)(CUdeviceptr, size_t);
    CUresult (*MemSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t);
    CUresult (*MemGetGranularity)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags);
    CUresult (*McCreate)(CUmemGenericAllocationHandle*, const CUmulticastObjectProp*);
    CUresult (*McAddDevice)(CUmemGenericAllocationHandle, CUdevice);
    CUresult (*McBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t,
                          unsigned long long);
    CUresult (*McGetGranularity)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags);
    CUresult (*DeviceGetAttribute)(int*, CUdevice_attribute, CUdevice);
    CUresult (*CtxGetDevice)(CUdevice*);
    bool ok = false;
};

static DriverApi g_drv;

static bool load_driver_api() {
    static std::once_flag once;
    std::call_once(once, [] {
        cudaFree(nullptr);  // make sure a context is bound
        bool ok = true;
        auto get = [&ok](const char* name, void** out) {
            cudaDriverEntryPointQueryResult q;
            if (cudaGetDriverEntryPoint(name, out, cudaEnableDefault, &q) != cudaSuccess ||
                q != cudaDriverEntryPointSuccess || *out == nullptr)
                ok = false;
        };
        get("cuMemCreate", reinterpret_cast<void**>(&g_drv.MemCreate));
        get("cuMemRelease", reinterpret_cast<void**>(&g_drv.MemRelease));
        get("cuMemExportToShareableHandle", reinterpret_cast<void**>(&g_drv.MemExport));
        get("cuMemImportFromShareableHandle", reinterpret_cast<void**>(&g_drv.MemImport));
        get("cuMemAddressReserve", reinterpret_cast<void**>(&g_drv.MemAddressReserve));
        get("cuMemAddressFree", reinterpret_cast<void**>(&g_drv.MemAddressFree));
        get("cuMemMap", reinterpret_cast<void**>(&g_drv.MemMap));
        get("cuMemUnmap", reinterpret_cast<void**>(&g_drv.MemUnmap));
        get("cuMemSetAccess", reinterpret_cast<void**>(&g_drv.MemSetAccess));
        get("cuMemGetAllocationGranularity", reinterpret_cast<void**>(&g_drv.MemGetGranularity));
        get("cuMulticastCreate", reinterpret_cast<void**>(&g_drv.McCreate));
        get("cuMulticastAddDevice", reinterpret_cast<void**>(&g_drv.McAddDevice));
        get("cuMulticastBindMem", reinterpret_cast<void**>(&g_drv.McBindMem));
        get("cuMulticastGetGranularity", reinterpret_cast<void**>(&g_drv.McGetGranularity));
        get("cuDeviceGetAttribute", reinterpret_cast<void**>(&g_drv.DeviceGetAttribute));
        get("cuCtxGetDevice", reinterpret_cast<void**>(&g_drv.CtxGetDevice));
        g_drv.ok = ok;
    });
    return g_drv.ok;
}

static int drv_fail(const char* what, CUresult r) {
    char buf[64];
    snprintf(buf, sizeof(buf), "CUresult %d", static_cast<int>(r));
    set_last_error(what, buf);
    return TONY_ERR_CUDA;
}

#define TONY_DRV_CHECK(call, name)                      \
    do {                                                \
        CUresult _r = (call);                           \
        if (_r != CUDA_SUCCESS) return drv_fail(name, _r); \
    } while (0)

static int current_device(CUdevice* dev) {
    TONY_DRV_CHECK(g_drv.CtxGetDevice(dev), "cuCtxGetDevice");
    return TONY_OK;
}

static CUmemAllocationProp alloc_prop(CUdevice dev) {
    CUmemAllocationProp p;
    memset(&p, 0, sizeof(p));
    p.type = CU_MEM_ALLOCATION_TYPE_PINNED;
    p.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    p.location.id = dev;
    p.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    return p;
}

static CUmulticastObjectProp mc_prop(int world, size_t bytes) {
    CUmulticastObjectProp p;
    memset(&p, 0, sizeof(p));
    p.numDevices = static_cast<unsigned int>(world);
    p.size = bytes;
    p.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    return p;
}

static int map_handle(CUmemGenericAllocationHandle h, size_t bytes, CUdevice dev, void** ptr) {
    CUdeviceptr va = 0;
    TONY_DRV_CHECK(g_drv.MemAddressReserve(&va, bytes, 0, 0, 0), "cuMemAddressReserve");
    CUresult r = g_drv.MemMap(va, bytes, 0, h, 0);
    if (r != CUDA_SUCCESS) {
        g_drv.MemAddressFree(va, bytes);
        return drv_fail("cuMemMap", r);
    }
    CUmemAccessDesc acc;
    memset(&acc, 0, sizeof(acc));
    acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    acc.location.id = dev;
    acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
    r = g_drv.MemSetAccess(va, bytes, &acc, 1);
    if (r != CUDA_SUCCESS) {
        g_drv.MemUnmap(va, bytes);
        g_drv.MemAddressFree(va, bytes);
        return drv_fail("cuMemSetAccess", r);
    }
    *ptr = reinterpret_cast<void*>(va);
    return TONY_OK;
}

}  // namespace tony

using namespace tony;

extern "C" int tony_nvls_supported(void) {
    if (!load_driver_api()) return 0;
    CUdevice dev;
    if (g_drv.CtxGetDevice(&dev) != CUDA_SUCCESS) return 0;
    int mc = 0, fd = 0;
    g_drv.DeviceGetAttribute(&mc, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev);
    g_drv.DeviceGetAttribute(&fd, CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR_SUPPORTED, dev);
    return (mc && fd) ? 1 : 0;
}

extern "C" int tony_vmm_granularity(int world, size_t bytes, size_t* padded_out) {
    if (!load_driver_api()) return TONY_ERR_UNSUPPORTED;
    if (!padded_out || world < 1) return TONY_ERR_INVALID;
    CUdevice dev;
    int rc = current_device(&dev);
    if (rc) return rc;
    CUmemAllocationProp ap = alloc_prop(dev);
    size_t g1 = 0, g2 = 0;
    TONY_DRV_CHECK(g_drv.MemGetGranularity(&g1, &ap, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED),
                   "cuMemGetAllocationGranularity");
    CUmulticastObjectProp mp = mc_prop(world, bytes);
    if (g_drv.McGetGranularity(&g2, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED) != CUDA_SUCCESS) g2 = g1;
    const size_t g = g1 > g2 ? g1 : g2;
    *padded_out = (bytes + g - 1) / g * g;
    return TONY_OK;
}

extern "C" int tony_vmm_alloc(size_t padded_bytes, unsigned long long* handle_out, int* fd_out) {
    if (!load_driver_api()) return TONY_ERR_UNSUPPORTED;
    if (!handle_out || !fd_out || padded_bytes == 0) return TONY_ERR_INVALID;
    CUdevice dev;
    int rc = current_device(&dev);
    if (rc) return rc;
    CUmemAllocationProp ap = alloc_prop(dev);
    CUmemGenericAllocationHandle h;
    TONY_DRV_CHECK(g_drv.MemCreate(&h, padded_bytes, &ap, 0), "cuMemCreate");
    int fd = -1;
    CUresult r = g_drv.MemExport(&fd, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
    if (r != CUDA_SUCCESS) {
        g_drv.MemRelease(h);
        return drv_fail("cuMemExportToShareableHandle", r);
    }
    *handle_out = static_cast<unsigned long long>(h);
    *fd_out = fd;
    return TONY_OK;
}

extern "C" int tony_vmm_import(int fd, unsigned long long* handle_out) {
    if (!load_driver_api()) return TONY_ERR_UNSUPPORTED;
    if (!handle_out || fd < 0) return TONY_ERR_INVALID;
    CUmemGenericAllocationHandle h;
    TONY_DRV_CHECK(g_drv.MemImport(&h, reinterpret_cast<void*>(static_cast<uintptr_t>(fd)),
                                   CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR),
                   "cuMemImportFromShareableHandle");
    *handle_out = static_cast<unsigned long long>(h);
    return TONY_OK;
}

extern "C" int tony_vmm_map(unsigned long long handle, size_t padded_bytes, void** ptr_out) {
    if (!load_driver_api()) return TONY_ERR_UNSUPPORTED;
    if (!ptr_out) return TONY_ERR_INVALID;
    CUdevice dev;
    int rc = current_device(&dev);
    if (rc) return rc;
    return map_handle(static_cast<CUmemGenericAllocationHandle>(handle), padded_bytes, dev, ptr_out);
}

extern "C" int tony_vmm_unmap(void* ptr, size_t padded_bytes) {
    if (!load_driver_api() || !ptr) return TONY_OK;
    g_drv.MemUnmap(reinterpret_cast<CUdeviceptr>(ptr), padded_bytes);
    g_drv.MemAddressFree(reinterpret_cast<CUdeviceptr>(ptr), padded_bytes);
    return TONY_OK;
}

extern "C" int tony_vmm_release(unsigned long long handle) {
    if (!load_driver_api()) return TONY_OK;
    g_drv.MemRelease(static_cast<CUmemGenericAllocationHandle>(handle));
    return TONY_OK;
}

extern "C" int tony_mc_create(size_t padded_bytes, int world, unsigned long long* handle_out, int* fd_out) {
    if (!load_driver_api()) return TONY_ERR_UNSUPPORTED;
    if (!handle_out || !fd_out) return TONY_ERR_INVALID;
    CUmulticastObjectProp mp = mc_prop(world, padded_bytes);
    CUmemGenericAllocationHandle h;
    TONY_DRV_CHECK(g_drv.McCreate(&h, &mp), "cuMulticastCreate");
    int fd = -1;
    CUresult r = g_drv.MemExport(&fd, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
    if (r != CUDA_SUCCESS) {
        g_drv.MemRelease(h);
        return drv_fail("cuMemExportToShareableHandle(multicast)", r);
    }
    *handle_out = static_cast<unsigned long long>(h);
    *fd_out = fd;
    return TONY_OK;
}

extern "C" int tony_mc_add_device(unsigned long long mc_handle) {
    if (!load_driver_api()) return TONY_ERR_UNSUPPORTED;
    CUdevice dev;
    int rc = current_device(&dev);
    if (rc) return rc;
    TONY_DRV_CHECK(g_drv.McAddDevice(static_cast<CUmemGenericAllocationHandle>(mc_handle), dev),
                   "cuMulticastAddDevice");
    return TONY_OK;
}

extern "C" int tony_mc_bind(unsigned long long mc_handle, unsigned long long mem_handle, size_t padded_bytes) {
    if (!load_driver_api()) return TONY_ERR_UNSUPPORTED;
    TONY_DRV_CHECK(g_drv.McBindMem(static_cast<CUmemGenericAllocationHandle>(mc_handle), 0,
                                   static_cast<CUmemGenericAllocationHandle>(mem_handle), 0, padded_bytes, 0),
                   "cuMulticastBindMem");
    return TONY_OK;
}
