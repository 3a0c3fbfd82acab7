from cuda import cuda
def ck(r):
    err = r[0]
    if err != cuda.CUresult.CUDA_SUCCESS:
        raise RuntimeError(str(err))
    return r[1:] if len(r) > 2 else (r[1] if len(r) == 2 else None)
ck(cuda.cuInit(0))
n = ck(cuda.cuDeviceGetCount())
print("devices", n)
for d in range(n):
    dev = ck(cuda.cuDeviceGet(d))
    mc = ck(cuda.cuDeviceGetAttribute(cuda.CUdevice_attribute.CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev))
    fab = ck(cuda.cuDeviceGetAttribute(cuda.CUdevice_attribute.CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_FABRIC_SUPPORTED, dev))
    fd = ck(cuda.cuDeviceGetAttribute(cuda.CUdevice_attribute.CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR_SUPPORTED, dev))  # noqa
    print(f"dev {d}: multicast_supported={mc} fabric={fab} posix_fd={fd}")
dev0 = ck(cuda.cuDeviceGet(0))
ctx = ck(cuda.cuDevicePrimaryCtxRetain(dev0))
ck(cuda.cuCtxSetCurrent(ctx))
prop = cuda.CUmulticastObjectProp()
prop.numDevices = n
prop.size = 2 << 20
prop.handleTypes = cuda.CUmemAllocationHandleType.CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR
prop.flags = 0
try:
    gran = ck(cuda.cuMulticastGetGranularity(prop, cuda.CUmulticastGranularity_flags.CU_MULTICAST_GRANULARITY_RECOMMENDED))
    print("granularity", gran)
    prop.size = max(prop.size, gran)
    h = ck(cuda.cuMulticastCreate(prop))
    print("cuMulticastCreate OK", h)
    for d in range(n):
        ck(cuda.cuMulticastAddDevice(h, ck(cuda.cuDeviceGet(d))))
    print("AddDevice OK for all")
    fdh = ck(cuda.cuMemExportToShareableHandle(h, cuda.CUmemAllocationHandleType.CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0))
    print("export fd OK", fdh)
except Exception as e:
    print("multicast FAILED:", e)
