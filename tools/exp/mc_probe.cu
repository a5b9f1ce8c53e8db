// mc_probe.cu -- does this box support NVLink multicast (NVLS)? Queries the attribute and, with >= 2 GPUs,
// tries cuMulticastCreate / AddDevice / BindMem / Map end to end.
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstring>
#include <vector>
#define CKD(x) do { CUresult r_ = (x); if (r_ != CUDA_SUCCESS) { const char *s; cuGetErrorString(r_, &s); printf("FAIL %s -> %s\n", #x, s); return 1; } } while (0)
int main() {
    CKD(cuInit(0));
    int n = 0; CKD(cuDeviceGetCount(&n));
    printf("devices %d\n", n);
    for (int d = 0; d < n; ++d) {
        CUdevice dev; CKD(cuDeviceGet(&dev, d));
        int mc = -1, fab = -1, vmm = -1;
        cuDeviceGetAttribute(&mc, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev);
        cuDeviceGetAttribute(&fab, CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_FABRIC_SUPPORTED, dev);
        cuDeviceGetAttribute(&vmm, CU_DEVICE_ATTRIBUTE_VIRTUAL_MEMORY_MANAGEMENT_SUPPORTED, dev);
        printf("dev %d multicast_supported=%d fabric_handle=%d vmm=%d\n", d, mc, fab, vmm);
    }
    if (n < 2) return 0;
    int nd = n;
    for (int d = 0; d < nd; ++d) { cudaSetDevice(d); cudaFree(0); }
    CUmulticastObjectProp prop; memset(&prop, 0, sizeof(prop));
    prop.numDevices = nd; prop.handleTypes = CU_MEM_HANDLE_TYPE_NONE; prop.flags = 0;
    size_t gran = 0, want = 64u << 20;
    prop.size = want;
    CKD(cuMulticastGetGranularity(&gran, &prop, CU_MULTICAST_GRANULARITY_RECOMMENDED));
    printf("multicast granularity %zu\n", gran);
    prop.size = ((want + gran - 1) / gran) * gran;
    CUmemGenericAllocationHandle mc;
    CKD(cuMulticastCreate(&mc, &prop));
    for (int d = 0; d < nd; ++d) { CUdevice dev; CKD(cuDeviceGet(&dev, d)); CKD(cuMulticastAddDevice(mc, dev)); }
    std::vector<CUmemGenericAllocationHandle> mem(nd);
    for (int d = 0; d < nd; ++d) {
        CUmemAllocationProp ap; memset(&ap, 0, sizeof(ap));
        ap.type = CU_MEM_ALLOCATION_TYPE_PINNED; ap.location.type = CU_MEM_LOCATION_TYPE_DEVICE; ap.location.id = d;
        size_t g2 = 0; CKD(cuMemGetAllocationGranularity(&g2, &ap, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
        CKD(cuMemCreate(&mem[d], prop.size, &ap, 0));
        CKD(cuMulticastBindMem(mc, 0, mem[d], 0, prop.size, 0));
    }
    cudaSetDevice(0);
    CUdeviceptr va; CKD(cuMemAddressReserve(&va, prop.size, gran, 0, 0));
    CKD(cuMemMap(va, prop.size, 0, mc, 0));
    CUmemAccessDesc ad; ad.location.type = CU_MEM_LOCATION_TYPE_DEVICE; ad.location.id = 0; ad.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
    CKD(cuMemSetAccess(va, prop.size, &ad, 1));
    printf("multicast object created, bound on %d devices and mapped: OK\n", nd);
    return 0;
}
