// N7 (VMM form) + N6 multicast_put: shareable physical allocations and NVLS multicast.
//
// The CUDA-IPC heap (heap.cu) is enough for peer loads and stores.  In-switch replication needs
// more: a multicast object (cuMulticastCreate) to which every GPU binds physical memory created
// with the virtual-memory-management API (cuMemCreate), exported between processes as POSIX file
// descriptors.  A store through the multicast mapping (multimem.st) is then replicated by
// NVSwitch into every GPU's bound memory: one write on the wire for a sample that every consumer
// needs.  The reference has no counterpart (its windows are per-instance, ddl/connection.py:88-139).
//
// Driver entry points are resolved at run time (cudaGetDriverEntryPoint), so the library loads
// without libcuda; everything here is feature-detected and reports BDDL_ERR_UNSUPPORTED when the
// platform lacks VMM, POSIX-FD handles or multicast.
#include <cuda.h>
#include <cuda_runtime.h>
#include <unistd.h>

#include "../include/ddl_b200.h"
#include "common.h"
#include "launch_count.h"

namespace {

struct Driver {
    bool ready = false;
    CUresult (*MemCreate)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long);
    CUresult (*MemRelease)(CUmemGenericAllocationHandle);
    CUresult (*MemAddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long);
    CUresult (*MemAddressFree)(CUdeviceptr, size_t);
    CUresult (*MemMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long);
    CUresult (*MemUnmap)(CUdeviceptr, size_t);
    CUresult (*MemSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t);
    CUresult (*MemExport)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long);
    CUresult (*MemImport)(CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType);
    CUresult (*MemGranularity)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags);
    CUresult (*McCreate)(CUmemGenericAllocationHandle*, const CUmulticastObjectProp*);
    CUresult (*McAddDevice)(CUmemGenericAllocationHandle, CUdevice);
    CUresult (*McBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t,
                          unsigned long long);
    CUresult (*McGranularity)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags);
    CUresult (*DeviceGetAttribute)(int*, CUdevice_attribute, CUdevice);
    CUresult (*DeviceGet)(CUdevice*, int);
} drv;

int resolve(const char* name, void** fn) {
    cudaDriverEntryPointQueryResult q;
    cudaError_t e = cudaGetDriverEntryPoint(name, fn, cudaEnableDefault, &q);
    if (e != cudaSuccess || q != cudaDriverEntryPointSuccess || !*fn) {
        cudaGetLastError();
        return bddl_fail(BDDL_ERR_UNSUPPORTED, "driver entry point %s unavailable", name);
    }
    return 0;
}

int load_driver() {
    if (drv.ready) return 0;
    int rc = 0;
#define R(field, name)                                   \
    if ((rc = resolve(name, (void**)&drv.field))) return rc;
    R(MemCreate, "cuMemCreate")
    R(MemRelease, "cuMemRelease")
    R(MemAddressReserve, "cuMemAddressReserve")
    R(MemAddressFree, "cuMemAddressFree")
    R(MemMap, "cuMemMap")
    R(MemUnmap, "cuMemUnmap")
    R(MemSetAccess, "cuMemSetAccess")
    R(MemExport, "cuMemExportToShareableHandle")
    R(MemImport, "cuMemImportFromShareableHandle")
    R(MemGranularity, "cuMemGetAllocationGranularity")
    R(McCreate, "cuMulticastCreate")
    R(McAddDevice, "cuMulticastAddDevice")
    R(McBindMem, "cuMulticastBindMem")
    R(McGranularity, "cuMulticastGetGranularity")
    R(DeviceGetAttribute, "cuDeviceGetAttribute")
    R(DeviceGet, "cuDeviceGet")
#undef R
    drv.ready = true;
    return 0;
}

#define DRV(call)                                                                                    \
    do {                                                                                             \
        CUresult r__ = (call);                                                                       \
        if (r__ != CUDA_SUCCESS) return bddl_fail((int)r__, "%s:%d %s -> CUresult %d", __FILE__, __LINE__, #call, (int)r__); \
    } while (0)

CUmemAllocationProp device_prop(int dev) {
    CUmemAllocationProp p = {};
    p.type = CU_MEM_ALLOCATION_TYPE_PINNED;
    p.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    p.location.id = dev;
    p.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    return p;
}

int map_rw(CUmemGenericAllocationHandle h, size_t size, int dev, void** ptr) {
    CUdeviceptr va = 0;
    DRV(drv.MemAddressReserve(&va, size, 0, 0, 0));
    CUresult r = drv.MemMap(va, size, 0, h, 0);
    if (r != CUDA_SUCCESS) {
        drv.MemAddressFree(va, size);
        return bddl_fail((int)r, "cuMemMap failed with CUresult %d", (int)r);
    }
    CUmemAccessDesc acc = {};
    acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    acc.location.id = dev;
    acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
    r = drv.MemSetAccess(va, size, &acc, 1);
    if (r != CUDA_SUCCESS) {
        drv.MemUnmap(va, size);
        drv.MemAddressFree(va, size);
        return bddl_fail((int)r, "cuMemSetAccess failed with CUresult %d", (int)r);
    }
    *ptr = (void*)va;
    return 0;
}

// One thread per 16 bytes: load from local memory, store once through the multicast mapping.
__global__ void __launch_bounds__(256) multicast_put_kernel(void* mc_dst, const uint4* __restrict__ src, int64_t n16) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += stride) {
        const uint4 v = src[i];
        char* p = (char*)mc_dst + i * 16;
        asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(__uint_as_float(v.x)),
                     "f"(__uint_as_float(v.y)), "f"(__uint_as_float(v.z)), "f"(__uint_as_float(v.w))
                     : "memory");
    }
    __threadfence_system();
}

}  // namespace

extern "C" {

int bddl_vmm_query(int32_t dev, int32_t* vmm, int32_t* posix_fd, int32_t* multicast) {
    *vmm = *posix_fd = *multicast = 0;
    if (load_driver()) return 0;  // not an error: the features are simply absent
    CUdevice d;
    if (drv.DeviceGet(&d, dev) != CUDA_SUCCESS) return 0;
    int v = 0;
    if (drv.DeviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_VIRTUAL_MEMORY_MANAGEMENT_SUPPORTED, d) == CUDA_SUCCESS) *vmm = v;
    if (drv.DeviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR_SUPPORTED, d) == CUDA_SUCCESS)
        *posix_fd = v;
    if (drv.DeviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, d) == CUDA_SUCCESS) *multicast = v;
    return 0;
}

// Physical allocation on `dev`, mapped read/write there; *fd is a POSIX handle other processes import.
int bddl_vmm_alloc(size_t bytes, int32_t dev, size_t granularity, void** ptr, uint64_t* handle, int32_t* fd,
                   size_t* rounded) {
    int rc = load_driver();
    if (rc) return rc;
    BDDL_CUDA(cudaSetDevice(dev));
    BDDL_CUDA(cudaFree(0));  // make sure the primary context exists
    CUmemAllocationProp prop = device_prop(dev);
    size_t gran = 0;
    DRV(drv.MemGranularity(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
    if (granularity > gran) gran = granularity;
    const size_t size = (bytes + gran - 1) / gran * gran;
    CUmemGenericAllocationHandle h;
    DRV(drv.MemCreate(&h, size, &prop, 0));
    int f = -1;
    CUresult r = drv.MemExport(&f, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
    if (r != CUDA_SUCCESS) {
        drv.MemRelease(h);
        return bddl_fail((int)r, "cuMemExportToShareableHandle failed with CUresult %d", (int)r);
    }
    rc = map_rw(h, size, dev, ptr);
    if (rc) {
        close(f);
        drv.MemRelease(h);
        return rc;
    }
    *handle = (uint64_t)h;
    *fd = f;
    *rounded = size;
    return 0;
}

int bddl_vmm_import(int32_t fd, size_t size, int32_t dev, void** ptr, uint64_t* handle) {
    int rc = load_driver();
    if (rc) return rc;
    BDDL_CUDA(cudaSetDevice(dev));
    BDDL_CUDA(cudaFree(0));
    CUmemGenericAllocationHandle h;
    DRV(drv.MemImport(&h, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
    rc = map_rw(h, size, dev, ptr);
    if (rc) {
        drv.MemRelease(h);
        return rc;
    }
    *handle = (uint64_t)h;
    return 0;
}

int bddl_vmm_free(void* ptr, size_t size, uint64_t handle) {
    int rc = load_driver();
    if (rc) return rc;
    if (ptr) {
        DRV(drv.MemUnmap((CUdeviceptr)ptr, size));
        DRV(drv.MemAddressFree((CUdeviceptr)ptr, size));
    }
    if (handle) DRV(drv.MemRelease((CUmemGenericAllocationHandle)handle));
    return 0;
}

int bddl_mc_granularity(size_t bytes, int32_t n_devices, size_t* gran) {
    int rc = load_driver();
    if (rc) return rc;
    CUmulticastObjectProp p = {};
    p.numDevices = (unsigned)n_devices;
    p.size = bytes;
    p.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    DRV(drv.McGranularity(gran, &p, CU_MULTICAST_GRANULARITY_RECOMMENDED));
    return 0;
}

// Rank 0: create the multicast object for n_devices GPUs and export it.
int bddl_mc_create(size_t size, int32_t n_devices, uint64_t* mc, int32_t* fd) {
    int rc = load_driver();
    if (rc) return rc;
    BDDL_CUDA(cudaFree(0));
    CUmulticastObjectProp p = {};
    p.numDevices = (unsigned)n_devices;
    p.size = size;
    p.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    CUmemGenericAllocationHandle h;
    DRV(drv.McCreate(&h, &p));
    int f = -1;
    CUresult r = drv.MemExport(&f, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
    if (r != CUDA_SUCCESS) {
        drv.MemRelease(h);
        return bddl_fail((int)r, "exporting the multicast object failed with CUresult %d", (int)r);
    }
    *mc = (uint64_t)h;
    *fd = f;
    return 0;
}

int bddl_mc_import(int32_t fd, uint64_t* mc) {
    int rc = load_driver();
    if (rc) return rc;
    BDDL_CUDA(cudaFree(0));
    CUmemGenericAllocationHandle h;
    DRV(drv.MemImport(&h, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
    *mc = (uint64_t)h;
    return 0;
}

int bddl_mc_add_device(uint64_t mc, int32_t dev) {
    int rc = load_driver();
    if (rc) return rc;
    CUdevice d;
    DRV(drv.DeviceGet(&d, dev));
    DRV(drv.McAddDevice((CUmemGenericAllocationHandle)mc, d));
    return 0;
}

// Bind this GPU's physical memory (a bddl_vmm_alloc handle) at offset 0 of the multicast object
// and map the object itself: stores through *mc_ptr reach every GPU's bound memory.
int bddl_mc_bind_and_map(uint64_t mc, uint64_t mem_handle, size_t size, int32_t dev, void** mc_ptr) {
    int rc = load_driver();
    if (rc) return rc;
    DRV(drv.McBindMem((CUmemGenericAllocationHandle)mc, 0, (CUmemGenericAllocationHandle)mem_handle, 0, size, 0));
    return map_rw((CUmemGenericAllocationHandle)mc, size, dev, mc_ptr);
}

// N6 multicast_put: replicate `bytes` (multiple of 16) from local memory into every GPU's copy.
int bddl_multicast_put(void* mc_dst, const void* d_src, int64_t bytes, void* stream) {
    if (!mc_dst || !d_src || bytes < 0 || (bytes & 15) || ((uintptr_t)mc_dst & 15) || ((uintptr_t)d_src & 15))
        return bddl_fail(BDDL_ERR_INVALID, "multicast_put: 16-byte aligned pointers and size required");
    if (bytes == 0) return 0;
    const int64_t n16 = bytes / 16;
    int64_t blocks = (n16 + 255) / 256;
    if (blocks > 148 * 8) blocks = 148 * 8;
    multicast_put_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(mc_dst, (const uint4*)d_src, n16);
    bddl_count_launch(1);
    BDDL_CUDA(cudaGetLastError());
    return 0;
}

}  // extern "C"
