// Virtual-memory-management allocations and NVLink multicast objects: the building blocks of the
// symmetric heap (cuda_mpi_scratch_b200/parallel/symmetric.py, SymmetricHeap).
//
// Why not cudaMalloc + cudaIpc* (round 1): one IPC mapping per plan per peer, no sub-allocation,
// and no way to bind the memory to a multicast object -- which in-switch reductions
// (multimem.ld_reduce / multimem.st, csrc/collectives/device_coll.cu) need.  Here
//   * every rank creates ONE physical allocation per heap arena (cuMemCreate, exportable as a
//     POSIX file descriptor), the descriptors travel over a Unix socket (SCM_RIGHTS, done by the
//     Python layer), every rank imports and maps all of them (cuMemMap + cuMemSetAccess): a
//     kernel addresses peer_base[r] + offset as before;
//   * one multicast object per arena (cuMulticastCreate on rank 0, imported by the others) to
//     which every rank adds its device and binds its allocation; mapped like memory, a store to
//     that mapping lands in EVERY rank's allocation and a multimem.ld_reduce returns the sum over
//     all of them, both done by the NVSwitch.
// Reference counterpart: none; the reference leaves device pointers to a CUDA-aware MPI
// (stencil2d/stencil2D.h:369-373), collectives to MPI on host scalars (mpicuda4.cu:368).
#include <cuda.h>
#include <cuda_runtime.h>
#include <unistd.h>

#include <cstring>
#include <string>

#include "b200comm.h"
#include "b200comm/error.hpp"
#include "core/capi_util.hpp"

namespace b200 {
namespace vmm {

// The library links the CUDA runtime statically and has no link-time dependency on libcuda: driver
// entry points are looked up at run time.
template <typename Fn>
static Fn driver_fn(const char* name) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    cudaError_t st = cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q);
    if (st != cudaSuccess || q != cudaDriverEntryPointSuccess) {
        cudaGetLastError();
        p = nullptr;
    }
    return reinterpret_cast<Fn>(p);
}

#define B200_DRV(name, fn_type)                                                                  \
    static fn_type name##_fn() {                                                                 \
        static fn_type f = driver_fn<fn_type>(#name);                                            \
        B200_REQUIRE(f != nullptr, ERR_CUDA, std::string(#name) + " is not available from the driver"); \
        return f;                                                                                \
    }

using MemCreateFn = CUresult (*)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long);
using MemReleaseFn = CUresult (*)(CUmemGenericAllocationHandle);
using MemExportFn = CUresult (*)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long);
using MemImportFn = CUresult (*)(CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType);
using MemReserveFn = CUresult (*)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long);
using MemAddrFreeFn = CUresult (*)(CUdeviceptr, size_t);
using MemMapFn = CUresult (*)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long);
using MemUnmapFn = CUresult (*)(CUdeviceptr, size_t);
using MemSetAccessFn = CUresult (*)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t);
using MemGranFn = CUresult (*)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags);
using McCreateFn = CUresult (*)(CUmemGenericAllocationHandle*, const CUmulticastObjectProp*);
using McAddDevFn = CUresult (*)(CUmemGenericAllocationHandle, CUdevice);
using McBindMemFn = CUresult (*)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t, unsigned long long);
using McGranFn = CUresult (*)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags);
using DevGetFn = CUresult (*)(CUdevice*, int);
using DevAttrFn = CUresult (*)(int*, CUdevice_attribute, CUdevice);

B200_DRV(cuMemCreate, MemCreateFn)
B200_DRV(cuMemRelease, MemReleaseFn)
B200_DRV(cuMemExportToShareableHandle, MemExportFn)
B200_DRV(cuMemImportFromShareableHandle, MemImportFn)
B200_DRV(cuMemAddressReserve, MemReserveFn)
B200_DRV(cuMemAddressFree, MemAddrFreeFn)
B200_DRV(cuMemMap, MemMapFn)
B200_DRV(cuMemUnmap, MemUnmapFn)
B200_DRV(cuMemSetAccess, MemSetAccessFn)
B200_DRV(cuMemGetAllocationGranularity, MemGranFn)
B200_DRV(cuMulticastCreate, McCreateFn)
B200_DRV(cuMulticastAddDevice, McAddDevFn)
B200_DRV(cuMulticastBindMem, McBindMemFn)
B200_DRV(cuMulticastGetGranularity, McGranFn)
B200_DRV(cuDeviceGet, DevGetFn)
B200_DRV(cuDeviceGetAttribute, DevAttrFn)

#define B200_CHECK_DRV(call)                                                                      \
    do {                                                                                          \
        CUresult r_ = (call);                                                                     \
        if (r_ != CUDA_SUCCESS)                                                                   \
            B200_FAIL(ERR_CUDA, std::string(#call) + " failed with CUresult " + std::to_string((int)r_)); \
    } while (0)

static CUmemAllocationProp alloc_prop(int device) {
    CUmemAllocationProp prop;
    std::memset(&prop, 0, sizeof(prop));
    prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
    prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    prop.location.id = device;
    prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    return prop;
}

static CUmulticastObjectProp mc_prop(int ndevices, size_t bytes) {
    CUmulticastObjectProp prop;
    std::memset(&prop, 0, sizeof(prop));
    prop.numDevices = (unsigned)ndevices;
    prop.size = bytes;
    prop.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    return prop;
}

}  // namespace vmm
}  // namespace b200

using namespace b200;
using namespace b200::vmm;

extern "C" {

// What `device` supports: *vmm = allocations exportable as file descriptors, *multicast = NVLink
// multicast objects (NVSwitch systems).  Never fails for a valid device; unsupported = 0.
B2C_EXPORT int b2c_vmm_support(int device, int* vmm, int* multicast) {
    B2C_TRY
    B200_REQUIRE(vmm && multicast, ERR_ARG, "null output");
    *vmm = 0;
    *multicast = 0;
    B200_CHECK_CUDA(cudaFree(nullptr));      // make sure the primary context exists
    CUdevice dev;
    B200_CHECK_DRV(cuDeviceGet_fn()(&dev, device));
    int v = 0;
    if (cuDeviceGetAttribute_fn()(&v, CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR_SUPPORTED, dev) == CUDA_SUCCESS)
        *vmm = v;
    v = 0;
    if (cuDeviceGetAttribute_fn()(&v, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev) == CUDA_SUCCESS)
        *multicast = v;
    B2C_CATCH
}

// Allocation granularity on `device`, and the multicast granularity for `ndevices` (0 if
// multicast is not available).  Arena sizes must be multiples of both.
B2C_EXPORT int b2c_vmm_granularity(int device, int ndevices, size_t* alloc_gran, size_t* mc_gran) {
    B2C_TRY
    B200_REQUIRE(alloc_gran && mc_gran, ERR_ARG, "null output");
    CUmemAllocationProp prop = alloc_prop(device);
    B200_CHECK_DRV(cuMemGetAllocationGranularity_fn()(alloc_gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
    *mc_gran = 0;
    int vm = 0, mc = 0;
    if (b2c_vmm_support(device, &vm, &mc) == 0 && mc && ndevices > 1) {
        CUmulticastObjectProp mp = mc_prop(ndevices, *alloc_gran);
        size_t g = 0;
        if (cuMulticastGetGranularity_fn()(&g, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED) == CUDA_SUCCESS) *mc_gran = g;
    }
    B2C_CATCH
}

// One physical allocation of `bytes` (a multiple of the granularity) on `device`; *fd is a file
// descriptor other processes can import (the caller closes it after it has been sent).
B2C_EXPORT int b2c_vmm_create(size_t bytes, int device, unsigned long long* handle, int* fd) {
    B2C_TRY
    B200_REQUIRE(handle && fd && bytes > 0, ERR_ARG, "bad argument");
    CUmemAllocationProp prop = alloc_prop(device);
    CUmemGenericAllocationHandle h;
    B200_CHECK_DRV(cuMemCreate_fn()(&h, bytes, &prop, 0));
    int f = -1;
    CUresult r = cuMemExportToShareableHandle_fn()(&f, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
    if (r != CUDA_SUCCESS) {
        cuMemRelease_fn()(h);
        B200_FAIL(ERR_CUDA, "cuMemExportToShareableHandle failed with CUresult " + std::to_string((int)r));
    }
    *handle = (unsigned long long)h;
    *fd = f;
    B2C_CATCH
}

// Import an allocation (or a multicast object) from a file descriptor received from its owner.
B2C_EXPORT int b2c_vmm_import(int fd, unsigned long long* handle) {
    B2C_TRY
    B200_REQUIRE(handle && fd >= 0, ERR_ARG, "bad argument");
    CUmemGenericAllocationHandle h;
    B200_CHECK_DRV(cuMemImportFromShareableHandle_fn()(&h, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
    *handle = (unsigned long long)h;
    B2C_CATCH
}

// Map `handle` (memory or multicast object) into this process and give `device` read/write access.
B2C_EXPORT int b2c_vmm_map(unsigned long long handle, size_t bytes, size_t align, int device, void** ptr) {
    B2C_TRY
    B200_REQUIRE(ptr && bytes > 0, ERR_ARG, "bad argument");
    CUdeviceptr va = 0;
    B200_CHECK_DRV(cuMemAddressReserve_fn()(&va, bytes, align, 0, 0));
    CUresult r = cuMemMap_fn()(va, bytes, 0, (CUmemGenericAllocationHandle)handle, 0);
    if (r != CUDA_SUCCESS) {
        cuMemAddressFree_fn()(va, bytes);
        B200_FAIL(ERR_CUDA, "cuMemMap failed with CUresult " + std::to_string((int)r));
    }
    CUmemAccessDesc acc;
    std::memset(&acc, 0, sizeof(acc));
    acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    acc.location.id = device;
    acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
    r = cuMemSetAccess_fn()(va, bytes, &acc, 1);
    if (r != CUDA_SUCCESS) {
        cuMemUnmap_fn()(va, bytes);
        cuMemAddressFree_fn()(va, bytes);
        B200_FAIL(ERR_CUDA, "cuMemSetAccess failed with CUresult " + std::to_string((int)r));
    }
    *ptr = reinterpret_cast<void*>(va);
    B2C_CATCH
}

B2C_EXPORT int b2c_vmm_unmap(void* ptr, size_t bytes) {
    B2C_TRY
    if (ptr) {
        B200_CHECK_DRV(cuMemUnmap_fn()((CUdeviceptr)(uintptr_t)ptr, bytes));
        B200_CHECK_DRV(cuMemAddressFree_fn()((CUdeviceptr)(uintptr_t)ptr, bytes));
    }
    B2C_CATCH
}

B2C_EXPORT int b2c_vmm_release(unsigned long long handle) {
    B2C_TRY
    B200_CHECK_DRV(cuMemRelease_fn()((CUmemGenericAllocationHandle)handle));
    B2C_CATCH
}

// Multicast object over `ndevices` GPUs for arenas of `bytes`; created by ONE rank, the others
// import *fd.  Every rank then adds its device, and -- after ALL have done so -- binds its memory.
B2C_EXPORT int b2c_mc_create(int ndevices, size_t bytes, unsigned long long* handle, int* fd) {
    B2C_TRY
    B200_REQUIRE(handle && fd && ndevices > 1 && bytes > 0, ERR_ARG, "bad argument");
    CUmulticastObjectProp prop = mc_prop(ndevices, bytes);
    CUmemGenericAllocationHandle h;
    B200_CHECK_DRV(cuMulticastCreate_fn()(&h, &prop));
    int f = -1;
    CUresult r = cuMemExportToShareableHandle_fn()(&f, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
    if (r != CUDA_SUCCESS) {
        cuMemRelease_fn()(h);
        B200_FAIL(ERR_CUDA, "export of the multicast object failed with CUresult " + std::to_string((int)r));
    }
    *handle = (unsigned long long)h;
    *fd = f;
    B2C_CATCH
}

B2C_EXPORT int b2c_mc_add_device(unsigned long long mc, int device) {
    B2C_TRY
    CUdevice dev;
    B200_CHECK_DRV(cuDeviceGet_fn()(&dev, device));
    B200_CHECK_DRV(cuMulticastAddDevice_fn()((CUmemGenericAllocationHandle)mc, dev));
    B2C_CATCH
}

B2C_EXPORT int b2c_mc_bind(unsigned long long mc, unsigned long long mem, size_t bytes) {
    B2C_TRY
    B200_CHECK_DRV(cuMulticastBindMem_fn()((CUmemGenericAllocationHandle)mc, 0, (CUmemGenericAllocationHandle)mem, 0, bytes, 0));
    B2C_CATCH
}

B2C_EXPORT int b2c_close_fd(int fd) {
    if (fd >= 0) ::close(fd);
    return 0;
}

}  // extern "C"
