// Core runtime: error ABI, rank->GPU binding, device/pinned memory, peer mapping.
//
// Capability parity with the reference's scattered set-up code:
//   * rank -> GPU: mpicuda2.cu:198-202 (task % device_count; round robin over
//     nodes under MPI_RROBIN_) and BindDevice(),
//     stencil2d/mpi-2d-stencil-subarray-cuda.cu:40-73 (launcher environment,
//     NUM_GPU_DEVICES).  Defect not reproduced: BindDevice uses local_rank
//     uninitialised when no variable is set -- here the
//     fallback is the global rank.
//   * pinned memory: host_allocator.h:58-93.
//   * device pointers crossing ranks: the reference hands them to a CUDA-aware
//     MPI (stencil2d/stencil2D.h:369-373); here peers map each other's
//     allocations (CUDA IPC between processes, peer access inside one) and
//     kernels store through the mapped addresses.
#include <cuda_runtime.h>

#include <cstdlib>
#include <cstring>
#include <string>

#include "b200comm.h"
#include "b200comm/error.hpp"
#include "core/capi_util.hpp"

using namespace b200;

static int env_int(const char* name, int dflt) {
    const char* v = std::getenv(name);
    if (!v || !*v) return dflt;
    char* end = nullptr;
    long x = std::strtol(v, &end, 10);
    if (end == v) return dflt;
    return (int)x;
}

extern "C" {

B2C_EXPORT const char* b2c_last_error(void) { return last_error(); }
B2C_EXPORT const char* b2c_error_class_string(int code) { return error_class_string(code); }
B2C_EXPORT int b2c_set_error_mode(int mode) {
    set_error_mode(mode == 1 ? ErrorMode::Abort : ErrorMode::Throw);
    return 0;
}
B2C_EXPORT int b2c_version(void) { return 100; }

B2C_EXPORT int b2c_device_count(int* n) {
    B2C_TRY
    B200_REQUIRE(n, ERR_ARG, "null output");
    int c = 0;
    cudaError_t st = cudaGetDeviceCount(&c);
    if (st != cudaSuccess) { cudaGetLastError(); c = 0; }
    int cap = env_int("NUM_GPU_DEVICES", -1);
    if (cap > 0 && cap < c) c = cap;
    *n = c;
    B2C_CATCH
}

B2C_EXPORT int b2c_bind_device(int policy, int rank, int node_count, int* device_out) {
    B2C_TRY
    int ndev = 0;
    int rc = b2c_device_count(&ndev);
    if (rc) return rc;
    B200_REQUIRE(ndev > 0, ERR_CUDA, "no CUDA device visible");
    int dev = 0;
    if (policy == 0) {
        int lr = env_int("LOCAL_RANK", -1);
        if (lr < 0) lr = env_int("OMPI_COMM_WORLD_LOCAL_RANK", -1);
        if (lr < 0) lr = env_int("MV2_COMM_WORLD_LOCAL_RANK", -1);
        if (lr < 0) lr = env_int("SLURM_LOCALID", -1);
        if (lr < 0) lr = rank;
        dev = lr % ndev;
    } else if (policy == 1) {
        dev = rank % ndev;
    } else if (policy == 2) {
        B200_REQUIRE(node_count > 0, ERR_ARG, "node_count must be positive");
        dev = (rank / node_count) % ndev;
    } else {
        B200_FAIL(ERR_ARG, "unknown binding policy");
    }
    B200_CHECK_CUDA(cudaSetDevice(dev));
    B200_CHECK_CUDA(cudaFree(nullptr));  // create the context now, before any transport set-up
    if (device_out) *device_out = dev;
    B2C_CATCH
}

B2C_EXPORT int b2c_set_device(int device) {
    B2C_TRY
    B200_CHECK_CUDA(cudaSetDevice(device));
    B2C_CATCH
}
B2C_EXPORT int b2c_get_device(int* device) {
    B2C_TRY
    B200_CHECK_CUDA(cudaGetDevice(device));
    B2C_CATCH
}
B2C_EXPORT int b2c_device_synchronize(void) {
    B2C_TRY
    B200_CHECK_CUDA(cudaDeviceSynchronize());
    B2C_CATCH
}
B2C_EXPORT int b2c_stream_synchronize(void* stream) {
    B2C_TRY
    B200_CHECK_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    B2C_CATCH
}
B2C_EXPORT int b2c_device_sm_count(int device, int* sms) {
    B2C_TRY
    B200_CHECK_CUDA(cudaDeviceGetAttribute(sms, cudaDevAttrMultiProcessorCount, device));
    B2C_CATCH
}

B2C_EXPORT int b2c_malloc(void** ptr, size_t bytes) {
    B2C_TRY
    B200_REQUIRE(ptr, ERR_ARG, "null output");
    void* p = nullptr;
    B200_CHECK_CUDA(cudaMalloc(&p, bytes ? bytes : 1));
    B200_CHECK_CUDA(cudaMemset(p, 0, bytes));
    *ptr = p;
    B2C_CATCH
}
B2C_EXPORT int b2c_free(void* ptr) {
    B2C_TRY
    if (ptr) B200_CHECK_CUDA(cudaFree(ptr));
    B2C_CATCH
}
B2C_EXPORT int b2c_memset_async(void* ptr, int value, size_t bytes, void* stream) {
    B2C_TRY
    B200_CHECK_CUDA(cudaMemsetAsync(ptr, value, bytes, (cudaStream_t)stream));
    B2C_CATCH
}
B2C_EXPORT int b2c_memcpy_async(void* dst, const void* src, size_t bytes, int kind, void* stream) {
    B2C_TRY
    B200_REQUIRE(kind >= 0 && kind <= 4, ERR_ARG, "bad cudaMemcpyKind");
    B200_CHECK_CUDA(cudaMemcpyAsync(dst, src, bytes, (cudaMemcpyKind)kind, (cudaStream_t)stream));
    B2C_CATCH
}
B2C_EXPORT int b2c_pinned_alloc(void** ptr, size_t bytes) {
    B2C_TRY
    B200_REQUIRE(ptr, ERR_ARG, "null output");
    void* p = nullptr;
    B200_CHECK_CUDA(cudaMallocHost(&p, bytes ? bytes : 1));
    *ptr = p;
    B2C_CATCH
}
B2C_EXPORT int b2c_pinned_free(void* ptr) {
    B2C_TRY
    if (ptr) B200_CHECK_CUDA(cudaFreeHost(ptr));
    B2C_CATCH
}

B2C_EXPORT int b2c_ipc_export(void* ptr, unsigned char handle[B2C_IPC_HANDLE_BYTES]) {
    B2C_TRY
    static_assert(sizeof(cudaIpcMemHandle_t) == B2C_IPC_HANDLE_BYTES, "handle size");
    B200_REQUIRE(ptr && handle, ERR_ARG, "null argument");
    cudaIpcMemHandle_t h;
    B200_CHECK_CUDA(cudaIpcGetMemHandle(&h, ptr));
    std::memcpy(handle, &h, sizeof(h));
    B2C_CATCH
}
B2C_EXPORT int b2c_ipc_open(const unsigned char handle[B2C_IPC_HANDLE_BYTES], void** ptr) {
    B2C_TRY
    B200_REQUIRE(ptr && handle, ERR_ARG, "null argument");
    cudaIpcMemHandle_t h;
    std::memcpy(&h, handle, sizeof(h));
    void* p = nullptr;
    B200_CHECK_CUDA(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    *ptr = p;
    B2C_CATCH
}
B2C_EXPORT int b2c_ipc_close(void* ptr) {
    B2C_TRY
    if (ptr) B200_CHECK_CUDA(cudaIpcCloseMemHandle(ptr));
    B2C_CATCH
}
B2C_EXPORT int b2c_can_access_peer(int device, int peer_device, int* yes) {
    B2C_TRY
    B200_REQUIRE(yes, ERR_ARG, "null output");
    if (device == peer_device) { *yes = 1; return 0; }
    B200_CHECK_CUDA(cudaDeviceCanAccessPeer(yes, device, peer_device));
    B2C_CATCH
}
B2C_EXPORT int b2c_enable_peer_access(int device, int peer_device) {
    B2C_TRY
    if (device == peer_device) return 0;
    int prev = 0;
    B200_CHECK_CUDA(cudaGetDevice(&prev));
    B200_CHECK_CUDA(cudaSetDevice(device));
    cudaError_t st = cudaDeviceEnablePeerAccess(peer_device, 0);
    if (st == cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); st = cudaSuccess; }
    cudaSetDevice(prev);
    if (st != cudaSuccess) fail_cuda((int)st, "cudaDeviceEnablePeerAccess", __FILE__, __LINE__);
    B2C_CATCH
}

}  // extern "C"
