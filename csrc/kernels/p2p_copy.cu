// Kernel-initiated P2P copy engine: put + device-side flags, and a
// device-resident ping-pong loop.
//
// Reference capability: the GPU ping-pong benchmarks
// test-benchmark/mpi-pingpong-gpu.cpp:24-83 (blocking MPI_Send/MPI_Recv of a
// device pointer) and mpi-pingpong-gpu-async.cpp:43-106 (Isend/Irecv/Wait,
// optional host staging) -- one host-timed, un-warmed round trip per process
// launch.  Here the sender's KERNEL writes the payload into the receiver's
// peer-mapped buffer and then publishes an epoch with a release store; the
// receiver's stream continues after a one-thread acquire kernel (or, for the
// latency floor, both sides stay inside one kernel for all iterations).
//
// Two data paths, selectable per call:
//   mode 0  LD/ST: 16-byte ld.global.nc + st.global, 4 in flight per thread
//   mode 1  TMA  : cp.async.bulk global->smem, cp.async.bulk smem->peer global,
//                  one elected thread per CTA drives a 4-stage ring
#include <cuda_runtime.h>

#include <atomic>
#include <stdint.h>

#include <algorithm>

#include "b200comm.h"
#include "b200comm/error.hpp"
#include "core/capi_util.hpp"
#include "core/trace.hpp"
#include "kernels/ptx.cuh"

namespace b200 {
namespace p2p {

constexpr int kThreads = 512;
constexpr int kUnroll = 4;
constexpr int kTmaStages = 4;
constexpr int kTmaChunk = 32 * 1024;

struct PutParams {
    char* dst;              // peer-mapped destination
    const char* src;        // local source
    size_t bytes;
    uint64_t* peer_flag;    // receiver's epoch word (nullptr: no signal)
    uint64_t flag_value;    // value to publish
    unsigned int* counter;  // LOCAL scratch: CTA arrival counter (self-resetting)
};

__device__ __forceinline__ void publish_when_last(const PutParams& P) {
    // Every CTA's stores must be visible at the receiver before the flag.
    __shared__ bool last;
    __syncthreads();
    if (threadIdx.x == 0) {
        ptx::fence_acq_rel_sys();
        last = atomicInc(P.counter, gridDim.x - 1) == gridDim.x - 1;
    }
    __syncthreads();
    if (last && threadIdx.x == 0 && P.peer_flag) ptx::st_release_sys(P.peer_flag, P.flag_value);
}

__global__ void __launch_bounds__(kThreads) put_ldst_kernel(const __grid_constant__ PutParams P) {
    const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t nthr = (size_t)gridDim.x * blockDim.x;
    const bool aligned = (((uintptr_t)P.dst | (uintptr_t)P.src) & 15) == 0;
    if (aligned) {
        const size_t nvec = P.bytes / 16;
        const int4* s = reinterpret_cast<const int4*>(P.src);
        int4* d = reinterpret_cast<int4*>(P.dst);
        size_t i = tid;
        for (; i + (kUnroll - 1) * nthr < nvec; i += kUnroll * nthr) {
            int4 v[kUnroll];
#pragma unroll
            for (int u = 0; u < kUnroll; ++u) v[u] = ptx::ld_nc_v4(s + i + u * nthr);
#pragma unroll
            for (int u = 0; u < kUnroll; ++u) ptx::st_v4(d + i + u * nthr, v[u]);
        }
        for (; i < nvec; i += nthr) ptx::st_v4(d + i, ptx::ld_nc_v4(s + i));
        for (size_t b = nvec * 16 + tid; b < P.bytes; b += nthr) P.dst[b] = P.src[b];
    } else {
        for (size_t b = tid; b < P.bytes; b += nthr) P.dst[b] = P.src[b];
    }
    publish_when_last(P);
}

__global__ void __launch_bounds__(128) put_tma_kernel(const __grid_constant__ PutParams P) {
    extern __shared__ __align__(128) unsigned char smem[];
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + kTmaStages * kTmaChunk);
    // requires 16-byte aligned src/dst and bytes % 16 == 0 (checked by the host)
    if (threadIdx.x == 0) {
        for (int s = 0; s < kTmaStages; ++s) ptx::mbar_init(&full[s], 1);
        ptx::fence_mbar_init();
        const size_t nchunks = (P.bytes + kTmaChunk - 1) / kTmaChunk;
        uint32_t phase[kTmaStages] = {0, 0, 0, 0};
        // software pipeline: issue loads kTmaStages ahead of the stores
        size_t issued = 0;
        auto issue = [&](size_t k) {
            const size_t c = (size_t)blockIdx.x + k * gridDim.x;
            const size_t off = c * kTmaChunk;
            const uint32_t n = (uint32_t)min((size_t)kTmaChunk, P.bytes - off);
            const int st = (int)(k % kTmaStages);
            ptx::mbar_arrive_expect_tx(&full[st], n);
            ptx::bulk_load_1d(smem + st * kTmaChunk, P.src + off, n, &full[st]);
        };
        size_t mine = 0;   // number of chunks this CTA owns
        if ((size_t)blockIdx.x < nchunks) mine = (nchunks - blockIdx.x + gridDim.x - 1) / gridDim.x;
        for (; issued < mine && issued < (size_t)kTmaStages; ++issued) issue(issued);
        for (size_t k = 0; k < mine; ++k) {
            const int st = (int)(k % kTmaStages);
            const size_t c = (size_t)blockIdx.x + k * gridDim.x;
            const size_t off = c * kTmaChunk;
            const uint32_t n = (uint32_t)min((size_t)kTmaChunk, P.bytes - off);
            ptx::mbar_wait(&full[st], phase[st]);
            phase[st] ^= 1u;
            ptx::bulk_store_1d(P.dst + off, smem + st * kTmaChunk, n);
            ptx::bulk_commit();
            if (issued < mine) {
                // the stage of chunk `issued` is the one whose store was issued
                // kTmaStages-1 groups ago... wait until its smem has been read
                ptx::bulk_wait_read<0>();
                issue(issued);
                ++issued;
            }
        }
        ptx::bulk_wait<0>();   // all peer writes complete
    }
    publish_when_last(P);
}

struct WaitParams {
    const uint64_t* flag;
    uint64_t value;
    unsigned int* status;
    uint64_t timeout_ns;
};

__global__ void flag_wait1_kernel(const __grid_constant__ WaitParams P) {
    const uint64_t t0 = ptx::globaltimer_ns();
    uint32_t spins = 0;
    while (ptx::ld_acquire_sys(P.flag) < P.value) {
        if ((++spins & 255u) == 0 && ptx::globaltimer_ns() - t0 > P.timeout_ns) {
            if (P.status) atomicOr(P.status, 4u);
            break;
        }
    }
}

__global__ void flag_set_kernel(uint64_t* flag, uint64_t value) {
    ptx::fence_acq_rel_sys();
    ptx::st_release_sys(flag, value);
}

struct PingPongParams {
    int role;                 // 0 initiator, 1 echo
    int iters;
    size_t bytes;
    const char* send_local;   // what I send (initiator: payload; echo: my receive buffer)
    char* recv_peer;          // the peer's receive buffer (peer mapped)
    const uint64_t* my_flag;
    uint64_t* peer_flag;
    uint64_t base_epoch;
    double* out_ns;           // initiator: total device time of the loop in ns
    unsigned int* status;
    uint64_t timeout_ns;
};

// Both GPUs run this kernel at the same time (one CTA each): the whole
// ping-pong loop stays on the devices, so the measured round trip is pure
// NVLink + memory-system latency with no launch or host in the path.
__global__ void __launch_bounds__(1024) pingpong_kernel(const __grid_constant__ PingPongParams P) {
    const size_t nvec = P.bytes / 16;
    const bool aligned = (((uintptr_t)P.send_local | (uintptr_t)P.recv_peer) & 15) == 0;
    uint64_t t0 = 0;
    __shared__ bool dead;
    if (threadIdx.x == 0) { dead = false; t0 = ptx::globaltimer_ns(); }
    __syncthreads();
    for (int it = 0; it < P.iters; ++it) {
        const uint64_t epoch = P.base_epoch + (uint64_t)it + 1;
        if (P.role == 1) {   // echo side waits first
            if (threadIdx.x == 0) {
                uint32_t spins = 0;
                const uint64_t w0 = ptx::globaltimer_ns();
                while (ptx::ld_acquire_sys(P.my_flag) < epoch)
                    if ((++spins & 1023u) == 0 && ptx::globaltimer_ns() - w0 > P.timeout_ns) { dead = true; break; }
            }
            __syncthreads();
            if (dead) break;
        }
        if (aligned) {
            const int4* s = reinterpret_cast<const int4*>(P.send_local);
            int4* d = reinterpret_cast<int4*>(P.recv_peer);
            for (size_t i = threadIdx.x; i < nvec; i += blockDim.x) {
                int4 v = __ldcv(s + i);   // never from a stale L1 line: the peer wrote this
                ptx::st_v4(d + i, v);
            }
            for (size_t b = nvec * 16 + threadIdx.x; b < P.bytes; b += blockDim.x)
                P.recv_peer[b] = *reinterpret_cast<const volatile char*>(P.send_local + b);
        } else {
            for (size_t b = threadIdx.x; b < P.bytes; b += blockDim.x)
                P.recv_peer[b] = *reinterpret_cast<const volatile char*>(P.send_local + b);
        }
        __syncthreads();
        if (threadIdx.x == 0) ptx::st_release_sys(P.peer_flag, epoch);   // release is cumulative over the barrier
        if (P.role == 0) {   // initiator waits for the echo
            if (threadIdx.x == 0) {
                uint32_t spins = 0;
                const uint64_t w0 = ptx::globaltimer_ns();
                while (ptx::ld_acquire_sys(P.my_flag) < epoch)
                    if ((++spins & 1023u) == 0 && ptx::globaltimer_ns() - w0 > P.timeout_ns) { dead = true; break; }
            }
            __syncthreads();
            if (dead) break;
        }
    }
    if (threadIdx.x == 0) {
        if (dead && P.status) atomicOr(P.status, 4u);
        if (P.out_ns) *P.out_ns = (double)(ptx::globaltimer_ns() - t0);
    }
}

}  // namespace p2p
}  // namespace b200

using namespace b200;
using namespace b200::p2p;

extern "C" {

// Copy `bytes` from local `src` to peer-mapped `dst`, then store `flag_value`
// to `peer_flag` with release semantics (skip with peer_flag == NULL).
// `counter` is 4 bytes of zero-initialised LOCAL device memory owned by the
// channel (CTA arrival counter).  mode 0 = LD/ST, 1 = TMA bulk.
B2C_EXPORT int b2c_p2p_put(void* dst, const void* src, size_t bytes, void* peer_flag,
                           unsigned long long flag_value, void* counter, int mode, int max_ctas,
                           void* stream) {
    B2C_TRY
    trace::Scope ts("p2p_put(enqueue)", "device", -1, -3, -1, (int64_t)bytes);
    B200_REQUIRE((dst && src) || bytes == 0, ERR_BUFFER, "null buffer");
    B200_REQUIRE(counter != nullptr, ERR_ARG, "null arrival counter");
    PutParams P{};
    P.dst = static_cast<char*>(dst); P.src = static_cast<const char*>(src); P.bytes = bytes;
    P.peer_flag = static_cast<uint64_t*>(peer_flag); P.flag_value = flag_value;
    P.counter = static_cast<unsigned int*>(counter);
    if (max_ctas <= 0) max_ctas = 64;
    const bool tma_ok = mode == 1 && bytes >= 16 && bytes % 16 == 0 &&
                        (((uintptr_t)dst | (uintptr_t)src) & 15) == 0;
    if (tma_ok) {
        const size_t chunks = (bytes + kTmaChunk - 1) / kTmaChunk;
        const int grid = (int)std::max<size_t>(1, std::min<size_t>((size_t)max_ctas, chunks));
        const size_t smem = (size_t)kTmaStages * kTmaChunk + kTmaStages * sizeof(uint64_t);
        // function attributes are per DEVICE: a process that drives several GPUs must set it on
        // each of them (setting it twice, from racing threads, is harmless)
        static std::atomic<bool> attr_done[64];
        int dev = 0;
        B200_CHECK_CUDA(cudaGetDevice(&dev));
        if (!attr_done[dev & 63]) {
            B200_CHECK_CUDA(cudaFuncSetAttribute(put_tma_kernel,
                                                 cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            attr_done[dev & 63] = true;
        }
        put_tma_kernel<<<grid, 128, smem, (cudaStream_t)stream>>>(P);
    } else {
        const size_t per_cta = (size_t)kThreads * 16 * kUnroll;
        const int grid = (int)std::max<size_t>(1, std::min<size_t>((size_t)max_ctas, (bytes + per_cta - 1) / per_cta));
        put_ldst_kernel<<<grid, kThreads, 0, (cudaStream_t)stream>>>(P);
    }
    B200_CHECK_LAUNCH();
    B2C_CATCH
}

// Stream-ordered wait until *flag >= value (device-side replacement of MPI_Wait / MPI_Recv).
B2C_EXPORT int b2c_flag_wait(const void* flag, unsigned long long value, void* status,
                             unsigned long long timeout_ms, void* stream) {
    B2C_TRY
    B200_REQUIRE(flag, ERR_ARG, "null flag");
    WaitParams P{static_cast<const uint64_t*>(flag), value, static_cast<unsigned int*>(status),
                 (timeout_ms ? timeout_ms : 20000ull) * 1000000ull};
    flag_wait1_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(P);
    B200_CHECK_LAUNCH();
    B2C_CATCH
}

B2C_EXPORT int b2c_flag_set(void* flag, unsigned long long value, void* stream) {
    B2C_TRY
    B200_REQUIRE(flag, ERR_ARG, "null flag");
    flag_set_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(static_cast<uint64_t*>(flag), value);
    B200_CHECK_LAUNCH();
    B2C_CATCH
}

// Device-resident ping-pong: `iters` round trips inside one kernel per side.
B2C_EXPORT int b2c_pingpong_device(int role, int iters, size_t bytes, const void* send_local,
                                   void* recv_peer, const void* my_flag, void* peer_flag,
                                   unsigned long long base_epoch, double* out_ns, void* status,
                                   unsigned long long timeout_ms, void* stream) {
    B2C_TRY
    B200_REQUIRE(role == 0 || role == 1, ERR_ARG, "role must be 0 or 1");
    B200_REQUIRE(send_local && recv_peer && my_flag && peer_flag, ERR_ARG, "null argument");
    PingPongParams P{};
    P.role = role; P.iters = iters; P.bytes = bytes;
    P.send_local = static_cast<const char*>(send_local); P.recv_peer = static_cast<char*>(recv_peer);
    P.my_flag = static_cast<const uint64_t*>(my_flag); P.peer_flag = static_cast<uint64_t*>(peer_flag);
    P.base_epoch = base_epoch; P.out_ns = out_ns; P.status = static_cast<unsigned int*>(status);
    P.timeout_ns = (timeout_ms ? timeout_ms : 20000ull) * 1000000ull;
    const int threads = bytes >= 16 * 1024 ? 1024 : (bytes >= 512 ? 256 : 32);
    pingpong_kernel<<<1, threads, 0, (cudaStream_t)stream>>>(P);
    B200_CHECK_LAUNCH();
    B2C_CATCH
}

}  // extern "C"
