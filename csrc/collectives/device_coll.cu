// Collectives on DEVICE tensors over peer-mapped symmetric memory.
//
// Reference capability: MPI_Reduce / MPI_Allreduce / MPI_Gather / MPI_Bcast /
// (MPI_Scatter) / barrier semantics (mpicuda2.cu:154,293, mpi9.cpp:51-54,
// mpi6.cpp:89-100, mpicuda3.cu:179,316) -- in the reference always on HOST
// scalars after a D2H copy.  Here the operands stay on the GPUs:
//   allreduce / reduce   one-shot  (every rank reads all P staged inputs, fixed
//                                   rank-order sum -> bit-identical everywhere)
//                        two-shot  (rank r reduces chunk r and stores it into all
//                                   peers' output staging; large messages)
//   bcast / gather / allgather / scatter
//                        push kernel (stores into peers' staging) + epoch flags
//   barrier              epoch flags only
// Every kernel signals with st.release.sys / waits with ld.acquire.sys on flags in
// the peers' workspaces; nothing goes through the host.  `phases` splits
// signal and wait for single-GPU emulation of several ranks.
//
// Workspace (symmetric, one per rank):
//   [0,512)      ready[P]     epoch of the staged input of rank p   (written by p)
//   [512,1024)   done[P]      epoch of rank p's reduced chunk       (written by p)
//   [1024,1536)  bar[P]       barrier epochs
//   [1536,2048)  data[P]      epoch of pushed data (bcast/gather/scatter)
//   [2048,2052)  CTA arrival counter (local use)    [2112,2116) status
//   [2560,3072)  consumed[P]  last operation rank p has completely finished (back-pressure)
//   [4096, ...)  IN[2][cap]  then OUT[2][cap]   (parity = epoch & 1)
#include <cuda_runtime.h>
#include <stdint.h>

#include <algorithm>
#include <new>
#include <string>

#include "b200comm.h"
#include "b200comm/error.hpp"
#include "core/capi_util.hpp"
#include "core/trace.hpp"
#include "kernels/ptx.cuh"

namespace b200 {
namespace coll {

constexpr int kMaxRanks = 64;
constexpr size_t kOffReady = 0, kOffDone = 512, kOffBar = 1024, kOffData = 1536, kOffCounter = 2048,
                 kOffStatus = 2112, kOffConsumed = 2560, kOffStaging = 4096;
constexpr int kThreads = 256;

struct Ws {
    unsigned char* mc;                // multicast mapping of the workspace (same offsets), or nullptr
    unsigned char* peer[kMaxRanks];   // every rank's workspace as mapped here
    int rank, nranks;
    size_t cap;                       // bytes of one staging half
    uint64_t timeout_ns;
};

__device__ __forceinline__ uint64_t* flag_at(const Ws& w, int owner, size_t region, int idx) {
    return reinterpret_cast<uint64_t*>(w.peer[owner] + region) + idx;
}
__device__ __forceinline__ unsigned char* in_half(const Ws& w, int owner, uint64_t epoch) {
    return w.peer[owner] + kOffStaging + (epoch & 1) * w.cap;
}
__device__ __forceinline__ unsigned char* out_half(const Ws& w, int owner, uint64_t epoch) {
    return w.peer[owner] + kOffStaging + 2 * w.cap + (epoch & 1) * w.cap;
}

// false: the flag did not arrive within the time-out (status bit 8 is set; b2c_coll_check raises).
__device__ __forceinline__ bool wait_flag(const Ws& w, const uint64_t* f, uint64_t epoch) {
    const uint64_t t0 = ptx::globaltimer_ns();
    uint32_t spins = 0;
    while (ptx::ld_acquire_sys(f) < epoch) {
        if ((++spins & 255u) == 0 && ptx::globaltimer_ns() - t0 > w.timeout_ns) {
            atomicOr(reinterpret_cast<unsigned int*>(w.peer[w.rank] + kOffStatus), 8u);
            return false;
        }
    }
    return true;
}

// Tell every rank (incl. myself) that my slot of `region` reached `epoch`.
__device__ __forceinline__ void signal_all(const Ws& w, size_t region, uint64_t epoch) {
    for (int r = threadIdx.x; r < w.nranks; r += blockDim.x)
        ptx::st_release_sys(flag_at(w, r, region, w.rank), epoch);
}
// Wait until every rank's slot of `region` in MY workspace reached `epoch`.
// The same answer in every thread of the CTA.  On false the callers RETURN: a kernel that lost a
// peer neither writes its output from stale data nor signals anybody (the peers then time out as
// well and every rank's check() reports it).  The handle is not usable after a time-out.
__device__ __forceinline__ bool wait_all(const Ws& w, size_t region, uint64_t epoch) {
    bool ok = true;
    if ((int)threadIdx.x < w.nranks) ok = wait_flag(w, flag_at(w, w.rank, region, threadIdx.x), epoch);
    return __syncthreads_and(ok) != 0;
}

template <typename T> __device__ __forceinline__ T apply(T a, T b, int op) {
    switch (op) {
        case 0: return a + b;
        case 1: return a * b;
        case 2: return a > b ? a : b;
        default: return a < b ? a : b;
    }
}

// 16-byte lanes: all bulk work of the reductions is done on int4 vectors of the STAGING
// buffers (256-byte aligned, padded past n), four independent vectors per thread in flight.
template <typename T> __device__ __forceinline__ int4 apply_vec(int4 a, const int4 b, int op) {
    constexpr int N = 16 / (int)sizeof(T);
    T* pa = reinterpret_cast<T*>(&a);
    const T* pb = reinterpret_cast<const T*>(&b);
#pragma unroll
    for (int k = 0; k < N; ++k) pa[k] = apply(pa[k], pb[k], op);
    return a;
}
constexpr int kVecUnroll = 4;

// acc[k] = op over ranks (fixed rank order: same bits on every rank) of vector v0 + k*stride
template <typename T>
__device__ __forceinline__ void reduce_vecs(const Ws& w, uint64_t epoch, int64_t v0, int64_t stride,
                                            int64_t vend, int op, int4 (&acc)[kVecUnroll]) {
    const int4* base0 = reinterpret_cast<const int4*>(in_half(w, 0, epoch));
#pragma unroll
    for (int k = 0; k < kVecUnroll; ++k)
        if (v0 + k * stride < vend) acc[k] = __ldcv(base0 + v0 + k * stride);
    for (int r = 1; r < w.nranks; ++r) {
        const int4* base = reinterpret_cast<const int4*>(in_half(w, r, epoch));
        int4 t[kVecUnroll];
#pragma unroll
        for (int k = 0; k < kVecUnroll; ++k)
            if (v0 + k * stride < vend) t[k] = __ldcv(base + v0 + k * stride);
#pragma unroll
        for (int k = 0; k < kVecUnroll; ++k)
            if (v0 + k * stride < vend) acc[k] = apply_vec<T>(acc[k], t[k], op);
    }
}

// Copy n elements from (16-byte aligned) staging to a user pointer of any alignment.
template <typename T>
__device__ __forceinline__ void copy_out(T* out, const T* src, int64_t n) {
    constexpr int E = 16 / (int)sizeof(T);
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, nthr = (int64_t)gridDim.x * blockDim.x;
    if ((reinterpret_cast<uintptr_t>(out) & 15) == 0) {
        const int64_t nv = n / E;
        const int4* s4 = reinterpret_cast<const int4*>(src);
        int4* d4 = reinterpret_cast<int4*>(out);
        for (int64_t v = tid; v < nv; v += nthr * kVecUnroll) {
            int4 t[kVecUnroll];
#pragma unroll
            for (int k = 0; k < kVecUnroll; ++k)
                if (v + k * nthr < nv) t[k] = __ldcv(s4 + v + k * nthr);
#pragma unroll
            for (int k = 0; k < kVecUnroll; ++k)
                if (v + k * nthr < nv) d4[v + k * nthr] = t[k];
        }
        for (int64_t i = nv * E + tid; i < n; i += nthr) out[i] = *reinterpret_cast<const volatile T*>(src + i);
    } else {
        for (int64_t i = tid; i < n; i += nthr) out[i] = *reinterpret_cast<const volatile T*>(src + i);
    }
}

// ---- one-shot allreduce / reduce ------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(kThreads) allreduce_oneshot_kernel(const __grid_constant__ Ws w, T* out,
                                                                     int64_t n, int op, int root,
                                                                     uint64_t epoch, int phases,
                                                                     const T* fused_in) {
    // fused_in != nullptr: SMALL message, ONE CTA, everything in this launch -- the CTA stages the
    // operand itself (no cudaMemcpyAsync before) and does the consumed-epoch exchange itself (no
    // tail kernel after): one launch instead of three, ~5 us less for an 8-byte allreduce.
    if (fused_in != nullptr) {
        unsigned char* stage = in_half(w, w.rank, epoch);
        const int64_t nb = n * (int64_t)sizeof(T);
        for (int64_t b = threadIdx.x; b < nb; b += blockDim.x)
            stage[b] = reinterpret_cast<const unsigned char*>(fused_in)[b];
        __syncthreads();
    }
    if (phases & 1) {
        // the input was staged by an earlier operation in this stream (or just above): publish it
        if (blockIdx.x == 0) signal_all(w, kOffReady, epoch);
    }
    if (!(phases & 2)) return;
    if (root >= 0 && w.rank != root) {
        if (fused_in != nullptr) {            // folded tail (see tail_kernel)
            signal_all(w, kOffConsumed, epoch);
            if (epoch > 1) wait_all(w, kOffConsumed, epoch - 1);
        }
        return;
    }
    if (!wait_all(w, kOffReady, epoch)) return;
    constexpr int E = 16 / (int)sizeof(T);
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, nthr = (int64_t)gridDim.x * blockDim.x;
    const bool aligned = (reinterpret_cast<uintptr_t>(out) & 15) == 0;
    const int64_t nv = aligned ? n / E : 0;
    for (int64_t v = tid; v < nv; v += nthr * kVecUnroll) {
        int4 acc[kVecUnroll];
        reduce_vecs<T>(w, epoch, v, nthr, nv, op, acc);
#pragma unroll
        for (int k = 0; k < kVecUnroll; ++k)
            if (v + k * nthr < nv) reinterpret_cast<int4*>(out)[v + k * nthr] = acc[k];
    }
    for (int64_t i = nv * E + tid; i < n; i += nthr) {     // tail, or everything if `out` is unaligned
        T acc = *reinterpret_cast<const volatile T*>(reinterpret_cast<const T*>(in_half(w, 0, epoch)) + i);
        for (int r = 1; r < w.nranks; ++r)
            acc = apply(acc, *reinterpret_cast<const volatile T*>(
                                 reinterpret_cast<const T*>(in_half(w, r, epoch)) + i), op);
        out[i] = acc;
    }
    if (fused_in != nullptr) {                // folded tail: every thread of the only CTA has read
        __syncthreads();
        signal_all(w, kOffConsumed, epoch);
        if (epoch > 1) wait_all(w, kOffConsumed, epoch - 1);
    }
}

// ---- two-shot allreduce: reduce-scatter into everybody's OUT, then copy out -----
// Works on whole 16-byte vectors of the padded staging buffers; rank r owns vectors
// [r*vchunk, (r+1)*vchunk).
template <typename T>
__global__ void __launch_bounds__(kThreads) allreduce_twoshot_kernel(const __grid_constant__ Ws w, T* out,
                                                                     int64_t n, int op, uint64_t epoch,
                                                                     int phases) {
    const int P = w.nranks;
    const int64_t nvec = (n * (int64_t)sizeof(T) + 15) / 16;
    const int64_t vchunk = (nvec + P - 1) / P;
    const int64_t lo = min(nvec, (int64_t)w.rank * vchunk), hi = min(nvec, lo + vchunk);
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, nthr = (int64_t)gridDim.x * blockDim.x;
    __shared__ bool last;
    if (phases & 1) {
        if (blockIdx.x == 0) signal_all(w, kOffReady, epoch);
    }
    if (phases & 2) {
        if (!wait_all(w, kOffReady, epoch)) return;
        for (int64_t v = lo + tid; v < hi; v += nthr * kVecUnroll) {
            int4 acc[kVecUnroll];
            reduce_vecs<T>(w, epoch, v, nthr, hi, op, acc);
            for (int r = 0; r < P; ++r) {
                int4* dst = reinterpret_cast<int4*>(out_half(w, r, epoch));
#pragma unroll
                for (int k = 0; k < kVecUnroll; ++k)
                    if (v + k * nthr < hi) dst[v + k * nthr] = acc[k];
            }
        }
        // my chunk is complete when the LAST CTA has stored it
        __syncthreads();
        if (threadIdx.x == 0) {
            ptx::fence_acq_rel_sys();
            unsigned int* cnt = reinterpret_cast<unsigned int*>(w.peer[w.rank] + kOffCounter);
            last = atomicInc(cnt, gridDim.x - 1) == gridDim.x - 1;
        }
        __syncthreads();
        if (last) signal_all(w, kOffDone, epoch);
    }
    if (phases & 4) {
        if (!wait_all(w, kOffDone, epoch)) return;
        copy_out<T>(out, reinterpret_cast<const T*>(out_half(w, w.rank, epoch)), n);
    }
}

// ---- NVLS: the same two steps with the NVSwitch doing the arithmetic and the fan-out --------
// multimem.ld_reduce on the MULTICAST address of IN[parity] returns the sum over all ranks' staged
// inputs (reduced inside the switch: each element crosses my link once instead of P-1 times), and
// multimem.st to the multicast address of OUT[parity] writes every rank's output staging with one
// store.  Rank r still owns vectors [r*vchunk, (r+1)*vchunk): one owner per element, so all ranks
// receive the same bits.  SUM only (what the switch implements for float / double).
template <typename T> struct Mm;
template <> struct Mm<float> {
    static __device__ __forceinline__ int4 ld_reduce(const void* mc) {
        int4 v;
        asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                     : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(mc) : "memory");
        return v;
    }
};
template <> struct Mm<double> {
    static __device__ __forceinline__ int4 ld_reduce(const void* mc) {
        union { int4 v; double d[2]; } u;
        asm volatile("multimem.ld_reduce.relaxed.sys.global.add.f64 %0, [%1];" : "=d"(u.d[0]) : "l"(mc) : "memory");
        asm volatile("multimem.ld_reduce.relaxed.sys.global.add.f64 %0, [%1];"
                     : "=d"(u.d[1]) : "l"(static_cast<const char*>(mc) + 8) : "memory");
        return u.v;
    }
};
template <> struct Mm<int> {
    static __device__ __forceinline__ int4 ld_reduce(const void* mc) {
        int4 v;
        const char* p = static_cast<const char*>(mc);
        asm volatile("multimem.ld_reduce.relaxed.sys.global.add.u32 %0, [%1];" : "=r"(v.x) : "l"(p) : "memory");
        asm volatile("multimem.ld_reduce.relaxed.sys.global.add.u32 %0, [%1];" : "=r"(v.y) : "l"(p + 4) : "memory");
        asm volatile("multimem.ld_reduce.relaxed.sys.global.add.u32 %0, [%1];" : "=r"(v.z) : "l"(p + 8) : "memory");
        asm volatile("multimem.ld_reduce.relaxed.sys.global.add.u32 %0, [%1];" : "=r"(v.w) : "l"(p + 12) : "memory");
        return v;
    }
};
template <> struct Mm<long long> {
    static __device__ __forceinline__ int4 ld_reduce(const void* mc) {
        union { int4 v; unsigned long long q[2]; } u;
        const char* p = static_cast<const char*>(mc);
        asm volatile("multimem.ld_reduce.relaxed.sys.global.add.u64 %0, [%1];" : "=l"(u.q[0]) : "l"(p) : "memory");
        asm volatile("multimem.ld_reduce.relaxed.sys.global.add.u64 %0, [%1];" : "=l"(u.q[1]) : "l"(p + 8) : "memory");
        return u.v;
    }
};
__device__ __forceinline__ void mm_st_v4(void* mc, const int4& v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc),
                 "f"(__int_as_float(v.x)), "f"(__int_as_float(v.y)), "f"(__int_as_float(v.z)),
                 "f"(__int_as_float(v.w)) : "memory");
}

template <typename T>
__global__ void __launch_bounds__(kThreads) allreduce_nvls_kernel(const __grid_constant__ Ws w, T* out,
                                                                  int64_t n, uint64_t epoch, int phases) {
    const int P = w.nranks;
    const int64_t nvec = (n * (int64_t)sizeof(T) + 15) / 16;
    const int64_t vchunk = (nvec + P - 1) / P;
    const int64_t lo = min(nvec, (int64_t)w.rank * vchunk), hi = min(nvec, lo + vchunk);
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, nthr = (int64_t)gridDim.x * blockDim.x;
    __shared__ bool last;
    if (phases & 1) {
        if (blockIdx.x == 0) signal_all(w, kOffReady, epoch);
    }
    if (phases & 2) {
        if (!wait_all(w, kOffReady, epoch)) return;
        const char* mc_in = reinterpret_cast<const char*>(w.mc + kOffStaging + (epoch & 1) * w.cap);
        char* mc_out = reinterpret_cast<char*>(w.mc + kOffStaging + 2 * w.cap + (epoch & 1) * w.cap);
        for (int64_t v = lo + tid; v < hi; v += nthr * kVecUnroll) {
            int4 acc[kVecUnroll];
#pragma unroll
            for (int k = 0; k < kVecUnroll; ++k)
                if (v + k * nthr < hi) acc[k] = Mm<T>::ld_reduce(mc_in + (v + k * nthr) * 16);
#pragma unroll
            for (int k = 0; k < kVecUnroll; ++k)
                if (v + k * nthr < hi) mm_st_v4(mc_out + (v + k * nthr) * 16, acc[k]);
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            ptx::fence_acq_rel_sys();
            unsigned int* cnt = reinterpret_cast<unsigned int*>(w.peer[w.rank] + kOffCounter);
            last = atomicInc(cnt, gridDim.x - 1) == gridDim.x - 1;
        }
        __syncthreads();
        if (last) signal_all(w, kOffDone, epoch);
    }
    if (phases & 4) {
        if (!wait_all(w, kOffDone, epoch)) return;
        copy_out<T>(out, reinterpret_cast<const T*>(out_half(w, w.rank, epoch)), n);
    }
}

// In place on a SYMMETRIC tensor (allocated from a multicast arena of the heap, same offset on
// every rank): no staging copy in, no copy out.  ready-barrier, my 1/P of the vectors reduced by the
// switch and stored back to every rank by the switch, done-barrier.
template <typename T>
__global__ void __launch_bounds__(kThreads) allreduce_nvls_inplace_kernel(const __grid_constant__ Ws w,
                                                                          unsigned char* mc_data, int64_t n,
                                                                          uint64_t epoch) {
    const int P = w.nranks;
    const int64_t nvec = (n * (int64_t)sizeof(T) + 15) / 16;      // blocks are 4 KiB granular: whole vectors
    const int64_t vchunk = (nvec + P - 1) / P;
    const int64_t lo = min(nvec, (int64_t)w.rank * vchunk), hi = min(nvec, lo + vchunk);
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, nthr = (int64_t)gridDim.x * blockDim.x;
    __shared__ bool last;
    if (blockIdx.x == 0) signal_all(w, kOffReady, epoch);      // my operand was complete at launch (stream order)
    if (!wait_all(w, kOffReady, epoch)) return;
    for (int64_t v = lo + tid; v < hi; v += nthr * kVecUnroll) {
        int4 acc[kVecUnroll];
#pragma unroll
        for (int k = 0; k < kVecUnroll; ++k)
            if (v + k * nthr < hi) acc[k] = Mm<T>::ld_reduce(mc_data + (v + k * nthr) * 16);
#pragma unroll
        for (int k = 0; k < kVecUnroll; ++k)
            if (v + k * nthr < hi) mm_st_v4(mc_data + (v + k * nthr) * 16, acc[k]);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        ptx::fence_acq_rel_sys();
        unsigned int* cnt = reinterpret_cast<unsigned int*>(w.peer[w.rank] + kOffCounter);
        last = atomicInc(cnt, gridDim.x - 1) == gridDim.x - 1;
    }
    __syncthreads();
    if (last) signal_all(w, kOffDone, epoch);
    if (!wait_all(w, kOffDone, epoch)) return;          // everybody's chunk has landed in my copy
    if (gridDim.x == 1) {                               // small tensor: the back-pressure tail rides along
        signal_all(w, kOffConsumed, epoch);
        if (epoch > 1) wait_all(w, kOffConsumed, epoch - 1);
    }
}

// bcast / allgather in place on a SYMMETRIC tensor: this rank's byte range [lo, hi) of the tensor
// (bcast: the whole tensor on the root, nothing elsewhere; allgather: block `rank`) is read from
// the local copy and stored ONCE to the multicast address -- the switch delivers it to every
// rank's copy.  One pass over the data, no staging.  ready-barrier (every rank has reached the
// operation, so its copy may be overwritten), stores, done-barrier (every range has landed here).
__global__ void __launch_bounds__(kThreads) spread_inplace_kernel(const __grid_constant__ Ws w,
                                                                  unsigned char* mc_data,
                                                                  const unsigned char* local_data,
                                                                  size_t lo, size_t hi, uint64_t epoch) {
    __shared__ bool last;
    if (blockIdx.x == 0) signal_all(w, kOffReady, epoch);
    if (!wait_all(w, kOffReady, epoch)) return;
    const int64_t v0 = (int64_t)(lo / 16), v1 = (int64_t)((hi + 15) / 16);
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, nthr = (int64_t)gridDim.x * blockDim.x;
    for (int64_t v = v0 + tid; v < v1; v += nthr * kVecUnroll) {
        int4 acc[kVecUnroll];
#pragma unroll
        for (int k = 0; k < kVecUnroll; ++k)
            if (v + k * nthr < v1) acc[k] = __ldcg(reinterpret_cast<const int4*>(local_data) + v + k * nthr);
#pragma unroll
        for (int k = 0; k < kVecUnroll; ++k)
            if (v + k * nthr < v1) mm_st_v4(mc_data + (v + k * nthr) * 16, acc[k]);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        ptx::fence_acq_rel_sys();
        unsigned int* cnt = reinterpret_cast<unsigned int*>(w.peer[w.rank] + kOffCounter);
        last = atomicInc(cnt, gridDim.x - 1) == gridDim.x - 1;
    }
    __syncthreads();
    if (last) signal_all(w, kOffDone, epoch);
    if (!wait_all(w, kOffDone, epoch)) return;
    if (gridDim.x == 1) {
        signal_all(w, kOffConsumed, epoch);
        if (epoch > 1) wait_all(w, kOffConsumed, epoch - 1);
    }
}

// bcast through the switch: the root stores its buffer ONCE to the multicast address of
// OUT[parity]; every rank (the root included) then copies out of its own staging.
__global__ void __launch_bounds__(kThreads) push_mc_kernel(const __grid_constant__ Ws w,
                                                           const unsigned char* src, size_t bytes,
                                                           uint64_t epoch) {
    __shared__ bool last;
    char* dst = reinterpret_cast<char*>(w.mc + kOffStaging + 2 * w.cap + (epoch & 1) * w.cap);
    const size_t nv = (bytes + 15) / 16;                 // staging is padded: whole vectors
    const bool al = (((uintptr_t)src) & 15) == 0;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < nv; i += (size_t)gridDim.x * blockDim.x) {
        int4 v;
        if (al && (i + 1) * 16 <= bytes) {
            v = ptx::ld_nc_v4(reinterpret_cast<const int4*>(src) + i);
        } else {
            unsigned char tmp[16];
            for (int b = 0; b < 16; ++b) tmp[b] = i * 16 + b < bytes ? src[i * 16 + b] : 0;
            v = *reinterpret_cast<int4*>(tmp);
        }
        mm_st_v4(dst + i * 16, v);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        ptx::fence_acq_rel_sys();
        unsigned int* cnt = reinterpret_cast<unsigned int*>(w.peer[w.rank] + kOffCounter);
        last = atomicInc(cnt, gridDim.x - 1) == gridDim.x - 1;
    }
    __syncthreads();
    if (last)
        for (int r = threadIdx.x; r < w.nranks; r += blockDim.x)
            ptx::st_release_sys(flag_at(w, r, kOffData, w.rank), epoch);
}

// ---- push: copy `bytes` of local `src` into OUT[parity] + dst_off of the ranks in
// `targets` (bit mask over ranks, up to 64), then publish data[my rank] there ------
struct PushArgs {
    const unsigned char* src;
    size_t bytes;
    size_t dst_off[kMaxRanks];   // per target: offset inside its OUT half
    size_t src_off[kMaxRanks];   // per target: offset inside src (scatter)
    unsigned long long targets;
};

__global__ void __launch_bounds__(kThreads) push_kernel(const __grid_constant__ Ws w,
                                                        const __grid_constant__ PushArgs a,
                                                        uint64_t epoch) {
    __shared__ bool last;
    for (int r = 0; r < w.nranks; ++r) {
        if (!(a.targets >> r & 1ull)) continue;
        unsigned char* dst = out_half(w, r, epoch) + a.dst_off[r];
        const unsigned char* src = a.src + a.src_off[r];
        const bool al = ((((uintptr_t)dst | (uintptr_t)src) & 15) == 0);
        const size_t nv = al ? a.bytes / 16 : 0;
        for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < nv;
             i += (size_t)gridDim.x * blockDim.x)
            ptx::st_v4(reinterpret_cast<int4*>(dst) + i, ptx::ld_nc_v4(reinterpret_cast<const int4*>(src) + i));
        for (size_t b = nv * 16 + (size_t)blockIdx.x * blockDim.x + threadIdx.x; b < a.bytes;
             b += (size_t)gridDim.x * blockDim.x)
            dst[b] = src[b];
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        ptx::fence_acq_rel_sys();
        unsigned int* cnt = reinterpret_cast<unsigned int*>(w.peer[w.rank] + kOffCounter);
        last = atomicInc(cnt, gridDim.x - 1) == gridDim.x - 1;
    }
    __syncthreads();
    if (last)
        for (int r = threadIdx.x; r < w.nranks; r += blockDim.x)
            if (a.targets >> r & 1ull) ptx::st_release_sys(flag_at(w, r, kOffData, w.rank), epoch);
}

// Wait for data[source] >= epoch for the ranks in `sources`, then copy `bytes` from
// my OUT half (+ src_off) to `dst`.
__global__ void __launch_bounds__(kThreads) pull_local_kernel(const __grid_constant__ Ws w,
                                                              unsigned char* dst, size_t src_off,
                                                              size_t bytes, unsigned long long sources,
                                                              uint64_t epoch) {
    bool ok = true;
    if ((int)threadIdx.x < w.nranks && (sources >> threadIdx.x & 1ull))
        ok = wait_flag(w, flag_at(w, w.rank, kOffData, threadIdx.x), epoch);
    if (!__syncthreads_and(ok)) return;       // a source never delivered: leave `dst` alone
    const unsigned char* src = out_half(w, w.rank, epoch) + src_off;
    const bool al = ((((uintptr_t)dst | (uintptr_t)src) & 15) == 0);
    const size_t nv = al ? bytes / 16 : 0;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < nv;
         i += (size_t)gridDim.x * blockDim.x)
        reinterpret_cast<int4*>(dst)[i] = __ldcv(reinterpret_cast<const int4*>(src) + i);
    for (size_t b = nv * 16 + (size_t)blockIdx.x * blockDim.x + threadIdx.x; b < bytes;
         b += (size_t)gridDim.x * blockDim.x)
        dst[b] = *reinterpret_cast<const volatile unsigned char*>(src + b);
}

// Back-pressure, appended to EVERY reduction / data-movement operation (one tiny kernel):
// publish that this rank has completely finished operation `epoch` -- it will not read
// IN[epoch&1] of any rank or its own OUT[epoch&1] again -- then wait until every rank has
// finished operation epoch-1.  The staging halves alternate with the operation number (ONE
// counter for reductions and data movement alike), so when this rank starts operation epoch+1
// and overwrites the halves last used by epoch-1, nobody is still reading them.  Without it a
// rank that only SENDS in rooted operations (reduce, gather, bcast root, scatter root) could run
// two operations ahead of a slow root and overwrite data the root had not read yet.  Waiting for
// epoch-1 (not epoch) keeps the wait off the critical path and makes the kernel safe for ranks
// emulated on one GPU (every rank's tail of epoch-1 was enqueued before any tail of epoch).
__global__ void tail_kernel(const __grid_constant__ Ws w, uint64_t epoch) {
    signal_all(w, kOffConsumed, epoch);
    if (epoch > 1) wait_all(w, kOffConsumed, epoch - 1);
}

__global__ void barrier_kernel(const __grid_constant__ Ws w, uint64_t epoch, int phases) {
    if (phases & 1) signal_all(w, kOffBar, epoch);
    if (phases & 2) wait_all(w, kOffBar, epoch);
}

struct Plan {
    Ws ws;
    int device, sms;
    uint64_t epoch;        // operation counter: reductions AND push-type collectives (the staging
                           // halves alternate with it, so the two families must share it)
    uint64_t bar_epoch;
    bool use_mc_bcast;     // bcast through multimem.st when the workspace has a multicast mapping
};

static size_t elem_size(int dtype) { return dtype == 8 || dtype == 5 ? 4 : 8; }

template <typename T>
static void launch_allreduce(Plan& p, void* out, int64_t n, int op, int root, int algo, int phases,
                             cudaStream_t st, const void* fused_in) {
    // two-shot CTAs wait (phase 4) for a flag that needs EVERY CTA of this grid to have
    // finished phase 2: the grid must be co-resident, so stay well below capacity
    const int per_sm = (algo >= 1 && root < 0) ? 2 : 4;
    const int64_t nvec = (n * (int64_t)sizeof(T) + 15) / 16;
    const int grid = (int)std::max<int64_t>(1, std::min<int64_t>((nvec + kThreads - 1) / kThreads, p.sms * per_sm));
    if (algo == 2 && root < 0)
        allreduce_nvls_kernel<T><<<grid, kThreads, 0, st>>>(p.ws, static_cast<T*>(out), n, p.epoch, phases);
    else if (algo == 1 && root < 0)
        allreduce_twoshot_kernel<T><<<grid, kThreads, 0, st>>>(p.ws, static_cast<T*>(out), n, op, p.epoch, phases);
    else
        allreduce_oneshot_kernel<T><<<fused_in ? 1 : grid, kThreads, 0, st>>>(
            p.ws, static_cast<T*>(out), n, op, root, p.epoch, phases, static_cast<const T*>(fused_in));
    B200_CHECK_LAUNCH();
}

}  // namespace coll
}  // namespace b200

using namespace b200;
using namespace b200::coll;

struct b2c_coll { Plan p; };

extern "C" {

B2C_EXPORT int b2c_coll_workspace_bytes(size_t max_bytes, size_t* bytes) {
    B2C_TRY
    B200_REQUIRE(bytes, ERR_ARG, "null output");
    const size_t cap = (max_bytes + 255) & ~(size_t)255;
    *bytes = kOffStaging + 4 * cap;
    B2C_CATCH
}

// `workspaces[r]` = rank r's workspace as mapped in this process; all zero-initialised.
B2C_EXPORT int b2c_coll_create(b2c_coll** out, int rank, int nranks, void* const* workspaces,
                               size_t max_bytes) {
    B2C_TRY
    B200_REQUIRE(out && workspaces, ERR_ARG, "null argument");
    B200_REQUIRE(nranks >= 1 && nranks <= kMaxRanks && rank >= 0 && rank < nranks, ERR_RANK, "bad rank");
    auto* h = new (std::nothrow) b2c_coll();
    B200_REQUIRE(h, ERR_NOMEM, "allocation failed");
    Plan& p = h->p;
    p = Plan{};
    for (int r = 0; r < nranks; ++r) p.ws.peer[r] = static_cast<unsigned char*>(workspaces[r]);
    p.ws.rank = rank; p.ws.nranks = nranks;
    p.ws.cap = (max_bytes + 255) & ~(size_t)255;
    p.ws.timeout_ns = 20ull * 1000 * 1000 * 1000;
    B200_CHECK_CUDA(cudaGetDevice(&p.device));
    B200_CHECK_CUDA(cudaDeviceGetAttribute(&p.sms, cudaDevAttrMultiProcessorCount, p.device));
    *out = h;
    B2C_CATCH
}
B2C_EXPORT int b2c_coll_destroy(b2c_coll* h) { B2C_TRY delete h; B2C_CATCH }

// `mc_base`: the multicast mapping of the workspace (SymmetricHeap block with multicast=True), or
// NULL.  With it, algo 2 of b2c_coll_allreduce reduces inside the NVSwitch and bcast uses one
// multicast store per vector (`mc_bcast` != 0).
B2C_EXPORT int b2c_coll_set_multicast(b2c_coll* h, void* mc_base, int mc_bcast) {
    B2C_TRY
    B200_REQUIRE(h, ERR_ARG, "null handle");
    h->p.ws.mc = static_cast<unsigned char*>(mc_base);
    h->p.use_mc_bcast = mc_base != nullptr && mc_bcast != 0;
    B2C_CATCH
}

// dtype: 5 int32, 6 int64, 8 float32, 9 float64.  op: 0 SUM 1 PROD 2 MAX 3 MIN.
// algo: 0 one-shot, 1 two-shot.  root < 0: allreduce.  `in` and `out` are LOCAL
// device pointers (they may alias).  phases: 7 = all in one call; for several
// emulated ranks on one GPU call with 1 (stage+publish) on every rank, then 2,
// then 4 (one-shot ignores 4).
B2C_EXPORT int b2c_coll_allreduce(b2c_coll* h, const void* in, void* out, int64_t count, int dtype,
                                  int op, int root, int algo, int phases, void* stream) {
    B2C_TRY
    trace::Scope ts("coll_allreduce(enqueue)", "device", -1, -3, -1, count);
    B200_REQUIRE(h && in && out && count >= 0, ERR_ARG, "bad argument");
    B200_REQUIRE(dtype == 5 || dtype == 6 || dtype == 8 || dtype == 9, ERR_TYPE,
                 "device reductions support int32, int64, float32, float64");
    B200_REQUIRE(op >= 0 && op <= 3, ERR_ARG, "device reductions support SUM, PROD, MAX, MIN");
    Plan& p = h->p;
    const size_t bytes = (size_t)count * elem_size(dtype);
    B200_REQUIRE(bytes <= p.ws.cap, ERR_COUNT,
                 "message of " + std::to_string(bytes) + " B exceeds the collective workspace (" +
                     std::to_string(p.ws.cap) + " B)");
    if (algo == 2)
        B200_REQUIRE(p.ws.mc != nullptr && op == 0 && root < 0, ERR_ARG,
                     "the in-switch (NVLS) allreduce needs a multicast-bound workspace and op SUM");
    cudaStream_t st = (cudaStream_t)stream;
    // small one-shot operations issued in one piece: a single launch does staging, reduction
    // and the consumed-epoch exchange
    const bool one_shot = !(algo >= 1 && root < 0);
    const void* fused_in = (one_shot && (phases & 3) == 3 && bytes <= 4096) ? in : nullptr;
    if (phases & 1) {
        p.epoch += 1;
        if (!fused_in) {
            unsigned char* stage = p.ws.peer[p.ws.rank] + kOffStaging + (p.epoch & 1) * p.ws.cap;
            B200_CHECK_CUDA(cudaMemcpyAsync(stage, in, bytes, cudaMemcpyDeviceToDevice, st));
        }
    }
    switch (dtype) {
        case 9: launch_allreduce<double>(p, out, count, op, root, algo, phases, st, fused_in); break;
        case 8: launch_allreduce<float>(p, out, count, op, root, algo, phases, st, fused_in); break;
        case 6: launch_allreduce<long long>(p, out, count, op, root, algo, phases, st, fused_in); break;
        default: launch_allreduce<int>(p, out, count, op, root, algo, phases, st, fused_in); break;
    }
    // last phase of this operation on this rank (two-shot: 4, one-shot / rooted: 2)
    const int final_phase = (algo >= 1 && root < 0) ? 4 : 2;
    if ((phases & final_phase) && !fused_in) {
        tail_kernel<<<1, 64, 0, st>>>(p.ws, p.epoch);
        B200_CHECK_LAUNCH();
    }
    B2C_CATCH
}

// In-place SUM allreduce of a symmetric tensor: `mc_data` is the MULTICAST address of the tensor
// (SymmetricBuffer.mc_ptr + offset; the tensor sits at the same offset of a multicast arena on
// every rank).  No staging: the switch reduces and redistributes.  The whole grid waits on
// flags, so it is sized to be co-resident.
B2C_EXPORT int b2c_coll_allreduce_symmetric(b2c_coll* h, void* mc_data, int64_t count, int dtype,
                                            void* stream) {
    B2C_TRY
    trace::Scope ts("coll_allreduce_symmetric(enqueue)", "device", -1, -3, -1, count);
    B200_REQUIRE(h && mc_data && count >= 0, ERR_ARG, "bad argument");
    B200_REQUIRE(dtype == 5 || dtype == 6 || dtype == 8 || dtype == 9, ERR_TYPE,
                 "device reductions support int32, int64, float32, float64");
    B200_REQUIRE(((uintptr_t)mc_data & 15) == 0, ERR_BUFFER, "symmetric tensors must be 16-byte aligned");
    Plan& p = h->p;
    cudaStream_t st = (cudaStream_t)stream;
    p.epoch += 1;
    const int64_t nvec = (count * (int64_t)elem_size(dtype) + 15) / 16;
    const int grid = (int)std::max<int64_t>(1, std::min<int64_t>((nvec / p.ws.nranks + kThreads - 1) / kThreads, p.sms * 2));
    unsigned char* mc = static_cast<unsigned char*>(mc_data);
    switch (dtype) {
        case 9: allreduce_nvls_inplace_kernel<double><<<grid, kThreads, 0, st>>>(p.ws, mc, count, p.epoch); break;
        case 8: allreduce_nvls_inplace_kernel<float><<<grid, kThreads, 0, st>>>(p.ws, mc, count, p.epoch); break;
        case 6: allreduce_nvls_inplace_kernel<long long><<<grid, kThreads, 0, st>>>(p.ws, mc, count, p.epoch); break;
        default: allreduce_nvls_inplace_kernel<int><<<grid, kThreads, 0, st>>>(p.ws, mc, count, p.epoch); break;
    }
    B200_CHECK_LAUNCH();
    if (grid > 1) {             // (a single CTA does the consumed-epoch exchange itself)
        tail_kernel<<<1, 64, 0, st>>>(p.ws, p.epoch);
        B200_CHECK_LAUNCH();
    }
    B2C_CATCH
}

// In place on a symmetric tensor (multicast address `mc_data`, this rank's own mapping
// `local_data`): store my byte range [lo, hi) into every rank's copy through the switch.
// bcast: lo = 0, hi = bytes on the root and lo = hi elsewhere; allgather: block `rank`.
// Ranges are 16-byte granular (symmetric blocks are 4 KiB granular, so rounding hi up is safe
// only for the LAST range: callers pass block sizes that are multiples of 16).
B2C_EXPORT int b2c_coll_spread_symmetric(b2c_coll* h, void* mc_data, const void* local_data, size_t lo,
                                         size_t hi, void* stream) {
    B2C_TRY
    trace::Scope ts("coll_spread_symmetric(enqueue)", "device", -1, -3, -1, (int64_t)(hi - lo));
    B200_REQUIRE(h && mc_data && local_data && lo <= hi, ERR_ARG, "bad argument");
    B200_REQUIRE((((uintptr_t)mc_data | (uintptr_t)local_data | lo) & 15) == 0, ERR_BUFFER,
                 "symmetric tensors and block offsets must be 16-byte aligned");
    Plan& p = h->p;
    cudaStream_t st = (cudaStream_t)stream;
    p.epoch += 1;
    const int64_t nvec = (int64_t)((hi - lo + 15) / 16);
    // the whole grid waits on flags: keep it co-resident
    const int grid = (int)std::max<int64_t>(1, std::min<int64_t>((nvec + kThreads * kVecUnroll - 1) / (kThreads * kVecUnroll), p.sms * 2));
    spread_inplace_kernel<<<grid, kThreads, 0, st>>>(p.ws, static_cast<unsigned char*>(mc_data),
                                                     static_cast<const unsigned char*>(local_data), lo, hi, p.epoch);
    B200_CHECK_LAUNCH();
    if (grid > 1) {
        tail_kernel<<<1, 64, 0, st>>>(p.ws, p.epoch);
        B200_CHECK_LAUNCH();
    }
    B2C_CATCH
}

// kind: 0 bcast (root pushes `bytes` to everybody), 1 gather (everybody pushes its block to
// root's slot `rank`), 2 allgather, 3 scatter (root pushes block r to rank r).
// phases: 1 = push, 2 = wait + copy out.  `buf` is the send buffer for pushers and the
// receive buffer for receivers (bcast: same buffer; gather: sendbuf/recvbuf given separately).
B2C_EXPORT int b2c_coll_move(b2c_coll* h, int kind, const void* sendbuf, void* recvbuf,
                             size_t block_bytes, int root, int phases, void* stream) {
    B2C_TRY
    trace::Scope ts("coll_move(enqueue)", "device", -1, -3, kind, (int64_t)block_bytes);
    B200_REQUIRE(h, ERR_ARG, "null handle");
    B200_REQUIRE(kind >= 0 && kind <= 3, ERR_ARG, "unknown collective kind");
    Plan& p = h->p;
    const int P = p.ws.nranks, me = p.ws.rank;
    B200_REQUIRE(kind == 2 || (root >= 0 && root < P), ERR_RANK, "invalid root");
    const size_t total = (kind == 0) ? block_bytes : block_bytes * (size_t)P;
    B200_REQUIRE(total <= p.ws.cap, ERR_COUNT, "message exceeds the collective workspace");
    cudaStream_t st = (cudaStream_t)stream;
    if (phases & 1) p.epoch += 1;
    const uint64_t epoch = p.epoch;
    const unsigned long long all = P == 64 ? ~0ull : ((1ull << P) - 1);
    if (phases & 1) {
        PushArgs a{};
        a.src = static_cast<const unsigned char*>(sendbuf);
        a.bytes = block_bytes;
        bool pusher = false;
        if (kind == 0 && me == root) { a.targets = all; pusher = true; }
        if (kind == 1) { a.targets = 1ull << root; a.dst_off[root] = (size_t)me * block_bytes; pusher = true; }
        if (kind == 2) { a.targets = all; for (int r = 0; r < P; ++r) a.dst_off[r] = (size_t)me * block_bytes; pusher = true; }
        if (kind == 3 && me == root) {
            a.targets = all;
            for (int r = 0; r < P; ++r) a.src_off[r] = (size_t)r * block_bytes;
            pusher = true;
        }
        if (pusher) {
            B200_REQUIRE(sendbuf != nullptr || block_bytes == 0, ERR_BUFFER, "null send buffer");
            const int grid = (int)std::max<size_t>(1, std::min<size_t>((block_bytes / 16 + kThreads - 1) / kThreads, (size_t)p.sms * 2));
            if (kind == 0 && p.ws.mc != nullptr && p.use_mc_bcast)
                push_mc_kernel<<<grid, kThreads, 0, st>>>(p.ws, a.src, block_bytes, epoch);
            else
                push_kernel<<<grid, kThreads, 0, st>>>(p.ws, a, epoch);
            B200_CHECK_LAUNCH();
        }
    }
    if (phases & 2) {
        unsigned long long sources = 0;
        size_t bytes = 0;
        if (kind == 0) { sources = 1ull << root; bytes = block_bytes; }
        if (kind == 1 && me == root) { sources = all; bytes = total; }
        if (kind == 2) { sources = all; bytes = total; }
        if (kind == 3) { sources = 1ull << root; bytes = block_bytes; }
        if (sources) {
            B200_REQUIRE(recvbuf != nullptr || bytes == 0, ERR_BUFFER, "null receive buffer");
            const int grid = (int)std::max<size_t>(1, std::min<size_t>((bytes / 16 + kThreads - 1) / kThreads, (size_t)p.sms * 2));
            pull_local_kernel<<<grid, kThreads, 0, st>>>(p.ws, static_cast<unsigned char*>(recvbuf), 0,
                                                         bytes, sources, epoch);
            B200_CHECK_LAUNCH();
        }
        tail_kernel<<<1, 64, 0, st>>>(p.ws, epoch);
        B200_CHECK_LAUNCH();
    }
    B2C_CATCH
}

B2C_EXPORT int b2c_coll_barrier(b2c_coll* h, int phases, void* stream) {
    B2C_TRY
    B200_REQUIRE(h, ERR_ARG, "null handle");
    Plan& p = h->p;
    if (phases & 1) p.bar_epoch += 1;
    barrier_kernel<<<1, 64, 0, (cudaStream_t)stream>>>(p.ws, p.bar_epoch, phases);
    B200_CHECK_LAUNCH();
    B2C_CATCH
}

// Synchronises the stream and reports device-side time-outs.
B2C_EXPORT int b2c_coll_check(b2c_coll* h, void* stream) {
    B2C_TRY
    B200_REQUIRE(h, ERR_ARG, "null handle");
    unsigned int st = 0;
    B200_CHECK_CUDA(cudaMemcpyAsync(&st, h->p.ws.peer[h->p.ws.rank] + kOffStatus, 4,
                                    cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    B200_CHECK_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    B200_REQUIRE((st & 8u) == 0, ERR_TIMEOUT, "device collective: a peer never arrived");
    B2C_CATCH
}

}  // extern "C"
