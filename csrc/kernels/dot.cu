// Distributed dot product: local reduction fused with a one-shot P2P allreduce.
//
// Reference capability: mpicuda2.cu / mpicuda3.cu / mpicuda4.cu
// -- per-rank kernel (atomicAdd variant :65-81, per-block partials :84-100,
// single-launch "last block finishes" mpicuda4.cu:157-185), then D2H of the
// scalar, then MPI_Reduce(SUM, root 0) on the host (mpicuda4.cu:354-368).
//
// Here ONE kernel per rank does everything:
//   1. grid-stride FMA over 16-byte vector loads, fp64 accumulation;
//   2. per-CTA partial -> scratch[blockIdx]; __threadfence; arrival counter;
//   3. the LAST CTA sums the partials in a fixed order (deterministic), then
//      -- instead of returning to the host -- stores the rank's partial into
//      slot[my_rank] of EVERY peer's symmetric slot table (8-byte peer stores
//      over NVLink) and release-stores the epoch next to it;
//   4. the same CTA acquires the epochs of all ranks in its own table and adds
//      the P values in rank order, so every rank gets bit-identical results.
// No D2H, no host MPI, no second launch.  `phases` lets single-GPU emulation of
// several ranks split push (bit 0) and gather (bit 1) into two launches, because
// kernels that wait on each other must not be queued on one GPU.
//
// The reference's other two local-reduce variants (atomicAdd per block, host
// finish of per-block partials) are kept in csrc/baseline/ as baselines.
#include <cuda_runtime.h>
#include <stdint.h>

#include <mutex>
#include <new>
#include <string>

#include "b200comm.h"
#include "b200comm/error.hpp"
#include "core/capi_util.hpp"
#include "core/trace.hpp"
#include "kernels/ptx.cuh"

namespace b200 {
namespace dot {

constexpr int kThreads = 256;
constexpr int kMaxRanks = 64;
constexpr int kUnroll = 4;

// One slot per (parity, source rank): {value, epoch}; 16 bytes.
struct Slot {
    double value;
    uint64_t epoch;
};

struct Params {
    const void* x;
    const void* y;
    int64_t n;
    double* partials;          // scratch: [gridDim.x]
    unsigned int* counter;     // scratch: arrival counter (self-resetting)
    Slot* my_slots;            // [2][nranks] in my symmetric block
    Slot* peer_slots[kMaxRanks];   // every rank's table (peer mapped), incl. mine
    double* out;               // device scalar result
    unsigned int* status;
    uint64_t timeout_ns;
    uint64_t epoch;            // 1, 2, 3, ... (same on every rank)
    // Back-pressure: before I overwrite slot [parity][me] in rank r's table, r must have finished
    // reading it in the last earlier epoch of the same parity in which r gathered (need_ack[r],
    // tracked by the host, 0 = never).  Readers publish the epoch they finished in ack[reader] of
    // every rank's table (the u64 array that follows the 2*nranks slots).
    uint64_t need_ack[kMaxRanks];
    int rank, nranks;
    int root;                  // -1: allreduce, else reduce to this rank
    int phases;                // bit0 local reduce + push, bit1 gather + sum
};

template <typename T> struct Vec;
template <> struct Vec<double> { using type = double2; static constexpr int N = 2; };
template <> struct Vec<float> { using type = float4; static constexpr int N = 4; };

__device__ __forceinline__ double dotv(const double2& a, const double2& b, double acc) {
    acc = fma(a.x, b.x, acc);
    return fma(a.y, b.y, acc);
}
__device__ __forceinline__ double dotv(const float4& a, const float4& b, double acc) {
    acc = fma((double)a.x, (double)b.x, acc);
    acc = fma((double)a.y, (double)b.y, acc);
    acc = fma((double)a.z, (double)b.z, acc);
    return fma((double)a.w, (double)b.w, acc);
}

__device__ __forceinline__ double block_sum(double v, double* sh) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0) sh[warp] = v;
    __syncthreads();
    double s = 0.0;
    if (warp == 0) {
        s = lane < (int)(blockDim.x >> 5) ? sh[lane] : 0.0;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    }
    __syncthreads();
    return s;   // valid in warp 0
}

// Steps 3 and 4 of the header comment, executed by one CTA.
__device__ void exchange_and_finish(const Params& P, double local, double* sh) {
    const int par = (int)(P.epoch & 1);
    if (P.phases & 1) {
        // push my partial to every rank that needs it (one 16-byte slot each)
        for (int r = threadIdx.x; r < P.nranks; r += blockDim.x) {
            if (P.root >= 0 && r != P.root) continue;
            if (P.need_ack[r] != 0) {
                // rank r may still be reading this slot (a rooted reduce lets the non-root ranks
                // run ahead of the root): wait for its acknowledgement of that epoch
                const uint64_t* ack = reinterpret_cast<const uint64_t*>(P.my_slots + 2 * (size_t)P.nranks) + r;
                const uint64_t t0 = ptx::globaltimer_ns();
                uint32_t spins = 0;
                while (ptx::ld_acquire_sys(ack) < P.need_ack[r]) {
                    if ((++spins & 255u) == 0 && ptx::globaltimer_ns() - t0 > P.timeout_ns) {
                        atomicOr(P.status, 2u);
                        break;
                    }
                }
            }
            Slot* s = P.peer_slots[r] + (size_t)par * P.nranks + P.rank;
            s->value = local;
            ptx::st_release_sys(&s->epoch, P.epoch);
        }
    }
    if (!(P.phases & 2)) return;
    if (P.root >= 0 && P.rank != P.root) return;
    // gather: thread r waits for rank r, then a fixed-order sum
    double v = 0.0;
    if ((int)threadIdx.x < P.nranks) {
        const Slot* s = P.my_slots + (size_t)par * P.nranks + threadIdx.x;
        const uint64_t t0 = ptx::globaltimer_ns();
        uint32_t spins = 0;
        while (ptx::ld_acquire_sys(&s->epoch) < P.epoch) {
            if ((++spins & 255u) == 0 && ptx::globaltimer_ns() - t0 > P.timeout_ns) {
                atomicOr(P.status, 2u);
                break;
            }
        }
        v = *reinterpret_cast<const volatile double*>(&s->value);
    }
    __shared__ double vals[kMaxRanks];
    if ((int)threadIdx.x < P.nranks) vals[threadIdx.x] = v;
    __syncthreads();
    // every slot of this epoch has been read: tell the pushers (ack[me] in THEIR tables)
    if ((int)threadIdx.x < P.nranks)
        ptx::st_release_sys(reinterpret_cast<uint64_t*>(P.peer_slots[threadIdx.x] + 2 * (size_t)P.nranks) + P.rank,
                            P.epoch);
    if (threadIdx.x == 0) {
        double tot = 0.0;
        for (int r = 0; r < P.nranks; ++r) tot += vals[r];   // rank order: same bits everywhere
        *P.out = tot;
    }
    (void)sh;
}

template <typename T>
__global__ void __launch_bounds__(kThreads, 4) dot_fused_kernel(const __grid_constant__ Params P) {
    using V = typename Vec<T>::type;
    constexpr int VN = Vec<T>::N;
    __shared__ double sh[kThreads / 32];
    __shared__ bool last;
    if (!(P.phases & 1)) {          // gather-only launch (emulation): one CTA
        if (blockIdx.x == 0) exchange_and_finish(P, 0.0, sh);
        return;
    }
    const T* x = static_cast<const T*>(P.x);
    const T* y = static_cast<const T*>(P.y);
    const int64_t nvec = P.n / VN;
    const V* xv = reinterpret_cast<const V*>(x);
    const V* yv = reinterpret_cast<const V*>(y);
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    double acc[kUnroll] = {0.0, 0.0, 0.0, 0.0};
    // main loop: kUnroll independent 16-byte loads of each operand in flight
    for (; i + (kUnroll - 1) * stride < nvec; i += kUnroll * stride) {
        V a[kUnroll], b[kUnroll];
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
            a[u] = __ldcs(xv + i + u * stride);
            b[u] = __ldcs(yv + i + u * stride);
        }
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) acc[u] = dotv(a[u], b[u], acc[u]);
    }
    for (; i < nvec; i += stride) acc[0] = dotv(__ldcs(xv + i), __ldcs(yv + i), acc[0]);
    // scalar tail
    for (int64_t k = nvec * VN + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < P.n; k += stride)
        acc[0] = fma((double)x[k], (double)y[k], acc[0]);
    double s = block_sum((acc[0] + acc[1]) + (acc[2] + acc[3]), sh);
    if (threadIdx.x == 0) {
        P.partials[blockIdx.x] = s;
        __threadfence();
        const unsigned int v = atomicInc(P.counter, gridDim.x - 1);   // wraps back to 0
        last = v == gridDim.x - 1;
    }
    __syncthreads();
    if (!last) return;
    __threadfence();
    // the last CTA: deterministic sum of the partials ...
    double t = 0.0;
    for (int k = threadIdx.x; k < (int)gridDim.x; k += blockDim.x)
        t += *reinterpret_cast<volatile double*>(P.partials + k);
    t = block_sum(t, sh);
    __shared__ double local_total;
    if (threadIdx.x == 0) local_total = t;
    __syncthreads();
    // ... and the cross-GPU exchange, in the same kernel
    exchange_and_finish(P, local_total, sh);
}

struct Plan {
    int device;
    int rank, nranks;
    int grid;
    double* partials;
    unsigned int* counter;
    double* out;
    unsigned int* status;
    Slot* my_slots;
    Slot* peer_slots[kMaxRanks];
    uint64_t epoch;
    uint64_t timeout_ns;
    bool owns_scratch;
    uint64_t last_read[kMaxRanks][2];   // last epoch of each parity in which rank r gathered
};

}  // namespace dot
}  // namespace b200

using namespace b200;
using namespace b200::dot;

struct b2c_dot { Plan p; };

static void launch_dot(const Plan& pl, const void* x, const void* y, int64_t n, int dtype, int root,
                       int phases, double* partials, unsigned int* counter, cudaStream_t st) {
    Params P{};
    P.x = x; P.y = y; P.n = n;
    P.partials = partials; P.counter = counter;
    P.my_slots = pl.my_slots;
    for (int r = 0; r < pl.nranks; ++r) P.peer_slots[r] = pl.peer_slots[r];
    P.out = pl.out; P.status = pl.status; P.timeout_ns = pl.timeout_ns;
    P.epoch = pl.epoch; P.rank = pl.rank; P.nranks = pl.nranks; P.root = root; P.phases = phases;
    for (int r = 0; r < pl.nranks; ++r) P.need_ack[r] = pl.last_read[r][pl.epoch & 1];
    const int grid = (phases & 1) ? pl.grid : 1;
    if (dtype == 9) dot_fused_kernel<double><<<grid, kThreads, 0, st>>>(P);
    else dot_fused_kernel<float><<<grid, kThreads, 0, st>>>(P);
    B200_CHECK_LAUNCH();
}

extern "C" {

// Bytes of symmetric memory a dot plan needs on every rank (slot tables).
B2C_EXPORT int b2c_dot_symmetric_bytes(int nranks, size_t* bytes) {
    B2C_TRY
    B200_REQUIRE(bytes && nranks >= 1 && nranks <= kMaxRanks, ERR_ARG, "bad rank count");
    // 2 x nranks slots {value, epoch}, then nranks acknowledgement words
    *bytes = (((size_t)2 * nranks * sizeof(Slot)) + (size_t)nranks * sizeof(uint64_t) + 255) & ~(size_t)255;
    B2C_CATCH
}

// `slot_tables[r]` = address of rank r's slot table as mapped in THIS process
// (slot_tables[rank] is the local one).  The tables must start zeroed.
B2C_EXPORT int b2c_dot_create(b2c_dot** out, int rank, int nranks, void* const* slot_tables) {
    B2C_TRY
    B200_REQUIRE(out && slot_tables, ERR_ARG, "null argument");
    B200_REQUIRE(nranks >= 1 && nranks <= kMaxRanks && rank >= 0 && rank < nranks, ERR_RANK,
                 "bad rank / rank count");
    auto* h = new (std::nothrow) b2c_dot();
    B200_REQUIRE(h, ERR_NOMEM, "plan allocation failed");
    Plan& p = h->p;
    p = Plan{};
    B200_CHECK_CUDA(cudaGetDevice(&p.device));
    int sms = 0;
    B200_CHECK_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, p.device));
    p.rank = rank; p.nranks = nranks;
    p.grid = sms * 4;     // 4 x 256 threads per SM, one wave, 128 B of loads in flight per thread
    for (int r = 0; r < nranks; ++r) p.peer_slots[r] = static_cast<Slot*>(slot_tables[r]);
    p.my_slots = p.peer_slots[rank];
    // scratch: per-CTA partials, arrival counter, result, status
    B200_CHECK_CUDA(cudaMalloc(&p.partials, (size_t)p.grid * sizeof(double)));
    B200_CHECK_CUDA(cudaMalloc(&p.counter, 256));
    B200_CHECK_CUDA(cudaMemset(p.counter, 0, 256));
    p.out = reinterpret_cast<double*>(reinterpret_cast<char*>(p.counter) + 64);
    p.status = reinterpret_cast<unsigned int*>(reinterpret_cast<char*>(p.counter) + 128);
    p.epoch = 0;
    p.timeout_ns = 20ull * 1000 * 1000 * 1000;
    p.owns_scratch = true;
    *out = h;
    B2C_CATCH
}

B2C_EXPORT int b2c_dot_destroy(b2c_dot* h) {
    B2C_TRY
    if (!h) return 0;
    if (h->p.owns_scratch) { cudaFree(h->p.partials); cudaFree(h->p.counter); }
    delete h;
    B2C_CATCH
}

// Enqueue one fused dot + allreduce (root < 0) or dot + reduce-to-root.
// dtype: 8 float32, 9 float64 (the basic-kind numbering of the datatype API).
// phases: 3 = everything in one launch (one rank per GPU);
//         1 then 2 = push and gather as separate launches (several emulated ranks on one GPU).
// The epoch advances when bit 0 is set.
B2C_EXPORT int b2c_dot_run(b2c_dot* h, const void* x, const void* y, int64_t n, int dtype, int root,
                           int phases, void* stream) {
    B2C_TRY
    trace::Scope ts("dot_run(enqueue)", "device", -1, -3, -1, n);
    B200_REQUIRE(h && (x || n == 0) && (y || n == 0), ERR_ARG, "null argument");
    B200_REQUIRE(dtype == 8 || dtype == 9, ERR_TYPE, "dot supports float32 and float64");
    B200_REQUIRE(n >= 0, ERR_COUNT, "negative length");
    B200_REQUIRE(((uintptr_t)x & 15) == 0 && ((uintptr_t)y & 15) == 0, ERR_BUFFER,
                 "dot operands must be 16-byte aligned");
    B200_REQUIRE(root < h->p.nranks, ERR_RANK, "root out of range");
    B200_REQUIRE((phases & 3) != 0, ERR_ARG, "phases must include push (1) and/or gather (2)");
    Plan& p = h->p;
    if (phases & 1) p.epoch += 1;
    launch_dot(p, x, y, n, dtype, root, phases, p.partials, p.counter, (cudaStream_t)stream);
    if (phases & 2)   // who reads its slot table in this epoch: the root, or everybody
        for (int r = 0; r < p.nranks; ++r)
            if (root < 0 || r == root) p.last_read[r][p.epoch & 1] = p.epoch;
    B2C_CATCH
}

// Device address of the result scalar (valid after the launch with bit 1 completed).
B2C_EXPORT int b2c_dot_result_ptr(b2c_dot* h, double** out) {
    B2C_TRY
    B200_REQUIRE(h && out, ERR_ARG, "null argument");
    *out = h->p.out;
    B2C_CATCH
}

// Synchronises `stream`, returns the result and raises ERR_TIMEOUT if a peer never delivered.
B2C_EXPORT int b2c_dot_result(b2c_dot* h, void* stream, double* value) {
    B2C_TRY
    B200_REQUIRE(h && value, ERR_ARG, "null argument");
    struct { double v; char pad[56]; unsigned int st; } host{};
    B200_CHECK_CUDA(cudaMemcpyAsync(&host.v, h->p.out, sizeof(double), cudaMemcpyDeviceToHost,
                                    (cudaStream_t)stream));
    B200_CHECK_CUDA(cudaMemcpyAsync(&host.st, h->p.status, sizeof(unsigned), cudaMemcpyDeviceToHost,
                                    (cudaStream_t)stream));
    B200_CHECK_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    B200_REQUIRE((host.st & 2u) == 0, ERR_TIMEOUT,
                 "dot allreduce: a peer's partial never arrived (device-side wait timed out)");
    *value = host.v;
    B2C_CATCH
}

B2C_EXPORT int b2c_dot_set_timeout_ms(b2c_dot* h, unsigned long long ms) {
    B2C_TRY
    B200_REQUIRE(h, ERR_ARG, "null plan");
    h->p.timeout_ns = ms * 1000000ull;
    B2C_CATCH
}

}  // extern "C"

// =============================================================================
// Local dot product for callers that must not allocate
// =============================================================================
//
// b2c_dot_create() above cudaMalloc's the kernel's scratch (per-CTA partial sums,
// arrival counter, slot table).  Some users of this library cannot do that at
// the place where they need a dot product:
//   * a solver step that runs inside another runtime's callback or a latency-
//     critical loop, where cudaMalloc is off limits because it synchronises the
//     whole device and stalls every other stream;
//   * a process whose GPU memory has been handed, in full, to a framework's
//     caching allocator or to a fixed arena -- a driver-level allocation can then
//     simply fail, or break the footprint the application promised its host;
//   * code that is only given raw device pointers and a stream, with no place to
//     keep a plan object alive between calls.
// For them the same kernel is offered with its scratch in a fixed device array
// that is part of the library image.  Every GPU on which the library is loaded
// carries these kNoAllocBytes (8448 bytes) until the process exits.
//
// Rules that follow from sharing one array:
//   * it keeps whatever the previous call left in it, so each call first clears
//     the control words that must start at zero (arrival counter, slot epochs,
//     status) with a stream-ordered memset; the partial sums are fully
//     overwritten and need no clearing;
//   * two calls whose execution overlaps on the same GPU (different streams, or
//     different host threads) use the same scratch at the same time and corrupt
//     each other's result -- serialise them on one stream;
//   * the grid is limited to what the array can hold (kNoAllocMaxCtas partial
//     sums); a request beyond that is rejected, never silently shrunk.
//
// The address of the array is looked up once per device, by the first call made
// while that device is current (cudaGetSymbolAddress; its return code is
// checked).  That first call does more than launch work, so make it before any
// stream capture is started.  The address is remembered per device ordinal for
// the life of the process: a program that calls cudaDeviceReset() on a device
// must not use this entry point on that device afterwards (the remembered
// address would belong to the destroyed context).
namespace b200 {
namespace dot {

constexpr int kNoAllocMaxCtas = 1024;                 // partial sums the scratch can hold
constexpr size_t kNoAllocCtrlBytes = 256;             // counter @0, status @64, slots @128
constexpr size_t kNoAllocBytes = kNoAllocCtrlBytes + (size_t)kNoAllocMaxCtas * sizeof(double);  // 8448

// File-scope device scratch.  In HOST code this name is only a handle for the
// symbol, not an address a kernel could use; the device address comes from
// cudaGetSymbolAddress below.
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocBytes];

struct NoAllocDevice {
    unsigned char* scratch = nullptr;   // device address of g_noalloc_scratch on this GPU
    int sms = 0;
};
static NoAllocDevice g_noalloc_dev[64];
static std::mutex g_noalloc_mu;

// One-time set-up for the CURRENT device: resolve the scratch address.
static const NoAllocDevice& noalloc_device() {
    int dev = 0;
    B200_CHECK_CUDA(cudaGetDevice(&dev));
    B200_REQUIRE(dev >= 0 && dev < 64, ERR_ARG, "device index out of range");
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    NoAllocDevice& d = g_noalloc_dev[dev];
    if (d.scratch == nullptr) {
        void* p = nullptr;
        cudaError_t st = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (st != cudaSuccess || p == nullptr)
            fail_cuda((int)st, "cudaGetSymbolAddress(g_noalloc_scratch)", __FILE__, __LINE__,
                      "cannot resolve the built-in dot scratch on this device:");
        int sms = 0;
        B200_CHECK_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
        d.sms = sms;
        d.scratch = static_cast<unsigned char*>(p);
    }
    return d;
}

}  // namespace dot
}  // namespace b200

extern "C" {

B2C_EXPORT int b2c_dot_local_noalloc_limits(int* max_ctas, size_t* scratch_bytes) {
    if (max_ctas) *max_ctas = kNoAllocMaxCtas;
    if (scratch_bytes) *scratch_bytes = kNoAllocBytes;
    return 0;
}

// result[0] = sum_i x[i] * y[i] for this GPU's vectors, computed by the fused dot
// kernel with the library's built-in scratch: no cudaMalloc, no cudaFree, no plan.
//   x, y    device vectors of n elements, 16-byte aligned (dtype 8 float32, 9 float64)
//   ctas    thread blocks to use; 0 = 4 per SM.  More than the scratch can hold
//           (b2c_dot_local_noalloc_limits) is an error.
//   result  device address of one double, written by the kernel
//   stream  the launch and the clearing memset are ordered on it; the caller
//           synchronises before reading `result`
// See the comment block above for the sharing rules (calls that overlap on a GPU
// race) and for the first-call set-up.
B2C_EXPORT int b2c_dot_local_noalloc(const void* x, const void* y, int64_t n, int dtype, int ctas,
                                     double* result, void* stream) {
    B2C_TRY
    B200_REQUIRE(result != nullptr, ERR_BUFFER, "null result pointer");
    B200_REQUIRE((x && y) || n == 0, ERR_BUFFER, "null input vector");
    B200_REQUIRE(dtype == 8 || dtype == 9, ERR_TYPE, "dot supports float32 (8) and float64 (9)");
    B200_REQUIRE(n >= 0, ERR_COUNT, "negative length");
    B200_REQUIRE(((uintptr_t)x & 15) == 0 && ((uintptr_t)y & 15) == 0, ERR_BUFFER,
                 "dot operands must be 16-byte aligned");
    B200_REQUIRE(ctas >= 0, ERR_ARG, "negative block count");
    const NoAllocDevice& d = noalloc_device();
    const int grid = ctas > 0 ? ctas : d.sms * 4;
    B200_REQUIRE(grid <= kNoAllocMaxCtas, ERR_ARG,
                 "requested " + std::to_string(grid) + " thread blocks but the built-in scratch holds "
                 "partial sums for at most " + std::to_string(kNoAllocMaxCtas) +
                 "; pass a smaller `ctas` or use b2c_dot_create(), which sizes its own scratch");
    cudaStream_t st = (cudaStream_t)stream;
    // what must start at zero: arrival counter, status word, slot epochs
    B200_CHECK_CUDA(cudaMemsetAsync(d.scratch, 0, kNoAllocCtrlBytes, st));
    Params P{};
    P.x = x; P.y = y; P.n = n;
    P.counter = reinterpret_cast<unsigned int*>(d.scratch);
    P.status = reinterpret_cast<unsigned int*>(d.scratch + 64);
    P.my_slots = reinterpret_cast<Slot*>(d.scratch + 128);      // [2 parities][1 rank]
    P.peer_slots[0] = P.my_slots;
    P.partials = reinterpret_cast<double*>(d.scratch + kNoAllocCtrlBytes);
    P.out = result;
    P.timeout_ns = 1000ull * 1000 * 1000;
    P.epoch = 1; P.rank = 0; P.nranks = 1; P.root = -1; P.phases = 3;
    if (dtype == 9) dot_fused_kernel<double><<<grid, kThreads, 0, st>>>(P);
    else dot_fused_kernel<float><<<grid, kThreads, 0, st>>>(P);
    B200_CHECK_LAUNCH();
    B2C_CATCH
}

}  // extern "C"
