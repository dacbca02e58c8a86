// Device-buffer support of the comm layer: classification of pointers, CUDA-IPC
// export / mapping of the sender's allocation, and the receiver-side "pull"
// that moves (and packs/unpacks) the payload with a kernel.
//
// Reference capability: MPI_Send / MPI_Recv / MPI_Isend / MPI_Irecv on DEVICE
// pointers with basic or derived datatypes through a CUDA-aware MPI
// (test-benchmark/mpi-pingpong-gpu.cpp:52-53,75-76, mpi-pingpong-gpu-async.cpp:
// 64-67, stencil2d/stencil2D.h:366-376).  Protocol here:
//   sender   isend(device buf)  -> REC_DEV_RTS {IPC handle, offset, runs}   (host fabric)
//   receiver match              -> map sender allocation (cached) and launch
//                                  the typed move kernel src(peer) -> dst(local)
//                               -> REC_DEV_FIN                              (host fabric)
//   sender   request completes on FIN.
// No payload byte touches host memory.
#include <cuda.h>
#include <cuda_runtime.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <ctime>
#include <cstring>
#include <map>
#include <mutex>
#include <tuple>
#include <vector>

#include "b200comm.h"
#include "b200comm/datatype.hpp"
#include "b200comm/error.hpp"
#include "comm/engine.hpp"
#include "core/capi_util.hpp"
#include "core/handles.hpp"
#include "kernels/ptx.cuh"

namespace b200 {
namespace packmove {
void launch_uniform(const void* src, void* dst, int64_t nruns, int64_t len, int64_t sstride,
                    int64_t dstride, cudaStream_t st);
void launch_segments(const void* src, void* dst, const Segment* dev_segs, int64_t nseg,
                     int64_t max_len, int64_t total_items_64k, cudaStream_t st,
                     const int64_t* dev_item_start);
}  // namespace packmove

namespace devp2p {

static std::mutex g_mu;

static bool has_cuda() {
    static int has = -1;
    if (has < 0) {
        int n = 0;
        has = (cudaGetDeviceCount(&n) == cudaSuccess && n > 0) ? 1 : 0;
        cudaGetLastError();
    }
    return has == 1;
}

using GetRangeFn = CUresult (*)(CUdeviceptr*, size_t*, CUdeviceptr);
static GetRangeFn get_range_fn() {
    static GetRangeFn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuMemGetAddressRange", &p, cudaEnableDefault, &q) == cudaSuccess)
            fn = reinterpret_cast<GetRangeFn>(p);
        cudaGetLastError();
    }
    return fn;
}

static bool classify(const void* p, int* device, uint64_t* base, uint64_t* bytes) {
    if (!p || !has_cuda()) return false;
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    if (a.type != cudaMemoryTypeDevice) return false;
    *device = a.device;
    *base = (uint64_t)(uintptr_t)p;
    *bytes = 0;
    if (GetRangeFn fn = get_range_fn()) {
        CUdeviceptr b = 0;
        size_t n = 0;
        if (fn(&b, &n, (CUdeviceptr)(uintptr_t)p) == CUDA_SUCCESS) {
            *base = (uint64_t)b;
            *bytes = (uint64_t)n;
        }
    }
    return true;
}

// Identity of the allocation behind an address (unique for the life of the process).
using PtrAttrFn = CUresult (*)(void*, CUpointer_attribute, CUdeviceptr);
static uint64_t buffer_id_of(uint64_t ptr) {
    static PtrAttrFn fn = [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuPointerGetAttribute", &p, cudaEnableDefault, &q) != cudaSuccess) p = nullptr;
        cudaGetLastError();
        return reinterpret_cast<PtrAttrFn>(p);
    }();
    unsigned long long id = 0;
    if (fn && fn(&id, CU_POINTER_ATTRIBUTE_BUFFER_ID, (CUdeviceptr)ptr) == CUDA_SUCCESS) return id;
    return 0;
}

static void export_handle(comm::DevDesc* d) {
    d->pid = (int64_t)getpid();
    d->buffer_id = buffer_id_of(d->base_ptr);
    std::memset(d->ipc_handle, 0, sizeof(d->ipc_handle));
    // keyed by address AND allocation identity: an address that was freed and handed out again
    // gets a fresh handle instead of the stale one
    static std::map<uint64_t, std::pair<uint64_t, cudaIpcMemHandle_t>> cache;
    std::lock_guard<std::mutex> lk(g_mu);
    auto it = cache.find(d->base_ptr);
    if (it != cache.end() && it->second.first != d->buffer_id) {
        cache.erase(it);
        it = cache.end();
    }
    if (it == cache.end()) {
        cudaIpcMemHandle_t h;
        if (cudaIpcGetMemHandle(&h, reinterpret_cast<void*>((uintptr_t)d->base_ptr)) != cudaSuccess) {
            cudaGetLastError();   // not IPC-capable: only same-process receivers can pull
            return;
        }
        it = cache.emplace(d->base_ptr, std::make_pair(d->buffer_id, h)).first;
    }
    std::memcpy(d->ipc_handle, &it->second.second, sizeof(cudaIpcMemHandle_t));
}

// ---- typed move with a small plan cache -----------------------------------------
struct MovePlan {
    bool uniform = false;
    int64_t nruns = 0, len = 0, sstride = 0, dstride = 0, s0 = 0, d0 = 0;
    Segment* dev_segs = nullptr;
    int64_t* dev_item_start = nullptr;   // prefix sums of the 64 KiB work items per segment
    int64_t nseg = 0, max_len = 0, items = 0, total = 0;
};

static bool as_uniform(const std::vector<Run>& r, int64_t want_len, int64_t* n, int64_t* len,
                       int64_t* stride, int64_t* d0) {
    if (r.empty()) return false;
    if (r.size() == 1) {   // one dense run can be cut into pieces of any length
        const int64_t l = want_len > 0 ? want_len : r[0].len;
        if (r[0].len % l) return false;
        *n = r[0].len / l; *len = l; *stride = l; *d0 = r[0].disp;
        return true;
    }
    const int64_t l = r[0].len, s = r[1].disp - r[0].disp;
    for (size_t i = 0; i < r.size(); ++i) {
        if (r[i].len != l) return false;
        if (i && r[i].disp - r[i - 1].disp != s) return false;
    }
    *n = (int64_t)r.size(); *len = l; *stride = s; *d0 = r[0].disp;
    return true;
}

static uint64_t hash_runs(const std::vector<Run>& r, uint64_t h) {
    for (const Run& x : r) {
        h = (h ^ (uint64_t)x.disp) * 1099511628211ull;
        h = (h ^ (uint64_t)x.len) * 1099511628211ull;
    }
    return h;
}

static const MovePlan& get_plan(const std::vector<Run>& s, const std::vector<Run>& d, int device) {
    // keyed by a hash of both run lists; the lists themselves are kept and compared on a hit, so
    // two layouts that collide get two plans instead of one silently wrong move
    struct Entry { std::vector<Run> s, d; MovePlan plan; };
    static std::multimap<std::pair<uint64_t, int>, Entry> cache;
    const uint64_t key = hash_runs(d, hash_runs(s, 1469598103934665603ull) ^ 0x9e3779b97f4a7c15ull);
    std::lock_guard<std::mutex> lk(g_mu);
    auto same = [](const std::vector<Run>& a, const std::vector<Run>& b) {
        if (a.size() != b.size()) return false;
        for (size_t i = 0; i < a.size(); ++i)
            if (a[i].disp != b[i].disp || a[i].len != b[i].len) return false;
        return true;
    };
    auto range = cache.equal_range({key, device});
    for (auto it = range.first; it != range.second; ++it)
        if (same(it->second.s, s) && same(it->second.d, d)) return it->second.plan;
    MovePlan p;
    int64_t sn, sl, ss, s0, dn, dl, ds, d0;
    const int64_t hint_s = d.size() > 1 ? d[0].len : 0, hint_d = s.size() > 1 ? s[0].len : 0;
    if (as_uniform(s, hint_s, &sn, &sl, &ss, &s0) && as_uniform(d, hint_d, &dn, &dl, &ds, &d0) &&
        sl == dl && sn == dn) {
        p.uniform = true;
        p.nruns = sn; p.len = sl; p.sstride = ss; p.dstride = ds; p.s0 = s0; p.d0 = d0;
        p.total = sn * sl;
    } else {
        std::vector<Segment> segs = merge_runs(s, d);
        p.nseg = (int64_t)segs.size();
        for (const Segment& g : segs) {
            p.max_len = std::max(p.max_len, g.len);
            p.items += (g.len + 65535) / 65536;
            p.total += g.len;
        }
        if (p.nseg) {
            B200_CHECK_CUDA(cudaMalloc(&p.dev_segs, segs.size() * sizeof(Segment)));
            B200_CHECK_CUDA(cudaMemcpy(p.dev_segs, segs.data(), segs.size() * sizeof(Segment),
                                       cudaMemcpyHostToDevice));
            if (p.max_len > 256) {
                std::vector<int64_t> start(segs.size() + 1, 0);
                for (size_t i = 0; i < segs.size(); ++i) start[i + 1] = start[i] + (segs[i].len + 65535) / 65536;
                B200_CHECK_CUDA(cudaMalloc(&p.dev_item_start, start.size() * sizeof(int64_t)));
                B200_CHECK_CUDA(cudaMemcpy(p.dev_item_start, start.data(), start.size() * sizeof(int64_t),
                                           cudaMemcpyHostToDevice));
            }
        }
    }
    return cache.emplace(std::make_pair(key, device), Entry{s, d, p})->second.plan;
}

static int64_t run_plan(const MovePlan& p, const void* src, void* dst, cudaStream_t st);

// Variant for callers that already hold a stable identity for the (src, dst)
// run lists: the plan is looked up by that identity, no hashing of the runs.
int64_t typed_move_cached(const void* src, const std::vector<Run>& sruns, void* dst,
                          const std::vector<Run>& druns, const void* identity, cudaStream_t st) {
    static std::map<const void*, MovePlan> by_id;
    const MovePlan* p;
    {
        std::lock_guard<std::mutex> lk(g_mu);
        auto it = by_id.find(identity);
        p = it == by_id.end() ? nullptr : &it->second;
    }
    if (!p) {
        int dev = 0;
        B200_CHECK_CUDA(cudaGetDevice(&dev));
        const MovePlan& built = get_plan(sruns, druns, dev);
        std::lock_guard<std::mutex> lk(g_mu);
        p = &by_id.emplace(identity, built).first->second;
    }
    return run_plan(*p, src, dst, st);
}

int64_t typed_move(const void* src, const std::vector<Run>& sruns, void* dst,
                   const std::vector<Run>& druns, cudaStream_t st) {
    int dev = 0;
    B200_CHECK_CUDA(cudaGetDevice(&dev));
    return run_plan(get_plan(sruns, druns, dev), src, dst, st);
}

static int64_t run_plan(const MovePlan& p, const void* src, void* dst, cudaStream_t st) {
    if (p.uniform) {
        packmove::launch_uniform(static_cast<const char*>(src) + p.s0, static_cast<char*>(dst) + p.d0,
                                 p.nruns, p.len, p.sstride, p.dstride, st);
    } else if (p.nseg) {
        packmove::launch_segments(src, dst, p.dev_segs, p.nseg, p.max_len, p.items, st, p.dev_item_start);
    }
    return p.total;
}

static cudaStream_t pull_stream(int device) {
    static thread_local std::map<int, cudaStream_t> streams;
    auto it = streams.find(device);
    if (it != streams.end()) return it->second;
    cudaStream_t s;
    B200_CHECK_CUDA(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
    streams[device] = s;
    return s;
}

static void* staging(int device, size_t bytes) {
    static thread_local std::map<int, std::pair<void*, size_t>> bufs;
    auto& b = bufs[device];
    if (b.second < bytes) {
        if (b.first) cudaFree(b.first);
        B200_CHECK_CUDA(cudaMalloc(&b.first, bytes));
        b.second = bytes;
    }
    return b.first;
}

static const void* map_source(const comm::DevDesc& d, int my_device) {
    if (d.pid == (int64_t)getpid()) {
        if (d.device != my_device) {
            cudaError_t st = cudaDeviceEnablePeerAccess(d.device, 0);
            if (st != cudaSuccess && st != cudaErrorPeerAccessAlreadyEnabled)
                fail_cuda((int)st, "cudaDeviceEnablePeerAccess", __FILE__, __LINE__);
            cudaGetLastError();
        }
        return reinterpret_cast<const void*>((uintptr_t)d.base_ptr);
    }
    // (pid, base) -> (allocation identity, mapping); a different identity at the same address means
    // the sender freed and re-allocated: drop the mapping of the dead allocation and map again
    static std::map<std::pair<int64_t, uint64_t>, std::pair<uint64_t, void*>> opened;
    std::lock_guard<std::mutex> lk(g_mu);
    auto it = opened.find({d.pid, d.base_ptr});
    if (it != opened.end()) {
        if (it->second.first == d.buffer_id) return it->second.second;
        cudaIpcCloseMemHandle(it->second.second);
        cudaGetLastError();
        opened.erase(it);
    }
    cudaIpcMemHandle_t h;
    std::memcpy(&h, d.ipc_handle, sizeof(h));
    bool zero = true;
    for (size_t i = 0; i < sizeof(h); ++i) zero = zero && reinterpret_cast<const char*>(&h)[i] == 0;
    B200_REQUIRE(!zero, ERR_PEER, "sender's device buffer is not IPC-exportable "
                                  "(allocate it with cudaMalloc / the default torch allocator)");
    void* p = nullptr;
    B200_CHECK_CUDA(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    opened[{d.pid, d.base_ptr}] = {d.buffer_id, p};
    return p;
}

static void pull(const comm::DevDesc& d, const Run* runs, void* buf, int64_t count,
                 const Datatype& type, bool dst_is_device, int64_t* moved) {
    std::vector<Run> sruns(runs, runs + d.nruns);
    const int64_t cap = type.size() * count;
    int dev = 0;
    if (dst_is_device) {
        cudaPointerAttributes a;
        B200_CHECK_CUDA(cudaPointerGetAttributes(&a, buf));
        dev = a.device;
    } else {
        B200_CHECK_CUDA(cudaGetDevice(&dev));
    }
    B200_CHECK_CUDA(cudaSetDevice(dev));
    const char* src = static_cast<const char*>(map_source(d, dev)) + d.offset;
    cudaStream_t st = pull_stream(dev);
    if (dst_is_device) {
        std::vector<Run> druns = type.flatten(count);
        *moved = typed_move(src, sruns, buf, druns, st);
        B200_CHECK_CUDA(cudaStreamSynchronize(st));
    } else {
        // device -> host: pack on the device, one D2H, unpack on the host
        const int64_t n = std::min<int64_t>(d.total_bytes, cap);
        void* stg = staging(dev, (size_t)std::max<int64_t>(d.total_bytes, 16));
        std::vector<Run> packed = {{0, d.total_bytes}};
        typed_move(src, sruns, stg, packed, st);
        std::vector<char> host((size_t)n);
        B200_CHECK_CUDA(cudaMemcpyAsync(host.data(), stg, (size_t)n, cudaMemcpyDeviceToHost, st));
        B200_CHECK_CUDA(cudaStreamSynchronize(st));
        type.unpack(host.data(), n, buf, count);
        *moved = n;
    }
    if (*moved > cap) *moved = cap;
}

static void unpack_to_device(const void* packed, int64_t bytes, void* buf, int64_t count,
                             const Datatype& type) {
    cudaPointerAttributes a;
    B200_CHECK_CUDA(cudaPointerGetAttributes(&a, buf));
    B200_CHECK_CUDA(cudaSetDevice(a.device));
    cudaStream_t st = pull_stream(a.device);
    if (type.contiguous_layout()) {
        B200_CHECK_CUDA(cudaMemcpyAsync(buf, packed, (size_t)bytes, cudaMemcpyHostToDevice, st));
    } else {
        void* stg = staging(a.device, (size_t)std::max<int64_t>(bytes, 16));
        B200_CHECK_CUDA(cudaMemcpyAsync(stg, packed, (size_t)bytes, cudaMemcpyHostToDevice, st));
        std::vector<Run> src = {{0, bytes}};
        typed_move(stg, src, buf, type.flatten(count), st);
    }
    B200_CHECK_CUDA(cudaStreamSynchronize(st));
}

// =====================================================================================
// Eager path: small contiguous device messages through a per-rank device MAILBOX
// =====================================================================================
// Reference capability: the latency end of MPI_Send / MPI_Recv / MPI_Isend / MPI_Irecv on device
// pointers (test-benchmark/mpi-pingpong-gpu.cpp:52-53,75-76, mpi-pingpong-gpu-async.cpp:64-67).
// The rendezvous above costs a descriptor with the whole run list, an IPC lookup, a pull kernel,
// a cudaStreamSynchronize and a FIN record per message.  Here every rank owns a mailbox in device
// memory, published once through the fabric's DevInfo table and mapped by its peers:
//   sender    put kernel: payload -> slot of (me -> dest) in dest's mailbox, st.release.sys of the
//             slot's flag, then a word in pinned host memory ("buffer read, reusable");
//             64-byte record {slot, seq, bytes} through the host ring.
//   receiver  on the match: ONE kernel that acquires the flag, copies slot -> user buffer and
//             writes a token into pinned host memory; the host only polls that word.
//             An ACK record hands the slot back.
// No cudaStreamSynchronize, no IPC work and no allocation per message.  If no slot is free or the
// peer has not published a mailbox yet, the caller falls back to the rendezvous.
namespace eager {

constexpr int kSlots = 4;                         // per (sender, receiver) pair
constexpr uint64_t kSlotBytes = 64 * 1024;        // largest eager message
constexpr int kPutThreads = 512;

// process-wide counters: which path device sends took (b2c_dev_eager_stats)
static std::atomic<unsigned long long> g_n_eager{0}, g_n_too_big{0}, g_n_no_mailbox{0}, g_n_no_slot{0},
    g_n_other_device{0};

// B200COMM_EAGER_TRACE=1: host timestamps (CLOCK_MONOTONIC, comparable across processes) of the
// steps of every eager message, printed to stderr -- where the microseconds of a small message go
static bool trace_on() {
    static const bool on = [] { const char* e = std::getenv("B200COMM_EAGER_TRACE"); return e && e[0] == '1'; }();
    return on;
}
static double now_us() {
    timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec * 1e6 + ts.tv_nsec * 1e-3;
}
#define EAGER_TRACE(...) do { if (trace_on()) std::fprintf(stderr, __VA_ARGS__); } while (0)

struct Mailbox {          // device memory of one rank
    // flags[src][slot] (u64 sequence numbers, zero initialised), then the slots
    static size_t ctrl_bytes(int n) { return ((size_t)n * kSlots * 8 + 4095) & ~(size_t)4095; }
    static size_t bytes(int n) { return ctrl_bytes(n) + (size_t)n * kSlots * kSlotBytes; }
    static size_t flag_off(int src, int slot) { return ((size_t)src * kSlots + slot) * 8; }
    static size_t slot_off(int n, int src, int slot) {
        return ctrl_bytes(n) + ((size_t)src * kSlots + slot) * kSlotBytes;
    }
};

// payload -> peer slot; then publish: flag in the PEER's memory (release), word in MY pinned memory
__global__ void __launch_bounds__(kPutThreads) eager_put_kernel(const unsigned char* __restrict__ src,
                                                                unsigned char* dst, uint64_t bytes,
                                                                uint64_t* flag, uint64_t seq,
                                                                volatile uint64_t* host_sent) {
    if (((((uintptr_t)src) | ((uintptr_t)dst)) & 15) == 0) {
        const uint64_t nv = bytes / 16;
        for (uint64_t i = threadIdx.x; i < nv; i += blockDim.x)
            reinterpret_cast<int4*>(dst)[i] = reinterpret_cast<const int4*>(src)[i];
        for (uint64_t b = nv * 16 + threadIdx.x; b < bytes; b += blockDim.x) dst[b] = src[b];
    } else {
        for (uint64_t b = threadIdx.x; b < bytes; b += blockDim.x) dst[b] = src[b];
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        ptx::fence_acq_rel_sys();
        ptx::st_relaxed_sys(flag, seq);
        ptx::st_release_sys(const_cast<uint64_t*>(host_sent), seq);
    }
}

// acquire the slot's flag, slot -> user buffer, token -> pinned host word
__global__ void __launch_bounds__(kPutThreads) eager_copyout_kernel(const uint64_t* flag, uint64_t seq,
                                                                    const unsigned char* slot,
                                                                    unsigned char* dst, uint64_t bytes,
                                                                    volatile uint64_t* host_token,
                                                                    uint64_t token, uint64_t timeout_ns) {
    if (threadIdx.x == 0 && flag != nullptr) {
        const uint64_t t0 = ptx::globaltimer_ns();
        uint32_t spins = 0;
        while (ptx::ld_acquire_sys(flag) < seq)
            if ((++spins & 255u) == 0 && ptx::globaltimer_ns() - t0 > timeout_ns) break;
    }
    __syncthreads();
    if (dst != nullptr) {
        if (((((uintptr_t)slot) | ((uintptr_t)dst)) & 15) == 0) {
            const uint64_t nv = bytes / 16;
            for (uint64_t i = threadIdx.x; i < nv; i += blockDim.x)
                reinterpret_cast<int4*>(dst)[i] = __ldcv(reinterpret_cast<const int4*>(slot) + i);
            for (uint64_t b = nv * 16 + threadIdx.x; b < bytes; b += blockDim.x)
                dst[b] = *reinterpret_cast<const volatile unsigned char*>(slot + b);
        } else {
            for (uint64_t b = threadIdx.x; b < bytes; b += blockDim.x)
                dst[b] = *reinterpret_cast<const volatile unsigned char*>(slot + b);
        }
    }
    __syncthreads();
    if (threadIdx.x == 0 && host_token != nullptr) {
        ptx::fence_acq_rel_sys();
        ptx::st_release_sys(const_cast<uint64_t*>(host_token), token);
    }
}

struct HostDst {          // receive into host memory / typed device layout: finished on the host side
    uint64_t token;
    const unsigned char* slot;
    void* buf;
    int64_t count;
    Datatype type;
    int64_t bytes;
    bool dst_is_device;
};

struct State {
    int me = -1, nranks = 0, device = -1;
    unsigned char* mailbox = nullptr;            // mine
    unsigned char* peer_box[comm::kMaxRanks] = {};   // peers' mailboxes as mapped here (lazily)
    bool peer_same_gpu[comm::kMaxRanks] = {};
    bool peer_tried[comm::kMaxRanks] = {};
    uint64_t* host_words = nullptr;              // pinned, device visible: sent[nranks], token
    uint64_t next_seq[comm::kMaxRanks] = {};     // per destination
    bool slot_busy[comm::kMaxRanks][kSlots] = {};
    uint64_t slot_seq[comm::kMaxRanks][kSlots] = {};   // last sequence number written to each slot
    uint64_t next_token = 0;
    uint64_t traced_sent[comm::kMaxRanks] = {};
    uint64_t traced_token = 0;
    std::vector<HostDst> host_dsts;
    unsigned char uuid[16] = {};
    cudaStream_t stream = nullptr;
};

static void device_uuid(int device, unsigned char out[16]) {
    cudaDeviceProp pr;
    B200_CHECK_CUDA(cudaGetDeviceProperties(&pr, device));
    std::memcpy(out, &pr.uuid, 16);
}

// publish a slot whose payload was written by EARLIER kernels of the same stream (typed gather)
__global__ void eager_flag_kernel(uint64_t* flag, uint64_t seq, volatile uint64_t* host_sent) {
    ptx::fence_acq_rel_sys();
    ptx::st_relaxed_sys(flag, seq);
    ptx::st_release_sys(const_cast<uint64_t*>(host_sent), seq);
}
__global__ void eager_token_kernel(volatile uint64_t* host_token, uint64_t token) {
    ptx::fence_acq_rel_sys();
    ptx::st_release_sys(const_cast<uint64_t*>(host_token), token);
}

// Flattened layout of (type, count), kept per process: the run list of a derived type is built
// once, the move plan that belongs to it is found by the ADDRESS of this entry.
struct TypedLayout {
    Datatype type;                 // keeps the committed type (and so its identity) alive
    std::vector<Run> runs;         // layout in user memory
    std::vector<Run> packed;       // {0, bytes}
    char send_id = 0, recv_id = 0; // two stable addresses = plan identities (gather / scatter)
};
static TypedLayout& typed_layout(const Datatype& t, int64_t count) {
    static std::map<std::pair<const void*, int64_t>, TypedLayout> cache;
    static std::mutex mu;
    std::lock_guard<std::mutex> lk(mu);
    auto it = cache.find({t.id(), count});
    if (it == cache.end()) {
        TypedLayout L;
        L.type = t;
        L.runs = t.flatten(count);
        L.packed = {{0, t.size() * count}};
        it = cache.emplace(std::make_pair(t.id(), count), std::move(L)).first;
    }
    return it->second;
}

struct DeviceGuard {       // run with `device` current, restore the caller's device afterwards
    int prev = -1;
    explicit DeviceGuard(int device) {
        B200_CHECK_CUDA(cudaGetDevice(&prev));
        if (prev != device) B200_CHECK_CUDA(cudaSetDevice(device));
        else prev = -1;
    }
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

// Create this rank's mailbox on `device` and publish it (first device send / receive).
static State& get_state(void** st, comm::Fabric* f, int me, int device) {
    if (*st) return *static_cast<State*>(*st);
    auto* S = new State();
    S->me = me;
    S->nranks = f->nranks();
    S->device = device;
    DeviceGuard g(device);
    const size_t bytes = Mailbox::bytes(S->nranks);
    B200_CHECK_CUDA(cudaMalloc(&S->mailbox, bytes));
    B200_CHECK_CUDA(cudaMemset(S->mailbox, 0, Mailbox::ctrl_bytes(S->nranks)));
    B200_CHECK_CUDA(cudaHostAlloc(&S->host_words, (size_t)(S->nranks + 1) * 8,
                                  cudaHostAllocMapped | cudaHostAllocPortable));
    std::memset(S->host_words, 0, (size_t)(S->nranks + 1) * 8);
    S->stream = pull_stream(device);
    device_uuid(device, S->uuid);
    B200_CHECK_CUDA(cudaDeviceSynchronize());
    comm::DevInfo* di = f->dev_info(me);
    di->device = device;
    di->pid = (int64_t)getpid();
    di->base = (uint64_t)(uintptr_t)S->mailbox;
    di->bytes = bytes;
    std::memcpy(di->gpu_uuid, S->uuid, 16);
    std::memset(di->ipc_handle, 0, sizeof(di->ipc_handle));
    cudaIpcMemHandle_t h;
    if (cudaIpcGetMemHandle(&h, S->mailbox) == cudaSuccess) std::memcpy(di->ipc_handle, &h, sizeof(h));
    else cudaGetLastError();           // not exportable: only peers inside this process can use it
    di->ready.store(1, std::memory_order_release);
    *st = S;
    return *S;
}

static void prepare(void** st, comm::Fabric* f, int me, int device) { get_state(st, f, me, device); }

// Rank `r`'s mailbox as addressable from my device, or nullptr (not published / not mappable).
static unsigned char* map_peer(State& S, comm::Fabric* f, int r) {
    if (S.peer_box[r]) return S.peer_box[r];
    if (S.peer_tried[r]) return nullptr;
    if (r == S.me) {
        S.peer_box[r] = S.mailbox;
        S.peer_same_gpu[r] = true;
        return S.mailbox;
    }
    comm::DevInfo* di = f->dev_info(r);
    if (di->ready.load(std::memory_order_acquire) != 1) return nullptr;   // ask again next time
    if ((size_t)di->bytes != Mailbox::bytes(S.nranks)) { S.peer_tried[r] = true; return nullptr; }
    S.peer_same_gpu[r] = std::memcmp(di->gpu_uuid, S.uuid, 16) == 0;
    void* p = nullptr;
    if (di->pid == (int64_t)getpid()) {
        if (di->device != S.device) {
            cudaError_t e = cudaDeviceEnablePeerAccess(di->device, 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) {
                cudaGetLastError();
                S.peer_tried[r] = true;
                return nullptr;
            }
            cudaGetLastError();
        }
        p = reinterpret_cast<void*>((uintptr_t)di->base);
    } else {
        cudaIpcMemHandle_t h;
        std::memcpy(&h, di->ipc_handle, sizeof(h));
        bool zero = true;
        for (size_t i = 0; i < sizeof(h); ++i) zero = zero && di->ipc_handle[i] == 0;
        if (zero || cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
            cudaGetLastError();
            S.peer_tried[r] = true;
            return nullptr;
        }
    }
    S.peer_box[r] = static_cast<unsigned char*>(p);
    return S.peer_box[r];
}

static bool send(void** st, comm::Fabric* f, int me, int dest, const void* buf, int device,
                 uint64_t bytes, const Datatype& type, int64_t count, comm::DevEagerMsg* m, bool* gated) {
    static const bool off = [] { const char* e = std::getenv("B200COMM_NO_EAGER"); return e && e[0] == '1'; }();
    if (off || bytes == 0 || bytes > kSlotBytes) { ++g_n_too_big; return false; }
    const double t_in = trace_on() ? now_us() : 0.0;
    State& S = get_state(st, f, me, device);
    if (device != S.device) { ++g_n_other_device; return false; }   // second GPU of a multi-device process
    DeviceGuard g(S.device);
    unsigned char* box = map_peer(S, f, dest);
    if (!box) { ++g_n_no_mailbox; return false; }
    int slot = -1;
    // slots are used round robin so that the receiver's flags only ever grow
    for (int k = 0; k < kSlots; ++k) {
        const int c = (int)((S.next_seq[dest] + k) % kSlots);
        if (!S.slot_busy[dest][c]) { slot = c; break; }
    }
    if (slot < 0) { ++g_n_no_slot; return false; }   // receiver is behind: rendezvous (no buffering limit)
    const uint64_t seq = ++S.next_seq[dest];
    S.slot_busy[dest][slot] = true;
    S.slot_seq[dest][slot] = seq;
    unsigned char* pslot = box + Mailbox::slot_off(S.nranks, me, slot);
    uint64_t* pflag = reinterpret_cast<uint64_t*>(box + Mailbox::flag_off(me, slot));
    if (type.contiguous_layout()) {
        eager_put_kernel<<<1, bytes <= 2048 ? 128 : kPutThreads, 0, S.stream>>>(
            static_cast<const unsigned char*>(buf), pslot, bytes, pflag, seq, S.host_words + dest);
        B200_CHECK_LAUNCH();
    } else {
        // derived type: gather by the sender's layout straight into the peer's slot (the packed
        // side is contiguous, so the NVLink stores are full lines), then publish
        TypedLayout& L = typed_layout(type, count);
        typed_move_cached(buf, L.runs, pslot, L.packed, &L.send_id, S.stream);
        eager_flag_kernel<<<1, 1, 0, S.stream>>>(pflag, seq, S.host_words + dest);
        B200_CHECK_LAUNCH();
    }
    m->slot = (uint32_t)slot;
    m->seq = seq;
    m->bytes = bytes;
    *gated = S.peer_same_gpu[dest];
    m->flags = *gated ? 1u : 0u;
    ++g_n_eager;
    EAGER_TRACE("[eager %d->%d seq %llu] send: enter %.1f launched %.1f\n", me, dest,
                (unsigned long long)seq, t_in, now_us());
    return true;
}

static bool sent(void* st, int dest, uint64_t seq) {
    State& S = *static_cast<State*>(st);
    const bool ok = reinterpret_cast<volatile uint64_t*>(S.host_words)[dest] >= seq;
    if (ok && trace_on() && S.traced_sent[dest] < seq) {
        S.traced_sent[dest] = seq;
        EAGER_TRACE("[eager %d->%d seq %llu] send: put kernel done seen %.1f\n", S.me, dest,
                    (unsigned long long)seq, now_us());
    }
    return ok;
}

static void ack(void* st, int dest, uint32_t slot) {
    State& S = *static_cast<State*>(st);
    if (dest >= 0 && dest < S.nranks && slot < (uint32_t)kSlots) S.slot_busy[dest][slot] = false;
}

static uint64_t recv_start(void** st, comm::Fabric* f, int me, int src, const comm::DevEagerMsg& m,
                           void* buf, int64_t count, const Datatype& type, int dst_device,
                           int64_t* moved) {
    const bool dst_is_device = dst_device >= 0;
    B200_REQUIRE(*st != nullptr, ERR_STATE, "eager device message but this rank has no mailbox");
    (void)f; (void)me;
    State& S = *static_cast<State*>(*st);
    const double t_in = trace_on() ? now_us() : 0.0;
    DeviceGuard g(S.device);
    const int64_t cap = type.size() * count;
    const int64_t n = std::min<int64_t>((int64_t)m.bytes, cap);
    const unsigned char* slot = S.mailbox + Mailbox::slot_off(S.nranks, src, (int)m.slot);
    // message announced after its put kernel had finished (same GPU): nothing to wait for
    const uint64_t* flag = (m.flags & 1u) ? nullptr
                                          : reinterpret_cast<const uint64_t*>(S.mailbox + Mailbox::flag_off(src, (int)m.slot));
    const uint64_t token = ++S.next_token;
    // contiguous destination on my mailbox's GPU: the copy-out kernel writes it directly
    const bool direct = dst_device == S.device && type.contiguous_layout();
    // derived-type destination on my GPU, whole message fits: wait for the flag, scatter by the
    // receiver's layout with the typed move kernel, token -- three launches, still no host sync
    const bool scatter = !direct && dst_device == S.device && (int64_t)m.bytes == cap;
    eager_copyout_kernel<<<1, n <= 2048 ? 128 : kPutThreads, 0, S.stream>>>(
        flag, m.seq, slot, direct ? static_cast<unsigned char*>(buf) : nullptr, (uint64_t)n,
        scatter ? nullptr : S.host_words + S.nranks, token, 20ull * 1000 * 1000 * 1000);
    B200_CHECK_LAUNCH();
    if (scatter) {
        TypedLayout& L = typed_layout(type, count);
        typed_move_cached(slot, L.packed, buf, L.runs, &L.recv_id, S.stream);
        eager_token_kernel<<<1, 1, 0, S.stream>>>(S.host_words + S.nranks, token);
        B200_CHECK_LAUNCH();
    } else if (!direct) {
        S.host_dsts.push_back(HostDst{token, slot, buf, count, type, n, dst_is_device});
    }
    EAGER_TRACE("[eager %d->%d seq %llu] recv: matched %.1f copy-out launched %.1f (token %llu)\n", src, S.me,
                (unsigned long long)m.seq, t_in, now_us(), (unsigned long long)token);
    *moved = n;
    return token;
}

static bool recv_done(void* st, uint64_t token) {
    State& S = *static_cast<State*>(st);
    if (reinterpret_cast<volatile uint64_t*>(S.host_words)[S.nranks] < token) return false;
    if (trace_on() && S.traced_token < token) {
        S.traced_token = token;
        EAGER_TRACE("[eager ->%d] recv: token %llu seen %.1f\n", S.me, (unsigned long long)token, now_us());
    }
    for (size_t i = 0; i < S.host_dsts.size(); ++i) {
        if (S.host_dsts[i].token != token) continue;
        // typed or host destination: the payload has landed in the slot; finish from here
        HostDst d = S.host_dsts[i];
        S.host_dsts.erase(S.host_dsts.begin() + (long)i);
        DeviceGuard g(S.device);
        if (d.dst_is_device) {
            std::vector<Run> srun = {{0, d.bytes}};
            typed_move(d.slot, srun, d.buf, d.type.flatten(d.count), S.stream);
            B200_CHECK_CUDA(cudaStreamSynchronize(S.stream));
        } else if (d.bytes > 0) {
            std::vector<char> host((size_t)d.bytes);
            B200_CHECK_CUDA(cudaMemcpyAsync(host.data(), d.slot, (size_t)d.bytes, cudaMemcpyDeviceToHost, S.stream));
            B200_CHECK_CUDA(cudaStreamSynchronize(S.stream));
            d.type.unpack(host.data(), d.bytes, d.buf, d.count);
        }
        break;
    }
    return true;
}

static void destroy(void* st) {
    State* S = static_cast<State*>(st);
    if (!S) return;
    int prev = -1;
    if (cudaGetDevice(&prev) == cudaSuccess && cudaSetDevice(S->device) == cudaSuccess) {
        cudaStreamSynchronize(S->stream);
        cudaFree(S->mailbox);
        cudaFreeHost(S->host_words);
        cudaSetDevice(prev);
    }
    cudaGetLastError();
    delete S;
}

void stats(unsigned long long out[5]) {
    out[0] = g_n_eager; out[1] = g_n_too_big; out[2] = g_n_no_mailbox; out[3] = g_n_no_slot;
    out[4] = g_n_other_device;
}

}  // namespace eager

struct Installer {
    Installer() {
        comm::DeviceHooks& h = comm::device_hooks();
        h.classify = classify;
        h.export_handle = export_handle;
        h.pull = pull;
        h.unpack_to_device = unpack_to_device;
        h.eager_prepare = eager::prepare;
        h.eager_send = eager::send;
        h.eager_sent = eager::sent;
        h.eager_ack = eager::ack;
        h.eager_recv_start = eager::recv_start;
        h.eager_recv_done = eager::recv_done;
        h.eager_destroy = eager::destroy;
    }
};
static Installer g_installer;

}  // namespace devp2p
}  // namespace b200

using namespace b200;

extern "C" {

// How the device sends of this process travelled so far: out[0] eager (receiver's device mailbox),
// then the reasons for the rendezvous instead: out[1] larger than a slot (or eager switched off),
// out[2] receiver has not published / could not map a mailbox, out[3] all slots in flight,
// out[4] buffer on another GPU than this rank's mailbox.
B2C_EXPORT int b2c_dev_eager_stats(unsigned long long out[5]) {
    B2C_TRY
    B200_REQUIRE(out != nullptr, ERR_ARG, "null output");
    devp2p::eager::stats(out);
    B2C_CATCH
}

// Typed device-to-device move: `scount` elements of `stype` at `src` into
// `dcount` elements of `dtype` at `dst` (either pointer may be peer mapped),
// packing and unpacking inside the moving kernel.  *moved = payload bytes.
B2C_EXPORT int b2c_typed_move(const void* src, b2c_type* stype, int64_t scount, void* dst,
                              b2c_type* dtype, int64_t dcount, void* stream, int64_t* moved) {
    B2C_TRY
    B200_REQUIRE(src && dst && stype && dtype, ERR_ARG, "null argument");
    B200_REQUIRE(stype->t.committed() && dtype->t.committed(), ERR_TYPE, "datatype not committed");
    // Flattening + intersecting the run lists is host work proportional to the
    // number of runs (4096 for one ghost column of a 4096^2 tile): do it once per
    // (types, counts, device) and keep the committed types alive with the entry.
    struct Key {
        const void* s; const void* d; int64_t sc, dc; int dev;
        bool operator<(const Key& o) const {
            return std::tie(s, d, sc, dc, dev) < std::tie(o.s, o.d, o.sc, o.dc, o.dev);
        }
    };
    struct Entry { Datatype st, dt; std::vector<Run> sruns, druns; };
    static std::map<Key, Entry> cache;
    static std::mutex mu;
    int dev = 0;
    B200_CHECK_CUDA(cudaGetDevice(&dev));
    const Entry* e;
    {
        std::lock_guard<std::mutex> lk(mu);
        Key k{stype->t.id(), dtype->t.id(), scount, dcount, dev};
        auto it = cache.find(k);
        if (it == cache.end())
            it = cache.emplace(k, Entry{stype->t, dtype->t, stype->t.flatten(scount),
                                        dtype->t.flatten(dcount)}).first;
        e = &it->second;
    }
    const int64_t n = devp2p::typed_move_cached(src, e->sruns, dst, e->druns, e, (cudaStream_t)stream);
    if (moved) *moved = n;
    B2C_CATCH
}

}  // extern "C"
