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

namespace b200 {
namespace packmove {
void launch_uniform(const void* src, void* dst, int64_t nruns, int64_t len, int64_t sstride,
                    int64_t dstride, cudaStream_t st);
void launch_segments(const void* src, void* dst, const Segment* dev_segs, int64_t nseg,
                     int64_t max_len, int64_t total_items_64k, cudaStream_t st);
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

static void export_handle(comm::DevDesc* d) {
    d->pid = (int64_t)getpid();
    std::memset(d->ipc_handle, 0, sizeof(d->ipc_handle));
    static std::map<uint64_t, cudaIpcMemHandle_t> cache;
    std::lock_guard<std::mutex> lk(g_mu);
    auto it = cache.find(d->base_ptr);
    if (it == cache.end()) {
        cudaIpcMemHandle_t h;
        if (cudaIpcGetMemHandle(&h, reinterpret_cast<void*>((uintptr_t)d->base_ptr)) != cudaSuccess) {
            cudaGetLastError();   // not IPC-capable: only same-process receivers can pull
            return;
        }
        it = cache.emplace(d->base_ptr, h).first;
    }
    std::memcpy(d->ipc_handle, &it->second, sizeof(cudaIpcMemHandle_t));
}

// ---- typed move with a small plan cache -----------------------------------------
struct MovePlan {
    bool uniform = false;
    int64_t nruns = 0, len = 0, sstride = 0, dstride = 0, s0 = 0, d0 = 0;
    Segment* dev_segs = nullptr;
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
    static std::map<std::pair<uint64_t, int>, MovePlan> cache;
    const uint64_t key = hash_runs(d, hash_runs(s, 1469598103934665603ull) ^ 0x9e3779b97f4a7c15ull);
    std::lock_guard<std::mutex> lk(g_mu);
    auto it = cache.find({key, device});
    if (it != cache.end()) return it->second;
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
        }
    }
    return cache.emplace(std::make_pair(key, device), p).first->second;
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
        packmove::launch_segments(src, dst, p.dev_segs, p.nseg, p.max_len, p.items, st);
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
    static std::map<std::pair<int64_t, uint64_t>, void*> opened;   // (pid, base) -> mapping
    std::lock_guard<std::mutex> lk(g_mu);
    auto it = opened.find({d.pid, d.base_ptr});
    if (it != opened.end()) return it->second;
    cudaIpcMemHandle_t h;
    std::memcpy(&h, d.ipc_handle, sizeof(h));
    bool zero = true;
    for (size_t i = 0; i < sizeof(h); ++i) zero = zero && reinterpret_cast<const char*>(&h)[i] == 0;
    B200_REQUIRE(!zero, ERR_PEER, "sender's device buffer is not IPC-exportable "
                                  "(allocate it with cudaMalloc / the default torch allocator)");
    void* p = nullptr;
    B200_CHECK_CUDA(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    opened[{d.pid, d.base_ptr}] = p;
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

struct Installer {
    Installer() {
        comm::DeviceHooks& h = comm::device_hooks();
        h.classify = classify;
        h.export_handle = export_handle;
        h.pull = pull;
        h.unpack_to_device = unpack_to_device;
    }
};
static Installer g_installer;

}  // namespace devp2p
}  // namespace b200

using namespace b200;

extern "C" {

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
