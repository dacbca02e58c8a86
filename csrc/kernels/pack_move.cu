// Datatype-driven move kernels: gather by the sender's layout and scatter by
// the receiver's layout in ONE pass, source and destination may both be
// peer-mapped -- "strided subarray and struct datatypes are packed and
// unpacked inside the same kernel that moves them".
//
// Reference capability: the reference hands device pointers + derived
// datatypes to a CUDA-aware MPI (stencil2d/stencil2D.h:369-373) and the MPI
// library packs on the GPU, moves, and unpacks with separate internal steps.
// Here the two flattened run lists are intersected on the host once
// (merge_runs) and a kernel copies segment by segment, or -- for regular
// layouts such as 2-D subarrays -- computes both addresses in closed form.
#include <cuda_runtime.h>
#include <stdint.h>

#include <algorithm>

#include "b200comm.h"
#include "b200comm/datatype.hpp"
#include "b200comm/error.hpp"
#include "core/capi_util.hpp"
#include "kernels/ptx.cuh"

namespace b200 {
namespace packmove {

// ---- closed form: n runs of `len` bytes, constant stride on both sides ------
struct UniformParams {
    const char* src;
    char* dst;
    int64_t nruns;
    int64_t len;          // bytes per run
    int64_t sstride, dstride;
};

template <typename V>
__global__ void uniform_move_kernel(const __grid_constant__ UniformParams P) {
    const int64_t wpr = P.len / (int64_t)sizeof(V);   // words per run
    const int64_t total = P.nruns * wpr;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = i / wpr, w = i - r * wpr;
        const V v = *reinterpret_cast<const V*>(P.src + r * P.sstride + w * (int64_t)sizeof(V));
        *reinterpret_cast<V*>(P.dst + r * P.dstride + w * (int64_t)sizeof(V)) = v;
    }
}

// ---- general: table of (src_off, dst_off, len) segments ----------------------
template <typename V>
__device__ __forceinline__ void copy_words(char* d, const char* s, int64_t n, int tid, int nthr) {
    const int64_t nw = n / (int64_t)sizeof(V);
    for (int64_t i = tid; i < nw; i += nthr)
        reinterpret_cast<V*>(d)[i] = reinterpret_cast<const V*>(s)[i];
}

__device__ __forceinline__ void copy_any(char* d, const char* s, int64_t n, int tid, int nthr) {
    const uintptr_t a = (uintptr_t)d | (uintptr_t)s | (uintptr_t)n;
    if ((a & 15) == 0) copy_words<int4>(d, s, n, tid, nthr);
    else if ((a & 7) == 0) copy_words<int2>(d, s, n, tid, nthr);
    else if ((a & 3) == 0) copy_words<int>(d, s, n, tid, nthr);
    else copy_words<char>(d, s, n, tid, nthr);
}

// Large segments: one CTA per (segment, 64 KiB chunk) work item.
__global__ void segment_move_cta_kernel(const char* src, char* dst, const Segment* segs,
                                        int64_t nseg, int64_t chunk, const int64_t* item_start) {
    for (int64_t item = blockIdx.x;; item += gridDim.x) {
        // work items are enumerated segment-major: item_start[s] = first item of segment s
        // (prefix sums built with the plan, item_start[nseg] = number of items) -> binary search.
        // Callers that bring their own table without prefix sums get the linear walk.
        int64_t k = item, sidx = 0;
        if (item_start != nullptr) {
            if (item >= item_start[nseg]) return;
            int64_t lo = 0, hi = nseg;            // last s with item_start[s] <= item
            while (hi - lo > 1) {
                const int64_t mid = (lo + hi) >> 1;
                if (item_start[mid] <= item) lo = mid; else hi = mid;
            }
            sidx = lo;
            k = item - item_start[lo];
        } else {
            for (; sidx < nseg; ++sidx) {
                const int64_t c = (segs[sidx].len + chunk - 1) / chunk;
                if (k < c) break;
                k -= c;
            }
        }
        if (sidx >= nseg) return;
        const Segment sg = segs[sidx];
        const int64_t off = k * chunk, n = min(chunk, sg.len - off);
        copy_any(dst + sg.dst_off + off, src + sg.src_off + off, n, threadIdx.x, blockDim.x);
    }
}

// Small segments: one thread per segment (consecutive threads -> consecutive
// segments, so packed destinations stay coalesced).
__global__ void segment_move_thread_kernel(const char* src, char* dst, const Segment* segs,
                                           int64_t nseg) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nseg;
         i += (int64_t)gridDim.x * blockDim.x) {
        const Segment sg = segs[i];
        copy_any(dst + sg.dst_off, src + sg.src_off, sg.len, 0, 1);
    }
}

static int widest(int64_t a) { return (a & 15) == 0 ? 16 : (a & 7) == 0 ? 8 : (a & 3) == 0 ? 4 : 1; }

// Launch a uniform move.  Returns false if the pattern is not uniform-compatible.
void launch_uniform(const void* src, void* dst, int64_t nruns, int64_t len, int64_t sstride,
                    int64_t dstride, cudaStream_t st) {
    UniformParams P{static_cast<const char*>(src), static_cast<char*>(dst), nruns, len, sstride, dstride};
    const int w = std::min({widest(len), widest(sstride), widest(dstride),
                            widest((int64_t)(uintptr_t)src), widest((int64_t)(uintptr_t)dst)});
    const int64_t total = nruns * (len / w);
    const int threads = 256;
    const int grid = (int)std::max<int64_t>(1, std::min<int64_t>((total + threads - 1) / threads, 148 * 8));
    if (w == 16) uniform_move_kernel<int4><<<grid, threads, 0, st>>>(P);
    else if (w == 8) uniform_move_kernel<int2><<<grid, threads, 0, st>>>(P);
    else if (w == 4) uniform_move_kernel<int><<<grid, threads, 0, st>>>(P);
    else uniform_move_kernel<char><<<grid, threads, 0, st>>>(P);
    B200_CHECK_LAUNCH();
}

void launch_segments(const void* src, void* dst, const Segment* dev_segs, int64_t nseg,
                     int64_t max_len, int64_t total_items_64k, cudaStream_t st,
                     const int64_t* dev_item_start) {
    if (max_len <= 256) {
        const int threads = 256;
        const int grid = (int)std::max<int64_t>(1, std::min<int64_t>((nseg + threads - 1) / threads, 148 * 8));
        segment_move_thread_kernel<<<grid, threads, 0, st>>>(static_cast<const char*>(src),
                                                             static_cast<char*>(dst), dev_segs, nseg);
    } else {
        const int grid = (int)std::max<int64_t>(1, std::min<int64_t>(total_items_64k, 148 * 4));
        segment_move_cta_kernel<<<grid, 256, 0, st>>>(static_cast<const char*>(src),
                                                      static_cast<char*>(dst), dev_segs, nseg, 65536,
                                                      dev_item_start);
    }
    B200_CHECK_LAUNCH();
}

}  // namespace packmove
}  // namespace b200

using namespace b200;

extern "C" {

// Move `nruns` runs of `len` bytes: src + r*sstride -> dst + r*dstride (either
// side may be peer-mapped).  The closed-form path of 2-D subarray transfers.
B2C_EXPORT int b2c_move_uniform(const void* src, void* dst, int64_t nruns, int64_t len,
                                int64_t sstride, int64_t dstride, void* stream) {
    B2C_TRY
    B200_REQUIRE(src && dst && nruns >= 0 && len >= 0, ERR_ARG, "bad argument");
    if (nruns && len) packmove::launch_uniform(src, dst, nruns, len, sstride, dstride, (cudaStream_t)stream);
    B2C_CATCH
}

// Move by an explicit segment table that already lives in device memory:
// triples (src_off, dst_off, len) of int64.
B2C_EXPORT int b2c_move_segments(const void* src, void* dst, const void* dev_segments, int64_t nseg,
                                 int64_t max_len, int64_t items_64k, void* stream) {
    B2C_TRY
    B200_REQUIRE(src && dst && dev_segments && nseg >= 0, ERR_ARG, "bad argument");
    if (nseg)
        packmove::launch_segments(src, dst, static_cast<const Segment*>(dev_segments), nseg, max_len,
                                  items_64k, (cudaStream_t)stream, nullptr);
    B2C_CATCH
}

}  // extern "C"
