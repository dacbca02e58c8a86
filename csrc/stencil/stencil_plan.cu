// Host side of the fused stencil: per-rank plan, layout, tensor maps, launches,
// plus the small unfused halo kernels (initial exchange, drain).
//
// Reference counterpart (capability only): the driver program
// stencil2d/mpi-2d-stencil-subarray-cuda.cu:126-172 (allocate tile + ghosts,
// build 16 TransferInfo, loop ExchangeData/Compute/TerminateCondition) and the
// plan builder stencil2d/stencil2D.h:381-437.  Here the "plan" is a table of 8
// peer base pointers + 8 epoch counters living in a symmetric allocation, and
// the loop body is one kernel.
#include <cuda.h>
#include <cuda_runtime.h>

#include <atomic>
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>

#include "b200comm.h"
#include "b200comm/error.hpp"
#include "core/capi_util.hpp"
#include "core/trace.hpp"
#include "stencil/stencil_kernel.cuh"
#include "stencil/stencil_strip_kernel.cuh"
#include "stencil/stencil_tstrip_kernel.cuh"

#include <algorithm>
#include <vector>

namespace b200 {
namespace stencil {

// ---- control pad layout (first kPadBytes of every rank's symmetric block) ----
constexpr size_t kPadBytes = 4096;
constexpr size_t kOffStepFlags = 0;     // 8 x u64: fused-step epoch counters
constexpr size_t kOffXchgFlags = 64;    // 8 x u64: unfused-exchange epoch counters
constexpr size_t kOffStatus = 128;      // u32 status word
constexpr size_t kOffStepCtr = 192;     // u64: fused steps completed (device-side mirror)
constexpr size_t kOffResid = 256;       // 3 doubles: last local residual (sum of squares) of a plain
                                        // step, then the two slots of the pipelined step_io path
constexpr size_t kOffRedSlots = 1024;   // scalar allreduce slots (see collectives)

constexpr int kLead = 16;  // core starts 128 B into each padded row

struct Layout {
    int W, H, R, g, lead, pitch, rows;
    size_t tile_bytes;   // one padded buffer
    size_t off_x[2];     // offsets of the two buffers inside the symmetric block
    size_t land_bytes;   // one column landing buffer
    size_t off_land[2][2];  // [buffer parity][0 west ghost | 1 east ghost]
    size_t total_bytes;
};

static Layout make_layout(int W, int H, int R) {
    Layout L{};
    L.W = W; L.H = H; L.R = R; L.g = R; L.lead = kLead;
    const int RA = (R + 1) & ~1;
    const int need = kLead + W + (R > RA ? R : RA);
    L.pitch = (need + 15) & ~15;
    L.rows = H + 2 * R;
    L.tile_bytes = ((size_t)L.pitch * L.rows * sizeof(double) + 1023) & ~(size_t)1023;
    L.off_x[0] = kPadBytes;
    L.off_x[1] = kPadBytes + L.tile_bytes;
    L.land_bytes = ((size_t)(L.rows + kLandLead) * R * sizeof(double) + 1023) & ~(size_t)1023;
    size_t off = kPadBytes + 2 * L.tile_bytes;
    for (int b = 0; b < 2; ++b)
        for (int sd = 0; sd < 2; ++sd) { L.off_land[b][sd] = off; off += L.land_bytes; }
    L.total_bytes = off;
    return L;
}

// ------------------------------ small kernels ------------------------------

struct XchgParams {
    const double* src;        // my current buffer
    double* peer[8];          // neighbours' current buffers (row directions N, S)
    double* peer_land[8];     // neighbours' landing buffers (directions with a column part)
    uint64_t* peer_flag[8];   // neighbours' xchg counters (slot opposite(d))
    int W, H, g, lead, pitch;
};

// Unfused halo push of the current buffer: region d of my core -> ghost region
// opposite(d) of neighbour d.  One CTA per direction, so the CTA itself is the
// last writer and can publish.  Edges are at most W*g or H*g elements (128 KiB
// for 16384 x 1): a set-up path, not worth a grid-wide arrival counter.
__global__ void halo_push1_kernel(const __grid_constant__ XchgParams P) {
    const int d = blockIdx.x;
    double* pb = P.peer[d];
    if (pb == nullptr) return;
    const int dy = dir_dy(d), dx = dir_dx(d);
    const int rows = dy == 0 ? P.H : P.g, cols = dx == 0 ? P.W : P.g;
    const int y0 = dy > 0 ? P.H - P.g : 0, x0 = dx > 0 ? P.W - P.g : 0;
    const long n = (long)rows * cols;
    for (long i = threadIdx.x; i < n; i += blockDim.x) {
        const int y = y0 + (int)(i / cols), x = x0 + (int)(i % cols);
        const double v = P.src[(size_t)(y + P.g) * P.pitch + P.lead + x];
        if (dx == 0)
            pb[(size_t)(y + P.g - dy * P.H) * P.pitch + P.lead + x] = v;
        else   // columns land contiguously: [padded row][k]
            P.peer_land[d][land_index(y + P.g - dy * P.H, P.g, P.g) + (x - x0)] = v;
    }
    __syncthreads();
    if (threadIdx.x == 0) ptx::red_add_release_sys(P.peer_flag[d], 1ull);
}

struct UnpackParams {
    double* grid;             // my current buffer
    const double* land[2];    // my west / east landing buffers of the same parity
    int active[8];            // which directions delivered data
    int W, H, g, lead, pitch;
};

// Landing buffers -> in-grid ghost columns (local, after a drain/exchange wait),
// so that the tile a user sees always carries complete ghost cells.
__global__ void land_unpack_kernel(const __grid_constant__ UnpackParams P) {
    const int rows = P.H + 2 * P.g;
    const int n = 2 * rows * P.g;
    for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < n; e += gridDim.x * blockDim.x) {
        const int side = e / (rows * P.g), rem = e % (rows * P.g);
        const int pr = rem / P.g, k = rem % P.g;
        const int vd = pr < P.g ? (side ? NE : NW)
                                : (pr >= P.H + P.g ? (side ? SE : SW) : (side ? E_ : W_));
        if (!P.active[vd]) continue;
        const int col = side ? P.lead + P.W + k : P.lead - P.g + k;
        P.grid[(size_t)pr * P.pitch + col] = P.land[side][land_index(pr, P.g, P.g) + k];
    }
}

struct WaitParams {
    const uint64_t* flags;     // 8 counters
    uint64_t epoch[8];         // required values (0: do not wait); multiplied by *step_ctr if set
    const uint64_t* step_ctr;  // device step counter or nullptr
    unsigned int* status;
    uint64_t timeout_ns;
};

// Device-side replacement of MPI_Waitall (stencil2d/stencil2D.h:375-376): one
// thread per direction acquires its counter.
__global__ void flag_wait_kernel(const __grid_constant__ WaitParams P) {
    const int d = threadIdx.x;
    if (d >= 8 || P.epoch[d] == 0) return;
    const uint64_t need = P.step_ctr ? (*P.step_ctr) * P.epoch[d] : P.epoch[d];
    const uint64_t t0 = ptx::globaltimer_ns();
    uint32_t spins = 0;
    while (ptx::ld_acquire_sys(P.flags + d) < need) {
        if ((++spins & 255u) == 0 && ptx::globaltimer_ns() - t0 > P.timeout_ns) {
            atomicOr(P.status, 1u);
            break;
        }
    }
}

__global__ void step_bump_kernel(uint64_t* ctr, uint64_t n) { *ctr += n; }

__global__ void sum_partials_kernel(const double* __restrict__ partials, int n, double* out) {
    __shared__ double sh[256];
    double s = 0.0;
    for (int i = threadIdx.x; i < n; i += blockDim.x) s += partials[i];
    sh[threadIdx.x] = s;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if ((int)threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0) *out = sh[0];
}

// Fill a rectangle of a padded tile (tiled launch; the reference's InitKernel
// uses one thread per block, stencil2d/mpi-2d-stencil-subarray-cuda.cu:17-28).
__global__ void fill_rect_kernel(double* base, int pitch, int x0, int y0, int w, int h, double v) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = blockIdx.y * blockDim.y + threadIdx.y;
    if (x < w && y < h) base[(size_t)(y0 + y) * pitch + x0 + x] = v;
}

// ------------------------------ plan ------------------------------

using EncodeTiledFn = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                   const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                   const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                   CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_tiled() {
    static EncodeTiledFn fn = nullptr;
    if (fn) return fn;
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    B200_CHECK_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q));
    B200_REQUIRE(p != nullptr && q == cudaDriverEntryPointSuccess, ERR_CUDA,
                 "cuTensorMapEncodeTiled not available from the driver");
    fn = reinterpret_cast<EncodeTiledFn>(p);
    return fn;
}

// Pipelined stepping (b2c_stencil_step_io): the coefficients of a sweep arrive from pinned host
// memory and its residual leaves to pinned host memory on two copy streams of the plan, so that
// the launching stream carries nothing but the sweep kernels.  Everything is double buffered:
// slot i & 1 belongs to call i, and call i + 2 waits (events) for call i to have let go of it.
struct IoState {
    cudaStream_t h2d = nullptr, d2h = nullptr;
    cudaEvent_t coef_ready[2] = {nullptr, nullptr};    // H2D of a coefficient slot has completed
    cudaEvent_t coef_free[2] = {nullptr, nullptr};     // every kernel that read the slot has finished
    cudaEvent_t sweep_done[2] = {nullptr, nullptr};    // residual slot written
    cudaEvent_t resid_copied[2] = {nullptr, nullptr};  // residual slot in host memory
    double* coef_stage = nullptr;    // pinned, 2 x 128
    double* resid_stage = nullptr;   // pinned, 2
    unsigned long long seq = 0;      // calls so far = next ticket
};

struct Plan {
    Layout L;
    int kind;          // 0 star, 1 box, 2 radius-1 star applied twice per sweep (L.R == 2)
    int push_corners;
    int device;
    unsigned char* base;         // my symmetric block
    unsigned char* nbr[8];       // neighbours' symmetric blocks (peer mapped) or nullptr
    CUtensorMap tmap[2];
    double* coef_dev;            // 2 x 128 doubles: two coefficient slots (see IoState)
    double* coef_cur;            // the slot the kernels read
    int ncoef;
    double* resid_partials;      // one partial per CTA, then the tail ticket of the temporal strip kernel
    unsigned* tail_ticket;
    struct IoState* io;          // pipelined step_io path, created on first use
    uint64_t steps_done;         // fused steps completed since creation
    uint64_t xchg_done;          // unfused exchanges completed
    int cur;                     // index of the buffer holding the current state
    int grid, nstages, ctas_per_sm;
    size_t smem;
    uint64_t timeout_ns;
    int tilesX, tilesY;
    int tile_h;                  // output rows per kernel tile (kTileH; 64 for 3-4 levels per sweep)
    int partial_levels;          // kind 2: levels of the NEXT sweep if fewer than R (0 = all)
    int debug;                   // B2C_STENCIL_DEBUG, read ONCE when the plan is created
    uint64_t contrib[8];         // signals per step a neighbour sends me per direction
    unsigned long long launches; // kernels launched through this plan (for bench accounting)
    // ---- strip kernel (algo 1) ----
    int algo;                    // 0 tile kernel, 1 strip kernel
    CUtensorMap strip_tmap[2];
    Segment* d_segs;
    int* d_seg_begin;
    int strip_grid, strip_slots;
    int strip_w;                 // cells per strip of this plan (even, <= kStripW)
    size_t strip_smem;
    // ---- temporal strip kernel (algo 2, kind 2 only): shares d_segs / d_seg_begin ----
    int ts_grid, ts_slots, ts_strip_w;
    size_t ts_smem;
    unsigned long long* ts_stamps;   // debug bit 32: [ts_grid][4] phase stamps of the last sweep
    double* tail_resid_out;          // per launch: where the folded tail puts the residual
    unsigned long long tail_bump;    // per launch: step-counter increment done by the last CTA
};

// Geometry of the temporal strip kernel: 6 columns per thread, 8 consumer warps.
constexpr int kTsCpt = 6, kTsWarps = 8;
using TsG = TsGeo<kTsCpt, kTsWarps>;

// Strip width such that (interior strips) x (row bands) uses as many of the G CTAs as
// possible: every CTA then owns one long segment and all CTAs of a band advance together.
static int choose_strip_width(int W, int H, int R, int G) {
    const int wmax = std::min(kStripW, W);
    int best_w = wmax & ~1;
    long best_score = -1;
    const int CY = (H + kChunkH - 1) / kChunkH;
    // Score = busy CTAs x strip width: narrow strips waste lanes, TMA efficiency and
    // sectors (68-cell strips on a 4096-wide tile ran at half the speed of 200-cell ones).
    // Evaluate every even width exactly the way build_strip_partition will use it.
    for (int sw = 32; sw <= (wmax & ~1); sw += 2) {
        const int NS = (W + sw - 1) / sw;
        const int last = W - (NS - 1) * sw;
        if (NS < 3 || (last != sw && last < R)) continue;
        const int NSi = NS - 2;
        const int NB = G / NSi;                       // row bands the partition will make
        if (NB < 1 || CY - 2 < 8 * NB) continue;
        // busy CTAs x per-CTA efficiency; the +40 models the fixed cost per chunk row
        // (halo columns, sector rounding, barrier traffic) in units of cells
        const long score = (long)NSi * NB * 1000L * sw / (sw + 40);
        if (score >= best_score) { best_score = score; best_w = sw; }
    }
    return best_w;
}

constexpr int kStripSlots = 5;   // default ring depth (B2C_STRIP_SLOTS = 3..6 | 8 | 10 | 12); deeper is SLOWER, see profiles/

// Work partition of the strip kernel: every CTA gets its boundary segments first
// (pieces of the west/east strips, single top/bottom chunks of the interior strips),
// then one contiguous run of interior chunks.  Also counts, per direction, how many
// segments signal it -- the neighbour (same geometry) sends me exactly as many.
static void build_strip_partition(Plan& p, int strip_w, int G, bool force_run,
                                  std::vector<Segment>& segs, std::vector<int>& begin) {
    const int W = p.L.W, H = p.L.H;
    const int NS = (W + strip_w - 1) / strip_w, CY = (H + kChunkH - 1) / kChunkH;
    std::vector<std::vector<Segment>> per((size_t)G);
    std::vector<long> load((size_t)G, 0);
    auto act = [&](int d) {
        const bool corner = dir_dy(d) != 0 && dir_dx(d) != 0;
        return p.nbr[d] != nullptr && (!corner || p.push_corners);
    };
    // Geometric edge mask of a segment; what it actually waits for / pushes to is the
    // subset with a neighbour.  The per-direction signal COUNT must be geometric: a
    // neighbour with the same tile geometry sends me that many signals even when my own
    // opposite side has no neighbour (non-periodic grids).
    for (int d = 0; d < 8; ++d) p.contrib[d] = 0;
    unsigned actmask = 0;
    for (int d = 0; d < 8; ++d) if (act(d)) actmask |= 1u << d;
    unsigned last_pin = 0;     // set by dirs_for: edges of that segment on the physical boundary
    auto dirs_for = [&](int strip, int j0, int j1) {
        const bool w = strip == 0, e = strip == NS - 1, n = j0 == 0, s = j1 == CY;
        unsigned m = 0;
        auto add = [&](bool c, int d) { if (c) m |= 1u << d; };
        add(n, N); add(s, S); add(w, W_); add(e, E_);
        add(n && w, NW); add(n && e, NE); add(s && w, SW); add(s && e, SE);
        for (int d = 0; d < 8; ++d) if (m >> d & 1u) p.contrib[d] += 1;
        last_pin = (n && !act(N) ? 1u : 0u) | (s && !act(S) ? 2u : 0u) | (w && !act(W_) ? 4u : 0u) |
                   (e && !act(E_) ? 8u : 0u);
        return m & actmask;
    };
    int rr = 0;   // round-robin cursor over CTAs
    auto give = [&](int strip, int j0, int j1) {
        const unsigned dm = dirs_for(strip, j0, j1);
        per[(size_t)rr].push_back(Segment{strip, j0, j1, dm, last_pin});
        load[(size_t)rr] += (j1 - j0) + 2;
        rr = (rr + 1) % G;
    };
    // west / east strips, cut into pieces of >= 4 chunks
    const int npieces = std::max(1, std::min(G, CY / 4));
    for (int b = 0; b < (NS > 1 ? 2 : 1); ++b) {
        const int strip = b == 0 ? 0 : NS - 1;
        for (int q = 0; q < npieces; ++q) {
            const int j0 = (int)((long)CY * q / npieces), j1 = (int)((long)CY * (q + 1) / npieces);
            if (j1 > j0) give(strip, j0, j1);
        }
    }
    // top / bottom chunk of every interior strip
    for (int strip = 1; strip + 1 < NS; ++strip) {
        give(strip, 0, 1);
        if (CY > 1) give(strip, CY - 1, CY);
    }
    // interior chunks: strips 1..NS-2, chunks 1..CY-2, strip-major, contiguous per CTA
    const long ich = (NS > 2 && CY > 2) ? (long)(NS - 2) * (CY - 2) : 0;
    const char* pe = std::getenv("B2C_STRIP_PART");
    const int NSi = NS - 2, CYi = CY - 2;
    const int NB = NSi > 0 ? G / NSi : 0;        // row bands so that bands x strips <= CTAs
    if (ich > 0 && NB >= 1 && CYi >= 8 * NB && !force_run && !(pe && std::string(pe) == "run")) {
        // band-major: CTA (band b, strip s) marches down rows of band b of strip s.  CTAs of
        // one band sit on ADJACENT strips and advance in lock-step, so at any instant DRAM
        // sees whole contiguous rows (like the x-fastest tile order), not scattered pieces.
        for (int b = 0; b < NB; ++b)
            for (int si = 0; si < NSi; ++si) {
                const int cta = b * NSi + si;
                const int j0 = 1 + (int)((long)CYi * b / NB), j1 = 1 + (int)((long)CYi * (b + 1) / NB);
                per[(size_t)cta].push_back(Segment{1 + si, j0, j1, 0u, 0u});
            }
    } else if (ich > 0) {
        long total = ich;
        for (long l : load) total += l;
        const long target = (total + G - 1) / G;
        long next = 0;
        for (int cta = 0; cta < G && next < ich; ++cta) {
            long want = target - load[(size_t)cta];
            if (cta == G - 1) want = ich - next;
            if (want <= 0) continue;
            long endi = std::min(ich, next + want);
            while (next < endi) {   // split at strip boundaries
                const int strip = 1 + (int)(next / (CY - 2));
                const int j0 = 1 + (int)(next % (CY - 2));
                const long room = (CY - 2) - (next % (CY - 2));
                const long n = std::min(room, endi - next);
                per[(size_t)cta].push_back(Segment{strip, j0, j0 + (int)n, 0u, 0u});
                next += n;
            }
        }
        // anything left (rounding) goes to the last CTA
        while (next < ich) {
            const int strip = 1 + (int)(next / (CY - 2));
            const int j0 = 1 + (int)(next % (CY - 2));
            const long n = std::min<long>((CY - 2) - (next % (CY - 2)), ich - next);
            per[(size_t)G - 1].push_back(Segment{strip, j0, j0 + (int)n, 0u, 0u});
            next += n;
        }
    }
    begin.assign(1, 0);
    for (int cta = 0; cta < G; ++cta) {
        for (const Segment& s : per[(size_t)cta]) segs.push_back(s);
        begin.push_back((int)segs.size());
    }
}

template <int R, int SLOTS>
static void launch_strip_s(const Plan& pl, const LaunchParams& lp, bool resid, cudaStream_t st);

template <int R>
static void launch_strip(const Plan& pl, const LaunchParams& lp, bool resid, cudaStream_t st) {
    switch (pl.strip_slots) {
        case 3: launch_strip_s<R, 3>(pl, lp, resid, st); return;
        case 4: launch_strip_s<R, 4>(pl, lp, resid, st); return;
        case 5: launch_strip_s<R, 5>(pl, lp, resid, st); return;
        case 6: launch_strip_s<R, 6>(pl, lp, resid, st); return;
        case 8: launch_strip_s<R, 8>(pl, lp, resid, st); return;
        case 10: launch_strip_s<R, 10>(pl, lp, resid, st); return;
        default: launch_strip_s<R, 12>(pl, lp, resid, st); return;
    }
}

template <int R, int SLOTS>
static void launch_strip_s(const Plan& pl, const LaunchParams& lp, bool resid, cudaStream_t st) {
    StripParams sp{};
    sp.lp = lp;
    sp.segs = pl.d_segs;
    sp.seg_begin = pl.d_seg_begin;
    sp.strip_w = pl.strip_w;
    sp.box_w = pl.strip_w + 4;
    auto k0 = stencil_strip_kernel<R, false, SLOTS>;
    auto k1 = stencil_strip_kernel<R, true, SLOTS>;
    static std::atomic<bool> attr_done[64];   // per device; setting it twice is harmless
    if (!attr_done[pl.device & 63]) {
        B200_CHECK_CUDA(cudaFuncSetAttribute(k0, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.strip_smem));
        B200_CHECK_CUDA(cudaFuncSetAttribute(k1, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.strip_smem));
        attr_done[pl.device & 63] = true;
    }
    if (resid) k1<<<pl.strip_grid, kStripThreads, pl.strip_smem, st>>>(pl.strip_tmap[pl.cur], sp);
    else k0<<<pl.strip_grid, kStripThreads, pl.strip_smem, st>>>(pl.strip_tmap[pl.cur], sp);
    B200_CHECK_LAUNCH();
}

template <int R, bool BOX, bool RESID, int NST, int TB = 0, int TH = kTileH, int LV = TB>
static void launch_one(const Plan& pl, const LaunchParams& lp, cudaStream_t st) {
    auto k = stencil_fused_kernel<R, BOX, RESID, NST, TB, TH, LV>;
    static std::atomic<bool> attr_done[64];   // per device; setting it twice is harmless
    if (!attr_done[pl.device & 63]) {
        B200_CHECK_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             (int)pl.smem));
        attr_done[pl.device & 63] = true;
    }
    k<<<pl.grid, kThreads, pl.smem, st>>>(pl.tmap[pl.cur], lp);
    B200_CHECK_LAUNCH();
}

template <int R, bool BOX>
static void launch_r(const Plan& pl, const LaunchParams& lp, bool resid, cudaStream_t st) {
    constexpr int NST = R <= 2 ? 3 : 4;
    if (R <= 2 && pl.nstages == 2) {   // shallower ring (B2C_TILE_STAGES=2), see DESIGN.md
        if (resid) launch_one<R, BOX, true, 2>(pl, lp, st);
        else launch_one<R, BOX, false, 2>(pl, lp, st);
        return;
    }
    if (resid) launch_one<R, BOX, true, NST>(pl, lp, st);
    else launch_one<R, BOX, false, NST>(pl, lp, st);
}

// kind 2: radius-1 star, R time steps per sweep, on the radius-R geometry
// Partial sweep (LV < T levels): rare (at most once per step() call), so only the 2-stage
// instantiation exists -- every plan allocates at least two stages of shared memory.
template <int T, int LV>
static void launch_tb_partial(const Plan& pl, const LaunchParams& lp, bool resid, cudaStream_t st) {
    if (pl.tile_h == 64) {
        if (resid) launch_one<T, false, true, 2, T, 64, LV>(pl, lp, st);
        else launch_one<T, false, false, 2, T, 64, LV>(pl, lp, st);
    } else {
        if (resid) launch_one<T, false, true, 2, T, kTileH, LV>(pl, lp, st);
        else launch_one<T, false, false, 2, T, kTileH, LV>(pl, lp, st);
    }
}

template <int T>
static void launch_tb(const Plan& pl, const LaunchParams& lp, bool resid, cudaStream_t st) {
    if (pl.partial_levels > 0 && pl.partial_levels < T) {
        if constexpr (T >= 2) if (pl.partial_levels == 1) { launch_tb_partial<T, 1>(pl, lp, resid, st); return; }
        if constexpr (T >= 3) if (pl.partial_levels == 2) { launch_tb_partial<T, 2>(pl, lp, resid, st); return; }
        if constexpr (T >= 4) if (pl.partial_levels == 3) { launch_tb_partial<T, 3>(pl, lp, resid, st); return; }
    }
    if (pl.tile_h == 64) {             // tall tiles: two stages is all that fits
        if (resid) launch_one<T, false, true, 2, T, 64>(pl, lp, st);
        else launch_one<T, false, false, 2, T, 64>(pl, lp, st);
        return;
    }
    switch (pl.nstages) {
        case 2:
            if (resid) launch_one<T, false, true, 2, T>(pl, lp, st);
            else launch_one<T, false, false, 2, T>(pl, lp, st);
            return;
        case 4:
            if (resid) launch_one<T, false, true, 4, T>(pl, lp, st);
            else launch_one<T, false, false, 4, T>(pl, lp, st);
            return;
        default:
            if (resid) launch_one<T, false, true, 3, T>(pl, lp, st);
            else launch_one<T, false, false, 3, T>(pl, lp, st);
    }
}

// kind 2, algo 2: the temporal strip kernel; LV < T is the partial sweep
template <int T, int LV>
static void launch_ts_lv(const Plan& pl, const LaunchParams& lp, bool resid, cudaStream_t st) {
    TsParams sp{};
    sp.lp = lp;
    sp.segs = pl.d_segs;
    sp.seg_begin = pl.d_seg_begin;
    sp.strip_w = pl.ts_strip_w;
    sp.nslots = pl.ts_slots;
    sp.resid = resid ? 1 : 0;
    sp.stamps = pl.ts_stamps;
    sp.tail_ticket = (resid || pl.tail_bump) ? pl.tail_ticket : nullptr;
    sp.resid_out = pl.tail_resid_out;
    sp.ctr_w = reinterpret_cast<unsigned long long*>(pl.base + kOffStepCtr);
    sp.bump = pl.tail_bump;
    if (pl.ts_stamps)      // the producer ACCUMULATES its waits: clear per sweep
        B200_CHECK_CUDA(cudaMemsetAsync(pl.ts_stamps, 0, (size_t)pl.ts_grid * 4 * sizeof(unsigned long long), st));
    auto k = stencil_tstrip_kernel<T, LV, kTsCpt, kTsWarps>;
    static std::atomic<bool> attr_done[64];   // per device; setting it twice is harmless
    if (!attr_done[pl.device & 63]) {
        B200_CHECK_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             (int)ts_smem_bytes<T, kTsCpt, kTsWarps>(6)));
        attr_done[pl.device & 63] = true;
    }
    k<<<pl.ts_grid, TsG::THREADS, pl.ts_smem, st>>>(sp);
    B200_CHECK_LAUNCH();
}

template <int T>
static void launch_ts(const Plan& pl, const LaunchParams& lp, bool resid, cudaStream_t st) {
    const int lv = (pl.partial_levels > 0 && pl.partial_levels < T) ? pl.partial_levels : T;
    if constexpr (T >= 3) if (lv == 2) { launch_ts_lv<T, 2>(pl, lp, resid, st); return; }
    if (lv == 1) { launch_ts_lv<T, 1>(pl, lp, resid, st); return; }
    launch_ts_lv<T, T>(pl, lp, resid, st);
}

static void launch_step(const Plan& pl, const LaunchParams& lp, bool resid, cudaStream_t st) {
    if (pl.kind == 2 && pl.algo == 2) {
        if (pl.L.R == 2) launch_ts<2>(pl, lp, resid, st);
        else launch_ts<3>(pl, lp, resid, st);
        return;
    }
    if (pl.kind == 2) {
        switch (pl.L.R) {
            case 2: launch_tb<2>(pl, lp, resid, st); return;
            case 3: launch_tb<3>(pl, lp, resid, st); return;
            case 4: launch_tb<4>(pl, lp, resid, st); return;
        }
    }
    if (pl.algo == 1) {
        if (pl.L.R == 1) launch_strip<1>(pl, lp, resid, st);
        else launch_strip<2>(pl, lp, resid, st);
        return;
    }
    if (pl.kind == 0) {
        switch (pl.L.R) {
            case 1: launch_r<1, false>(pl, lp, resid, st); return;
            case 2: launch_r<2, false>(pl, lp, resid, st); return;
            case 3: launch_r<3, false>(pl, lp, resid, st); return;
            case 4: launch_r<4, false>(pl, lp, resid, st); return;
        }
    } else {
        switch (pl.L.R) {
            case 1: launch_r<1, true>(pl, lp, resid, st); return;
            case 2: launch_r<2, true>(pl, lp, resid, st); return;
        }
    }
    B200_FAIL(ERR_ARG, "unsupported stencil kind/radius");
}

// Signals per step that arrive on my counter `d` (= what the neighbour in direction d,
// which has the same tile geometry, sends towards me).
static uint64_t contributions(const Plan& pl, int d) { return pl.contrib[dir_opposite(d)]; }

static void tile_contributions(Plan& pl) {
    for (int d = 0; d < 8; ++d) {
        const int dy = dir_dy(d), dx = dir_dx(d);
        pl.contrib[d] = (dy != 0 && dx != 0) ? 1 : (dy != 0 ? (uint64_t)pl.tilesX : (uint64_t)pl.tilesY);
    }
}

static bool dir_active(const Plan& pl, int d) {
    const bool corner = dir_dy(d) != 0 && dir_dx(d) != 0;
    return pl.nbr[d] != nullptr && (!corner || pl.push_corners);
}

}  // namespace stencil
}  // namespace b200

using namespace b200;
using namespace b200::stencil;

struct b2c_stencil { Plan p; };

extern "C" {

B2C_EXPORT int b2c_stencil_layout(int W, int H, int R, b2c_stencil_layout_t* out) {
    B2C_TRY
    B200_REQUIRE(out != nullptr, ERR_ARG, "null output");
    B200_REQUIRE(W >= 2 && H >= 1 && (W % 2) == 0, ERR_ARG, "tile width must be even and >= 2");
    B200_REQUIRE(R >= 1 && R <= kMaxRadius, ERR_ARG, "radius must be in [1,4]");
    B200_REQUIRE(W >= R && H >= R, ERR_ARG, "tile smaller than the stencil radius");
    // an edge strip of width R must not straddle two kernel tiles (128 x 32 cells)
    B200_REQUIRE((W % kTileW == 0 || W % kTileW >= R) && (H % kTileH == 0 || H % kTileH >= R),
                 ERR_ARG, "W mod 128 and H mod 32 must be 0 or >= radius");
    Layout L = make_layout(W, H, R);
    out->W = W; out->H = H; out->radius = R; out->ghost = L.g; out->lead = L.lead;
    out->pitch = L.pitch; out->rows = L.rows; out->tile_bytes = L.tile_bytes;
    out->off_buf0 = L.off_x[0]; out->off_buf1 = L.off_x[1]; out->total_bytes = L.total_bytes;
    out->pad_bytes = kPadBytes; out->off_red_slots = kOffRedSlots;
    B2C_CATCH
}

B2C_EXPORT int b2c_stencil_create(b2c_stencil** out, int W, int H, int R, int kind,
                                  void* local_base, void* const nbr_base[8], int push_corners) {
    B2C_TRY
    B200_REQUIRE(out && local_base, ERR_ARG, "null argument");
    B200_REQUIRE(kind == 0 || (kind == 1 && R <= 2) || (kind == 2 && R >= 2), ERR_ARG,
                 "kind must be 0 (star, R<=4), 1 (box, R<=2) or 2 (radius-1 star x R time steps "
                 "per sweep, R = 2..4 = the ghost width)");
    b2c_stencil_layout_t tmp;
    int rc = b2c_stencil_layout(W, H, R, &tmp);
    if (rc) return rc;
    auto* h = new (std::nothrow) b2c_stencil();
    B200_REQUIRE(h, ERR_NOMEM, "plan allocation failed");
    Plan& p = h->p;
    std::memset(&p, 0, sizeof(p));
    p.L = make_layout(W, H, R);
    p.kind = kind;
    p.push_corners = (kind != 0) ? 1 : (push_corners != 0);
    B200_CHECK_CUDA(cudaGetDevice(&p.device));
    p.base = static_cast<unsigned char*>(local_base);
    for (int d = 0; d < 8; ++d) p.nbr[d] = static_cast<unsigned char*>(nbr_base ? nbr_base[d] : nullptr);
    p.tile_h = kTileH;
    if (kind == 2 && kTileH == 32) {
        // 3 or 4 time levels per sweep are compute bound: 64-row tiles (16 rows per thread) do
        // 7.0 instead of 8.0 cell updates per output cell for 3 levels (846 vs 790 Gcell/s)
        const char* he = std::getenv("B2C_TB_TILE_H");
        p.tile_h = he ? (std::atoi(he) == 64 ? 64 : 32) : (R >= 3 ? 64 : 32);
        if (!(H % p.tile_h == 0 || H % p.tile_h >= R)) p.tile_h = 32;   // edge strip must not straddle tiles
    }
    p.tilesX = (W + kTileW - 1) / kTileW;
    p.tilesY = (H + p.tile_h - 1) / p.tile_h;
    p.timeout_ns = 20ull * 1000 * 1000 * 1000;
    { const char* dbg = std::getenv("B2C_STENCIL_DEBUG"); p.debug = dbg ? std::atoi(dbg) : 0; }

    // shared memory / grid
    const int RA = (R + 1) & ~1;
    const size_t stage = (((size_t)(kTileW + 2 * RA) * (p.tile_h + 2 * R) * 8) + 127) & ~(size_t)127;
    p.nstages = R <= 2 ? 3 : 4;
    p.ctas_per_sm = R <= 2 ? 2 : 1;
    {
        // How much to keep in flight is THE tuning parameter of this kernel on B200
        // (profiles/README.md): with many tiles per SM, one CTA per SM and a 2-stage ring
        // (36-72 KB in flight per SM ~ HBM latency x bandwidth) runs at copy-level DRAM
        // efficiency (16384^2: 370 vs 340 Gcell/s with 2 CTAs x 3 stages); with few tiles per
        // SM the deeper, wider configuration hides the start-up better (4096^2: 310 vs 296).
        cudaDeviceProp pr;
        B200_CHECK_CUDA(cudaGetDeviceProperties(&pr, p.device));
        const long tiles_per_sm = (long)p.tilesX * p.tilesY / std::max(1, pr.multiProcessorCount);
        if (R <= 2 && tiles_per_sm >= 64) { p.nstages = 2; p.ctas_per_sm = 1; }
        // Two time steps per sweep: one CTA per SM (the two-level register window is wide) and
        // a 3-stage ring -- the compute phase per tile is ~2x longer than a single step's, so
        // more has to be in flight to cover it (measured 16384^2: 2/3 stages = 696/783 Gcell/s).
        if (kind == 2) { p.ctas_per_sm = 1; p.nstages = 3; }
    }
    if (const char* te = std::getenv("B2C_TILE_STAGES")) {
        const int v = std::atoi(te);
        if ((R <= 2 && (v == 2 || v == 3)) || (kind == 2 && v >= 2 && v <= 4)) p.nstages = v;
    }
    if (const char* ce = std::getenv("B2C_TILE_CTAS_PER_SM"))
        if (R <= 2 && kind != 2 && (std::atoi(ce) == 1 || std::atoi(ce) == 2)) p.ctas_per_sm = std::atoi(ce);
    if (p.tile_h == 64) p.nstages = 2;
    p.smem = p.nstages * stage + 2 * p.nstages * sizeof(uint64_t);
    cudaDeviceProp prop;
    B200_CHECK_CUDA(cudaGetDeviceProperties(&prop, p.device));
    const int ntiles = p.tilesX * p.tilesY;
    p.grid = prop.multiProcessorCount * p.ctas_per_sm;
    if (p.grid > ntiles) p.grid = ntiles;

    // tensor maps over the two padded buffers
    EncodeTiledFn enc = get_encode_tiled();
    for (int b = 0; b < 2; ++b) {
        cuuint64_t gdim[2] = {(cuuint64_t)p.L.pitch, (cuuint64_t)p.L.rows};
        cuuint64_t gstr[1] = {(cuuint64_t)p.L.pitch * sizeof(double)};
        cuuint32_t box[2] = {(cuuint32_t)(kTileW + 2 * RA), (cuuint32_t)(p.tile_h + 2 * R)};
        cuuint32_t estr[2] = {1, 1};
        const char* pe = std::getenv("B2C_STENCIL_L2PROMO");
        const int promo = pe ? std::atoi(pe) : 128;
        const CUtensorMapL2promotion l2p = promo >= 256 ? CU_TENSOR_MAP_L2_PROMOTION_L2_256B
                                         : promo >= 128 ? CU_TENSOR_MAP_L2_PROMOTION_L2_128B
                                         : promo >= 64  ? CU_TENSOR_MAP_L2_PROMOTION_L2_64B
                                                        : CU_TENSOR_MAP_L2_PROMOTION_NONE;
        CUresult r = enc(&p.tmap[b], CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 2, p.base + p.L.off_x[b], gdim,
                         gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                         l2p, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        B200_REQUIRE(r == CUDA_SUCCESS, ERR_CUDA,
                     "cuTensorMapEncodeTiled failed with CUresult " + std::to_string((int)r));
    }
    tile_contributions(p);
    // ---- strip kernel: star stencils of radius <= 2 whose edge strips do not straddle
    // a strip / chunk boundary.  B2C_STENCIL_ALGO=tile|strip overrides the choice.
    {
        const char* ae = std::getenv("B2C_STENCIL_ALGO");
        const char* ge = std::getenv("B2C_STRIP_CTAS_PER_SM");
        const int per_sm = ge ? std::max(1, std::min(kStripCtasPerSm, std::atoi(ge))) : 2;
        const long chunks_all = (long)((W + kStripW - 1) / kStripW) * ((H + kChunkH - 1) / kChunkH);
        p.strip_grid = (int)std::max<long>(1, std::min<long>(prop.multiProcessorCount * per_sm, (chunks_all + 7) / 8));
        const char* we = std::getenv("B2C_STRIP_WIDTH");
        p.strip_w = we ? (std::max(2, std::min(kStripW, std::atoi(we))) & ~1)
                       : choose_strip_width(W, H, R, p.strip_grid);
        const bool eligible = kind == 0 && R <= 2 &&
                              (W % p.strip_w == 0 || W % p.strip_w >= R) &&
                              (H % kChunkH == 0 || H % kChunkH >= R);
        // Measured (profiles/stencil_strip_sweeps.txt): with its ring depth tuned the tile
        // kernel ties the strip kernel on 16384^2 cells per GPU (369 vs 365-368 Gcell/s on one
        // GPU, 716 vs 716 on two) and wins on every smaller tile (4096^2: 308 vs 240; long
        // sequential segments pay their pipeline fill), so the tile kernel is the default and
        // the strip kernel (6 % less DRAM traffic) is opt-in: B2C_STENCIL_ALGO=strip.
        bool use = false;
        if (ae && std::string(ae) == "strip") use = eligible;
        if (ae && std::string(ae) == "tile") use = false;
        if (ae && std::string(ae) == "strip" && kind == 0)
            B200_REQUIRE(eligible, ERR_ARG, "strip kernel not applicable to this stencil/tile");
        p.algo = use ? 1 : 0;
    }
    if (p.algo == 1) {
        {
            const char* se = std::getenv("B2C_STRIP_SLOTS");
            const int s = se ? std::atoi(se) : kStripSlots;
            p.strip_slots = ((s >= 3 && s <= 6) || s == 8 || s == 10) ? s : 12;
        }
        p.strip_smem = (size_t)p.strip_slots * kChunkBytes + 128 + 2 * p.strip_slots * sizeof(uint64_t);
        std::vector<Segment> segs;
        std::vector<int> begin;
        build_strip_partition(p, p.strip_w, p.strip_grid, false, segs, begin);
        B200_CHECK_CUDA(cudaMalloc(&p.d_segs, std::max<size_t>(1, segs.size()) * sizeof(Segment)));
        B200_CHECK_CUDA(cudaMalloc(&p.d_seg_begin, begin.size() * sizeof(int)));
        B200_CHECK_CUDA(cudaMemcpy(p.d_segs, segs.data(), segs.size() * sizeof(Segment), cudaMemcpyHostToDevice));
        B200_CHECK_CUDA(cudaMemcpy(p.d_seg_begin, begin.data(), begin.size() * sizeof(int), cudaMemcpyHostToDevice));
        for (int b = 0; b < 2; ++b) {
            cuuint64_t gdim[2] = {(cuuint64_t)p.L.pitch, (cuuint64_t)p.L.rows};
            cuuint64_t gstr[1] = {(cuuint64_t)p.L.pitch * sizeof(double)};
            cuuint32_t box[2] = {(cuuint32_t)(p.strip_w + 4), (cuuint32_t)kChunkH};
            cuuint32_t estr[2] = {1, 1};
            CUresult r = enc(&p.strip_tmap[b], CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 2, p.base + p.L.off_x[b],
                             gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                             CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                             CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
            B200_REQUIRE(r == CUDA_SUCCESS, ERR_CUDA,
                         "cuTensorMapEncodeTiled (strip) failed with CUresult " + std::to_string((int)r));
        }
    }
    // ---- temporal strip kernel (kind 2): strips of <= 1536 columns marched top to bottom, the
    // time levels in registers (stencil_tstrip_kernel.cuh).  Default whenever the tile gives every
    // CTA a long enough march; B2C_TB_ALGO=strip|tile overrides.
    if (kind == 2) {
        const char* ae = std::getenv("B2C_TB_ALGO");
        const int NS = (W + TsG::SWMAX - 1) / TsG::SWMAX;
        int sw = (((W + NS - 1) / NS) + kTsCpt - 1) / kTsCpt * kTsCpt;
        if (const char* we = std::getenv("B2C_TS_WIDTH")) {
            const int v = std::atoi(we) / kTsCpt * kTsCpt;
            if (v >= kTsCpt && v <= TsG::SWMAX) sw = v;
        }
        const int NS2 = (W + sw - 1) / sw;
        const int last = W - (NS2 - 1) * sw;
        const int CY = (H + kChunkH - 1) / kChunkH;
        const bool eligible = R <= 3 && H >= 16 && last >= R && (H % kChunkH == 0 || H % kChunkH >= R);
        const long steps_per_cta = (long)NS2 * CY / std::max(1, prop.multiProcessorCount);
        // strong scaling at 8 GPUs (8192 x 4096 tiles, 41 chunk rows per CTA): 6192 Gcell/s with
        // this kernel vs 5160 with the tile kernel (profiles/r2/bench_8gpu_strong*.json)
        bool use = eligible && steps_per_cta >= 16;
        if (ae && std::string(ae) == "strip") use = eligible;   // whenever it applies, however small the tile
        if (ae && std::string(ae) == "tile") use = false;
        // A tile on the edge of a NON-PERIODIC grid has no neighbour there: its ghost cells are
        // boundary data, and the intermediate time levels must stay fixed beyond that edge.  Only
        // the strip kernel implements that (Segment::pin).
        bool boundary = false;
        for (int d : {(int)N, (int)S, (int)W_, (int)E_}) boundary = boundary || p.nbr[d] == nullptr;
        if (boundary) {
            B200_REQUIRE(eligible, ERR_ARG,
                         "temporal blocking on a non-periodic grid needs 2 or 3 time steps per sweep, a tile "
                         "of at least 16 rows and (rows mod 4) == 0 or >= the number of steps");
            use = true;
        }
        if (use) {
            p.algo = 2;
            p.ts_strip_w = sw;
            p.ts_grid = (int)std::max<long>(1, std::min<long>(prop.multiProcessorCount, (long)NS2 * CY / 4));
            // 2-row slots of 24.1 KiB.  Like the other two kernels this one is FASTER with a shallow
            // ring (profiles/r2/tstrip_knobs.txt: 2/3/4/5/6 slots = 940/1063/1061/1034/1029 Gcell/s):
            // ~50 KB in flight per SM is the latency x bandwidth product, more only adds queueing.
            p.ts_slots = 3;
            if (const char* se = std::getenv("B2C_TS_SLOTS")) {
                const int v = std::atoi(se);
                if (v >= 2 && v <= 6) p.ts_slots = v;
            }
            p.ts_smem = R == 2 ? ts_smem_bytes<2, kTsCpt, kTsWarps>(p.ts_slots)
                               : ts_smem_bytes<3, kTsCpt, kTsWarps>(p.ts_slots);
            std::vector<Segment> segs;
            std::vector<int> begin;
            build_strip_partition(p, sw, p.ts_grid, true, segs, begin);
            B200_CHECK_CUDA(cudaMalloc(&p.d_segs, std::max<size_t>(1, segs.size()) * sizeof(Segment)));
            B200_CHECK_CUDA(cudaMalloc(&p.d_seg_begin, begin.size() * sizeof(int)));
            B200_CHECK_CUDA(cudaMemcpy(p.d_segs, segs.data(), segs.size() * sizeof(Segment), cudaMemcpyHostToDevice));
            B200_CHECK_CUDA(cudaMemcpy(p.d_seg_begin, begin.data(), begin.size() * sizeof(int), cudaMemcpyHostToDevice));
            if (p.debug & 32)
                B200_CHECK_CUDA(cudaMalloc(&p.ts_stamps, (size_t)p.ts_grid * 4 * sizeof(unsigned long long)));
        }
    }
    p.ncoef = kind == 2 ? 5 : (kind == 0 ? 4 * R + 1 : (2 * R + 1) * (2 * R + 1));
    B200_CHECK_CUDA(cudaMalloc(&p.coef_dev, 2 * 128 * sizeof(double)));
    p.coef_cur = p.coef_dev;
    {
        const size_t np = (size_t)std::max(std::max(p.grid, p.strip_grid), p.ts_grid);
        B200_CHECK_CUDA(cudaMalloc(&p.resid_partials, (np + 1) * sizeof(double)));
        B200_CHECK_CUDA(cudaMemset(p.resid_partials, 0, (np + 1) * sizeof(double)));
        p.tail_ticket = reinterpret_cast<unsigned*>(p.resid_partials + np);
    }
    // default: Jacobi average of the 4R (or (2R+1)^2-1) neighbours
    double hc[128] = {0};
    const int nn = p.ncoef - 1;
    if (kind != 1) { for (int k = 1; k < p.ncoef; ++k) hc[k] = 1.0 / nn; }
    else { for (int k = 0; k < p.ncoef; ++k) hc[k] = (k == p.ncoef / 2) ? 0.0 : 1.0 / nn; }
    B200_CHECK_CUDA(cudaMemcpy(p.coef_dev, hc, sizeof(double) * p.ncoef, cudaMemcpyHostToDevice));
    *out = h;
    B2C_CATCH
}

B2C_EXPORT int b2c_stencil_destroy(b2c_stencil* h) {
    B2C_TRY
    if (!h) return 0;
    if (IoState* io = h->p.io) {
        cudaStreamSynchronize(io->h2d);
        cudaStreamSynchronize(io->d2h);
        for (int i = 0; i < 2; ++i) {
            cudaEventDestroy(io->coef_ready[i]); cudaEventDestroy(io->coef_free[i]);
            cudaEventDestroy(io->sweep_done[i]); cudaEventDestroy(io->resid_copied[i]);
        }
        cudaStreamDestroy(io->h2d);
        cudaStreamDestroy(io->d2h);
        cudaFreeHost(io->coef_stage);
        cudaFreeHost(io->resid_stage);
        delete io;
    }
    cudaFree(h->p.coef_dev);
    cudaFree(h->p.resid_partials);
    if (h->p.d_segs) cudaFree(h->p.d_segs);
    if (h->p.d_seg_begin) cudaFree(h->p.d_seg_begin);
    if (h->p.ts_stamps) cudaFree(h->p.ts_stamps);
    delete h;
    B2C_CATCH
}

B2C_EXPORT int b2c_stencil_set_coef(b2c_stencil* h, const double* host_coef, int n, void* stream) {
    B2C_TRY
    B200_REQUIRE(h && host_coef, ERR_ARG, "null argument");
    B200_REQUIRE(n == h->p.ncoef, ERR_COUNT,
                 "coefficient count must be " + std::to_string(h->p.ncoef));
    B200_CHECK_CUDA(cudaMemcpyAsync(h->p.coef_cur, host_coef, sizeof(double) * n,
                                    cudaMemcpyHostToDevice, (cudaStream_t)stream));
    B2C_CATCH
}

B2C_EXPORT int b2c_stencil_set_timeout_ms(b2c_stencil* h, unsigned long long ms) {
    B2C_TRY
    B200_REQUIRE(h, ERR_ARG, "null plan");
    h->p.timeout_ns = ms * 1000000ull;
    B2C_CATCH
}

B2C_EXPORT int b2c_stencil_info(b2c_stencil* h, b2c_stencil_info_t* o) {
    B2C_TRY
    B200_REQUIRE(h && o, ERR_ARG, "null argument");
    const Plan& p = h->p;
    o->cur = p.cur; o->steps_done = p.steps_done;
    o->grid = p.algo == 2 ? p.ts_grid : (p.algo == 1 ? p.strip_grid : p.grid);
    o->threads = p.algo == 2 ? TsG::THREADS : (p.algo == 1 ? kStripThreads : kThreads);
    o->smem_bytes = p.algo == 2 ? p.ts_smem : (p.algo == 1 ? p.strip_smem : p.smem);
    o->nstages = p.algo == 2 ? p.ts_slots : (p.algo == 1 ? p.strip_slots : p.nstages); o->tiles_x = p.tilesX; o->tiles_y = p.tilesY;
    o->ncoef = p.ncoef; o->launches_per_step = 1;
    o->buf[0] = p.base + p.L.off_x[0]; o->buf[1] = p.base + p.L.off_x[1];
    o->resid = reinterpret_cast<double*>(p.base + kOffResid);
    o->launches = p.launches;
    B2C_CATCH
}

static int drain_impl(Plan& p, bool xchg, cudaStream_t st) {
    WaitParams wp{};
    wp.flags = reinterpret_cast<const uint64_t*>(p.base + (xchg ? kOffXchgFlags : kOffStepFlags));
    wp.status = reinterpret_cast<unsigned*>(p.base + kOffStatus);
    wp.timeout_ns = p.timeout_ns;
    bool any = false;
    // fused steps: the kernel multiplies the DEVICE step counter (graph replays advance it);
    // unfused exchanges: plain host-side epochs
    wp.step_ctr = xchg ? nullptr : reinterpret_cast<const uint64_t*>(p.base + kOffStepCtr);
    for (int d = 0; d < 8; ++d) {
        wp.epoch[d] = 0;
        if (!dir_active(p, d)) continue;
        wp.epoch[d] = xchg ? p.xchg_done : (p.steps_done ? contributions(p, d) : 0);
        any = any || wp.epoch[d] != 0;
    }
    if (!any) return 0;
    flag_wait_kernel<<<1, 32, 0, st>>>(wp);
    B200_CHECK_LAUNCH();
    // refresh the in-grid ghost columns of the current buffer from the landing buffers
    UnpackParams up{};
    up.grid = reinterpret_cast<double*>(p.base + p.L.off_x[p.cur]);
    up.land[0] = reinterpret_cast<const double*>(p.base + p.L.off_land[p.cur][0]);
    up.land[1] = reinterpret_cast<const double*>(p.base + p.L.off_land[p.cur][1]);
    bool cols = false;
    for (int d = 0; d < 8; ++d) {
        up.active[d] = dir_active(p, d) && dir_dx(d) != 0 && wp.epoch[d] != 0;
        cols = cols || up.active[d];
    }
    if (!cols) return 1;
    up.W = p.L.W; up.H = p.L.H; up.g = p.L.g; up.lead = p.L.lead; up.pitch = p.L.pitch;
    const int n = 2 * p.L.rows * p.L.g;
    land_unpack_kernel<<<(n + 255) / 256 > 64 ? 64 : (n + 255) / 256, 256, 0, st>>>(up);
    B200_CHECK_LAUNCH();
    return 2;
}

// Launch `nsteps` fused steps.  flags: bit0 = compute the residual of the LAST
// step into the pad, bit1 = append the drain (wait until the neighbours' halo
// pushes of the last step have landed, so ghosts of the current buffer are
// complete when the stream is synchronised).
static void step_impl(Plan& p, int nsteps, int flags, cudaStream_t st, double* resid_out) {
    // The temporal strip kernel does the tail of a sweep itself (last CTA: residual sum, step
    // counter); the other kernels get one small kernel for each.
    const bool fold = p.algo == 2;
    for (int s = 0; s < nsteps; ++s) {
        LaunchParams lp{};
        const int in = p.cur, outb = p.cur ^ 1;
        lp.in = reinterpret_cast<const double*>(p.base + p.L.off_x[in]);
        lp.out = reinterpret_cast<double*>(p.base + p.L.off_x[outb]);
        lp.coef = p.coef_cur;
        lp.my_flags = reinterpret_cast<const uint64_t*>(p.base + kOffStepFlags);
        lp.status = reinterpret_cast<unsigned*>(p.base + kOffStatus);
        lp.timeout_ns = p.timeout_ns;
        lp.resid_partials = p.resid_partials;
        for (int d = 0; d < 8; ++d) {
            if (dir_active(p, d)) {
                lp.peer_out[d] = reinterpret_cast<double*>(p.nbr[d] + p.L.off_x[outb]);
                // my west edge becomes the neighbour's EAST ghost and vice versa
                lp.peer_land[d] = dir_dx(d) == 0 ? nullptr :
                    reinterpret_cast<double*>(p.nbr[d] + p.L.off_land[outb][dir_dx(d) < 0 ? 1 : 0]);
                lp.peer_flag[d] =
                    reinterpret_cast<uint64_t*>(p.nbr[d] + kOffStepFlags) + dir_opposite(d);
                lp.signals_per_step[d] = contributions(p, d);
            } else {
                lp.peer_out[d] = nullptr; lp.peer_land[d] = nullptr; lp.peer_flag[d] = nullptr;
                lp.signals_per_step[d] = 0;
            }
        }
        // epochs are derived on the device from the step counter in the pad (+ the offset of
        // this step inside the batch), so the same launch parameters work for every replay of
        // a captured graph
        lp.step_ctr = reinterpret_cast<const uint64_t*>(p.base + kOffStepCtr);
        lp.step_off = (uint64_t)s;
        lp.my_land[0] = reinterpret_cast<const double*>(p.base + p.L.off_land[in][0]);
        lp.my_land[1] = reinterpret_cast<const double*>(p.base + p.L.off_land[in][1]);
        lp.W = p.L.W; lp.H = p.L.H; lp.g = p.L.g; lp.lead = p.L.lead; lp.pitch = p.L.pitch;
        lp.tilesX = p.tilesX; lp.tilesY = p.tilesY; lp.push_corners = p.push_corners;
        lp.debug = p.debug;
        const bool resid = (flags & 1) && s == nsteps - 1;
        p.tail_resid_out = resid_out;
        p.tail_bump = (fold && s == nsteps - 1) ? (unsigned long long)nsteps : 0ull;
        launch_step(p, lp, resid, st);
        p.launches += (resid && !fold) ? 2 : 1;
        if (resid && !fold) {
            sum_partials_kernel<<<1, 256, 0, st>>>(p.resid_partials,
                                                   p.algo == 1 ? p.strip_grid : p.grid, resid_out);
            B200_CHECK_LAUNCH();
        }
        p.cur = outb;
    }
    if (nsteps > 0) {
        if (!fold) {
            step_bump_kernel<<<1, 1, 0, st>>>(reinterpret_cast<uint64_t*>(p.base + kOffStepCtr), (uint64_t)nsteps);
            B200_CHECK_LAUNCH();
            p.launches += 1;
        }
        p.steps_done += (uint64_t)nsteps;
    }
    if (flags & 2) p.launches += drain_impl(p, false, st);
}

B2C_EXPORT int b2c_stencil_step(b2c_stencil* h, int nsteps, int flags, void* stream) {
    B2C_TRY
    trace::Scope ts("stencil_step(enqueue)", "device", -1, -3, nsteps);
    B200_REQUIRE(h && nsteps >= 0, ERR_ARG, "bad argument");
    step_impl(h->p, nsteps, flags, (cudaStream_t)stream, reinterpret_cast<double*>(h->p.base + kOffResid));
    B2C_CATCH
}

static IoState& ensure_io(Plan& p) {
    if (p.io) return *p.io;
    IoState* io = new IoState();
    B200_CHECK_CUDA(cudaStreamCreateWithFlags(&io->h2d, cudaStreamNonBlocking));
    B200_CHECK_CUDA(cudaStreamCreateWithFlags(&io->d2h, cudaStreamNonBlocking));
    for (int i = 0; i < 2; ++i) {
        B200_CHECK_CUDA(cudaEventCreateWithFlags(&io->coef_ready[i], cudaEventDisableTiming));
        B200_CHECK_CUDA(cudaEventCreateWithFlags(&io->coef_free[i], cudaEventDisableTiming));
        B200_CHECK_CUDA(cudaEventCreateWithFlags(&io->sweep_done[i], cudaEventDisableTiming));
        B200_CHECK_CUDA(cudaEventCreateWithFlags(&io->resid_copied[i], cudaEventDisableTiming));
    }
    B200_CHECK_CUDA(cudaMallocHost(&io->coef_stage, 2 * 128 * sizeof(double)));
    B200_CHECK_CUDA(cudaMallocHost(&io->resid_stage, 2 * sizeof(double)));
    io->resid_stage[0] = io->resid_stage[1] = 0.0;
    p.io = io;
    return *io;
}

// One call of a pipelined loop: `nsteps` time steps whose coefficients come from host memory and
// whose residual (flags bit0) goes back to host memory, with both copies on the plan's own copy
// streams -- the H2D of call i+1 and the D2H of call i-1 run while the kernels of call i do, and
// `stream` carries the kernels only.  host_coef may be any host memory (it is staged through
// pinned slots of the plan) or NULL (coefficients unchanged).  *ticket identifies the call for
// b2c_stencil_io_wait; results of the last TWO calls are kept.
B2C_EXPORT int b2c_stencil_step_io(b2c_stencil* h, int nsteps, int flags, const double* host_coef, int n,
                                   void* stream, unsigned long long* ticket) {
    B2C_TRY
    B200_REQUIRE(h && nsteps >= 1 && ticket, ERR_ARG, "bad argument");
    Plan& p = h->p;
    IoState& io = ensure_io(p);
    cudaStream_t st = (cudaStream_t)stream;
    if (host_coef) {
        B200_REQUIRE(n == p.ncoef, ERR_COUNT, "coefficient count must be " + std::to_string(p.ncoef));
        const int cur = p.coef_cur == p.coef_dev ? 0 : 1, tgt = cur ^ 1;
        // staging slot tgt: its previous H2D (two switches ago) has long completed
        B200_CHECK_CUDA(cudaEventSynchronize(io.coef_ready[tgt]));
        std::memcpy(io.coef_stage + 128 * tgt, host_coef, sizeof(double) * n);
        // device slot tgt: free once the kernels that read it (before the last switch) are done
        B200_CHECK_CUDA(cudaStreamWaitEvent(io.h2d, io.coef_free[tgt], 0));
        B200_CHECK_CUDA(cudaMemcpyAsync(p.coef_dev + 128 * tgt, io.coef_stage + 128 * tgt, sizeof(double) * n,
                                        cudaMemcpyHostToDevice, io.h2d));
        B200_CHECK_CUDA(cudaEventRecord(io.coef_ready[tgt], io.h2d));
        // everything queued so far reads slot cur
        B200_CHECK_CUDA(cudaEventRecord(io.coef_free[cur], st));
        B200_CHECK_CUDA(cudaStreamWaitEvent(st, io.coef_ready[tgt], 0));
        p.coef_cur = p.coef_dev + 128 * tgt;
    }
    const int rs = (int)(io.seq & 1);
    double* rdev = reinterpret_cast<double*>(p.base + kOffResid) + 1 + rs;
    // the residual slot is rewritten by this call: its copy-out of two calls ago must be over
    B200_CHECK_CUDA(cudaStreamWaitEvent(st, io.resid_copied[rs], 0));
    const int T = p.kind == 2 ? p.L.R : 1;
    const int full = nsteps / T, rem = nsteps % T;
    if (T == 1) {
        step_impl(p, nsteps, flags, st, rdev);
    } else {
        if (full) step_impl(p, full, rem ? 0 : flags, st, rdev);
        if (rem) {
            p.partial_levels = rem;
            step_impl(p, 1, flags, st, rdev);
            p.partial_levels = 0;
        }
    }
    B200_CHECK_CUDA(cudaEventRecord(io.sweep_done[rs], st));
    B200_CHECK_CUDA(cudaStreamWaitEvent(io.d2h, io.sweep_done[rs], 0));
    if (flags & 1)
        B200_CHECK_CUDA(cudaMemcpyAsync(io.resid_stage + rs, rdev, sizeof(double), cudaMemcpyDeviceToHost, io.d2h));
    B200_CHECK_CUDA(cudaEventRecord(io.resid_copied[rs], io.d2h));
    *ticket = io.seq++;
    B2C_CATCH
}

// Block until call `ticket` of b2c_stencil_step_io has finished on the device and its residual
// (local sum of squared updates of the last time step) is in host memory; returns it.
B2C_EXPORT int b2c_stencil_io_wait(b2c_stencil* h, unsigned long long ticket, double* resid) {
    B2C_TRY
    B200_REQUIRE(h && h->p.io, ERR_ARG, "no pipelined step was issued");
    IoState& io = *h->p.io;
    B200_REQUIRE(ticket < io.seq && io.seq - ticket <= 2, ERR_ARG,
                 "only the last two pipelined steps can be waited for");
    B200_CHECK_CUDA(cudaEventSynchronize(io.resid_copied[ticket & 1]));
    if (resid) *resid = io.resid_stage[ticket & 1];
    B2C_CATCH
}

// kind 2 only: ONE sweep that computes `levels` (1 .. R) time levels instead of all R -- the
// tail of a run whose step count is not a multiple of R.  Same exchange, same epochs.
B2C_EXPORT int b2c_stencil_step_levels(b2c_stencil* h, int levels, int flags, void* stream) {
    B2C_TRY
    B200_REQUIRE(h, ERR_ARG, "null plan");
    B200_REQUIRE(h->p.kind == 2 && levels >= 1 && levels <= h->p.L.R, ERR_ARG,
                 "levels must be in [1, levels per sweep] of a temporally blocked plan");
    h->p.partial_levels = levels;
    const int rc = b2c_stencil_step(h, 1, flags, stream);
    h->p.partial_levels = 0;
    return rc;
    B2C_CATCH
}

// Host bookkeeping for replays of a captured graph of `nsteps` steps (the device-side
// counter is advanced by the graph itself): moves the plan's step count and buffer parity.
B2C_EXPORT int b2c_stencil_advance(b2c_stencil* h, long long nsteps) {
    B2C_TRY
    B200_REQUIRE(h, ERR_ARG, "null plan");
    Plan& p = h->p;
    B200_REQUIRE(nsteps >= 0 || (uint64_t)(-nsteps) <= p.steps_done, ERR_ARG, "cannot rewind past step 0");
    p.steps_done = (uint64_t)((long long)p.steps_done + nsteps);
    if (nsteps & 1) p.cur ^= 1;
    B2C_CATCH
}

B2C_EXPORT int b2c_stencil_drain(b2c_stencil* h, void* stream) {
    B2C_TRY
    trace::Scope ts("stencil_drain(enqueue)", "device");
    B200_REQUIRE(h, ERR_ARG, "null plan");
    h->p.launches += drain_impl(h->p, false, (cudaStream_t)stream);
    B2C_CATCH
}

// Unfused halo exchange of the CURRENT buffer, phase 1: push my core edges
// into the neighbours' ghost cells and bump their exchange counters.
B2C_EXPORT int b2c_stencil_exchange_push(b2c_stencil* h, void* stream) {
    B2C_TRY
    trace::Scope ts("stencil_exchange_push(enqueue)", "device");
    B200_REQUIRE(h, ERR_ARG, "null plan");
    Plan& p = h->p;
    XchgParams xp{};
    xp.src = reinterpret_cast<const double*>(p.base + p.L.off_x[p.cur]);
    xp.W = p.L.W; xp.H = p.L.H; xp.g = p.L.g; xp.lead = p.L.lead; xp.pitch = p.L.pitch;
    for (int d = 0; d < 8; ++d) {
        if (dir_active(p, d)) {
            xp.peer[d] = reinterpret_cast<double*>(p.nbr[d] + p.L.off_x[p.cur]);
            xp.peer_land[d] = dir_dx(d) == 0 ? nullptr :
                reinterpret_cast<double*>(p.nbr[d] + p.L.off_land[p.cur][dir_dx(d) < 0 ? 1 : 0]);
            xp.peer_flag[d] = reinterpret_cast<uint64_t*>(p.nbr[d] + kOffXchgFlags) + dir_opposite(d);
        }
    }
    halo_push1_kernel<<<8, 512, 0, (cudaStream_t)stream>>>(xp);
    B200_CHECK_LAUNCH();
    p.launches += 1;
    p.xchg_done += 1;
    B2C_CATCH
}

// Phase 2: wait until every neighbour's push of the same exchange has landed.
B2C_EXPORT int b2c_stencil_exchange_wait(b2c_stencil* h, void* stream) {
    B2C_TRY
    trace::Scope ts("stencil_exchange_wait(enqueue)", "device");
    B200_REQUIRE(h, ERR_ARG, "null plan");
    h->p.launches += drain_impl(h->p, true, (cudaStream_t)stream);
    B2C_CATCH
}

B2C_EXPORT int b2c_stencil_fill(b2c_stencil* h, int which, int region, double value, void* stream) {
    // region: 0 = whole padded tile, 1 = core only
    B2C_TRY
    B200_REQUIRE(h && (which == 0 || which == 1), ERR_ARG, "bad argument");
    Plan& p = h->p;
    double* b = reinterpret_cast<double*>(p.base + p.L.off_x[which]);
    int x0, y0, w, hh;
    if (region == 0) { x0 = 0; y0 = 0; w = p.L.pitch; hh = p.L.rows; }
    else { x0 = p.L.lead; y0 = p.L.g; w = p.L.W; hh = p.L.H; }
    dim3 blk(64, 4), grd((w + 63) / 64, (hh + 3) / 4);
    fill_rect_kernel<<<grd, blk, 0, (cudaStream_t)stream>>>(b, p.L.pitch, x0, y0, w, hh, value);
    B200_CHECK_LAUNCH();
    p.launches += 1;
    B2C_CATCH
}

// Returns the device status word (0 = healthy) after synchronising `stream`.
B2C_EXPORT int b2c_stencil_status(b2c_stencil* h, void* stream, unsigned int* status) {
    B2C_TRY
    B200_REQUIRE(h && status, ERR_ARG, "null argument");
    B200_CHECK_CUDA(cudaMemcpyAsync(status, h->p.base + kOffStatus, sizeof(unsigned),
                                    cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    B200_CHECK_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    B2C_CATCH
}

// Phase timing of the LAST sweep of the temporal strip kernel (plan created with
// B2C_STENCIL_DEBUG=32; synchronises `stream`).  All in microseconds over the CTAs of this rank:
// out[0] mean and out[1] max time from kernel start to the end of a CTA's boundary segments,
// out[2] mean and out[3] max CTA lifetime, out[4] mean and out[5] max time a producer warp spent
// waiting for halo flags, out[6] first-start to last-end of the grid.
B2C_EXPORT int b2c_stencil_phase_stats(b2c_stencil* h, void* stream, double out[7]) {
    B2C_TRY
    B200_REQUIRE(h && out, ERR_ARG, "null argument");
    Plan& p = h->p;
    B200_REQUIRE(p.ts_stamps != nullptr, ERR_STATE,
                 "no phase stamps: create the plan with B2C_STENCIL_DEBUG=32 (temporal strip kernel only)");
    std::vector<unsigned long long> v((size_t)p.ts_grid * 4);
    B200_CHECK_CUDA(cudaMemcpyAsync(v.data(), p.ts_stamps, v.size() * sizeof(unsigned long long),
                                    cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    B200_CHECK_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    double sb = 0, mb = 0, sl = 0, ml = 0, sw = 0, mw = 0;
    unsigned long long t0 = ~0ull, t1 = 0;
    for (int c = 0; c < p.ts_grid; ++c) {
        const double b = (double)(v[4 * c + 1] - v[4 * c + 0]) * 1e-3, l = (double)(v[4 * c + 2] - v[4 * c + 0]) * 1e-3,
                     w = (double)v[4 * c + 3] * 1e-3;
        sb += b; sl += l; sw += w;
        mb = std::max(mb, b); ml = std::max(ml, l); mw = std::max(mw, w);
        t0 = std::min(t0, v[4 * c + 0]); t1 = std::max(t1, v[4 * c + 2]);
    }
    out[0] = sb / p.ts_grid; out[1] = mb; out[2] = sl / p.ts_grid; out[3] = ml;
    out[4] = sw / p.ts_grid; out[5] = mw; out[6] = (double)(t1 - t0) * 1e-3;
    B2C_CATCH
}

}  // extern "C"
