// Fused stencil update + halo exchange for one rank's tile (sm_100a).
//
// What the reference does per time step (stencil2d/mpi-2d-stencil-subarray-cuda.cu:169-172):
//   ExchangeData()  -> 8 MPI_Irecv + 8 MPI_Isend of subarray types on a device
//                      pointer + MPI_Waitall (stencil2d/stencil2D.h:363-377)
//   Compute()       -> empty stub (:32-33)
// i.e. host-driven, no overlap, and no stencil at all.
//
// What this kernel does per time step, in ONE launch per rank and with no host
// or MPI involvement:
//   * a persistent grid (<= 2 CTAs per SM) walks the tile list with the
//     boundary ring FIRST;
//   * one producer thread per CTA feeds a 3-stage shared-memory ring with 2-D
//     TMA box loads (cp.async.bulk.tensor) completed on mbarriers;
//   * 8 consumer warps apply the radius-R star (or 3x3 box) stencil from
//     shared memory with a register sliding window and write 16-byte aligned
//     rows of the output tile;
//   * consumers of a boundary tile also store the freshly computed edge cells
//     into the NEIGHBOUR GPU's memory (peer addresses over NVLink): edge ROWS
//     go straight into the neighbour's ghost rows (1 KiB contiguous per tile
//     row); edge COLUMNS are staged in shared memory and written as one
//     coalesced burst into a contiguous per-neighbour "landing" buffer
//     [padded row][g] -- strided 8-byte stores over NVLink were measured at
//     ~8 ns EACH (0.28 ms per step for 2 x 16384 cells), so columns never travel
//     strided.  One thread then release-increments the neighbour's
//     per-direction epoch counter (red.release.sys);
//   * the consumer of a west/east boundary tile patches the ghost columns of its
//     TMA box in shared memory from its landing buffer (the in-grid ghost
//     columns are refreshed from the landing buffers only when a run drains);
//   * before a boundary tile is loaded, the producer acquires (ld.acquire.sys)
//     the epoch counters of the directions whose ghost cells the tile reads --
//     the device-side replacement of MPI_Waitall.  The same acquire also
//     protects the write-after-read hazard on the neighbour's ghost cells of
//     the double-buffered grid (see DESIGN.md "halo protocol").
// Interior tiles never wait, so the halo traffic is in flight while ~98 % of
// the tile list is being computed.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "kernels/ptx.cuh"

namespace b200 {
namespace stencil {

constexpr int kTileW = 128;        // output cells per tile row (1 KiB of fp64)
#ifndef B2C_TILE_H
#define B2C_TILE_H 32
#endif
constexpr int kTileH = B2C_TILE_H;         // output rows per tile
constexpr int kConsumers = 256;    // 8 consumer warps
constexpr int kThreads = kConsumers + 32;  // + 1 producer warp
constexpr int kRowsPerThread = kTileH / 4;  // 4 row bands, 64 column pairs
constexpr int kMaxRadius = 4;

// Neighbour directions, numbered like the reference's MPIGridCellID
// (stencil2d/stencil2D.h:86-88): NW N NE W E SW S SE.
enum Dir : int { NW = 0, N = 1, NE = 2, W_ = 3, E_ = 4, SW = 5, S = 6, SE = 7 };
__host__ __device__ constexpr int dir_dy(int d) { return d < 3 ? -1 : (d < 5 ? 0 : 1); }
__host__ __device__ constexpr int dir_dx(int d) {
    return (d == 0 || d == 3 || d == 5) ? -1 : ((d == 1 || d == 6) ? 0 : 1);
}
__host__ __device__ constexpr int dir_opposite(int d) { return 7 - d; }
// Column landing buffers are indexed [kLandLead + padded_row - g][R] so that
// the burst of core row 0 starts on a 128-byte line.
constexpr int kLandLead = 16;
__host__ __device__ constexpr size_t land_index(int padded_row, int g, int R) {
    return (size_t)(padded_row - g + kLandLead) * R;
}
// directions with a column (east/west) component: NW NE W E SW SE
constexpr unsigned kColDirs = (1u << 0) | (1u << 2) | (1u << 3) | (1u << 4) | (1u << 5) | (1u << 7);

template <int R, int TH = kTileH>
struct Geo {
    static constexpr int RA = (R + 1) & ~1;  // column halo rounded up to 16 B
    static constexpr int BOXW = kTileW + 2 * RA;
    static constexpr int BOXH = TH + 2 * R;
    static constexpr int BOX_BYTES = BOXW * BOXH * 8;
    static constexpr int STAGE_BYTES = (BOX_BYTES + 127) & ~127;
};

struct LaunchParams {
    double* out;               // local padded output buffer
    const double* in;          // local padded input buffer (same layout)
    const double* coef;        // device coefficients (layout: see Stencil2D docs)
    double* peer_out[8];       // neighbour's output buffer per direction (nullptr: none)
    double* peer_land[8];      // dx != 0 only: neighbour's column landing buffer (out parity)
    const double* my_land[2];  // my west / east column landing buffers (in parity)
    uint64_t* peer_flag[8];    // neighbour's epoch counter that I bump per direction
    const uint64_t* my_flags;  // my 8 epoch counters (indexed by the direction data comes from)
    // my_flags[d] must reach (*step_ctr + step_off) * signals_per_step[d] before the ghost
    // cells of direction d are read.  The step count lives in DEVICE memory (bumped by a
    // one-thread kernel after each batch of steps) so that a captured CUDA graph of N
    // steps can be replayed: its kernel parameters never change, the counter does.
    const uint64_t* step_ctr;
    uint64_t step_off;
    uint64_t signals_per_step[8];
    double* resid_partials;    // [gridDim.x] per-CTA sum of (out-in)^2, or nullptr
    unsigned int* status;      // device status word (bit0: flag wait timed out)
    uint64_t timeout_ns;       // spin deadline per wait
    int W, H;                  // core size in cells
    int g;                     // ghost width (== stencil radius)
    int lead;                  // column of the first core cell inside a padded row
    int pitch;                 // elements per padded row
    int tilesX, tilesY;
    int push_corners;          // also exchange the 4 corner blocks
    int debug;                 // bit0: skip column pushes, bit1: skip flag waits (experiments only)
};

__device__ __forceinline__ void decode_tile(int i, int nbx, int nby, int& tx, int& ty) {
    // Boundary ring first, then the interior with x fastest: tiles i+-1 (x
    // neighbours) and i+-(nbx-2) (y neighbours) are all in flight within about
    // one round of the persistent grid, so halo sectors are re-read from L2,
    // not from DRAM (ncu: y-fastest order re-read 23 % of the grid from DRAM).
    if (nby == 1) { tx = i; ty = 0; return; }
    if (i < nbx) { tx = i; ty = 0; return; }
    i -= nbx;
    if (i < nbx) { tx = i; ty = nby - 1; return; }
    i -= nbx;
    const int mid = nby - 2;
    if (i < mid) { tx = 0; ty = 1 + i; return; }
    i -= mid;
    if (nbx > 1) {
        if (i < mid) { tx = nbx - 1; ty = 1 + i; return; }
        i -= mid;
    }
    const int midx = nbx - 2;
    tx = 1 + i % midx;
    ty = 1 + i / midx;
}

// Bit d set <=> the tile touches the tile edge facing direction d.
__device__ __forceinline__ unsigned edge_mask(int tx, int ty, int nbx, int nby) {
    const bool n = ty == 0, s = ty == nby - 1, w = tx == 0, e = tx == nbx - 1;
    unsigned m = 0;
    if (n) m |= 1u << N;
    if (s) m |= 1u << S;
    if (w) m |= 1u << W_;
    if (e) m |= 1u << E_;
    if (n && w) m |= 1u << NW;
    if (n && e) m |= 1u << NE;
    if (s && w) m |= 1u << SW;
    if (s && e) m |= 1u << SE;
    return m;
}

__device__ __forceinline__ unsigned active_dirs(const LaunchParams& P, unsigned m) {
    unsigned a = 0;
#pragma unroll
    for (int d = 0; d < 8; ++d) {
        const bool corner = (d == NW || d == NE || d == SW || d == SE);
        if ((m >> d & 1u) && P.peer_out[d] != nullptr && (!corner || P.push_corners)) a |= 1u << d;
    }
    return a;
}

// TB (temporal blocking, TB = R time steps per sweep): the geometry, halo width and the whole
// exchange protocol are those of radius R with corners, but the arithmetic is a radius-1 star
// applied R times.  Each thread slides three input rows through registers and forms time level
// k for its two columns plus R-k columns on either side (rows likewise), level k+1 from the
// registers of level k; nothing intermediate touches shared or global memory and no barrier
// separates the levels.  One sweep reads and writes the grid once for R updates, so the DRAM
// roofline is 16/R bytes per cell update.  Every value is computed with the same operation order
// as single steps would use, on this GPU or on the neighbour that owns the cell, so results stay
// bit-identical.
// TH: output rows per tile (a multiple of 4; every consumer thread owns TH/4 rows of two
// columns).  Taller tiles cut the redundant rows of the temporally blocked variants
// (TB = 3: 8.0 -> 7.0 cell updates per output cell) at the price of a shallower ring.
// LV: time levels actually computed by this sweep (1 <= LV <= TB).  LV < TB is the partial sweep
// that ends a run whose step count is not a multiple of TB: same geometry and exchange, fewer
// levels.
template <int R, bool BOX, bool RESID, int NSTAGES, int TB = 0, int TH = B2C_TILE_H, int LV = TB>
__global__ void __launch_bounds__(kThreads, (NSTAGES <= 3 && R <= 2 && TB == 0) ? 2 : 1)
stencil_fused_kernel(const __grid_constant__ CUtensorMap tmap_in,
                     const __grid_constant__ LaunchParams P) {
    static_assert(TB == 0 || (TB == R && R >= 2 && !BOX), "temporal blocking: radius-1 star, R levels");
    static_assert(TB == 0 ? LV == 0 : (LV >= 1 && LV <= TB), "levels per sweep");
    static_assert(TH % 4 == 0 && TH >= 4 * kMaxRadius, "tile height: 4 row bands");
    constexpr int kTileH = TH;               // shadow the namespace-level defaults
    constexpr int kRowsPerThread = TH / 4;
    using G = Geo<R, TH>;
    extern __shared__ __align__(128) unsigned char smem[];
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + NSTAGES * G::STAGE_BYTES);
    uint64_t* empty = full + NSTAGES;
    __shared__ double red_smem[kConsumers / 32];
    __shared__ double colstage[2][2][kTileH][R];   // [parity][west|east][tile row][k]
    int bparity = 0;                                // alternates per boundary tile

    const int tid = threadIdx.x;
    const int warp = tid >> 5, lane = tid & 31;
    const int nbx = P.tilesX, nby = P.tilesY;
    const int ntiles = nbx * nby;

    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < NSTAGES; ++s) {
            ptx::mbar_init(&full[s], 1);
            ptx::mbar_init(&empty[s], kConsumers / 32);
        }
        ptx::fence_mbar_init();
    }
    __syncthreads();

    if (warp == kConsumers / 32) {
        // ===================== producer: flag acquire + TMA =====================
        if (lane == 0) {
            ptx::prefetch_tensormap(&tmap_in);
            int stage = 0;
            uint32_t phase = 0;
            for (int i = blockIdx.x; i < ntiles; i += gridDim.x) {
                int tx, ty;
                decode_tile(i, nbx, nby, tx, ty);
                const unsigned act = active_dirs(P, edge_mask(tx, ty, nbx, nby));
                if (act && !(P.debug & 2)) {
                    const uint64_t t0 = ptx::globaltimer_ns();
                    for (int d = 0; d < 8; ++d) {
                        if (!(act >> d & 1u)) continue;
                        uint32_t spins = 0;
                        const uint64_t need = (*P.step_ctr + P.step_off) * P.signals_per_step[d];
                        while (ptx::ld_acquire_sys(P.my_flags + d) < need) {
                            if ((++spins & 1023u) == 0 &&
                                ptx::globaltimer_ns() - t0 > P.timeout_ns) {
                                atomicOr(P.status, 1u);
                                break;
                            }
                        }
                    }
                    // ghost cells were written by another GPU's generic proxy;
                    // the TMA below reads them through the async proxy.
                    ptx::fence_proxy_async();
                }
                ptx::mbar_wait(&empty[stage], phase ^ 1u);
                ptx::mbar_arrive_expect_tx(&full[stage], G::BOX_BYTES);
                if (P.debug & 8)
                    ptx::tma_load_2d_hint(smem + stage * G::STAGE_BYTES, &tmap_in,
                                          P.lead + tx * kTileW - G::RA, ty * kTileH + P.g - R,
                                          &full[stage], (P.debug & 16) ? ptx::l2_policy_evict_last()
                                                                       : ptx::l2_policy_evict_first());
                else
                    ptx::tma_load_2d(smem + stage * G::STAGE_BYTES, &tmap_in,
                                     P.lead + tx * kTileW - G::RA, ty * kTileH + P.g - R, &full[stage]);
                if (++stage == NSTAGES) { stage = 0; phase ^= 1u; }
            }
        }
        return;
    }

    // =============================== consumers ===============================
    constexpr int NCOEF = TB ? 5 : (BOX ? (2 * R + 1) * (2 * R + 1) : 4 * R + 1);
    double c[NCOEF];
#pragma unroll
    for (int k = 0; k < NCOEF; ++k) c[k] = __ldg(P.coef + k);

    const int cpair = tid & 63;       // column pair inside the tile
    const int band = tid >> 6;        // row band (4 bands of kRowsPerThread rows)
    const int c0 = 2 * cpair;         // first of my two columns (tile coords)
    const int r0 = band * kRowsPerThread;
    double racc = 0.0;

    int stage = 0;
    uint32_t phase = 0;
    for (int i = blockIdx.x; i < ntiles; i += gridDim.x) {
        int tx, ty;
        decode_tile(i, nbx, nby, tx, ty);
        const unsigned act = active_dirs(P, edge_mask(tx, ty, nbx, nby));

        ptx::mbar_wait(&full[stage], phase);
        const double* box = reinterpret_cast<const double*>(smem + stage * G::STAGE_BYTES);
        if (act & kColDirs) {
            // Ghost columns of this box come from the landing buffers, not from the
            // grid: one element per thread, rows = padded rows ty*TH .. +BOXH.
            double* wbox = reinterpret_cast<double*>(smem + stage * G::STAGE_BYTES);
            for (int e = tid; e < 2 * G::BOXH * R; e += kConsumers) {
                const int side = e / (G::BOXH * R), rem = e % (G::BOXH * R);
                const int j = rem / R, k = rem % R;
                const int pr = ty * kTileH + j;                 // padded row (g == R)
                if (pr >= P.H + 2 * R) continue;
                const int vd = pr < R ? (side ? NE : NW) : (pr >= P.H + R ? (side ? SE : SW)
                                                                          : (side ? E_ : W_));
                if (!(act >> vd & 1u)) continue;
                const double v = P.my_land[side][land_index(pr, R, R) + k];
                const int col = side ? G::RA + (P.W - tx * kTileW) + k : G::RA - R + k;
                wbox[j * G::BOXW + col] = v;
            }
            ptx::fence_proxy_async();   // generic smem writes vs the next TMA into this stage
            ptx::named_bar_sync(3, kConsumers);
        }
        // box(row, col): row = tile row + R, col = tile col + RA
        const double* mycol = box + c0 + G::RA;

        double2 res[kRowsPerThread];
        if constexpr (TB != 0) {
            // lv[0]: input rows (aligned columns c0-RA .. c0+1+RA); lv[k], k = 1..LV-1: time level k
            // at columns c0-(LV-k) .. c0+1+(LV-k).  Index [2] is the newest row of each window.
            // Input row j (tile row r0 - LV + j) lets level k produce its row r0 - LV + j - k once
            // j >= 2k; level LV (the output) starts at j = 2 LV with tile row r0.
            constexpr int LW = 2 * G::RA + 2;
            double lv[LV][3][LW];
#pragma unroll
            for (int j = 0; j < kRowsPerThread + 2 * LV; ++j) {
                {
                    const double* p = mycol + (r0 + (R - LV) + j) * G::BOXW - G::RA;
#pragma unroll
                    for (int k = 0; k < LW / 2; ++k) {
                        const double2 v = *reinterpret_cast<const double2*>(p + 2 * k);
                        lv[0][2][2 * k] = v.x;
                        lv[0][2][2 * k + 1] = v.y;
                    }
                }
#pragma unroll
                for (int k = 1; k <= LV; ++k) {
                    if (j < 2 * k) continue;
                    // column i of level k is tile column c0 - (LV-k) + i; its west neighbour in
                    // level k-1 has index i + off, centre i + off + 1, east i + off + 2
                    const int off = (k == 1) ? G::RA - LV : 0;
                    const int ncol = 2 + 2 * (LV - k);
                    double outv[2 + 2 * LV];
#pragma unroll
                    for (int i = 0; i < ncol; ++i) {
                        double a = c[0] * lv[k - 1][1][i + off + 1];
                        a = fma(c[1], lv[k - 1][1][i + off], a);          // west
                        a = fma(c[2], lv[k - 1][1][i + off + 2], a);      // east
                        a = fma(c[3], lv[k - 1][0][i + off + 1], a);      // north
                        a = fma(c[4], lv[k - 1][2][i + off + 1], a);      // south
                        outv[i] = a;
                    }
                    if (k < LV) {
#pragma unroll
                        for (int i = 0; i < ncol; ++i) lv[k][2][i] = outv[i];
                    } else {
                        const int y = j - 2 * LV;
                        res[y] = make_double2(outv[0], outv[1]);
                        if constexpr (RESID) {   // residual of the LAST of the LV updates
                            const int gy = ty * kTileH + r0 + y, gx = tx * kTileW + c0;
                            if (gy < P.H && gx < P.W) {
                                const int o1 = (LV == 1) ? G::RA - LV : 0;
                                const double dx_ = outv[0] - lv[LV - 1][1][o1 + 1];
                                const double dy_ = outv[1] - lv[LV - 1][1][o1 + 2];
                                racc = fma(dx_, dx_, racc);
                                racc = fma(dy_, dy_, racc);
                            }
                        }
                    }
                }
                // slide every window that received a row in this iteration
#pragma unroll
                for (int k = 0; k < LV; ++k) {
                    if (j < 2 * k) continue;
#pragma unroll
                    for (int i = 0; i < LW; ++i) { lv[k][0][i] = lv[k][1][i]; lv[k][1][i] = lv[k][2][i]; }
                }
            }
        } else if constexpr (!BOX) {
            double2 win[2 * R + 1];
#pragma unroll
            for (int k = 0; k < 2 * R; ++k)
                win[k] = *reinterpret_cast<const double2*>(mycol + (r0 + k) * G::BOXW);
#pragma unroll
            for (int y = 0; y < kRowsPerThread; ++y) {
                win[2 * R] =
                    *reinterpret_cast<const double2*>(mycol + (r0 + y + 2 * R) * G::BOXW);
                // horizontal strip of the centre row: cols c0-RA .. c0+1+RA
                double h[2 * G::RA + 2];
                const double* crow = mycol + (r0 + y + R) * G::BOXW - G::RA;
#pragma unroll
                for (int k = 0; k < G::RA + 1; ++k) {
                    if (2 * k == G::RA) {
                        h[2 * k] = win[R].x;
                        h[2 * k + 1] = win[R].y;
                    } else {
                        const double2 v = *reinterpret_cast<const double2*>(crow + 2 * k);
                        h[2 * k] = v.x;
                        h[2 * k + 1] = v.y;
                    }
                }
                double ax = c[0] * win[R].x, ay = c[0] * win[R].y;
#pragma unroll
                for (int k = 1; k <= R; ++k) {
                    const int b = 1 + 4 * (k - 1);
                    ax = fma(c[b + 0], h[G::RA - k], ax);       // west
                    ay = fma(c[b + 0], h[G::RA + 1 - k], ay);
                    ax = fma(c[b + 1], h[G::RA + k], ax);       // east
                    ay = fma(c[b + 1], h[G::RA + 1 + k], ay);
                    ax = fma(c[b + 2], win[R - k].x, ax);       // north
                    ay = fma(c[b + 2], win[R - k].y, ay);
                    ax = fma(c[b + 3], win[R + k].x, ax);       // south
                    ay = fma(c[b + 3], win[R + k].y, ay);
                }
                res[y] = make_double2(ax, ay);
                if constexpr (RESID) {
                    const int gy = ty * kTileH + r0 + y, gx = tx * kTileW + c0;
                    if (gy < P.H && gx < P.W) {
                        const double dx_ = ax - win[R].x, dy_ = ay - win[R].y;
                        racc = fma(dx_, dx_, racc);
                        racc = fma(dy_, dy_, racc);
                    }
                }
#pragma unroll
                for (int k = 0; k < 2 * R; ++k) win[k] = win[k + 1];
            }
        } else {
#pragma unroll
            for (int y = 0; y < kRowsPerThread; ++y) {
                double ax = 0.0, ay = 0.0, cx = 0.0, cy = 0.0;
#pragma unroll
                for (int dy = -R; dy <= R; ++dy) {
                    double h[2 * G::RA + 2];
                    const double* crow = mycol + (r0 + y + R + dy) * G::BOXW - G::RA;
#pragma unroll
                    for (int k = 0; k < G::RA + 1; ++k) {
                        const double2 v = *reinterpret_cast<const double2*>(crow + 2 * k);
                        h[2 * k] = v.x;
                        h[2 * k + 1] = v.y;
                    }
                    if (dy == 0) { cx = h[G::RA]; cy = h[G::RA + 1]; }
#pragma unroll
                    for (int dx = -R; dx <= R; ++dx) {
                        const double w = c[(dy + R) * (2 * R + 1) + (dx + R)];
                        ax = fma(w, h[G::RA + dx], ax);
                        ay = fma(w, h[G::RA + 1 + dx], ay);
                    }
                }
                res[y] = make_double2(ax, ay);
                if constexpr (RESID) {
                    const int gy = ty * kTileH + r0 + y, gx = tx * kTileW + c0;
                    if (gy < P.H && gx < P.W) {
                        racc = fma(ax - cx, ax - cx, racc);
                        racc = fma(ay - cy, ay - cy, racc);
                    }
                }
            }
        }
        // all shared-memory reads of this stage are done: hand it back to TMA
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(&empty[stage]);

        // ------------------------------ stores ------------------------------
        const int gx = tx * kTileW + c0;  // core coords of my first column
        if (gx < P.W) {
#pragma unroll
            for (int y = 0; y < kRowsPerThread; ++y) {
                const int gy = ty * kTileH + r0 + y;
                if (gy < P.H) {
                    double* dst = P.out + (size_t)(gy + P.g) * P.pitch + P.lead + gx;
                    if (P.debug & 4) ptx::st_cs_v2f64(dst, res[y].x, res[y].y);   // streaming store
                    else *reinterpret_cast<double2*>(dst) = res[y];
                }
            }
        }
        if (act) {
            // -------- halo push --------
            if (gx < P.W) {
                // edge rows: straight into the neighbour's ghost rows (contiguous)
#pragma unroll 1
                for (int d = N; d <= S; d += S - N) {
                    if (!(act >> d & 1u)) continue;
                    const int dy = dir_dy(d);
                    double* pb = P.peer_out[d];
#pragma unroll
                    for (int y = 0; y < kRowsPerThread; ++y) {
                        const int gy = ty * kTileH + r0 + y;
                        if (gy >= P.H) continue;
                        if (dy < 0 ? gy >= P.g : gy < P.H - P.g) continue;
                        double* prow = pb + (size_t)(gy + P.g - dy * P.H) * P.pitch + P.lead;
                        *reinterpret_cast<double2*>(prow + gx) = res[y];
                    }
                }
                // edge columns: stage in shared memory
                if (act & kColDirs) {
#pragma unroll
                    for (int y = 0; y < kRowsPerThread; ++y) {
#pragma unroll
                        for (int q = 0; q < 2; ++q) {
                            const int x = gx + q;
                            const double v = q ? res[y].y : res[y].x;
                            if (x < P.g) colstage[bparity][0][r0 + y][x] = v;
                            if (x >= P.W - P.g && x < P.W) colstage[bparity][1][r0 + y][x - (P.W - P.g)] = v;
                        }
                    }
                }
            }
            // every consumer's peer stores / staging writes happen-before warp 0's work
            ptx::named_bar_sync(1, kConsumers);
            if (warp == 0) {
                if ((act & kColDirs) && !(P.debug & 1)) {
                    // coalesced bursts into the neighbours' landing buffers
                    const int rows_here = min(kTileH, P.H - ty * kTileH);
                    for (int d = 0; d < 8; ++d) {
                        const int dy = dir_dy(d), dx = dir_dx(d);
                        if (dx == 0 || !(act >> d & 1u)) continue;
                        const int side = dx < 0 ? 0 : 1;
                        // tile rows [ya, yb) take part in direction d
                        int ya = 0, yb = rows_here;
                        if (dy < 0) yb = min(yb, P.g - ty * kTileH);
                        if (dy > 0) ya = max(ya, P.H - P.g - ty * kTileH);
                        const int n = (yb - ya) * R;
                        double* dst = P.peer_land[d] +
                                      land_index(ty * kTileH + ya + P.g - dy * P.H, R, R);
                        const double* src = &colstage[bparity][side][ya][0];
                        for (int e = lane; e < n; e += 32) dst[e] = src[e];
                    }
                }
                __syncwarp();
                if (lane == 0) {
                    for (int d = 0; d < 8; ++d)
                        if (act >> d & 1u) ptx::red_add_release_sys(P.peer_flag[d], 1ull);
                }
            }
            bparity ^= 1;
        }
        if (++stage == NSTAGES) { stage = 0; phase ^= 1u; }
    }

    if constexpr (RESID) {
        // deterministic per-CTA partial: lanes -> warps -> thread 0
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) racc += __shfl_xor_sync(0xffffffffu, racc, o);
        if (lane == 0) red_smem[warp] = racc;
        ptx::named_bar_sync(2, kConsumers);
        if (tid == 0) {
            double s = 0.0;
#pragma unroll
            for (int w = 0; w < kConsumers / 32; ++w) s += red_smem[w];
            P.resid_partials[blockIdx.x] = s;
        }
    }
}

}  // namespace stencil
}  // namespace b200
