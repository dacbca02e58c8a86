// Second-generation fused stencil + halo kernel: column strips with a
// shared-memory ROW RING (star stencils, radius <= 2).
//
// Why (ncu on the tile kernel, profiles/README.md): the 128x32 tiles re-read
// their y halo (2 of 34 rows) and pay a full 32-byte sector for each 16-byte x
// halo, 14 % more DRAM reads than the grid holds, in 1 KiB row pieces.  Here
//   * a CTA owns a SEGMENT = (strip of 252 columns) x (range of 8-row chunks)
//     and marches down it: every input row is loaded ONCE (a TMA box of
//     256 x 8 fp64 = 2 KiB contiguous per row) into a ring of NSLOTS chunks;
//   * each consumer thread owns two columns and carries the 2R+1 input rows of
//     its columns in registers from chunk to chunk (no vertical re-reads at
//     all); only the chunk above and below a segment are loaded extra;
//   * x over-fetch drops to 4 of 256 columns.
// The halo protocol is the one of the tile kernel (per-direction epoch
// counters, edge rows pushed straight into the neighbour's ghost rows, edge
// columns through contiguous landing buffers, ghost columns of the loaded
// chunks patched from the landing buffers).  The host partitions the work so
// that every CTA processes its boundary segments FIRST (west/east strips are
// cut into one piece per CTA, north/south chunk rows are dealt round robin),
// then a contiguous run of the interior; each segment waits/signals once.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "kernels/ptx.cuh"
#include "stencil/stencil_kernel.cuh"

namespace b200 {
namespace stencil {

constexpr int kStripW = 252;          // MAX cells per strip (box = 252 + 2*2 = 256 columns); the
                                      // actual width is chosen per plan so that bands x strips = CTAs
constexpr int kChunkH = 4;            // rows per TMA chunk (2 and 8 measured slower, see profiles/)
constexpr int kStripCtasPerSm = 4;
constexpr int kStripBoxW = 256;
constexpr int kStripConsumers = 128;  // 4 consumer warps, 2 columns per thread
constexpr int kStripThreads = kStripConsumers + 32;
constexpr int kChunkBytes = kStripBoxW * kChunkH * 8;   // 16 KiB
constexpr int kColStageRows = 32;     // column halo rows staged before one burst

struct Segment {
    int strip;        // strip index (x0 = strip * kStripW)
    int j0, j1;       // chunks [j0, j1) are computed; chunks j0-1 .. j1 are loaded
    unsigned dirs;    // directions this segment reads ghosts of / pushes to (bit d)
    unsigned pin;     // tile edges it touches that have NO neighbour (the physical boundary of a
                      // non-periodic grid): 1 north, 2 south, 4 west, 8 east.  The temporally blocked
                      // strip kernel keeps intermediate time levels fixed beyond those edges.
};

struct StripParams {
    LaunchParams lp;          // same fields as the tile kernel (tilesX/tilesY unused)
    const Segment* segs;      // all segments, grouped per CTA
    const int* seg_begin;     // [gridDim.x + 1] offsets into segs
    int strip_w;              // cells per strip (even, <= kStripW)
    int box_w;                // strip_w + 4: columns per TMA box row == shared-memory row pitch
};

template <int R, bool RESID, int NSLOTS>
__global__ void __launch_bounds__(kStripThreads, kStripCtasPerSm)
stencil_strip_kernel(const __grid_constant__ CUtensorMap tmap_in,
                     const __grid_constant__ StripParams SP) {
    constexpr int RA = 2;
    static_assert(R >= 1 && R <= 2, "strip kernel supports radius 1 and 2");
    const LaunchParams& P = SP.lp;
    extern __shared__ __align__(128) unsigned char smem[];
    // 128 bytes of slack: idle threads (columns >= 252) read a little past the last row
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + NSLOTS * kChunkBytes + 128);
    uint64_t* empty = full + NSLOTS;
    __shared__ double red_smem[kStripConsumers / 32];
    __shared__ double colstage[2][kColStageRows][R];   // [west|east][row][k]

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int sbeg = SP.seg_begin[blockIdx.x], send = SP.seg_begin[blockIdx.x + 1];

    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < NSLOTS; ++s) {
            ptx::mbar_init(&full[s], 1);
            ptx::mbar_init(&empty[s], kStripConsumers / 32);
        }
        ptx::fence_mbar_init();
    }
    __syncthreads();

    if (warp == kStripConsumers / 32) {
        // ===================== producer =====================
        if (lane == 0) {
            ptx::prefetch_tensormap(&tmap_in);
            uint32_t k = 0;   // running chunk counter of this CTA (slot = k % NSLOTS)
            for (int s = sbeg; s < send; ++s) {
                const Segment sg = SP.segs[s];
                const unsigned act = sg.dirs;
                if (act && !(P.debug & 2)) {
                    const uint64_t t0 = ptx::globaltimer_ns();
                    for (int d = 0; d < 8; ++d) {
                        if (!(act >> d & 1u)) continue;
                        uint32_t spins = 0;
                        const uint64_t need = (*P.step_ctr + P.step_off) * P.signals_per_step[d];
                        while (ptx::ld_acquire_sys(P.my_flags + d) < need) {
                            if ((++spins & 1023u) == 0 && ptx::globaltimer_ns() - t0 > P.timeout_ns) {
                                atomicOr(P.status, 1u);
                                break;
                            }
                        }
                    }
                    ptx::fence_proxy_async();
                }
                const int x = P.lead + sg.strip * SP.strip_w - RA;
                const uint32_t chunk_tx = (uint32_t)(SP.box_w * kChunkH * 8);
                for (int j = sg.j0 - 1; j <= sg.j1; ++j, ++k) {
                    const int slot = (int)(k % NSLOTS);
                    const uint32_t phase = (k / NSLOTS) & 1u;
                    ptx::mbar_wait(&empty[slot], phase ^ 1u);
                    ptx::mbar_arrive_expect_tx(&full[slot], chunk_tx);
                    ptx::tma_load_2d(smem + slot * kChunkBytes, &tmap_in, x, P.g + j * kChunkH,
                                     &full[slot]);
                }
            }
        }
        return;
    }

    // ===================== consumers =====================
    constexpr int NCOEF = 4 * R + 1;
    double c[NCOEF];
#pragma unroll
    for (int i = 0; i < NCOEF; ++i) c[i] = __ldg(P.coef + i);
    const int c0 = 2 * tid;            // my first column inside the strip
    double racc = 0.0;
    uint32_t k = 0;                    // mirrors the producer's chunk counter

    auto wait_chunk = [&](uint32_t kk) {
        ptx::mbar_wait(&full[kk % NSLOTS], (kk / NSLOTS) & 1u);
    };
    auto release_chunk = [&](uint32_t kk) {
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(&empty[kk % NSLOTS]);
    };
    auto chunk_ptr = [&](uint32_t kk) {
        return reinterpret_cast<double*>(smem + (kk % NSLOTS) * kChunkBytes);
    };

    for (int s = sbeg; s < send; ++s) {
        const Segment sg = SP.segs[s];
        const unsigned act = sg.dirs;
        const int x0 = sg.strip * SP.strip_w;
        const int sw = min(SP.strip_w, P.W - x0);              // valid columns of this strip
        const int bw = SP.box_w;                               // shared-memory row pitch (elements)
        const bool mine = c0 < sw;                             // I own two valid columns (W is even)
        const bool west = x0 == 0, east = x0 + sw == P.W;
        const bool patch = (act & kColDirs) != 0;
        int staged0 = -1;                                      // first core row held in colstage

        // Ghost columns of a freshly loaded chunk come from the landing buffers.
        auto patch_chunk = [&](uint32_t kk, int j) {
            double* ch = chunk_ptr(kk);
            for (int e = tid; e < 2 * kChunkH * R; e += kStripConsumers) {
                const int side = e / (kChunkH * R), rem = e % (kChunkH * R);
                const int r = rem / R, q = rem % R;
                if ((side == 0 && !west) || (side == 1 && !east)) continue;
                const int pr = P.g + j * kChunkH + r;           // padded row
                if (pr < 0 || pr >= P.H + 2 * R) continue;
                const int vd = pr < R ? (side ? NE : NW) : (pr >= P.H + R ? (side ? SE : SW)
                                                                          : (side ? E_ : W_));
                if (!(act >> vd & 1u)) continue;
                const int col = side ? RA + sw + q : RA - R + q;
                ch[r * bw + col] = P.my_land[side][land_index(pr, R, R) + q];
            }
            ptx::fence_proxy_async();
            ptx::named_bar_sync(3, kStripConsumers);
        };
        // Burst the staged edge columns [staged0, upto) into the neighbours' landing buffers.
        auto flush_cols = [&](int upto) {
            if (staged0 < 0 || upto <= staged0) return;
            ptx::named_bar_sync(4, kStripConsumers);
            if (warp == 0 && !(P.debug & 1)) {
                for (int d = 0; d < 8; ++d) {
                    const int dy = dir_dy(d), dx = dir_dx(d);
                    if (dx == 0 || !(act >> d & 1u)) continue;
                    if ((dx < 0 && !west) || (dx > 0 && !east)) continue;
                    int ya = staged0, yb = upto;                  // core rows
                    if (dy < 0) yb = min(yb, P.g);
                    if (dy > 0) ya = max(ya, P.H - P.g);
                    if (dy == 0) { /* all rows */ }
                    const int n = (yb - ya) * R;
                    if (n <= 0) continue;
                    double* dst = P.peer_land[d] + land_index(ya + P.g - dy * P.H, R, R);
                    const double* src = &colstage[dx < 0 ? 0 : 1][ya - staged0][0];
                    for (int e = lane; e < n; e += 32) dst[e] = src[e];
                }
            }
            ptx::named_bar_sync(4, kStripConsumers);
            staged0 = -1;
        };

        // ---- prologue: chunk j0-1 (rows above) and chunk j0
        double2 win[2 * R + 1];
        wait_chunk(k);
        if (patch) patch_chunk(k, sg.j0 - 1);
        wait_chunk(k + 1);
        if (patch) patch_chunk(k + 1, sg.j0);
        {
            const double* above = chunk_ptr(k) + RA + c0;
            const double* cur = chunk_ptr(k + 1) + RA + c0;
#pragma unroll
            for (int i = 0; i < R; ++i)
                win[i] = *reinterpret_cast<const double2*>(above + (kChunkH - R + i) * bw);
#pragma unroll
            for (int i = 0; i < R; ++i)
                win[R + i] = *reinterpret_cast<const double2*>(cur + i * bw);
        }
        release_chunk(k);          // the chunk above lives on in registers only
        ++k;                       // k now indexes chunk j0

        for (int j = sg.j0; j < sg.j1; ++j, ++k) {
            wait_chunk(k + 1);                                   // look-ahead rows
            if (patch) patch_chunk(k + 1, j + 1);
            const double* cur = chunk_ptr(k) + RA + c0;
            const double* nxt = chunk_ptr(k + 1) + RA + c0;
            const int ybase = j * kChunkH;
            const int gx = x0 + c0;
            if (patch && staged0 < 0) staged0 = ybase;
#pragma unroll
            for (int r = 0; r < kChunkH; ++r) {
                const int y = ybase + r;
                // input row y + R
                win[2 * R] = (r + R < kChunkH)
                                 ? *reinterpret_cast<const double2*>(cur + (r + R) * bw)
                                 : *reinterpret_cast<const double2*>(nxt + (r + R - kChunkH) * bw);
                const double* crow = cur + r * bw;
                const double2 lft = *reinterpret_cast<const double2*>(crow - 2);
                const double2 rgt = *reinterpret_cast<const double2*>(crow + 2);
                // h[] = columns c0-2 .. c0+3 of the centre row
                const double h0 = lft.x, h1 = lft.y, h2 = win[R].x, h3 = win[R].y, h4 = rgt.x, h5 = rgt.y;
                double ax = c[0] * h2, ay = c[0] * h3;
                ax = fma(c[1], h1, ax); ay = fma(c[1], h2, ay);          // west 1
                ax = fma(c[2], h3, ax); ay = fma(c[2], h4, ay);          // east 1
                ax = fma(c[3], win[R - 1].x, ax); ay = fma(c[3], win[R - 1].y, ay);   // north 1
                ax = fma(c[4], win[R + 1].x, ax); ay = fma(c[4], win[R + 1].y, ay);   // south 1
                if constexpr (R == 2) {
                    ax = fma(c[5], h0, ax); ay = fma(c[5], h1, ay);      // west 2
                    ax = fma(c[6], h4, ax); ay = fma(c[6], h5, ay);      // east 2
                    ax = fma(c[7], win[0].x, ax); ay = fma(c[7], win[0].y, ay);       // north 2
                    ax = fma(c[8], win[4].x, ax); ay = fma(c[8], win[4].y, ay);       // south 2
                }
                if (mine && y < P.H) {
                    if constexpr (RESID) {
                        const double dx_ = ax - h2, dy_ = ay - h3;
                        racc = fma(dx_, dx_, racc);
                        racc = fma(dy_, dy_, racc);
                    }
                    double* dst = P.out + (size_t)(y + P.g) * P.pitch + P.lead + gx;
                    if (P.debug & 4) ptx::st_cs_v2f64(dst, ax, ay);
                    else *reinterpret_cast<double2*>(dst) = make_double2(ax, ay);
                    if (act) {
                        // edge rows: straight into the neighbour's ghost rows
                        if ((act >> N & 1u) && y < P.g)
                            *reinterpret_cast<double2*>(P.peer_out[N] + (size_t)(y + P.g + P.H) * P.pitch +
                                                        P.lead + gx) = make_double2(ax, ay);
                        if ((act >> S & 1u) && y >= P.H - P.g)
                            *reinterpret_cast<double2*>(P.peer_out[S] + (size_t)(y + P.g - P.H) * P.pitch +
                                                        P.lead + gx) = make_double2(ax, ay);
                        // edge columns: stage for the burst
                        if (patch) {
                            const int sr = y - staged0;
#pragma unroll
                            for (int q = 0; q < 2; ++q) {
                                const int xx = gx + q;
                                const double v = q ? ay : ax;
                                if (west && xx < P.g) colstage[0][sr][xx] = v;
                                if (east && xx >= P.W - P.g) colstage[1][sr][xx - (P.W - P.g)] = v;
                            }
                        }
                    }
                }
#pragma unroll
                for (int i = 0; i < 2 * R; ++i) win[i] = win[i + 1];
            }
            release_chunk(k);
            if (patch && (ybase + kChunkH - staged0 >= kColStageRows || j + 1 == sg.j1))
                flush_cols(min(ybase + kChunkH, P.H));
        }
        // the look-ahead chunk j1 was only read: hand it back as well
        release_chunk(k);
        ++k;
        if (act) {
            ptx::named_bar_sync(1, kStripConsumers);
            if (tid == 0)
                for (int d = 0; d < 8; ++d)
                    if (act >> d & 1u) ptx::red_add_release_sys(P.peer_flag[d], 1ull);
        }
    }

    if constexpr (RESID) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) racc += __shfl_xor_sync(0xffffffffu, racc, o);
        if (lane == 0) red_smem[warp] = racc;
        ptx::named_bar_sync(2, kStripConsumers);
        if (tid == 0) {
            double t = 0.0;
#pragma unroll
            for (int w = 0; w < kStripConsumers / 32; ++w) t += red_smem[w];
            P.resid_partials[blockIdx.x] = t;
        }
    }
}

}  // namespace stencil
}  // namespace b200
