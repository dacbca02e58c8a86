// Temporally blocked stencil + halo kernel, third generation: column STRIPS marched from top to
// bottom with the time levels carried in REGISTERS (radius-1 star, T = 2..4 time steps per sweep).
//
// Why (ncu + bench of the tile kernel with 3 levels per sweep, profiles/README.md): every thread
// of the tile kernel owns 2 columns x 16 rows of a 128 x 64 tile and recomputes the intermediate
// levels of its whole neighbourhood: 7.0 cell updates per output cell instead of 3.  That made the
// kernel FP64/LDS bound (0.69 of its HBM roofline) and, sustained, POWER bound (1000 W cap,
// clocks down to 1845 MHz).  Here
//   * a CTA owns a strip of up to NWARPS*32*CPT columns and marches down it; every input row is
//     loaded ONCE (one 1-D bulk copy per row into a ring of 4-row chunks);
//   * a thread owns CPT (6) adjacent columns and keeps, for every time level k < LV, the three
//     newest rows of a window that is (LV-k) columns wider on either side.  Each new input row
//     yields one row of every level; the output row lags LV rows behind the input row.  Nothing
//     intermediate touches shared or global memory, no barrier separates the levels, and there is
//     NO vertical redundancy: (10+8+6)/6 = 4.0 cell updates per output cell for 3 levels;
//   * 6 columns = 48 bytes per lane makes the 16-byte shared-memory loads of a quarter warp hit
//     eight different bank groups (3*lane mod 8 is a permutation): no bank conflicts without any
//     swizzle (4 columns per lane would be 2-way conflicted).
// The halo protocol is the one of the other two kernels (per-direction epoch counters, acquire
// before a boundary segment is loaded, ghost columns patched from the landing buffers, release
// after the pushes), and so is the work partition of the strip kernel: every CTA does its boundary
// segments FIRST, then a balanced run of interior chunks.  Edge rows and edge columns are staged in
// shared memory and leave as full 128-byte lines (rows) / contiguous bursts (columns).
// Bit-identical to T single steps: every value is formed with the operation order of the
// single-step kernel.
//
// Reference counterpart: none -- the reference's Compute() is an empty stub
// (stencil2d/mpi-2d-stencil-subarray-cuda.cu:32-33) and its exchange is host-driven
// (stencil2d/stencil2D.h:363-377).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <type_traits>

#include "kernels/ptx.cuh"
#include "stencil/stencil_kernel.cuh"
#include "stencil/stencil_strip_kernel.cuh"   // Segment

namespace b200 {
namespace stencil {

constexpr int kTsChunkH = 4;         // rows per chunk == granularity of the segments
constexpr int kTsSlotRows = 2;       // rows per ring slot (finer than a chunk: slots are handed back,
                                     // and refilled, twice per chunk)
constexpr int kTsRA = 4;             // halo columns loaded on either side (keeps 16-byte alignment)
constexpr int kTsColStageRows = 32;  // edge-column rows staged before one burst
constexpr int kTsOutSlots = 6;       // output rows staged in shared memory (one row per slot)
static_assert(kTsChunkH == kChunkH, "the strip partition builder is shared");

template <int CPT, int NWARPS>
struct TsGeo {
    static constexpr int CONS = NWARPS * 32;             // consumer threads
    // + a producer WARPGROUP (only its first lane works).  Registers are handed out per SM
    // sub-partition, so a 9th warp would cap every thread at 168 registers; with a whole extra
    // warpgroup the producer side shrinks to 72 registers (setmaxnreg.dec) and the consumers grow
    // to 216 (setmaxnreg.inc) -- the windows of three time levels need ~200.
    static constexpr int THREADS = CONS + 128;
    static constexpr int PRODUCER_REGS = 72, CONSUMER_REGS = 216;
    static constexpr int SWMAX = CONS * CPT;             // widest strip (cells)
    static constexpr int PITCH = SWMAX + 2 * kTsRA;      // shared-memory row pitch (elements)
    static constexpr int SLOT_BYTES = PITCH * kTsSlotRows * 8;
    static_assert(SLOT_BYTES % 128 == 0, "slot alignment");
};

struct TsParams {
    LaunchParams lp;          // same fields as the tile kernel (tilesX/tilesY unused)
    const Segment* segs;      // all segments, grouped per CTA
    const int* seg_begin;     // [gridDim.x + 1] offsets into segs
    int strip_w;              // cells per strip (a multiple of CPT, <= SWMAX)
    int nslots;               // ring depth
    int resid;                // accumulate the residual of the last level
    // debug bit 32 (B2C_STENCIL_DEBUG): per-CTA globaltimer stamps of the LAST sweep,
    // [cta][0] start, [1] end of the boundary segments, [2] end, [3] ns the producer spent
    // waiting for halo flags; nullptr = off
    unsigned long long* stamps;
    // Tail work of the sweep, done by the LAST CTA to finish (ticket counter that wraps to 0, the
    // single-launch reduction of mpicuda4.cu:157-185 in the reference): sum the per-CTA residual
    // partials into *resid_out (resid != 0) and advance the device step counter by `bump`.
    // tail_ticket == nullptr: neither (the host launches separate kernels for both).
    unsigned* tail_ticket;
    double* resid_out;
    unsigned long long* ctr_w;
    unsigned long long bump;
};

// Dynamic shared memory: nslots ring slots, then the output stage (kTsOutSlots rows of SWMAX
// cells), then the mbarriers (2 per ring slot, 2 per output slot, 1 "segment done").
template <int T, int CPT, int NWARPS>
constexpr size_t ts_smem_bytes(int nslots) {
    return (size_t)nslots * TsGeo<CPT, NWARPS>::SLOT_BYTES +
           (size_t)kTsOutSlots * TsGeo<CPT, NWARPS>::SWMAX * 8 +
           (2 * (size_t)nslots + 2 * kTsOutSlots + 1) * sizeof(uint64_t);
}

// T: ghost width of the plan (= levels of a full sweep); LV: levels computed by THIS sweep
// (LV < T: the partial sweep that ends a run whose step count is not a multiple of T).
template <int T, int LV, int CPT, int NWARPS>
__global__ void __launch_bounds__(TsGeo<CPT, NWARPS>::THREADS, 1)
stencil_tstrip_kernel(const __grid_constant__ TsParams SP) {
    using G = TsGeo<CPT, NWARPS>;
    constexpr int CH = kTsChunkH, SR = kTsSlotRows, RA = kTsRA;
    static_assert(CH % SR == 0 && SR % 2 == 0, "slot rows");
    static_assert(T >= 2 && T <= 4 && LV >= 1 && LV <= T, "levels");
    static_assert(CPT % 2 == 0 && CPT >= T && RA >= T, "columns per thread");
    const LaunchParams& P = SP.lp;
    extern __shared__ __align__(128) unsigned char smem[];
    const int nslots = SP.nslots;
    double* ostage = reinterpret_cast<double*>(smem + (size_t)nslots * G::SLOT_BYTES);
    uint64_t* full = reinterpret_cast<uint64_t*>(ostage + kTsOutSlots * G::SWMAX);
    uint64_t* empty = full + nslots;
    uint64_t* ofull = empty + nslots;          // output slot written by all consumer warps
    uint64_t* oempty = ofull + kTsOutSlots;    // output slot read out by the store warp
    uint64_t* segdone = oempty + kTsOutSlots;  // consumers finished a boundary segment
    constexpr int NOS = kTsOutSlots;
    __shared__ double red_smem[NWARPS];
    __shared__ double colstage[2][kTsColStageRows][T];   // [west|east][row][k]

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int sbeg = SP.seg_begin[blockIdx.x], send = SP.seg_begin[blockIdx.x + 1];

    if (tid == 0) {
        for (int s = 0; s < nslots; ++s) {
            ptx::mbar_init(&full[s], 1);
            ptx::mbar_init(&empty[s], NWARPS);
        }
        for (int s = 0; s < NOS; ++s) {
            ptx::mbar_init(&ofull[s], NWARPS);
            ptx::mbar_init(&oempty[s], 1);
        }
        ptx::mbar_init(segdone, 1);
        ptx::fence_mbar_init();
    }
    __syncthreads();

    if (warp >= NWARPS) {
        // ============ producer warp: flag acquire, one bulk copy per row, ghost-column patch ============
        ptx::setmaxnreg_dec<G::PRODUCER_REGS>();
        if (warp == NWARPS) {
            int slot = 0;
            uint32_t phase = 0;
            const int rows_total = P.H + 2 * P.g;
            for (int s = sbeg; s < send; ++s) {
                const Segment sg = SP.segs[s];
                const unsigned act = sg.dirs;
                if (act && !(P.debug & 2)) {
                    const uint64_t tw0 = SP.stamps ? ptx::globaltimer_ns() : 0;
                    if (lane < 8 && (act >> lane & 1u)) {      // one lane per direction
                        const uint64_t t0 = ptx::globaltimer_ns();
                        uint32_t spins = 0;
                        const uint64_t need = (*P.step_ctr + P.step_off) * P.signals_per_step[lane];
                        while (ptx::ld_acquire_sys(P.my_flags + lane) < need) {
                            if ((++spins & 1023u) == 0 && ptx::globaltimer_ns() - t0 > P.timeout_ns) {
                                atomicOr(P.status, 1u);
                                break;
                            }
                        }
                    }
                    // ghost rows were written by another GPU's generic proxy; the bulk copies
                    // below read them through the async proxy
                    ptx::fence_proxy_async();
                    __syncwarp();
                    if (SP.stamps && lane == 0)
                        SP.stamps[4 * blockIdx.x + 3] += ptx::globaltimer_ns() - tw0;
                }
                const int x0 = sg.strip * SP.strip_w;
                const int sw = min(SP.strip_w, P.W - x0);
                const bool west = x0 == 0, east = x0 + sw == P.W;
                const bool patch = (act & kColDirs) != 0;
                // Ghost COLUMNS of an edge strip do not come from the grid but from the landing
                // buffers: the row copies leave them out and this warp fills them in with plain
                // stores before it arrives on the slot's barrier (the consumers see no difference).
                const int lo = (patch && west) ? RA : 0, hi = (patch && east) ? RA : 0;
                const uint32_t row_bytes = (uint32_t)(sw + 2 * RA - lo - hi) * 8u;
                const double* col0 = P.in + P.lead + x0 - RA + lo;
                for (int j = sg.j0 - 1; j <= sg.j1; ++j) {
#pragma unroll 1
                    for (int h = 0; h < CH / SR; ++h) {
                        ptx::mbar_wait(&empty[slot], phase ^ 1u);
                        // rows of the slot that exist in the padded buffer (the chunk above the
                        // first and below the last core chunk stick out by 4 - T rows)
                        const int pr0 = P.g + j * CH + h * SR;
                        const int ra = max(0, -pr0), rb = min(SR, rows_total - pr0);
                        const int nrows = max(0, rb - ra);
                        double* ch = reinterpret_cast<double*>(smem + (size_t)slot * G::SLOT_BYTES);
                        if (lane == 0) {
                            ptx::mbar_expect_tx(&full[slot], (uint32_t)nrows * row_bytes);
                            for (int r = ra; r < rb; ++r)
                                ptx::bulk_load_1d(ch + (size_t)r * G::PITCH + lo,
                                                  col0 + (size_t)(pr0 + r) * P.pitch, row_bytes, &full[slot]);
                        }
                        if (patch) {
                            const int e = lane;
                            if (e < 2 * SR * T) {
                                const int side = e / (SR * T), rem = e % (SR * T);
                                const int r = rem / T, q = rem % T;
                                const int pr = pr0 + r;                 // padded row
                                if (((side == 0 && west) || (side == 1 && east)) && pr >= 0 && pr < rows_total) {
                                    const int vd = pr < T ? (side ? NE : NW)
                                                          : (pr >= P.H + T ? (side ? SE : SW) : (side ? E_ : W_));
                                    const double v = (act >> vd & 1u)
                                        ? P.my_land[side][land_index(pr, T, T) + q]
                                        : P.in[(size_t)pr * P.pitch + P.lead + (side ? P.W + q : q - T)];
                                    ch[r * G::PITCH + (side ? RA + sw + q : RA - T + q)] = v;
                                }
                            }
                            // plain smem writes vs the bulk copy that refills this slot later
                            ptx::fence_proxy_async_smem();
                            __syncwarp();
                        }
                        if (lane == 0) ptx::mbar_arrive(&full[slot]);
                        if (++slot == nslots) { slot = 0; phase ^= 1u; }
                    }
                }
            }
        }
        if (warp == NWARPS + 1 && lane == 0) {
            // ============ store thread: staged output rows -> global memory (one bulk copy per
            // row; edge rows also into the neighbour's ghost rows), then the halo signals ============
            // The consumers never issue a global store: plain STG made the next row's arithmetic
            // wait for the store to read its registers, per-warp bulk copies back-pressured the
            // issuing warps (ncu, profiles/README.md).
            uint64_t n = 0, freed = 0;          // rows issued / output slots handed back
            uint32_t segph = 0;
            unsigned pend[8] = {0, 0, 0, 0, 0, 0, 0, 0};   // signals owed per direction
            int npend = 0;
            auto free_upto = [&](uint64_t m) {
                for (; freed < m; ++freed) ptx::mbar_arrive(&oempty[freed % NOS]);
            };
            for (int s = sbeg; s < send; ++s) {
                const Segment sg = SP.segs[s];
                const unsigned act = sg.dirs;
                const int x0 = sg.strip * SP.strip_w;
                const int sw = min(SP.strip_w, P.W - x0);
                const int ylo = sg.j0 * CH, yhi = min(sg.j1 * CH, P.H);
                const size_t col = (size_t)P.lead + x0;
                for (int y = ylo; y < yhi; ++y, ++n) {
                    const int os = (int)(n % NOS);
                    ptx::mbar_wait(&ofull[os], (uint32_t)(n / NOS) & 1u);
                    const double* src = ostage + (size_t)os * G::SWMAX;
                    ptx::bulk_store_1d(P.out + (size_t)(y + P.g) * P.pitch + col, src, sw * 8);
                    if ((act >> N & 1u) && y < T)
                        ptx::bulk_store_1d(P.peer_out[N] + (size_t)(y + P.g + P.H) * P.pitch + col, src, sw * 8);
                    if ((act >> S & 1u) && y >= P.H - T)
                        ptx::bulk_store_1d(P.peer_out[S] + (size_t)(y + P.g - P.H) * P.pitch + col, src, sw * 8);
                    ptx::bulk_commit();
                    ptx::bulk_wait_read<1>();   // every row but the newest has left shared memory
                    free_upto(n);
                }
                if (act) {
                    // the consumers' own peer stores (edge columns) of this segment are done
                    ptx::mbar_wait(segdone, segph);
                    segph ^= 1u;
                    for (int d = 0; d < 8; ++d) pend[d] += act >> d & 1u;
                    npend += 1;
                }
                // Signal ONCE per run of boundary segments (they come first in every CTA's list),
                // not per segment: waiting for the remote writes of a segment to complete costs an
                // NVLink round trip during which no output row leaves the stage.  The neighbours
                // need these signals a whole sweep later.
                const bool next_act = s + 1 < send && SP.segs[s + 1].dirs != 0;
                if (npend && !next_act) {
                    ptx::bulk_wait<0>();        // my bulk copies into the neighbours' ghost rows
                    free_upto(n);
                    ptx::fence_acq_rel_sys();
                    for (int d = 0; d < 8; ++d)
                        if (pend[d]) { ptx::red_add_relaxed_sys(P.peer_flag[d], (uint64_t)pend[d]); pend[d] = 0; }
                    npend = 0;
                }
            }
            ptx::bulk_wait<0>();      // the output stage must outlive its last bulk copies
        }
        return;
    }

    // =============================== consumers ===============================
    ptx::setmaxnreg_inc<G::CONSUMER_REGS>();
    constexpr int W0 = CPT + 2 * RA;    // level-0 window: columns c0-RA .. c0+CPT+RA-1
    double c[5];
#pragma unroll
    for (int k = 0; k < 5; ++k) c[k] = __ldg(P.coef + k);
    const int c0 = CPT * tid;           // my first column inside the strip
    // residual: NRA accumulators -- a single one made every output row end in a chain of CPT
    // dependent DFMAs that two warps per scheduler could not hide (+6 % per sweep); one per column
    // spills
    constexpr int NRA = 3;
    double racc[NRA];
#pragma unroll
    for (int q = 0; q < NRA; ++q) racc[q] = 0.0;
    int slot = 0;                       // mirrors the producer's ring position
    uint32_t phase = 0;
    int os = 0;                         // mirrors the store thread's output slot
    uint32_t ophase = 0;
    // lv[k][row % 3][col]: the three newest rows of time level k.  Level k covers columns
    // c0-(LV-k) .. c0+CPT-1+(LV-k) (level 0: the aligned window above).  All indices are
    // compile-time constants after unrolling, so the windows live in registers.
    double lv[LV][3][W0];
#pragma unroll
    for (int k = 0; k < LV; ++k)
#pragma unroll
        for (int a = 0; a < 3; ++a)
#pragma unroll
            for (int b = 0; b < W0; ++b) lv[k][a][b] = 0.0;

    if (SP.stamps && tid == 0) {
        SP.stamps[4 * blockIdx.x + 0] = ptx::globaltimer_ns();
        SP.stamps[4 * blockIdx.x + 1] = 0;
    }
    for (int s = sbeg; s < send; ++s) {
        const Segment sg = SP.segs[s];
        const unsigned act = sg.dirs;
        if (SP.stamps && tid == 0 && act == 0 && SP.stamps[4 * blockIdx.x + 1] == 0)
            SP.stamps[4 * blockIdx.x + 1] = ptx::globaltimer_ns();      // boundary segments done
        const int x0 = sg.strip * SP.strip_w;
        const int sw = min(SP.strip_w, P.W - x0);              // valid columns of this strip (even)
        const bool west = x0 == 0, east = x0 + sw == P.W;
        const bool patch = (act & kColDirs) != 0;
        const int ylo = sg.j0 * CH, yhi = min(sg.j1 * CH, P.H);  // output rows of this segment
        int staged0 = -1;                                      // first core row held in colstage

        // Burst the staged edge columns [staged0, upto) into the neighbours' landing buffers.
        auto flush_cols = [&](int upto) {
            if (staged0 < 0 || upto <= staged0) return;
            ptx::named_bar_sync(4, G::CONS);
            if (warp == 0 && !(P.debug & 1)) {
                for (int d = 0; d < 8; ++d) {
                    const int dy = dir_dy(d), dx = dir_dx(d);
                    if (dx == 0 || !(act >> d & 1u)) continue;
                    if ((dx < 0 && !west) || (dx > 0 && !east)) continue;
                    int ya = staged0, yb = upto;                  // core rows
                    if (dy < 0) yb = min(yb, P.g);
                    if (dy > 0) ya = max(ya, P.H - P.g);
                    const int n = (yb - ya) * T;
                    if (n <= 0) continue;
                    double* dst = P.peer_land[d] + land_index(ya + P.g - dy * P.H, T, T);
                    const double* src = &colstage[dx < 0 ? 0 : 1][ya - staged0][0];
                    for (int e = lane; e < n; e += 32) dst[e] = src[e];
                }
            }
            ptx::named_bar_sync(4, G::CONS);
            staged0 = -1;
        };
        // One input row: load it, advance every level by one row, emit output row y.
        // PH = (input row) mod 3 selects the register set of every window (compile time).  ACT:
        // the segment touches the tile edge (staging code present); RS: residual wanted.
        auto row = [&](auto phc, auto actc, auto rsc, auto pinc, const double* srow, double* orow_s, int y) {
            constexpr int PH = decltype(phc)::value;
            constexpr bool ACT = decltype(actc)::value, RS = decltype(rsc)::value, PIN = decltype(pinc)::value;
#pragma unroll
            for (int m = 0; m < W0 / 2; ++m) {
                // the outermost pairs are only half used (columns c0-LV .. c0+CPT-1+LV are needed):
                // keep them 16-byte loads, see ptx::lds_v2f64
                const bool half_used = (2 * m < RA - LV) || (2 * m + 1 >= W0 - (RA - LV));
                const double2 v = half_used ? ptx::lds_v2f64(srow + 2 * m)
                                            : *reinterpret_cast<const double2*>(srow + 2 * m);
                lv[0][PH][2 * m] = v.x;
                lv[0][PH][2 * m + 1] = v.y;
            }
            double outv[CPT];
            double ctr[CPT];
#pragma unroll
            for (int k = 1; k <= LV; ++k) {
                // level k-1 rows i-k-1 (north), i-k (centre), i-k+1 (south) -> level k row i-k
                const int sN = (PH + 30 - (k + 1)) % 3, sM = (PH + 30 - k) % 3, sS = (PH + 30 - (k - 1)) % 3;
                // column q of level k is c0-(LV-k)+q; in level k-1 its west neighbour has index
                // q+off, itself q+off+1, its east neighbour q+off+2
                const int off = (k == 1) ? RA - LV : 0;
                const int ncol = CPT + 2 * (LV - k);
                // PIN: beyond an edge of the PHYSICAL domain (non-periodic grid, no neighbour) the
                // ghost cells are boundary data, fixed in time: an intermediate level there keeps
                // the value of the level below instead of being updated.  Row i-k of level k:
                bool pin_row = false;
                if (PIN && k < LV) {
                    const int yk = y + LV - k;                         // core row of this level's new row
                    pin_row = ((sg.pin & 1u) && yk < 0) || ((sg.pin & 2u) && yk >= P.H);
                }
#pragma unroll
                for (int q = 0; q < ncol; ++q) {
                    double a = c[0] * lv[k - 1][sM][q + off + 1];
                    a = fma(c[1], lv[k - 1][sM][q + off], a);          // west
                    a = fma(c[2], lv[k - 1][sM][q + off + 2], a);      // east
                    a = fma(c[3], lv[k - 1][sN][q + off + 1], a);      // north
                    a = fma(c[4], lv[k - 1][sS][q + off + 1], a);      // south
                    if (PIN && k < LV) {
                        const int xk = x0 + c0 - (LV - k) + q;         // tile column of this value
                        if (pin_row || ((sg.pin & 4u) && xk < 0) || ((sg.pin & 8u) && xk >= P.W))
                            a = lv[k - 1][sM][q + off + 1];
                    }
                    if (k < LV) {
                        lv[k < LV ? k : 0][sM][q] = a;
                    } else {
                        outv[q < CPT ? q : 0] = a;
                        if (RS) ctr[q < CPT ? q : 0] = lv[k - 1][sM][q + off + 1];
                    }
                }
            }
            if (y < ylo || y >= yhi) return;              // warm-up / look-ahead rows (CTA-uniform)
            // The output row goes to this warp's stage in shared memory (48 bytes per lane: no
            // bank conflicts) and leaves as ONE bulk copy per warp row, issued by lane 0 (see
            // step()).  Plain STG.128 here made the next row's arithmetic wait for the stores to
            // read their registers (ncu: long-scoreboard stalls on the first DMUL after each STG).
#pragma unroll
            for (int p = 0; p < CPT / 2; ++p)
                *reinterpret_cast<double2*>(orow_s + 2 * p) = make_double2(outv[2 * p], outv[2 * p + 1]);
            if (RS) {
#pragma unroll
                for (int q = 0; q < CPT; ++q)
                    if (c0 + q < sw) {
                        const double d0 = outv[q] - ctr[q];
                        racc[q % NRA] = fma(d0, d0, racc[q % NRA]);
                    }
            }
            if (ACT) {
                if (patch) {
                    const int sr = y - staged0;
#pragma unroll
                    for (int q = 0; q < CPT; ++q) {
                        const int xx = c0 + q;
                        if (west && xx < T) colstage[0][sr][xx] = outv[q];
                        if (east && xx >= sw - T && xx < sw) colstage[1][sr][xx - (sw - T)] = outv[q];
                    }
                }
            }
        };

        // The march: input rows (j0-1)*4 .. (j1+1)*4-1, three per loop trip (the period of the
        // 3-row windows, so every register index is a compile-time constant and the loop body
        // stays small enough for the instruction cache); ring slots are taken and handed back
        // every fourth row.
        const int i_begin = (sg.j0 - 1) * CH, i_end = (sg.j1 + 1) * CH;
        const double* base = nullptr;
        auto step = [&](auto phc, auto actc, auto rsc, auto pinc, int i) {
            constexpr bool ACT = decltype(actc)::value;
            const int r = (i - i_begin) & (CH - 1);
            if ((r & (SR - 1)) == 0) {
                ptx::mbar_wait(&full[slot], phase);
                base = reinterpret_cast<const double*>(smem + (size_t)slot * G::SLOT_BYTES) + c0;
            }
            if (ACT) {
                if (r == 0 && patch && staged0 < 0 && i - LV + CH > ylo) staged0 = max(i - LV, ylo);
            }
            const int y = i - LV;
            const bool valid = y >= ylo && y < yhi;        // CTA-uniform
            if (valid) ptx::mbar_wait(&oempty[os], ophase ^ 1u);
            row(phc, actc, rsc, pinc, base + (r & (SR - 1)) * G::PITCH, ostage + (size_t)os * G::SWMAX + c0, y);
            if (valid) {
                // my part of the output row is staged: order it before the store thread's bulk copy
                ptx::fence_proxy_async_smem();
                __syncwarp();
                if (lane == 0) ptx::mbar_arrive(&ofull[os]);
                if (++os == NOS) { os = 0; ophase ^= 1u; }
            }
            if ((r & (SR - 1)) == SR - 1) {
                // all shared-memory reads of this slot are done: hand it back to the producer
                __syncwarp();
                if (lane == 0) ptx::mbar_arrive(&empty[slot]);
                if (++slot == nslots) { slot = 0; phase ^= 1u; }
            }
            if (r == CH - 1) {
                if (ACT) {
                    const int yend = min(i - LV + 1, yhi);
                    if (patch && staged0 >= 0 &&
                        (yend - staged0 + CH > kTsColStageRows || i + 1 == i_end))
                        flush_cols(yend);
                }
            }
        };
        auto march = [&](auto actc, auto rsc, auto pinc) {
            for (int i = i_begin; i < i_end; i += 3) {
                step(std::integral_constant<int, 0>{}, actc, rsc, pinc, i);
                if (i + 1 >= i_end) break;
                step(std::integral_constant<int, 1>{}, actc, rsc, pinc, i + 1);
                if (i + 2 >= i_end) break;
                step(std::integral_constant<int, 2>{}, actc, rsc, pinc, i + 2);
            }
        };
        if (sg.pin) {          // touches the physical boundary (the boundary code path copes with act == 0)
            if (SP.resid) march(std::true_type{}, std::true_type{}, std::true_type{});
            else march(std::true_type{}, std::false_type{}, std::true_type{});
        } else if (act) {
            if (SP.resid) march(std::true_type{}, std::true_type{}, std::false_type{});
            else march(std::true_type{}, std::false_type{}, std::false_type{});
        } else {
            if (SP.resid) march(std::false_type{}, std::true_type{}, std::false_type{});
            else march(std::false_type{}, std::false_type{}, std::false_type{});
        }

        if (act) {
            // every consumer's peer stores (edge columns) happen-before the store thread's release
            ptx::named_bar_sync(1, G::CONS);
            if (tid == 0) ptx::mbar_arrive(segdone);
        }
    }

    if (SP.stamps && tid == 0) {
        const unsigned long long t = ptx::globaltimer_ns();
        if (SP.stamps[4 * blockIdx.x + 1] == 0) SP.stamps[4 * blockIdx.x + 1] = t;
        SP.stamps[4 * blockIdx.x + 2] = t;
    }
    if (SP.resid) {
        // deterministic per-CTA partial: lanes -> warps -> thread 0
#pragma unroll
        double rsum = 0.0;
#pragma unroll
        for (int q = 0; q < NRA; ++q) rsum += racc[q];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) rsum += __shfl_xor_sync(0xffffffffu, rsum, o);
        if (lane == 0) red_smem[warp] = rsum;
        ptx::named_bar_sync(2, G::CONS);
        if (tid == 0) {
            double t = 0.0;
#pragma unroll
            for (int w = 0; w < NWARPS; ++w) t += red_smem[w];
            P.resid_partials[blockIdx.x] = t;
        }
    }
    if (SP.tail_ticket) {
        static_assert(G::CONS == 256, "the tail reduction mirrors sum_partials_kernel<<<1, 256>>>");
        __shared__ int is_last;
        if (tid == 0) {
            __threadfence();                      // my partial before my ticket
            const unsigned last = gridDim.x - 1;
            is_last = atomicInc(SP.tail_ticket, last) == last;
        }
        ptx::named_bar_sync(2, G::CONS);
        if (is_last) {
            // Every other CTA has taken its ticket, so its consumers are done, so its producer
            // has read the step counter for the last time: the counter may move now.
            __threadfence();
            if (SP.resid) {
                // same summation tree as sum_partials_kernel: the two paths agree bit for bit.
                // The ring slots are idle by now (every load was consumed).
                double* sh = reinterpret_cast<double*>(smem);
                double a = 0.0;
                for (int i = tid; i < (int)gridDim.x; i += G::CONS) a += __ldcg(P.resid_partials + i);
                sh[tid] = a;
                ptx::named_bar_sync(2, G::CONS);
#pragma unroll 1
                for (int o = G::CONS / 2; o > 0; o >>= 1) {
                    if (tid < o) sh[tid] += sh[tid + o];
                    ptx::named_bar_sync(2, G::CONS);
                }
                if (tid == 0) *SP.resid_out = sh[0];
            }
            if (tid == 0 && SP.bump) *SP.ctr_w += SP.bump;
        }
    }
}

}  // namespace stencil
}  // namespace b200
