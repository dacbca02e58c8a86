// C++ face of the device-side components: symmetric (peer-mapped) memory, the fused
// stencil + halo plan, the fused dot + allreduce plan.  Header-only over the C ABI
// (b200comm.h) and the communicator API (comm.hpp): a C++ program gets the flagship kernels
// without Python.
//
// Reference capability: the GPU programs of the reference are C++ main()s that hand device
// pointers to MPI (stencil2d/mpi-2d-stencil-subarray-cuda.cu:77-179, mpicuda4.cu:189-395);
// these classes are what such a program is written against here.  One rank per GPU: every
// rank's kernel of step s may wait for a neighbour's kernel of step s-1, so ranks that share a
// GPU must not use FusedStencil with more than one rank (use the phase-split calls of the C ABI
// for that, as Python's StencilGroup does).
#pragma once
#include <unistd.h>

#include <cmath>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "b200comm.h"
#include "b200comm/comm.hpp"
#include "b200comm/stencil2d.hpp"

namespace b200 {
namespace device {

// Throws (or aborts, per error mode) with the library's message when a C ABI call failed.
inline void check(int rc, const char* what) {
    if (rc != 0) ::b200::fail_comm(rc, std::string(what) + ": " + b2c_last_error(), __FILE__, __LINE__);
}
#define B200_DEVICE_OK(call) ::b200::device::check((call), #call)

// One device allocation per rank, mapped into every rank of `c` (CUDA IPC between processes,
// plain pointers between thread ranks of one process).  Collective.
class SymmetricBlock {
public:
    SymmetricBlock() = default;
    SymmetricBlock(const comm::Comm& c, size_t bytes) : comm_(c), bytes_(bytes) {
        const int n = c.size(), me = c.rank();
        B200_DEVICE_OK(b2c_malloc(&local_, bytes));
        struct Rec { unsigned char handle[B2C_IPC_HANDLE_BYTES]; unsigned long long ptr; long long pid; } mine{};
        B200_DEVICE_OK(b2c_ipc_export(local_, mine.handle));
        mine.ptr = reinterpret_cast<unsigned long long>(local_);
        mine.pid = (long long)::getpid();
        std::vector<Rec> all((size_t)n);
        const Datatype byte = Datatype::basic(Basic::BYTE);
        c.allgather(&mine, (int64_t)sizeof(Rec), byte, all.data(), (int64_t)sizeof(Rec), byte);
        peers_.assign((size_t)n, nullptr);
        opened_.assign((size_t)n, false);
        for (int r = 0; r < n; ++r) {
            if (r == me) { peers_[r] = local_; continue; }
            if (all[r].pid == mine.pid) {            // thread rank of this process: same address space
                peers_[r] = reinterpret_cast<void*>(all[r].ptr);
            } else {
                B200_DEVICE_OK(b2c_ipc_open(all[r].handle, &peers_[r]));
                opened_[r] = true;
            }
        }
        c.barrier();
    }
    SymmetricBlock(const SymmetricBlock&) = delete;
    SymmetricBlock& operator=(const SymmetricBlock&) = delete;
    SymmetricBlock(SymmetricBlock&& o) noexcept { *this = std::move(o); }
    SymmetricBlock& operator=(SymmetricBlock&& o) noexcept {
        if (this != &o) {
            release();
            comm_ = o.comm_; bytes_ = o.bytes_; local_ = o.local_;
            peers_ = std::move(o.peers_); opened_ = std::move(o.opened_);
            o.local_ = nullptr; o.peers_.clear(); o.opened_.clear();
        }
        return *this;
    }
    ~SymmetricBlock() { release(); }

    void* local() const { return local_; }
    void* peer(int rank) const { return rank < 0 ? nullptr : peers_[(size_t)rank]; }
    size_t bytes() const { return bytes_; }
    const comm::Comm& comm() const { return comm_; }

    // Collective: nobody unmaps or frees while a peer may still be storing into the block.
    void release() {
        if (!local_) return;
        if (!comm_.is_null()) comm_.barrier();
        for (size_t r = 0; r < peers_.size(); ++r)
            if (opened_[r]) b2c_ipc_close(peers_[r]);
        if (!comm_.is_null()) comm_.barrier();
        b2c_free(local_);
        local_ = nullptr;
    }

private:
    comm::Comm comm_;
    size_t bytes_ = 0;
    void* local_ = nullptr;
    std::vector<void*> peers_;
    std::vector<bool> opened_;
};

// Fused 2-D stencil + halo exchange on a 2-D Cartesian communicator (dimension 0 = rows).
// kind: Star (radius 1..4), Box (radius 1..2), or temporal blocking (`temporal` = 2..4 time
// steps per kernel sweep of the radius-1 star; 4 needs a fully periodic grid).
class FusedStencil {
public:
    enum Kind { Star = 0, Box = 1 };

    FusedStencil(const comm::Comm& cart, int H, int W, int radius = 1, Kind kind = Star, int temporal = 1,
                 void* stream = nullptr)
        : cart_(cart), H_(H), W_(W), radius_(radius), temporal_(temporal), stream_(stream) {
        if (!cart.is_cart() || cart.cart_dims().size() != 2)
            ::b200::fail_comm(ERR_ARG, "FusedStencil needs a 2-D Cartesian communicator", __FILE__, __LINE__);
        if (temporal != 1) {
            bool periodic = true;
            for (int p : cart.cart_periods()) periodic = periodic && p != 0;
            // 2 or 3 levels also run on non-periodic grids (ghost cells beyond the global edge are
            // boundary data, kept fixed through the intermediate levels by the strip kernel)
            if (temporal < 2 || temporal > 4 || radius != 1 || kind != Star || (temporal == 4 && !periodic))
                ::b200::fail_comm(ERR_ARG, "temporal blocking: 2..4 levels of the radius-1 star; 4 levels "
                                           "need a fully periodic grid", __FILE__, __LINE__);
        }
        const int g = radius * temporal;
        B200_DEVICE_OK(b2c_stencil_layout(W, H, g, &lay_));
        block_ = SymmetricBlock(cart, lay_.total_bytes);
        void* nbr[8];
        for (int d = 0; d < 8; ++d) {
            int dx, dy;
            stencil2d::cell_offset((stencil2d::GridCellID)d, &dx, &dy);
            const int r = stencil2d::offset_rank(cart, dx, dy);
            nbr[d] = r == comm::PROC_NULL ? nullptr : block_.peer(r);
        }
        B200_DEVICE_OK(b2c_stencil_create(&h_, W, H, g, temporal > 1 ? 2 : (int)kind, block_.local(), nbr, 1));
    }
    FusedStencil(const FusedStencil&) = delete;
    FusedStencil& operator=(const FusedStencil&) = delete;
    ~FusedStencil() {
        if (h_) {
            b2c_stream_synchronize(stream_);
            b2c_stencil_destroy(h_);
        }
    }

    const b2c_stencil_layout_t& layout() const { return lay_; }
    int ghost() const { return lay_.ghost; }
    long long steps_done() const { return steps_; }

    // Both buffers: whole padded tile = ghost_value, then core = core_value.
    void fill(double core_value, double ghost_value) {
        for (int b = 0; b < 2; ++b) {
            B200_DEVICE_OK(b2c_stencil_fill(h_, b, 0, ghost_value, stream_));
            B200_DEVICE_OK(b2c_stencil_fill(h_, b, 1, core_value, stream_));
        }
    }
    // Copy an H x W row-major host array into / out of the core of the current buffer.
    void upload_core(const double* host) { copy_core(const_cast<double*>(host), true); }
    void download_core(double* host) { copy_core(host, false); }
    // (H+2g) x (W+2g) row-major: core + ghost ring of the current buffer.
    void download_tile(double* host) {
        const b2c_stencil_info_t inf = info();
        const int g = lay_.ghost, tw = W_ + 2 * g;
        std::vector<double> padded((size_t)lay_.rows * lay_.pitch);
        B200_DEVICE_OK(b2c_memcpy_async(padded.data(), inf.buf[inf.cur], padded.size() * sizeof(double), 2, stream_));
        B200_DEVICE_OK(b2c_stream_synchronize(stream_));
        for (int y = 0; y < H_ + 2 * g; ++y)
            std::memcpy(host + (size_t)y * tw, padded.data() + (size_t)y * lay_.pitch + lay_.lead - g, sizeof(double) * tw);
    }
    // Inverse of download_tile: core AND ghost ring of the current buffer from a host array.
    void upload_tile(const double* host) {
        const b2c_stencil_info_t inf = info();
        const int g = lay_.ghost, tw = W_ + 2 * g;
        std::vector<double> padded((size_t)lay_.rows * lay_.pitch);
        const size_t nbytes = padded.size() * sizeof(double);
        B200_DEVICE_OK(b2c_memcpy_async(padded.data(), inf.buf[inf.cur], nbytes, 2, stream_));
        B200_DEVICE_OK(b2c_stream_synchronize(stream_));
        for (int y = 0; y < H_ + 2 * g; ++y)
            std::memcpy(padded.data() + (size_t)y * lay_.pitch + lay_.lead - g, host + (size_t)y * tw, sizeof(double) * tw);
        B200_DEVICE_OK(b2c_memcpy_async(inf.buf[inf.cur], padded.data(), nbytes, 1, stream_));
        B200_DEVICE_OK(b2c_stream_synchronize(stream_));
    }
    // Checkpoint / resume of this rank's state: geometry, step count, coefficients, the tile with
    // its ghost ring (one binary file per rank).  The reference only has the per-rank text dump
    // (stencil2d/mpi-2d-stencil-subarray-cuda.cu:114-119,164-176).  After load_checkpoint() on a
    // fresh plan of the same geometry call exchange() on all ranks; the continuation is
    // bit-identical.  Returns the step count of the checkpoint.
    void save_checkpoint(const std::string& path) {
        const int g = lay_.ghost;
        std::vector<double> tile((size_t)(H_ + 2 * g) * (W_ + 2 * g));
        download_tile(tile.data());
        const long long hdr[8] = {kCheckpointMagic, H_, W_, radius_, temporal_, steps_, (long long)coef_.size(),
                                  (long long)cart_.rank()};
        std::FILE* f = std::fopen(path.c_str(), "wb");
        bool ok = f != nullptr;
        ok = ok && std::fwrite(hdr, sizeof(hdr), 1, f) == 1;
        ok = ok && (coef_.empty() || std::fwrite(coef_.data(), sizeof(double), coef_.size(), f) == coef_.size());
        ok = ok && std::fwrite(tile.data(), sizeof(double), tile.size(), f) == tile.size();
        if (f) ok = (std::fclose(f) == 0) && ok;
        if (!ok) ::b200::fail_comm(ERR_STATE, "cannot write checkpoint " + path, __FILE__, __LINE__);
    }
    long long load_checkpoint(const std::string& path) {
        std::FILE* f = std::fopen(path.c_str(), "rb");
        if (!f) ::b200::fail_comm(ERR_STATE, "cannot open checkpoint " + path, __FILE__, __LINE__);
        long long hdr[8] = {0};
        bool ok = std::fread(hdr, sizeof(hdr), 1, f) == 1;
        if (ok && (hdr[0] != kCheckpointMagic || hdr[1] != H_ || hdr[2] != W_ || hdr[3] != radius_ ||
                   hdr[4] != temporal_ || hdr[7] != cart_.rank())) {
            std::fclose(f);
            ::b200::fail_comm(ERR_ARG, "checkpoint does not match this plan (tile size, radius, levels or rank)",
                              __FILE__, __LINE__);
        }
        std::vector<double> c((size_t)(ok ? hdr[6] : 0));
        ok = ok && (c.empty() || std::fread(c.data(), sizeof(double), c.size(), f) == c.size());
        const int g = lay_.ghost;
        std::vector<double> tile((size_t)(H_ + 2 * g) * (W_ + 2 * g));
        ok = ok && std::fread(tile.data(), sizeof(double), tile.size(), f) == tile.size();
        std::fclose(f);
        if (!ok) ::b200::fail_comm(ERR_STATE, "short read from checkpoint " + path, __FILE__, __LINE__);
        if (!c.empty()) set_coeffs(c);
        upload_tile(tile.data());
        return hdr[5];
    }
    void set_coeffs(const std::vector<double>& c) {
        coef_ = c;   // keeps the host copy alive until the asynchronous copy has been consumed
        B200_DEVICE_OK(b2c_stencil_set_coef(h_, coef_.data(), (int)coef_.size(), stream_));
        B200_DEVICE_OK(b2c_stream_synchronize(stream_));
    }
    void set_timeout_ms(unsigned long long ms) { B200_DEVICE_OK(b2c_stencil_set_timeout_ms(h_, ms)); }

    // Unfused halo exchange of the current state (needed once after the core was written).
    void exchange() {
        B200_DEVICE_OK(b2c_stream_synchronize(stream_));
        cart_.barrier();
        B200_DEVICE_OK(b2c_stencil_exchange_push(h_, stream_));
        B200_DEVICE_OK(b2c_stencil_exchange_wait(h_, stream_));
        B200_DEVICE_OK(b2c_stream_synchronize(stream_));
        cart_.barrier();
    }
    // Enqueue n fused time steps (any n; with temporal blocking a remainder is one partial sweep).
    void step(int n, bool residual = false, bool drain = true) {
        const int flags = (residual ? 1 : 0) | (drain ? 2 : 0);
        if (temporal_ == 1) {
            B200_DEVICE_OK(b2c_stencil_step(h_, n, flags, stream_));
        } else {
            const int full = n / temporal_, rem = n % temporal_;
            if (rem == 0) {
                B200_DEVICE_OK(b2c_stencil_step(h_, full, flags, stream_));
            } else {
                if (full) B200_DEVICE_OK(b2c_stencil_step(h_, full, 0, stream_));
                B200_DEVICE_OK(b2c_stencil_step_levels(h_, rem, flags, stream_));
            }
        }
        steps_ += n;
    }
    // Pipelined variant for loops that feed coefficients in and read the residual out EVERY call:
    // the copies run on the plan's copy streams next to the neighbouring calls' kernels.  Returns
    // a ticket for wait_io(); keep at most two calls in flight.
    unsigned long long step_io(int n, const std::vector<double>* coeffs = nullptr, bool residual = true,
                               bool drain = true) {
        const int flags = (residual ? 1 : 0) | (drain ? 2 : 0);
        unsigned long long ticket = 0;
        B200_DEVICE_OK(b2c_stencil_step_io(h_, n, flags, coeffs ? coeffs->data() : nullptr,
                                           coeffs ? (int)coeffs->size() : 0, stream_, &ticket));
        steps_ += n;
        return ticket;
    }
    // Blocks until that call has run; returns its LOCAL sum of squared updates.
    double wait_io(unsigned long long ticket) {
        double r = 0.0;
        B200_DEVICE_OK(b2c_stencil_io_wait(h_, ticket, &r));
        return r;
    }
    void synchronize() { B200_DEVICE_OK(b2c_stream_synchronize(stream_)); }
    // Global L2 norm of the last update (after step(..., residual = true)); collective.
    double residual() {
        double local = 0.0, sum = 0.0;
        B200_DEVICE_OK(b2c_memcpy_async(&local, info().resid, sizeof(double), 2, stream_));
        B200_DEVICE_OK(b2c_stream_synchronize(stream_));
        cart_.allreduce(&local, &sum, 1, Datatype::basic(Basic::DOUBLE), comm::Op::SUM);
        return std::sqrt(sum);
    }
    // Synchronises; throws ERR_TIMEOUT if a device-side halo wait gave up.
    void check() {
        unsigned st = 0;
        B200_DEVICE_OK(b2c_stencil_status(h_, stream_, &st));
        if (st & 1u) ::b200::fail_comm(ERR_TIMEOUT, "halo flag wait timed out on the device", __FILE__, __LINE__);
    }
    b2c_stencil_info_t info() const {
        b2c_stencil_info_t inf{};
        B200_DEVICE_OK(b2c_stencil_info(h_, &inf));
        return inf;
    }
    b2c_stencil* handle() const { return h_; }

private:
    void copy_core(double* host, bool to_device) {
        const b2c_stencil_info_t inf = info();
        std::vector<double> padded((size_t)lay_.rows * lay_.pitch);
        const size_t nbytes = padded.size() * sizeof(double);
        B200_DEVICE_OK(b2c_memcpy_async(padded.data(), inf.buf[inf.cur], nbytes, 2, stream_));
        B200_DEVICE_OK(b2c_stream_synchronize(stream_));
        for (int y = 0; y < H_; ++y) {
            double* row = padded.data() + (size_t)(y + lay_.ghost) * lay_.pitch + lay_.lead;
            if (to_device) std::memcpy(row, host + (size_t)y * W_, sizeof(double) * W_);
            else std::memcpy(host + (size_t)y * W_, row, sizeof(double) * W_);
        }
        if (to_device) {
            B200_DEVICE_OK(b2c_memcpy_async(inf.buf[inf.cur], padded.data(), nbytes, 1, stream_));
            B200_DEVICE_OK(b2c_stream_synchronize(stream_));
        }
    }

    static constexpr long long kCheckpointMagic = 0x6232635f636b7074ll;   // "b2c_ckpt"
    comm::Comm cart_;
    int H_, W_, radius_, temporal_;
    void* stream_;
    b2c_stencil_layout_t lay_{};
    SymmetricBlock block_;
    b2c_stencil* h_ = nullptr;
    std::vector<double> coef_;
    long long steps_ = 0;
};

// Dot product whose local reduction is fused with a one-shot P2P allreduce: one kernel per rank,
// the result on every rank (or on `root`).  One rank per GPU (see the note at the top).
class FusedDot {
public:
    explicit FusedDot(const comm::Comm& c, void* stream = nullptr) : comm_(c), stream_(stream) {
        size_t bytes = 0;
        B200_DEVICE_OK(b2c_dot_symmetric_bytes(c.size(), &bytes));
        block_ = SymmetricBlock(c, bytes);
        std::vector<void*> tables((size_t)c.size());
        for (int r = 0; r < c.size(); ++r) tables[(size_t)r] = block_.peer(r);
        B200_DEVICE_OK(b2c_dot_create(&h_, c.rank(), c.size(), tables.data()));
    }
    FusedDot(const FusedDot&) = delete;
    FusedDot& operator=(const FusedDot&) = delete;
    ~FusedDot() {
        if (h_) {
            b2c_stream_synchronize(stream_);
            b2c_dot_destroy(h_);
        }
    }
    // x, y: device pointers to n elements each (this rank's slice).
    double dot(const double* x, const double* y, int64_t n, int root = -1) { return run(x, y, n, 9, root); }
    double dot(const float* x, const float* y, int64_t n, int root = -1) { return run(x, y, n, 8, root); }

private:
    double run(const void* x, const void* y, int64_t n, int dtype, int root) {
        double v = 0.0;
        B200_DEVICE_OK(b2c_dot_run(h_, x, y, n, dtype, root, 3, stream_));
        B200_DEVICE_OK(b2c_dot_result(h_, stream_, &v));
        return v;
    }
    comm::Comm comm_;
    void* stream_;
    SymmetricBlock block_;
    b2c_dot* h_ = nullptr;
};

// Collectives on DEVICE buffers over a SymmetricBlock workspace: allreduce (one-shot for small
// messages -- a single launch up to 4 KiB -- two-shot for large ones), reduce, bcast, gather,
// allgather, scatter, barrier; all stream-ordered kernels with epoch flags, no host in the data
// path.  Reference capability: MPI_Allreduce mpi9.cpp:51-54, MPI_Reduce mpicuda4.cu:368, MPI_Bcast
// mpicuda2.cu:154, MPI_Gather mpi6.cpp:89-100 -- on host scalars there.  (The in-switch variants
// need a multicast arena of the Python SymmetricHeap; see cuda_mpi_scratch_b200.ops.collectives.)
class DeviceCollectives {
public:
    enum Op { Sum = 0, Prod = 1, Max = 2, Min = 3 };
    enum Type { Int32 = 5, Int64 = 6, Float32 = 8, Float64 = 9 };

    DeviceCollectives(const comm::Comm& c, size_t max_bytes, void* stream = nullptr)
        : comm_(c), stream_(stream), max_bytes_(max_bytes) {
        size_t bytes = 0;
        B200_DEVICE_OK(b2c_coll_workspace_bytes(max_bytes, &bytes));
        block_ = SymmetricBlock(c, bytes);
        std::vector<void*> ws((size_t)c.size());
        for (int r = 0; r < c.size(); ++r) ws[(size_t)r] = block_.peer(r);
        B200_DEVICE_OK(b2c_coll_create(&h_, c.rank(), c.size(), ws.data(), max_bytes));
    }
    DeviceCollectives(const DeviceCollectives&) = delete;
    DeviceCollectives& operator=(const DeviceCollectives&) = delete;
    ~DeviceCollectives() {
        if (h_) {
            b2c_stream_synchronize(stream_);
            b2c_coll_destroy(h_);
        }
    }
    // in == out allowed.  Large messages (>= 512 KiB) take the two-shot algorithm.
    void allreduce(const void* in, void* out, int64_t count, Type t, Op op = Sum) {
        const size_t esz = (t == Int32 || t == Float32) ? 4 : 8;
        const int algo = (size_t)count * esz >= (512u << 10) && comm_.size() > 1 ? 1 : 0;
        B200_DEVICE_OK(b2c_coll_allreduce(h_, in, out, count, (int)t, (int)op, -1, algo, 7, stream_));
    }
    void reduce(const void* in, void* out, int64_t count, Type t, int root, Op op = Sum) {
        B200_DEVICE_OK(b2c_coll_allreduce(h_, in, out, count, (int)t, (int)op, root, 0, 7, stream_));
    }
    void bcast(void* buf, size_t bytes, int root) { B200_DEVICE_OK(b2c_coll_move(h_, 0, buf, buf, bytes, root, 3, stream_)); }
    void gather(const void* send, void* recv, size_t block_bytes, int root) {
        B200_DEVICE_OK(b2c_coll_move(h_, 1, send, recv, block_bytes, root, 3, stream_));
    }
    void allgather(const void* send, void* recv, size_t block_bytes) {
        B200_DEVICE_OK(b2c_coll_move(h_, 2, send, recv, block_bytes, 0, 3, stream_));
    }
    void scatter(const void* send, void* recv, size_t block_bytes, int root) {
        B200_DEVICE_OK(b2c_coll_move(h_, 3, send, recv, block_bytes, root, 3, stream_));
    }
    void barrier() { B200_DEVICE_OK(b2c_coll_barrier(h_, 3, stream_)); }
    // synchronises the stream; throws if a peer never arrived (device-side time-out)
    void check() { B200_DEVICE_OK(b2c_coll_check(h_, stream_)); }

private:
    comm::Comm comm_;
    void* stream_;
    size_t max_bytes_;
    SymmetricBlock block_;
    b2c_coll* h_ = nullptr;
};

}  // namespace device
}  // namespace b200
