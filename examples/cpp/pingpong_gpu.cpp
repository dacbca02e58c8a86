// GPU ping-pong between rank 0 and rank 1 through the MPI-like API, N doubles in DEVICE memory:
//
//   python -m cuda_mpi_scratch_b200.launch -n 2 ./pingpong_gpu N [--async] [--host-copy] [--pinned] [--iters K]
//
// Covers both of the reference's benchmarks: test-benchmark/mpi-pingpong-gpu.cpp:24-83
// (blocking Send/Recv of device pointers) and mpi-pingpong-gpu-async.cpp:43-106 (Isend/Irecv/
// Wait; -DHOST_COPY = stage through host memory, -DPAGE_LOCKED = pinned staging buffers from the
// allocator of host_allocator.h, here b200::PinnedAllocator).  Differences on purpose: one
// untimed warm-up round trip and K timed ones (the reference times a single cold iteration,
// SURVEY 2.7 #12), every request is completed, and BOTH ranks verify what they received.
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <string>
#include <vector>

#include "b200comm.h"
#include "b200comm/comm.hpp"
#include "b200comm/pinned_allocator.hpp"

using namespace b200;
using namespace b200::comm;

#define B2C_OK(call)                                                       \
    do {                                                                   \
        if ((call) != 0) {                                                 \
            std::cerr << #call << ": " << b2c_last_error() << std::endl;   \
            std::exit(1);                                                  \
        }                                                                  \
    } while (0)

enum { H2D = 1, D2H = 2 };   // cudaMemcpyKind values

template <class HostVec>
static int run(const Comm& world, size_t n, bool async, bool host_copy, int iters) {
    const int me = world.rank(), peer = 1 - me;
    const Datatype dbl = Datatype::basic(Basic::DOUBLE);
    const size_t bytes = n * sizeof(double);
    double *dsend = nullptr, *drecv = nullptr;
    B2C_OK(b2c_malloc((void**)&dsend, bytes));
    B2C_OK(b2c_malloc((void**)&drecv, bytes));
    HostVec hsend(n, me == 0 ? 3.0 : 0.0), hrecv(n, 0.0);
    B2C_OK(b2c_memcpy_async(dsend, hsend.data(), bytes, H2D, nullptr));
    B2C_OK(b2c_stream_synchronize(nullptr));

    // one direction of the exchange: `out` leaves this rank, `in` arrives
    auto xfer_send = [&](const double* dev) {
        const void* p = dev;
        if (host_copy) {
            B2C_OK(b2c_memcpy_async(hsend.data(), dev, bytes, D2H, nullptr));
            B2C_OK(b2c_stream_synchronize(nullptr));
            p = hsend.data();
        }
        if (async) {
            Request r = world.isend(p, (int64_t)n, dbl, peer, 0x10);
            r.wait();
        } else {
            world.send(p, (int64_t)n, dbl, peer, 0x10);
        }
    };
    auto xfer_recv = [&](double* dev) {
        void* p = host_copy ? (void*)hrecv.data() : (void*)dev;
        if (async) {
            Request r = world.irecv(p, (int64_t)n, dbl, peer, 0x10);
            r.wait();
        } else {
            world.recv(p, (int64_t)n, dbl, peer, 0x10);
        }
        if (host_copy) {
            B2C_OK(b2c_memcpy_async(dev, hrecv.data(), bytes, H2D, nullptr));
            B2C_OK(b2c_stream_synchronize(nullptr));
        }
    };

    double best = 1e30, total = 0.0;
    for (int it = -1; it < iters; ++it) {      // it == -1 is the warm-up
        world.barrier();
        const double t0 = wtime();
        if (me == 0) {
            xfer_send(dsend);
            xfer_recv(drecv);
        } else {
            xfer_recv(drecv);
            xfer_send(drecv);                  // bounce what arrived
        }
        const double dt = wtime() - t0;
        if (it >= 0) {
            best = std::min(best, dt);
            total += dt;
        }
    }
    // verification on both ranks: everybody must hold rank 0's pattern
    std::vector<double> check(n, -1.0);
    B2C_OK(b2c_memcpy_async(check.data(), drecv, bytes, D2H, nullptr));
    B2C_OK(b2c_stream_synchronize(nullptr));
    const bool ok = std::all_of(check.begin(), check.end(), [](double v) { return v == 3.0; });
    int good = ok ? 1 : 0, all_good = 0;
    world.allreduce(&good, &all_good, 1, Datatype::basic(Basic::INT32), Op::MIN);
    if (me == 0) {
        const double avg = total / iters;
        std::cout << "pingpong " << (async ? "async" : "blocking") << (host_copy ? " host-staged" : " device")
                  << " bytes " << bytes << " round-trip avg " << avg * 1e6 << " us best " << best * 1e6
                  << " us one-way " << 2.0 * bytes / best / 1e9 << " GB/s\n"
                  << (all_good ? "PASSED" : "FAILED") << std::endl;
        unsigned long long st[5] = {0, 0, 0, 0, 0};
        b2c_dev_eager_stats(st);
        std::cout << "device sends of rank 0: " << st[0] << " eager (mailbox), rendezvous: " << st[1]
                  << " too large, " << st[2] << " no mailbox yet, " << st[3] << " no free slot" << std::endl;
    }
    world.barrier();
    B2C_OK(b2c_free(dsend));
    B2C_OK(b2c_free(drecv));
    return all_good ? 0 : 1;
}

int main(int argc, char** argv) {
    if (argc < 2) {
        std::cerr << "usage: " << argv[0] << " <number of doubles> [--async] [--host-copy] [--pinned] [--iters K]\n";
        return 1;     // (the reference falls through to atoi(NULL) here)
    }
    const size_t n = std::strtoull(argv[1], nullptr, 10);
    bool async = false, host_copy = false, pinned = false;
    int iters = 10;
    for (int i = 2; i < argc; ++i) {
        const std::string a = argv[i];
        if (a == "--async") async = true;
        else if (a == "--host-copy") host_copy = true;
        else if (a == "--pinned") pinned = true;
        else if (a == "--iters" && i + 1 < argc) iters = std::atoi(argv[++i]);
        else { std::cerr << "unknown option " << a << "\n"; return 1; }
    }
    if (n == 0 || iters < 1) { std::cerr << "need N > 0 and K > 0\n"; return 1; }
    Comm world = init_from_env();
    if (world.size() != 2) {
        if (world.rank() == 0) std::cerr << "run with exactly 2 ranks\n";
        finalize(world);
        return 1;
    }
    int device = -1;
    B2C_OK(b2c_bind_device(0, world.rank(), 1, &device));
    int rc;
    if (pinned) rc = run<std::vector<double, PinnedAllocator<double>>>(world, n, async, host_copy, iters);
    else rc = run<std::vector<double>>(world, n, async, host_copy, iters);
    finalize(world);
    return rc;
}
