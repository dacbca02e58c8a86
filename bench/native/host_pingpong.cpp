// Native host ping-pong over the built-in transport (no Python on the path):
//   g++ -std=c++17 -O2 -Iinclude bench/native/host_pingpong.cpp -Lcuda_mpi_scratch_b200/lib -lb200comm -o host_pingpong
//   python -m cuda_mpi_scratch_b200.launch -n 2 ./host_pingpong
// Blocking send/recv of host buffers between rank 0 and rank 1 (the pattern of the reference's
// mpi4.cpp:24-44), round-trip time per size after warm-up.
#include <cstdio>
#include <vector>

#include "b200comm/comm.hpp"

using namespace b200;
using namespace b200::comm;

int main() {
    Comm world = init_from_env();
    const int me = world.rank();
    const Datatype byte = Datatype::basic(Basic::BYTE);
    for (size_t n : {(size_t)8, (size_t)1024, (size_t)16384, (size_t)262144, (size_t)4194304, (size_t)16777216}) {
        std::vector<char> buf(n, 1);
        const int iters = n <= 16384 ? 20000 : (n <= 262144 ? 2000 : 50);
        double t0 = 0;
        for (int it = 0; it < iters + 100; ++it) {
            if (it == 100) { world.barrier(); t0 = wtime(); }
            if (me == 0) { world.send(buf.data(), (int64_t)n, byte, 1, 1); world.recv(buf.data(), (int64_t)n, byte, 1, 1); }
            else if (me == 1) { world.recv(buf.data(), (int64_t)n, byte, 0, 1); world.send(buf.data(), (int64_t)n, byte, 0, 1); }
        }
        const double rtt = (wtime() - t0) / iters;
        if (me == 0) std::printf("bytes %zu rtt_us %.2f one_way_GBps %.3f\n", n, rtt * 1e6, 2.0 * n / rtt / 1e9);
    }
    finalize(world);
    return 0;
}
