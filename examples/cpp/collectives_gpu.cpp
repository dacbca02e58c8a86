// Device collectives from C++ (one rank per GPU):
//
//   python -m cuda_mpi_scratch_b200.launch -n 8 ./collectives_gpu
//
// What the reference does with MPI_Allreduce / MPI_Reduce / MPI_Bcast / MPI_Gather on HOST scalars
// after a D2H copy (mpi9.cpp:51-54, mpicuda4.cu:354-368, mpicuda2.cu:154, mpi6.cpp:89-100), done on
// device buffers by kernels that read and write the peers' memory: b200::device::DeviceCollectives.
// The ranks must sit on DIFFERENT GPUs: the kernels of two ranks wait for each other.
#include <cstdlib>
#include <iostream>
#include <set>
#include <vector>

#include "b200comm.h"
#include "b200comm/comm.hpp"
#include "b200comm/device.hpp"

using namespace b200;
using namespace b200::comm;

#define B2C_OK(call)                                                       \
    do {                                                                   \
        if ((call) != 0) {                                                 \
            std::cerr << #call << ": " << b2c_last_error() << std::endl;   \
            std::exit(1);                                                  \
        }                                                                  \
    } while (0)

int main() {
    Comm world = init_from_env();
    const int me = world.rank(), np = world.size();
    int dev = -1;
    B2C_OK(b2c_bind_device(0, me, 1, &dev));
    std::vector<int> devs((size_t)np);
    const Datatype i32 = Datatype::basic(Basic::INT32);
    world.allgather(&dev, 1, i32, devs.data(), 1, i32);
    if (std::set<int>(devs.begin(), devs.end()).size() < (size_t)np) {
        if (me == 0) std::cout << "ranks share a GPU: skipped (device collectives need one rank per GPU)\n";
        finalize(world);
        return 0;
    }
    int good = 1;
    {
        device::DeviceCollectives coll(world, 8u << 20);
        const int64_t n = 1 << 18;
        std::vector<double> h((size_t)n, (double)(me + 1));
        double* d = nullptr;
        B2C_OK(b2c_malloc((void**)&d, n * sizeof(double)));
        B2C_OK(b2c_memcpy_async(d, h.data(), n * sizeof(double), 1, nullptr));
        B2C_OK(b2c_stream_synchronize(nullptr));
        coll.allreduce(d, d, n, device::DeviceCollectives::Float64);          // two-shot (2 MiB)
        coll.allreduce(d, d, 1, device::DeviceCollectives::Float64);          // one launch (8 bytes)
        coll.bcast(d, 1024, np - 1);
        coll.barrier();
        coll.check();
        B2C_OK(b2c_memcpy_async(h.data(), d, n * sizeof(double), 2, nullptr));
        B2C_OK(b2c_stream_synchronize(nullptr));
        const double s = np * (np + 1) / 2.0;
        // element 0: sum of sums (then broadcast from the last rank, which holds the same value)
        good = h[0] == s * np && h[(size_t)n - 1] == s;
        B2C_OK(b2c_free(d));
    }
    int all_good = 0;
    world.allreduce(&good, &all_good, 1, i32, Op::MIN);
    if (me == 0) std::cout << "device collectives on " << np << " GPUs: " << (all_good ? "PASSED" : "FAILED") << std::endl;
    finalize(world);
    return all_good ? 0 : 1;
}
