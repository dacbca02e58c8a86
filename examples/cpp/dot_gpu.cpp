// Block-decomposed dot product of two vectors of ones over all ranks (one rank per PROCESS):
//
//   python -m cuda_mpi_scratch_b200.launch -n 4 ./dot_gpu [log2 N] [--float] [--rrobin] [--variant 0|1|2] [--iters K]
//
// The program of the reference's mpicuda2.cu:105-312 / mpicuda3.cu / mpicuda4.cu:189-395:
// rank -> GPU binding (rank % devices, or round-robin over nodes after counting the distinct
// host names), host vectors of ones -> H2D, local reduction on the GPU, global sum, wall-clock
// time from the earliest begin to the latest end over all ranks.  Two paths are timed:
//   fused     ONE kernel per rank: local reduction + one-shot P2P allreduce over peer-mapped
//             slot tables (b2c_dot_run) -- no D2H before the reduction, no host reduction;
//   reference the reference's structure: kernel (variant 0 atomicAdd per block, 1 per-block
//             partials finished on the host, 2 last-block-finishes) -> D2H -> reduce(SUM, root 0).
// The slot tables show the symmetric-memory recipe in plain C++: allocate, export a CUDA-IPC
// handle, allgather the handles through the communicator, open the peers' blocks.
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <numeric>
#include <set>
#include <string>
#include <vector>

#include "b200comm.h"
#include "b200comm/comm.hpp"

using namespace b200;
using namespace b200::comm;

#define B2C_OK(call)                                                       \
    do {                                                                   \
        if ((call) != 0) {                                                 \
            std::cerr << #call << ": " << b2c_last_error() << std::endl;   \
            std::exit(1);                                                  \
        }                                                                  \
    } while (0)

enum { H2D = 1, D2H = 2, DT_F32 = 8, DT_F64 = 9 };

// Number of distinct hosts in the job (mpicuda2.cu:118-156): names to root, count, broadcast.
static int count_nodes(const Comm& world) {
    char name[256] = {0};
    std::strncpy(name, processor_name().c_str(), sizeof(name) - 1);
    const Datatype ch = Datatype::basic(Basic::CHAR);
    std::vector<char> all(world.rank() == 0 ? (size_t)256 * world.size() : 0);
    world.gather(name, 256, ch, all.data(), 256, ch, 0);
    int nodes = 0;
    if (world.rank() == 0) {
        std::set<std::string> uniq;
        for (int r = 0; r < world.size(); ++r) uniq.insert(std::string(&all[(size_t)256 * r]));
        nodes = (int)uniq.size();
    }
    world.bcast(&nodes, 1, Datatype::basic(Basic::INT32), 0);
    return nodes;
}

// Earliest begin / latest end over all ranks (mpicuda3.cu:176-179,315-325), on rank 0.
static double global_elapsed(const Comm& world, double begin, double end) {
    const Datatype dbl = Datatype::basic(Basic::DOUBLE);
    std::vector<double> b(world.size()), e(world.size());
    world.gather(&begin, 1, dbl, b.data(), 1, dbl, 0);
    world.gather(&end, 1, dbl, e.data(), 1, dbl, 0);
    if (world.rank() != 0) return 0.0;
    return *std::max_element(e.begin(), e.end()) - *std::min_element(b.begin(), b.end());
}

template <typename real>
static int run(const Comm& world, int64_t n_total, int variant, int iters) {
    const int me = world.rank(), np = world.size();
    const int dtype = sizeof(real) == 8 ? DT_F64 : DT_F32;
    const int64_t n = n_total / np;
    const size_t bytes = (size_t)n * sizeof(real);

    // ---- data: ones on the host, copied to the device (mpicuda2.cu:209-237)
    std::vector<real> ones((size_t)n, (real)1);
    real *x = nullptr, *y = nullptr;
    B2C_OK(b2c_malloc((void**)&x, bytes));
    B2C_OK(b2c_malloc((void**)&y, bytes));
    B2C_OK(b2c_memcpy_async(x, ones.data(), bytes, H2D, nullptr));
    B2C_OK(b2c_memcpy_async(y, ones.data(), bytes, H2D, nullptr));
    B2C_OK(b2c_stream_synchronize(nullptr));

    // ---- symmetric slot tables: every rank maps every rank's table
    size_t sym_bytes = 0;
    B2C_OK(b2c_dot_symmetric_bytes(np, &sym_bytes));
    void* mine = nullptr;
    B2C_OK(b2c_malloc(&mine, sym_bytes));
    unsigned char handle[B2C_IPC_HANDLE_BYTES];
    B2C_OK(b2c_ipc_export(mine, handle));
    std::vector<unsigned char> handles((size_t)np * B2C_IPC_HANDLE_BYTES);
    const Datatype byte = Datatype::basic(Basic::BYTE);
    world.allgather(handle, B2C_IPC_HANDLE_BYTES, byte, handles.data(), B2C_IPC_HANDLE_BYTES, byte);
    std::vector<void*> tables(np, nullptr);
    for (int r = 0; r < np; ++r) {
        if (r == me) tables[r] = mine;
        else B2C_OK(b2c_ipc_open(&handles[(size_t)r * B2C_IPC_HANDLE_BYTES], &tables[r]));
    }
    b2c_dot* plan = nullptr;
    B2C_OK(b2c_dot_create(&plan, me, np, tables.data()));
    world.barrier();

    // Do several ranks share a GPU (more ranks than devices)?  Then a kernel must never wait
    // for a kernel of another rank that may not be resident: run the push half, meet on the
    // host, run the gather half.  One rank per GPU uses the single fused launch.
    int my_dev = -1;
    B2C_OK(b2c_get_device(&my_dev));
    std::vector<int> devs(np);
    const Datatype i32 = Datatype::basic(Basic::INT32);
    world.allgather(&my_dev, 1, i32, devs.data(), 1, i32);
    const bool shared_gpu = std::set<int>(devs.begin(), devs.end()).size() < (size_t)np;
    if (me == 0 && shared_gpu) std::cout << "ranks share a GPU: push and gather as two launches\n";

    // ---- fused path
    double fused = 0.0, t_fused = 0.0;
    for (int it = -1; it < iters; ++it) {
        world.barrier();
        const double t0 = wtime();
        if (!shared_gpu) {
            B2C_OK(b2c_dot_run(plan, x, y, n, dtype, /*root: all ranks*/ -1, /*one launch*/ 3, nullptr));
        } else {
            B2C_OK(b2c_dot_run(plan, x, y, n, dtype, -1, /*local reduction + push*/ 1, nullptr));
            B2C_OK(b2c_stream_synchronize(nullptr));
            world.barrier();
            B2C_OK(b2c_dot_run(plan, x, y, n, dtype, -1, /*gather + sum*/ 2, nullptr));
        }
        B2C_OK(b2c_dot_result(plan, nullptr, &fused));
        const double t1 = wtime();
        const double el = global_elapsed(world, t0, t1);
        if (it >= 0) t_fused += el;
    }

    // ---- reference-structured path: kernel -> D2H -> host reduce to rank 0
    const int blocks = b2c_baseline_dot_blocks(n);
    real* dout = nullptr;
    B2C_OK(b2c_malloc((void**)&dout, sizeof(real) * blocks));
    std::vector<real> partial(blocks);
    const Datatype rt = Datatype::basic(sizeof(real) == 8 ? Basic::DOUBLE : Basic::FLOAT);
    real staged = 0;
    double t_ref = 0.0;
    for (int it = -1; it < iters; ++it) {
        world.barrier();
        const double t0 = wtime();
        if (variant == 0) B2C_OK(b2c_memset_async(dout, 0, sizeof(real), nullptr));   // atomicAdd target
        B2C_OK(b2c_baseline_dot(variant, x, y, n, dtype, dout, nullptr));
        real local;
        if (variant == 1) {
            B2C_OK(b2c_memcpy_async(partial.data(), dout, sizeof(real) * blocks, D2H, nullptr));
            B2C_OK(b2c_stream_synchronize(nullptr));
            local = std::accumulate(partial.begin(), partial.end(), (real)0);
        } else {
            B2C_OK(b2c_memcpy_async(&local, dout, sizeof(real), D2H, nullptr));
            B2C_OK(b2c_stream_synchronize(nullptr));
        }
        world.reduce(&local, &staged, 1, rt, Op::SUM, 0);
        const double t1 = wtime();
        const double el = global_elapsed(world, t0, t1);
        if (it >= 0) t_ref += el;
    }

    int rc = 0;
    if (me == 0) {
        const double want = (double)n_total;
        const double tol = sizeof(real) == 8 ? 0.0 : 1e-6 * want;   // fp32 partial sums round
        const bool ok = fused == want && std::fabs((double)staged - want) <= tol;
        std::cout << "ranks " << np << " elements " << n_total << " (" << sizeof(real) * 8 << "-bit)\n"
                  << "dot product result (fused kernel + P2P allreduce): " << fused << "  time "
                  << t_fused / iters * 1e3 << " ms\n"
                  << "dot product result (variant " << variant << " kernel + D2H + host reduce): "
                  << (double)staged << "  time " << t_ref / iters * 1e3 << " ms\n"
                  << (ok ? "PASSED" : "FAILED") << std::endl;
        rc = ok ? 0 : 1;
    }
    world.bcast(&rc, 1, Datatype::basic(Basic::INT32), 0);
    world.barrier();      // no rank unmaps while a peer may still be storing into its table
    B2C_OK(b2c_dot_destroy(plan));
    for (int r = 0; r < np; ++r)
        if (r != me) B2C_OK(b2c_ipc_close(tables[r]));
    world.barrier();
    B2C_OK(b2c_free(mine));
    B2C_OK(b2c_free(dout));
    B2C_OK(b2c_free(x));
    B2C_OK(b2c_free(y));
    return rc;
}

int main(int argc, char** argv) {
    int log2n = 24, variant = 2, iters = 5;
    bool use_float = false, rrobin = false;
    for (int i = 1; i < argc; ++i) {
        const std::string a = argv[i];
        if (a == "--float") use_float = true;
        else if (a == "--rrobin") rrobin = true;
        else if (a == "--variant" && i + 1 < argc) variant = std::atoi(argv[++i]);
        else if (a == "--iters" && i + 1 < argc) iters = std::atoi(argv[++i]);
        else if (a[0] != '-') log2n = std::atoi(a.c_str());
        else { std::cerr << "unknown option " << a << "\n"; return 1; }
    }
    if (log2n < 10 || log2n > 34 || variant < 0 || variant > 2 || iters < 1) {
        std::cerr << "usage: dot_gpu [log2 N in 10..34] [--float] [--rrobin] [--variant 0|1|2] [--iters K]\n";
        return 1;
    }
    Comm world = init_from_env();
    const int64_t n_total = (int64_t)1 << log2n;
    if (n_total % world.size() != 0) {     // mpicuda2.cu:161-165
        if (world.rank() == 0) std::cerr << "N must be divisible by the number of ranks\n";
        finalize(world);
        return 1;
    }
    int device = -1;
    if (rrobin) B2C_OK(b2c_bind_device(/*(rank / nodes) % devices*/ 2, world.rank(), count_nodes(world), &device));
    else B2C_OK(b2c_bind_device(/*rank % devices*/ 1, world.rank(), 1, &device));
    const int rc = use_float ? run<float>(world, n_total, variant, iters)
                             : run<double>(world, n_total, variant, iters);
    finalize(world);
    return rc;
}
