// The flagship from plain C++: fused 5-point Jacobi stencil + halo exchange, one rank per GPU.
//
//   python -m cuda_mpi_scratch_b200.launch -n 8 ./stencil_fused_gpu [tile side] [steps] [temporal]
//
// The loop of the reference's stencil2d/mpi-2d-stencil-subarray-cuda.cu:169-172
// (do { ExchangeData; Compute } while (!Terminate)) with a real Compute and with the exchange
// inside the compute kernel: periodic Py x Px grid, every rank owns a side x side fp64 tile,
// `steps` time steps (`temporal` of them per kernel sweep), wall-clock time from the earliest
// begin to the latest end over all ranks (mpicuda3.cu:315-325), and -- for grids small enough to
// redo on the host -- a comparison of every rank's tile with a host computation of the GLOBAL grid.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <iostream>
#include <vector>

#include "b200comm/device.hpp"

using namespace b200;
using namespace b200::comm;
using b200::device::FusedStencil;

static std::vector<double> host_steps(std::vector<double> g, int GH, int GW, int steps, const double* c,
                                      double* last_sumsq) {
    std::vector<double> n(g.size());
    for (int s = 0; s < steps; ++s) {
        double acc = 0.0;
        for (int y = 0; y < GH; ++y) {
            const int yn = (y + GH - 1) % GH, ys = (y + 1) % GH;
            for (int x = 0; x < GW; ++x) {
                const int xw = (x + GW - 1) % GW, xe = (x + 1) % GW;
                double a = c[0] * g[(size_t)y * GW + x];       // same operation order as the kernel
                a = std::fma(c[1], g[(size_t)y * GW + xw], a);
                a = std::fma(c[2], g[(size_t)y * GW + xe], a);
                a = std::fma(c[3], g[(size_t)yn * GW + x], a);
                a = std::fma(c[4], g[(size_t)ys * GW + x], a);
                n[(size_t)y * GW + x] = a;
                const double d = a - g[(size_t)y * GW + x];
                acc += d * d;
            }
        }
        g.swap(n);
        *last_sumsq = acc;
    }
    return g;
}

int main(int argc, char** argv) {
    const int side = argc > 1 ? std::atoi(argv[1]) : 1024;
    const int steps = argc > 2 ? std::atoi(argv[2]) : 30;
    const int temporal = argc > 3 ? std::atoi(argv[3]) : 3;
    if (side < 32 || side % 2 || steps < 1 || temporal < 1 || temporal > 4) {
        std::cerr << "usage: stencil_fused_gpu [even tile side >= 32] [steps >= 1] [temporal 1..4]\n";
        return 1;
    }
    Comm world = init_from_env();
    const int n = world.size(), me = world.rank();
    int ndev = 0, device = -1;
    if (b2c_device_count(&ndev) != 0 || ndev < 1) {
        if (me == 0) std::cerr << "no CUDA device\n";
        finalize(world);
        return 1;
    }
    if (n > ndev) {       // see the note in b200comm/device.hpp
        if (me == 0) std::cerr << "needs one GPU per rank (" << n << " ranks, " << ndev << " GPUs)\n";
        finalize(world);
        return 1;
    }
    device::check(b2c_bind_device(/*rank % devices*/ 1, me, 1, &device), "bind_device");
    const std::vector<int> dims = dims_create(n, 2);
    Comm cart = world.cart_create(dims, {1, 1});
    const std::vector<int> co = cart.cart_coords(cart.rank());
    int rc = 0;
    {
        FusedStencil st(cart, side, side, 1, FusedStencil::Star, temporal);
        const std::vector<double> coef = {0.2, 0.15, 0.25, 0.1, 0.3};
        st.set_coeffs(coef);
        const int GH = dims[0] * side, GW = dims[1] * side;
        auto init = [](int gy, int gx) { return std::sin(1e-2 * gy) * std::cos(2e-2 * gx); };
        std::vector<double> tile((size_t)side * side);
        for (int y = 0; y < side; ++y)
            for (int x = 0; x < side; ++x) tile[(size_t)y * side + x] = init(co[0] * side + y, co[1] * side + x);
        st.fill(0.0, 0.0);
        st.upload_core(tile.data());
        st.exchange();

        const int warm = 3;
        st.step(warm);
        st.synchronize();
        cart.barrier();
        const double t0 = wtime();
        st.step(steps, /*residual=*/true);
        st.synchronize();
        const double t1 = wtime();
        st.check();
        const double res = st.residual();
        const Datatype dbl = Datatype::basic(Basic::DOUBLE);
        double b = t0, e = t1, bmin = 0, emax = 0;
        cart.allreduce(&b, &bmin, 1, dbl, Op::MIN);
        cart.allreduce(&e, &emax, 1, dbl, Op::MAX);
        const double cells = (double)GH * GW;
        if (me == 0)
            std::cout << "grid " << GH << " x " << GW << " on " << dims[0] << " x " << dims[1] << " GPUs, " << steps
                      << " steps (" << temporal << " per sweep): " << (emax - bmin) / steps * 1e3 << " ms/step, "
                      << cells * steps / (emax - bmin) / 1e9 << " Gcell/s, residual " << res << std::endl;

        // The same steps as a loop that hands in coefficients and reads the residual back EVERY
        // call (step_io / wait_io: the copies run next to the neighbouring calls' kernels).
        const int io_calls = 4, extra = io_calls * temporal;
        double io_local = 0.0, io_sum = 0.0;
        {
            unsigned long long pend = 0;
            for (int i = 0; i < io_calls; ++i) {
                const unsigned long long t = st.step_io(temporal, &coef, /*residual=*/true, /*drain=*/i == io_calls - 1);
                if (i) io_local = st.wait_io(pend);
                pend = t;
            }
            io_local = st.wait_io(pend);
            st.check();
            cart.allreduce(&io_local, &io_sum, 1, dbl, Op::SUM);
        }

        if ((long long)GH * GW <= (1 << 22)) {         // small enough to redo on the host
            st.download_core(tile.data());
            std::vector<double> all(me == 0 ? (size_t)n * tile.size() : 0);
            cart.gather(tile.data(), (int64_t)tile.size(), dbl, all.data(), (int64_t)tile.size(), dbl, 0);
            if (me == 0) {
                std::vector<double> g((size_t)GH * GW);
                for (int y = 0; y < GH; ++y)
                    for (int x = 0; x < GW; ++x) g[(size_t)y * GW + x] = init(y, x);
                double sumsq = 0.0;
                g = host_steps(std::move(g), GH, GW, warm + steps, coef.data(), &sumsq);
                double sumsq_io = 0.0;
                g = host_steps(std::move(g), GH, GW, extra, coef.data(), &sumsq_io);
                double err = 0.0;
                for (int r = 0; r < n; ++r) {
                    const std::vector<int> rc2 = cart.cart_coords(r);
                    for (int y = 0; y < side; ++y)
                        for (int x = 0; x < side; ++x)
                            err = std::max(err, std::fabs(all[(size_t)r * tile.size() + (size_t)y * side + x] -
                                                          g[(size_t)(rc2[0] * side + y) * GW + rc2[1] * side + x]));
                }
                const bool ok = err == 0.0 && std::fabs(res - std::sqrt(sumsq)) <= 1e-10 * std::sqrt(sumsq) &&
                                std::fabs(io_sum - sumsq_io) <= 1e-10 * sumsq_io;
                std::cout << "max |device - host| = " << err << ", host residual " << std::sqrt(sumsq)
                          << ", pipelined loop residual " << std::sqrt(io_sum) << " (host " << std::sqrt(sumsq_io) << ")\n"
                          << (ok ? "PASSED" : "FAILED") << std::endl;
                rc = ok ? 0 : 1;
            }
            world.bcast(&rc, 1, Datatype::basic(Basic::INT32), 0);

            // Checkpoint / resume: a fresh plan restored from this rank's file continues with
            // the same bits as the plan that wrote it.
            const std::string ck = "/tmp/stencil_fused_gpu_" + std::to_string((long long)getpid()) + "_" +
                                   std::to_string(me) + ".ckpt";
            st.save_checkpoint(ck);
            st.step(2 * temporal + 1);
            st.synchronize();
            std::vector<double> cont(tile.size()), again(tile.size());
            st.download_core(cont.data());
            {
                FusedStencil st2(cart, side, side, 1, FusedStencil::Star, temporal);
                st2.fill(0.0, 0.0);
                const long long at = st2.load_checkpoint(ck);
                st2.exchange();
                st2.step(2 * temporal + 1);
                st2.synchronize();
                st2.check();
                st2.download_core(again.data());
                int same = (at == warm + steps + extra && cont == again) ? 1 : 0, all_same = 0;
                cart.allreduce(&same, &all_same, 1, Datatype::basic(Basic::INT32), Op::MIN);
                if (me == 0) std::cout << "checkpoint at step " << at << ", resumed run "
                                       << (all_same ? "bit-identical" : "DIFFERS") << std::endl;
                if (!all_same) rc = 1;
            }
            std::remove(ck.c_str());
        }
    }   // the plan and its symmetric block are released here (collective)
    finalize(world);
    return rc;
}
