// CPU halo-exchange driver against the C++ stencil2d library, one rank per PROCESS
// (start with `python -m cuda_mpi_scratch_b200.launch -n 4 ./halo_exchange_cpu OUTDIR`).
// Same run as the reference's stencil2d/mpi-2d-stencil-subarray.cpp:35-100: periodic
// sqrt(N) x sqrt(N) grid, 16x16 tile, 5x5 stencil, tile = -1, core = rank, ONE exchange,
// per-rank file "<row>_<col>" in the reference's dump format.
#include <cmath>
#include <fstream>
#include <iostream>
#include <string>
#include <vector>

#include "b200comm/stencil2d.hpp"

using namespace b200;
using namespace b200::comm;
using namespace b200::stencil2d;

int main(int argc, char** argv) {
    const std::string outdir = argc > 1 ? argv[1] : ".";
    Comm world = init_from_env();
    const int n = world.size();
    const int dim = (int)std::lround(std::sqrt((double)n));
    std::vector<int> dims = dim * dim == n ? std::vector<int>{dim, dim} : dims_create(n, 2);
    Comm cart = world.cart_create(dims, {1, 1});
    std::vector<int> co = cart.cart_coords(cart.rank());
    const int tile = 16, stencil = 5, tot = tile + 2 * (stencil / 2);
    std::vector<double> data((size_t)tot * tot, -1.0);
    Array2D layout(tot, tot, tot);
    HaloPlan plan(cart, layout, stencil, stencil, Datatype::basic(Basic::DOUBLE));
    Array2DAccessor<double> core(data.data(), plan.core());
    for (int y = 0; y < plan.core().height; ++y)
        for (int x = 0; x < plan.core().width; ++x) core(x, y) = (double)cart.rank();
    std::ofstream os(outdir + "/" + std::to_string(co[0]) + "_" + std::to_string(co[1]));
    os << "Rank:  " << cart.rank() << "\nCoord: " << co[0] << ", " << co[1] << "\n\nCompute grid\n";
    print_cart_grid(os, cart);
    os << "\n" << tile << " x " << tile << " grid size\n" << tot << " x " << tot
       << " total(with ghost/halo regions) grid size\n" << stencil << " x " << stencil << " stencil\n\nArray\n";
    print_array(data.data(), layout, os);
    os << "\n";
    plan.exchange(data.data());          // do { ExchangeData; Compute } while(!Terminate) runs once
    os << "Array after exchange\n";
    print_array(data.data(), layout, os);
    finalize(world);
    return 0;
}
