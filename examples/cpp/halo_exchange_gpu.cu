// GPU halo-exchange driver, one rank per PROCESS, tile in DEVICE memory:
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -Iinclude examples/cpp/halo_exchange_gpu.cu \
//        -Lcuda_mpi_scratch_b200/lib -lb200comm -o halo_exchange_gpu
//   python -m cuda_mpi_scratch_b200.launch -n 9 ./halo_exchange_gpu [tile [stencil [outdir]]]
//
// The run of the reference's stencil2d/mpi-2d-stencil-subarray-cuda.cu:77-179: bind a GPU,
// periodic Py x Px grid, tile = -1 everywhere, core = rank (set by a kernel through the
// device-usable Array2DAccessor), ONE exchange of the 8 halo regions with the DEVICE pointer
// handed to the comm layer (subarray types are applied on the GPU by the pack/move kernel),
// then the tile is copied back and written to "<row>_<col>" in the reference's dump format.
// Unlike the reference (2.7 #2, #8) the stencil height follows the stencil width argument and
// the copy-back happens before the communicator is shut down.
#include <cuda_runtime.h>

#include <cmath>
#include <cstdlib>
#include <fstream>
#include <iostream>
#include <string>
#include <vector>

#include "b200comm.h"
#include "b200comm/stencil2d.hpp"

using namespace b200;
using namespace b200::comm;
using namespace b200::stencil2d;

#define B2C_OK(call)                                                                  \
    do {                                                                              \
        if ((call) != 0) {                                                            \
            std::cerr << #call << ": " << b2c_last_error() << std::endl;              \
            std::exit(1);                                                             \
        }                                                                             \
    } while (0)

// One thread per cell of `region` (the reference launches one BLOCK per cell,
// mpi-2d-stencil-subarray-cuda.cu:25-28).
template <typename T>
__global__ void init_region(T* data, Array2D region, T value) {
    Array2DAccessor<T> a(data, region);
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = blockIdx.y * blockDim.y + threadIdx.y;
    if (x < region.width && y < region.height) a(x, y) = value;
}

template <typename T>
static void init_array(T* data, const Array2D& region, T value) {
    dim3 block(32, 8), grid((region.width + 31) / 32, (region.height + 7) / 8);
    init_region<T><<<grid, block>>>(data, region, value);
    if (cudaGetLastError() != cudaSuccess || cudaDeviceSynchronize() != cudaSuccess) {
        std::cerr << "init kernel failed" << std::endl;
        std::exit(1);
    }
}

int main(int argc, char** argv) {
    const int tile = argc > 1 ? std::atoi(argv[1]) : 16;
    const int stencil = argc > 2 ? std::atoi(argv[2]) : 5;
    const std::string outdir = argc > 3 ? argv[3] : ".";
    if (tile < 1 || stencil < 1 || stencil % 2 == 0 || stencil / 2 > tile) {
        std::cerr << "usage: halo_exchange_gpu [tile side] [odd stencil width <= 2*tile+1] [outdir]\n";
        return 1;
    }
    Comm world = init_from_env();
    int device = -1;
    B2C_OK(b2c_bind_device(/*policy: local rank*/ 0, world.rank(), 1, &device));
    const int n = world.size();
    const int side = (int)std::lround(std::sqrt((double)n));
    std::vector<int> dims = side * side == n ? std::vector<int>{side, side} : dims_create(n, 2);
    Comm cart = world.cart_create(dims, {1, 1});
    std::vector<int> co = cart.cart_coords(cart.rank());

    const int tot = tile + 2 * (stencil / 2);
    Array2D layout(tot, tot, tot);
    HaloPlan plan(cart, layout, stencil, stencil, element_type<double>());
    double* dev = nullptr;
    B2C_OK(b2c_malloc((void**)&dev, sizeof(double) * tot * tot));
    init_array(dev, layout, -1.0);
    init_array(dev, plan.core(), (double)cart.rank());

    std::vector<double> host((size_t)tot * tot);
    B2C_OK(b2c_memcpy_async(host.data(), dev, host.size() * sizeof(double), cudaMemcpyDeviceToHost, nullptr));
    B2C_OK(b2c_stream_synchronize(nullptr));
    std::ofstream os(outdir + "/" + std::to_string(co[0]) + "_" + std::to_string(co[1]));
    os << "Rank:  " << cart.rank() << "\nCoord: " << co[0] << ", " << co[1] << "\n\nCUDA device id: "
       << device << "\n\nCompute grid\n";
    print_cart_grid(os, cart);
    os << "\n" << tile << " x " << tile << " grid size\n" << tot << " x " << tot
       << " total(with ghost/halo regions) grid size\n" << stencil << " x " << stencil << " stencil\n\nArray\n";
    print_array(host.data(), layout, os);
    os << "\n";

    plan.exchange(dev);   // device pointer + derived datatypes, as with a CUDA-aware MPI

    B2C_OK(b2c_memcpy_async(host.data(), dev, host.size() * sizeof(double), cudaMemcpyDeviceToHost, nullptr));
    B2C_OK(b2c_stream_synchronize(nullptr));
    os << "Array after exchange\n";
    print_array(host.data(), layout, os);
    world.barrier();      // nobody unmaps a tile a peer may still be reading
    B2C_OK(b2c_free(dev));
    finalize(world);
    return 0;
}
