// Reference-STYLE local dot-product kernels, kept only as baselines for the
// fused kernel in csrc/kernels/dot.cu.  They reproduce the reference's three
// strategies re-written for sm_100a with the reference's
// defects fixed (missing barriers, int indices, fp64 atomics compiled out,
// SURVEY.md 2.7 #10):
//   variant 0  one atomicAdd per block into a single scalar
//              (dot_product_kernel, mpicuda2.cu:65-81)
//   variant 1  per-block partials, the HOST finishes the sum
//              (partial_dot_product_kernel, mpicuda2.cu:84-100, :277-279)
//   variant 2  per-block partials + "last block finishes" on the device
//              (dot_product_full_kernel, mpicuda4.cu:157-185)
// All use the reference's launch shape: 512 threads, min(N/512, 65535) blocks,
// scalar loads, a shared-memory tree.  The cross-rank step stays on the host
// (D2H + allreduce through the comm layer or NCCL), exactly the path the fused
// kernel removes.
#include <cuda_runtime.h>
#include <stdint.h>

#include "b200comm.h"
#include "b200comm/error.hpp"
#include "core/capi_util.hpp"

namespace {

constexpr int kBlock = 512;

template <typename T>
__device__ T block_tree(T v, T* cache) {
    cache[threadIdx.x] = v;
    __syncthreads();
    for (int i = kBlock / 2; i > 0; i >>= 1) {
        if ((int)threadIdx.x < i) cache[threadIdx.x] += cache[threadIdx.x + i];
        __syncthreads();
    }
    return cache[0];
}

template <typename T>
__device__ T strided_dot(const T* a, const T* b, int64_t n) {
    T acc = 0;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
         i += (int64_t)gridDim.x * blockDim.x)
        acc += a[i] * b[i];
    return acc;
}

template <typename T>
__global__ void dot_atomic_kernel(const T* a, const T* b, int64_t n, T* out) {
    __shared__ T cache[kBlock];
    T r = block_tree(strided_dot(a, b, n), cache);
    if (threadIdx.x == 0) atomicAdd(out, r);
}

template <typename T>
__global__ void dot_partials_kernel(const T* a, const T* b, int64_t n, T* out) {
    __shared__ T cache[kBlock];
    T r = block_tree(strided_dot(a, b, n), cache);
    if (threadIdx.x == 0) out[blockIdx.x] = r;
}

// The reference's DEMONSTRATION of a data race (ref_parallel-dot-product-atomics.cu:26-32,
// -DNO_SYNC): every block adds its partial to the scalar with a plain read-modify-write
// instead of atomicAdd, so concurrent blocks lose updates and the result comes out too
// small.  Kept as variant 3 for the same teaching purpose (and as a target for
// `compute-sanitizer --tool racecheck`); never used by the product.
template <typename T>
__global__ void dot_racy_kernel(const T* a, const T* b, int64_t n, T* out) {
    __shared__ T cache[kBlock];
    T r = block_tree(strided_dot(a, b, n), cache);
    if (threadIdx.x == 0) *reinterpret_cast<volatile T*>(out) = *reinterpret_cast<volatile T*>(out) + r;
}

__device__ unsigned int g_count = 0;

template <typename T>
__global__ void dot_lastblock_kernel(const T* a, const T* b, int64_t n, T* out) {
    __shared__ T cache[kBlock];
    __shared__ bool last;
    T r = block_tree(strided_dot(a, b, n), cache);
    if (threadIdx.x == 0) {
        out[blockIdx.x] = r;
        __threadfence();
        last = atomicInc(&g_count, gridDim.x - 1) == gridDim.x - 1;
    }
    __syncthreads();
    if (!last) return;
    T s = 0;
    for (int i = threadIdx.x; i < (int)gridDim.x; i += blockDim.x)
        s += *reinterpret_cast<volatile T*>(out + i);
    __syncthreads();
    s = block_tree(s, cache);
    if (threadIdx.x == 0) out[0] = s;
}

template <typename T>
void launch(int variant, const T* a, const T* b, int64_t n, T* out, int blocks, cudaStream_t st) {
    if (variant == 0) {
        cudaMemsetAsync(out, 0, sizeof(T), st);
        dot_atomic_kernel<T><<<blocks, kBlock, 0, st>>>(a, b, n, out);
    } else if (variant == 1) {
        dot_partials_kernel<T><<<blocks, kBlock, 0, st>>>(a, b, n, out);
    } else if (variant == 3) {
        cudaMemsetAsync(out, 0, sizeof(T), st);
        dot_racy_kernel<T><<<blocks, kBlock, 0, st>>>(a, b, n, out);
    } else {
        dot_lastblock_kernel<T><<<blocks, kBlock, 0, st>>>(a, b, n, out);
    }
}

}  // namespace

extern "C" {

// Blocks the reference would launch for `n` elements (mpicuda4.cu:338-339), at least 1.
B2C_EXPORT int b2c_baseline_dot_blocks(int64_t n) {
    int64_t b = n / kBlock;
    if (b > 65535) b = 65535;
    if (b < 1) b = 1;
    return (int)b;
}

// `out` must hold b2c_baseline_dot_blocks(n) elements.  variant 0/2 leave the
// local result in out[0]; variant 1 leaves per-block partials for the host.
B2C_EXPORT int b2c_baseline_dot(int variant, const void* a, const void* b, int64_t n, int dtype,
                                void* out, void* stream) {
    B2C_TRY
    using namespace b200;
    B200_REQUIRE(variant >= 0 && variant <= 3, ERR_ARG, "variant must be 0, 1, 2 or 3 (racy demo)");
    B200_REQUIRE(dtype == 8 || dtype == 9, ERR_TYPE, "float32 or float64 only");
    const int blocks = b2c_baseline_dot_blocks(n);
    if (dtype == 9)
        launch<double>(variant, (const double*)a, (const double*)b, n, (double*)out, blocks, (cudaStream_t)stream);
    else
        launch<float>(variant, (const float*)a, (const float*)b, n, (float*)out, blocks, (cudaStream_t)stream);
    B200_CHECK_LAUNCH();
    B2C_CATCH
}

}  // extern "C"
