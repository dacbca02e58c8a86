// Standard-conforming allocator over page-locked (pinned) host memory.
//
// Capability parity with the reference's host_allocator<T> (host_allocator.h:58-93:
// cudaMallocHost / cudaFreeHost behind the C++03 allocator interface, POD
// construct/destroy specialisations :11-54, used by the async ping-pong under
// -DPAGE_LOCKED, test-benchmark/mpi-pingpong-gpu-async.cpp:43-49).  This is the
// C++17 form: allocator_traits supplies construct/destroy/rebind, so only
// allocate/deallocate and equality are needed.  Header-only; the page-locked
// memory comes from the library (b2c_pinned_alloc), so a plain C++ program needs
// neither the CUDA headers nor its own copy of the CUDA runtime.
#pragma once
#include <cstddef>
#include <limits>
#include <new>
#include <type_traits>

#include "b200comm.h"

namespace b200 {

template <class T>
class PinnedAllocator {
public:
    using value_type = T;
    using is_always_equal = std::true_type;
    PinnedAllocator() noexcept = default;
    template <class U>
    PinnedAllocator(const PinnedAllocator<U>&) noexcept {}

    T* allocate(std::size_t n) {
        if (n > std::numeric_limits<std::size_t>::max() / sizeof(T)) throw std::bad_array_new_length();
        void* p = nullptr;
        if (b2c_pinned_alloc(&p, n * sizeof(T)) != 0 || p == nullptr) throw std::bad_alloc();
        return static_cast<T*>(p);
    }
    void deallocate(T* p, std::size_t) noexcept { b2c_pinned_free(p); }
};
template <class T, class U>
bool operator==(const PinnedAllocator<T>&, const PinnedAllocator<U>&) noexcept { return true; }
template <class T, class U>
bool operator!=(const PinnedAllocator<T>&, const PinnedAllocator<U>&) noexcept { return false; }

}  // namespace b200
