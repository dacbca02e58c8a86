// Derived datatypes: a committed type is a flat, ordered list of
// (signed byte displacement, byte length) runs plus size / extent / signature,
// which is exactly what a device kernel needs to gather or scatter it.
//
// Capability parity with the MPI datatype calls the reference exercises
// (SURVEY.md 2.3): MPI_Type_indexed (mpi7.cpp:32-41), MPI_Type_create_struct +
// MPI_Type_extent (mpi8.cpp:38-54), MPI_Type_create_subarray 1-D and 2-D in C
// order (mpi-complex-types.cpp:33-36, stencil2d/stencil2D.h:210-228),
// MPI_Type_create_hindexed over a subarray with byte displacements between
// separately allocated buffers (mpi-complex-types.cpp:38-50), commit / free.
// Displacements are signed 64-bit (the reference subtracts
// unrelated heap addresses).
#pragma once
#include <cstdint>
#include <memory>
#include <string>
#include <vector>

#pragma GCC visibility push(default)   // public C++ API: exported from the library
namespace b200 {

struct Run {
    int64_t disp;  // byte displacement from the buffer address (may be negative)
    int64_t len;   // bytes
};

enum class Basic : int {
    BYTE = 0, CHAR = 1, INT8 = 2, UINT8 = 3, INT16 = 4, INT32 = 5, INT64 = 6, UINT64 = 7,
    FLOAT = 8, DOUBLE = 9, UINT32 = 10, NUM_BASIC = 11
};
size_t basic_size(Basic b);
const char* basic_name(Basic b);

enum class Order : int { C = 0, FORTRAN = 1 };

class Datatype {
public:
    Datatype();  // invalid (null) type
    // ---- constructors (MPI_Type_*) ----
    static Datatype basic(Basic b);
    static Datatype contiguous(int count, const Datatype& old);
    static Datatype vector(int count, int blocklen, int stride, const Datatype& old);
    static Datatype hvector(int count, int blocklen, int64_t stride_bytes, const Datatype& old);
    static Datatype indexed(const std::vector<int>& blocklens, const std::vector<int>& disps,
                            const Datatype& old);
    static Datatype hindexed(const std::vector<int>& blocklens,
                             const std::vector<int64_t>& byte_disps, const Datatype& old);
    static Datatype create_struct(const std::vector<int>& blocklens,
                                  const std::vector<int64_t>& byte_disps,
                                  const std::vector<Datatype>& types);
    static Datatype subarray(const std::vector<int>& sizes, const std::vector<int>& subsizes,
                             const std::vector<int>& starts, Order order, const Datatype& old);
    static Datatype resized(const Datatype& old, int64_t lb, int64_t extent);

    Datatype& commit();          // MPI_Type_commit: merges adjacent runs, freezes the type
    void free();                 // MPI_Type_free: drops this handle
    bool valid() const { return (bool)impl_; }
    bool committed() const;
    bool is_basic() const;
    int64_t size() const;        // payload bytes of one element (MPI_Type_size)
    int64_t extent() const;      // MPI_Type_extent / MPI_Type_get_extent
    int64_t lb() const;
    const std::vector<Run>& runs() const;   // flat runs of ONE element
    uint64_t signature_hash() const;        // order-sensitive hash of the basic-type sequence
    bool contiguous_layout() const;         // one run at displacement 0 with len == extent
    // true if one element is `n` values of a single basic kind (what reductions need)
    bool homogeneous(Basic* kind, int64_t* n) const;
    std::string describe() const;

    // Runs of `count` consecutive elements (element i shifted by i * extent),
    // adjacent runs merged.
    std::vector<Run> flatten(int64_t count) const;
    // Regular pattern: `n` runs of `len` bytes, `stride` bytes apart, first at `disp0`.
    bool uniform(int64_t count, int64_t* n, int64_t* len, int64_t* stride, int64_t* disp0) const;

    // Host pack / unpack (the oracle for the device pack/move kernels).
    void pack(const void* buf, int64_t count, void* packed) const;
    void unpack(const void* packed, int64_t packed_bytes, void* buf, int64_t count) const;

    bool operator==(const Datatype& o) const { return impl_ == o.impl_; }
    const void* id() const { return impl_.get(); }   // identity of the underlying type

    struct Impl;
private:
    explicit Datatype(std::shared_ptr<Impl> p);
    std::shared_ptr<Impl> impl_;
};

// One copy segment of a typed transfer: packed-stream order, both sides resolved.
struct Segment {
    int64_t src_off;
    int64_t dst_off;
    int64_t len;
};
// Intersect the run lists of sender and receiver in packed-stream coordinates:
// the result moves `min(total_src, total_dst)` bytes from the source layout
// straight into the destination layout with no intermediate packed buffer.
std::vector<Segment> merge_runs(const std::vector<Run>& src, const std::vector<Run>& dst);

}  // namespace b200
#pragma GCC visibility pop
