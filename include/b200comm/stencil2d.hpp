// 2-D halo-exchange mini-library: layout descriptors, the 13 regions of a tile,
// the 8-neighbour exchange plan and the (unfused) exchange itself.
//
// Capability parity with stencil2d/stencil2D.h of the reference:
//   Array2D / Array2DAccessor (:30-75), RegionID (:79-82), MPIGridCellID
//   (:86-88), SubArrayRegion (:107-201), CreateMPISubArrayType (:210-228),
//   OffsetTaskId / MPIOffsetRegion (:232-299), TransferInfo (:303-311),
//   CreateReceiveInfo / CreateSendInfo (:319-359), CreateSendRecvArrays
//   (:381-437), ExchangeData (:363-377), Print (:92-102), PrintCartesianGrid
//   (:513-530), TestSubRegionExtraction (:441-510).
//
// Differences by design: ONE axis convention -- x is the fast
// axis, y the slow axis, cart dimension 0 is the row (y) -- instead of the
// reference's two cancelling transpositions (#1); rectangular tiles, stencils
// and process grids; the plan owns and frees its datatypes (#4); works on host
// AND device buffers: a device tile is moved by the pack/move kernel straight
// into the neighbour's ghost region.
#pragma once
#include <ostream>
#include <string>
#include <vector>

#include "b200comm/comm.hpp"
#include "b200comm/datatype.hpp"

#if defined(__CUDACC__)
#define B200_HD __host__ __device__
#else
#define B200_HD
#endif

#pragma GCC visibility push(default)   // public C++ API: exported from the library
namespace b200 {
namespace stencil2d {

// Pure layout descriptor; the data pointer is kept separately (stencil2d/README.md:10-13).
struct Array2D {
    int width = 0, height = 0;     // elements
    int xOffset = 0, yOffset = 0;  // position inside the enclosing allocation
    int rowStride = 0;             // elements between consecutive rows of the allocation
    B200_HD Array2D() {}
    B200_HD Array2D(int w, int h, int rs, int xoff = 0, int yoff = 0)
        : width(w), height(h), xOffset(xoff), yOffset(yoff), rowStride(rs) {}
};
std::ostream& operator<<(std::ostream& os, const Array2D& a);

template <typename T>
class Array2DAccessor {
public:
    B200_HD Array2DAccessor() : data_(nullptr) {}
    B200_HD Array2DAccessor(T* data, const Array2D& layout) : data_(data), layout_(layout) {}
    B200_HD T& operator()(int x, int y) const {
        return data_[(size_t)(layout_.yOffset + y) * layout_.rowStride + layout_.xOffset + x];
    }
    B200_HD const Array2D& layout() const { return layout_; }
private:
    T* data_;
    Array2D layout_;
};

// 9 grid cells + 4 full edges (same numbering as the reference, :79-82).
enum RegionID {
    TOP_LEFT = 0, TOP_CENTER, TOP_RIGHT, CENTER_LEFT, CENTER, CENTER_RIGHT,
    BOTTOM_LEFT, BOTTOM_CENTER, BOTTOM_RIGHT, TOP, LEFT, BOTTOM, RIGHT, NUM_REGIONS
};
// 8 neighbour directions (:86-88).
enum GridCellID {
    CELL_TOP_LEFT = 0, CELL_TOP_CENTER, CELL_TOP_RIGHT, CELL_CENTER_LEFT, CELL_CENTER_RIGHT,
    CELL_BOTTOM_LEFT, CELL_BOTTOM_CENTER, CELL_BOTTOM_RIGHT, NUM_CELLS
};
const char* region_name(RegionID r);

// Sub-layout of region `rid` of `g`; ghost width = stencilWidth/2, ghost height
// = stencilHeight/2.  The result keeps g's rowStride so it addresses the same
// allocation.
Array2D sub_array_region(const Array2D& g, int stencilWidth, int stencilHeight, RegionID rid);

// (dx, dy) of a neighbour cell: x = column offset, y = row offset, row 0 on top.
void cell_offset(GridCellID c, int* dx, int* dy);
// Rank of the neighbour at (dx, dy) on a 2-D Cartesian communicator whose
// dimension 0 is the row axis; PROC_NULL outside a non-periodic grid.
int offset_rank(const comm::Comm& cart, int dx, int dy);

// Customisation point: the datatype of one array element.  The reference declares
// CreateArrayElementType<T>() and lets client code specialise it (stencil2D.h:206,
// mpi-2d-stencil-subarray.cpp:10 for double); here the arithmetic types are provided
// and user types are added by specialising this template the same way.
template <typename T> Datatype element_type();
template <> inline Datatype element_type<double>() { return Datatype::basic(Basic::DOUBLE); }
template <> inline Datatype element_type<float>() { return Datatype::basic(Basic::FLOAT); }
template <> inline Datatype element_type<int>() { return Datatype::basic(Basic::INT32); }
template <> inline Datatype element_type<long long>() { return Datatype::basic(Basic::INT64); }
template <> inline Datatype element_type<char>() { return Datatype::basic(Basic::CHAR); }
template <> inline Datatype element_type<unsigned char>() { return Datatype::basic(Basic::UINT8); }

// Subarray datatype selecting `sub` inside the allocation that `g` describes.
Datatype make_subarray_type(const Array2D& g, const Array2D& sub, const Datatype& elem);

struct TransferInfo {
    int peer = comm::PROC_NULL;   // source (receive) or destination (send) rank
    int tag = 0;
    RegionID region = CENTER;
    Array2D layout;               // the region inside the tile allocation
    Datatype type;                // committed subarray type
};

// One entry of the plan (CreateReceiveInfo / CreateSendInfo, stencil2D.h:319-359):
// `make_recv_info` describes ghost region `local_target` of the full tile, filled by the
// neighbour in direction `remote_source`; `make_send_info` describes region `local_source`
// of the CORE, sent to the neighbour in direction `remote_target`.
TransferInfo make_recv_info(const comm::Comm& cart, GridCellID remote_source, RegionID local_target,
                            const Array2D& tile, int stencilWidth, int stencilHeight, int tag,
                            const Datatype& elem);
TransferInfo make_send_info(const comm::Comm& cart, GridCellID remote_target, RegionID local_source,
                            const Array2D& tile, int stencilWidth, int stencilHeight, int tag,
                            const Datatype& elem);

// The 8 (send, receive) pairs of one halo exchange.  Tag = id of the SEND
// region, so messages stay distinct even when several directions map to the
// same peer (1x1, 1x2, 2x2 periodic grids) -- stencil2D.h:422.
class HaloPlan {
public:
    HaloPlan() = default;
    HaloPlan(const comm::Comm& cart, const Array2D& tile, int stencilWidth, int stencilHeight,
             const Datatype& elem);
    // All receives, then all sends, one waitall (ExchangeData, :363-377).
    void exchange(void* tile_data) const;
    const std::vector<TransferInfo>& recvs() const { return recv_; }
    const std::vector<TransferInfo>& sends() const { return send_; }
    const Array2D& tile() const { return tile_; }
    Array2D core() const { return sub_array_region(tile_, sw_, sh_, CENTER); }
    int64_t bytes_sent_per_exchange() const;
private:
    comm::Comm cart_;
    Array2D tile_;
    int sw_ = 0, sh_ = 0;
    std::vector<TransferInfo> recv_, send_;
};

// Row-major dump, every value followed by a blank (Print, :92-102).
template <typename T>
void print_array(const T* data, const Array2D& g, std::ostream& os) {
    Array2DAccessor<const T> a(data, g);
    for (int row = 0; row < g.height; ++row) {
        for (int col = 0; col < g.width; ++col) os << a(col, row) << ' ';
        os << '\n';
    }
}
// Rank layout of a 2-D Cartesian communicator (PrintCartesianGrid, :513-530).
void print_cart_grid(std::ostream& os, const comm::Comm& cart);
// The reference's print-only self test (:441-510) as text, with the grid sized
// correctly (w + 2*(stencil/2); the reference forgets the factor 2, 2.7 #9).
std::string region_report(int w, int h, int stencilWidth, int stencilHeight);

}  // namespace stencil2d
}  // namespace b200
#pragma GCC visibility pop
