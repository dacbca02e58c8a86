// Halo-exchange mini-library implementation -- see include/b200comm/stencil2d.hpp.
#include <sstream>

#include "b200comm.h"
#include "b200comm/stencil2d.hpp"
#include "core/capi_util.hpp"
#include "core/trace.hpp"
#include "core/handles.hpp"

namespace b200 {
namespace stencil2d {

std::ostream& operator<<(std::ostream& os, const Array2D& a) {
    return os << "width:  " << a.width << ", height: " << a.height << ", x offset: " << a.xOffset
              << ", y offset: " << a.yOffset;
}

const char* region_name(RegionID r) {
    static const char* n[] = {"top left", "top center", "top right", "center left", "center",
                              "center right", "bottom left", "bottom center", "bottom right",
                              "top", "left", "bottom", "right"};
    return (r >= 0 && r < NUM_REGIONS) ? n[r] : "?";
}

namespace {
// A region is a (column band, row band) pair.
enum Band { LOW, MID, HIGH, FULL };
struct BandPair { Band col, row; };
const BandPair kBands[NUM_REGIONS] = {
    {LOW, LOW},  {MID, LOW},  {HIGH, LOW},    // top left / center / right
    {LOW, MID},  {MID, MID},  {HIGH, MID},    // center left / center / right
    {LOW, HIGH}, {MID, HIGH}, {HIGH, HIGH},   // bottom left / center / right
    {FULL, LOW}, {LOW, FULL}, {FULL, HIGH}, {HIGH, FULL}};   // top, left, bottom, right

void band_extent(Band b, int total, int ghost, int* off, int* len) {
    switch (b) {
        case LOW: *off = 0; *len = ghost; break;
        case MID: *off = ghost; *len = total - 2 * ghost; break;
        case HIGH: *off = total - ghost; *len = ghost; break;
        default: *off = 0; *len = total; break;
    }
}

// Send region (slice of the core) and receive region (ghost cells of the tile)
// associated with a neighbour direction.
const RegionID kSendRegion[NUM_CELLS] = {TOP_LEFT, TOP, TOP_RIGHT, LEFT, RIGHT,
                                         BOTTOM_LEFT, BOTTOM, BOTTOM_RIGHT};
const RegionID kGhostRegion[NUM_CELLS] = {TOP_LEFT, TOP_CENTER, TOP_RIGHT, CENTER_LEFT,
                                          CENTER_RIGHT, BOTTOM_LEFT, BOTTOM_CENTER, BOTTOM_RIGHT};
}  // namespace

Array2D sub_array_region(const Array2D& g, int stencilWidth, int stencilHeight, RegionID rid) {
    B200_REQUIRE(rid >= 0 && rid < NUM_REGIONS, ERR_ARG, "unknown region id");
    const int gw = stencilWidth / 2, gh = stencilHeight / 2;
    B200_REQUIRE(gw >= 0 && gh >= 0 && g.width >= 2 * gw && g.height >= 2 * gh, ERR_ARG,
                 "array smaller than twice the ghost width");
    int xo, w, yo, h;
    band_extent(kBands[rid].col, g.width, gw, &xo, &w);
    band_extent(kBands[rid].row, g.height, gh, &yo, &h);
    return Array2D(w, h, g.rowStride, g.xOffset + xo, g.yOffset + yo);
}

void cell_offset(GridCellID c, int* dx, int* dy) {
    static const int ox[NUM_CELLS] = {-1, 0, 1, -1, 1, -1, 0, 1};
    static const int oy[NUM_CELLS] = {-1, -1, -1, 0, 0, 1, 1, 1};
    B200_REQUIRE(c >= 0 && c < NUM_CELLS, ERR_ARG, "unknown grid cell id");
    *dx = ox[c];
    *dy = oy[c];
}

int offset_rank(const comm::Comm& cart, int dx, int dy) {
    std::vector<int> co = cart.cart_coords(cart.rank());
    B200_REQUIRE(co.size() == 2, ERR_TOPOLOGY, "halo exchange needs a 2-D Cartesian communicator");
    co[0] += dy;   // dimension 0 = rows
    co[1] += dx;   // dimension 1 = columns
    return cart.cart_rank(co);
}

Datatype make_subarray_type(const Array2D& g, const Array2D& sub, const Datatype& elem) {
    const int rows = g.yOffset + g.height;
    Datatype t = Datatype::subarray({rows, g.rowStride}, {sub.height, sub.width},
                                    {sub.yOffset, sub.xOffset}, Order::C, elem);
    t.commit();
    return t;
}

TransferInfo make_recv_info(const comm::Comm& cart, GridCellID remote_source, RegionID local_target,
                            const Array2D& tile, int stencilWidth, int stencilHeight, int tag,
                            const Datatype& elem) {
    int dx, dy;
    cell_offset(remote_source, &dx, &dy);
    TransferInfo r;
    r.region = local_target;
    r.layout = sub_array_region(tile, stencilWidth, stencilHeight, local_target);
    r.type = make_subarray_type(tile, r.layout, elem);
    r.peer = offset_rank(cart, dx, dy);
    r.tag = tag;
    return r;
}

TransferInfo make_send_info(const comm::Comm& cart, GridCellID remote_target, RegionID local_source,
                            const Array2D& tile, int stencilWidth, int stencilHeight, int tag,
                            const Datatype& elem) {
    int dx, dy;
    cell_offset(remote_target, &dx, &dy);
    const Array2D core_l = sub_array_region(tile, stencilWidth, stencilHeight, CENTER);
    TransferInfo s;
    s.region = local_source;
    s.layout = sub_array_region(core_l, stencilWidth, stencilHeight, local_source);
    s.type = make_subarray_type(tile, s.layout, elem);
    s.peer = offset_rank(cart, dx, dy);
    s.tag = tag;
    return s;
}

HaloPlan::HaloPlan(const comm::Comm& cart, const Array2D& tile, int stencilWidth,
                   int stencilHeight, const Datatype& elem)
    : cart_(cart), tile_(tile), sw_(stencilWidth), sh_(stencilHeight) {
    B200_REQUIRE(cart.is_cart(), ERR_TOPOLOGY, "halo plan needs a Cartesian communicator");
    for (int c = 0; c < NUM_CELLS; ++c) {
        // send: the slice of my core that faces neighbour c, to neighbour c; the tag is
        // the id of the send region
        const RegionID sreg = kSendRegion[c];
        send_.push_back(make_send_info(cart, (GridCellID)c, sreg, tile_, sw_, sh_, (int)sreg, elem));
        // receive: my ghost cells on the OPPOSITE side, from the neighbour there; that
        // neighbour sends the region facing direction c, hence the same tag
        const int opp = NUM_CELLS - 1 - c;
        recv_.push_back(make_recv_info(cart, (GridCellID)opp, kGhostRegion[opp], tile_, sw_, sh_,
                                       (int)sreg, elem));
    }
}

void HaloPlan::exchange(void* data) const {
    trace::Scope ts("halo_exchange", "halo", -1, -3, -1, bytes_sent_per_exchange());
    std::vector<comm::Request> reqs;
    reqs.reserve(recv_.size() + send_.size());
    for (const TransferInfo& t : recv_) reqs.push_back(cart_.irecv(data, 1, t.type, t.peer, t.tag));
    for (const TransferInfo& t : send_) reqs.push_back(cart_.isend(data, 1, t.type, t.peer, t.tag));
    comm::waitall(reqs);
}

int64_t HaloPlan::bytes_sent_per_exchange() const {
    int64_t n = 0;
    for (const TransferInfo& t : send_)
        if (t.peer != comm::PROC_NULL) n += t.type.size();
    return n;
}

void print_cart_grid(std::ostream& os, const comm::Comm& cart) {
    std::vector<int> d = cart.cart_dims();
    B200_REQUIRE(d.size() == 2, ERR_TOPOLOGY, "need a 2-D Cartesian communicator");
    for (int r = 0; r < d[0]; ++r) {
        for (int c = 0; c < d[1]; ++c) os << cart.cart_rank({r, c}) << ' ';
        os << '\n';
    }
}

std::string region_report(int w, int h, int stencilWidth, int stencilHeight) {
    std::ostringstream os;
    const int tw = w + 2 * (stencilWidth / 2), th = h + 2 * (stencilHeight / 2);
    Array2D grid(tw, th, tw);
    os << "GRID\nWidth: " << tw << ", Height: " << th << "\nStencil: " << stencilWidth << ", "
       << stencilHeight << "\n";
    for (int r = TOP_LEFT; r <= BOTTOM_RIGHT; ++r)
        os << region_name((RegionID)r) << ": " << sub_array_region(grid, stencilWidth, stencilHeight, (RegionID)r) << "\n";
    Array2D core = sub_array_region(grid, stencilWidth, stencilHeight, CENTER);
    os << "SUBGRID (core)\nWidth: " << core.width << ", Height: " << core.height << "\n";
    for (int r = TOP_LEFT; r < NUM_REGIONS; ++r)
        os << region_name((RegionID)r) << ": " << sub_array_region(core, stencilWidth, stencilHeight, (RegionID)r) << "\n";
    return os.str();
}

}  // namespace stencil2d
}  // namespace b200

// ------------------------------------------------------------------ C ABI
using namespace b200;
using namespace b200::stencil2d;

struct b2c_halo_plan { HaloPlan p; };

extern "C" {

B2C_EXPORT int b2c_sub_array_region(const b2c_array2d* g, int stencil_w, int stencil_h, int region,
                                    b2c_array2d* out) {
    B2C_TRY
    B200_REQUIRE(g && out, ERR_ARG, "null argument");
    Array2D a(g->width, g->height, g->row_stride, g->x_offset, g->y_offset);
    Array2D s = sub_array_region(a, stencil_w, stencil_h, (RegionID)region);
    out->width = s.width; out->height = s.height; out->x_offset = s.xOffset;
    out->y_offset = s.yOffset; out->row_stride = s.rowStride;
    B2C_CATCH
}

B2C_EXPORT int b2c_halo_plan_create(b2c_comm* cart, const b2c_array2d* tile, int stencil_w,
                                    int stencil_h, b2c_type* elem, b2c_halo_plan** out) {
    B2C_TRY
    B200_REQUIRE(cart && tile && elem && out, ERR_ARG, "null argument");
    Array2D a(tile->width, tile->height, tile->row_stride, tile->x_offset, tile->y_offset);
    *out = new b2c_halo_plan{HaloPlan(cart->c, a, stencil_w, stencil_h, elem->t)};
    B2C_CATCH
}
B2C_EXPORT int b2c_halo_plan_destroy(b2c_halo_plan* p) {
    B2C_TRY delete p; B2C_CATCH
}
B2C_EXPORT int b2c_halo_exchange(b2c_halo_plan* p, void* tile_data) {
    B2C_TRY
    B200_REQUIRE(p && tile_data, ERR_ARG, "null argument");
    p->p.exchange(tile_data);
    B2C_CATCH
}
// which: 0 = receives, 1 = sends; entry i in [0,8)
B2C_EXPORT int b2c_halo_plan_entry(b2c_halo_plan* p, int which, int i, int* peer, int* tag,
                                   int* region, b2c_array2d* layout, int64_t* bytes) {
    B2C_TRY
    B200_REQUIRE(p && i >= 0 && i < NUM_CELLS, ERR_ARG, "bad argument");
    const TransferInfo& t = which ? p->p.sends()[(size_t)i] : p->p.recvs()[(size_t)i];
    if (peer) *peer = t.peer;
    if (tag) *tag = t.tag;
    if (region) *region = (int)t.region;
    if (bytes) *bytes = t.type.size();
    if (layout) {
        layout->width = t.layout.width; layout->height = t.layout.height;
        layout->x_offset = t.layout.xOffset; layout->y_offset = t.layout.yOffset;
        layout->row_stride = t.layout.rowStride;
    }
    B2C_CATCH
}
B2C_EXPORT int b2c_region_report(int w, int h, int stencil_w, int stencil_h, char* buf, int cap) {
    B2C_TRY
    std::string s = region_report(w, h, stencil_w, stencil_h);
    B200_REQUIRE(cap > 0, ERR_ARG, "empty buffer");
    snprintf(buf, (size_t)cap, "%s", s.c_str());
    B2C_CATCH
}

}  // extern "C"
