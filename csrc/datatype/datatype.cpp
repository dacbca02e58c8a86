// Datatype engine (host side): type constructors, flattening, commit, host
// pack/unpack, run-list merging.  See include/b200comm/datatype.hpp for the
// reference call sites each constructor covers.
#include "b200comm/datatype.hpp"

#include <algorithm>
#include <cstring>
#include <sstream>

#include "b200comm/error.hpp"

namespace b200 {

size_t basic_size(Basic b) {
    switch (b) {
        case Basic::BYTE: case Basic::CHAR: case Basic::INT8: case Basic::UINT8: return 1;
        case Basic::INT16: return 2;
        case Basic::INT32: case Basic::UINT32: case Basic::FLOAT: return 4;
        case Basic::INT64: case Basic::UINT64: case Basic::DOUBLE: return 8;
        default: return 0;
    }
}
const char* basic_name(Basic b) {
    static const char* n[] = {"BYTE", "CHAR", "INT8", "UINT8", "INT16", "INT32", "INT64",
                              "UINT64", "FLOAT", "DOUBLE", "UINT32"};
    int i = (int)b;
    return (i >= 0 && i < (int)Basic::NUM_BASIC) ? n[i] : "?";
}

// A signature is the sequence of basic types in packed order, run-length
// encoded: (kind, repeat).
struct SigItem { int kind; int64_t rep; };

struct Datatype::Impl {
    std::vector<Run> runs;         // one element, in packed (type-map) order
    std::vector<SigItem> sig;      // matching packed order
    int64_t size = 0, lb = 0, extent = 0;
    bool committed = false, basic = false;
    std::string name;
};

static void push_run(std::vector<Run>& v, int64_t disp, int64_t len) {
    if (len <= 0) return;
    if (!v.empty() && v.back().disp + v.back().len == disp) v.back().len += len;
    else v.push_back({disp, len});
}
static void push_sig(std::vector<SigItem>& s, int kind, int64_t rep) {
    if (rep <= 0) return;
    if (!s.empty() && s.back().kind == kind) s.back().rep += rep;
    else s.push_back({kind, rep});
}
// Append `count` copies of `old` (each shifted by k*extent) at byte offset `base`.
static void append_block(Datatype::Impl& dst, const Datatype::Impl& old, int64_t base,
                         int64_t count) {
    if (count <= 0) return;
    if (old.runs.size() == 1 && old.runs[0].len == old.extent) {
        // dense element: the block is one run
        push_run(dst.runs, base + old.runs[0].disp, old.extent * count);
    } else {
        for (int64_t k = 0; k < count; ++k)
            for (const Run& r : old.runs) push_run(dst.runs, base + k * old.extent + r.disp, r.len);
    }
    if (old.sig.size() == 1) push_sig(dst.sig, old.sig[0].kind, old.sig[0].rep * count);
    else
        for (int64_t k = 0; k < count; ++k)
            for (const SigItem& s : old.sig) push_sig(dst.sig, s.kind, s.rep);
    dst.size += old.size * count;
}
static void finish_bounds(Datatype::Impl& t, int64_t lb, int64_t ub, bool any) {
    t.lb = any ? lb : 0;
    t.extent = any ? ub - lb : 0;
}

Datatype::Datatype() = default;
Datatype::Datatype(std::shared_ptr<Impl> p) : impl_(std::move(p)) {}

static const Datatype::Impl& need(const std::shared_ptr<Datatype::Impl>& p) {
    B200_REQUIRE((bool)p, ERR_TYPE, "null datatype");
    return *p;
}

Datatype Datatype::basic(Basic b) {
    B200_REQUIRE((int)b >= 0 && (int)b < (int)Basic::NUM_BASIC, ERR_TYPE, "unknown basic type");
    auto p = std::make_shared<Impl>();
    const int64_t s = (int64_t)basic_size(b);
    p->runs = {{0, s}};
    p->sig = {{(int)b, 1}};
    p->size = s; p->lb = 0; p->extent = s;
    p->committed = true; p->basic = true;
    p->name = basic_name(b);
    return Datatype(p);
}

Datatype Datatype::contiguous(int count, const Datatype& old) {
    const Impl& o = need(old.impl_);
    B200_REQUIRE(count >= 0, ERR_COUNT, "negative count");
    auto p = std::make_shared<Impl>();
    append_block(*p, o, 0, count);
    finish_bounds(*p, o.lb, o.lb + o.extent * count, count > 0);
    p->name = "contiguous(" + std::to_string(count) + "," + o.name + ")";
    return Datatype(p);
}

Datatype Datatype::hvector(int count, int blocklen, int64_t stride_bytes, const Datatype& old) {
    const Impl& o = need(old.impl_);
    B200_REQUIRE(count >= 0 && blocklen >= 0, ERR_COUNT, "negative count/blocklen");
    auto p = std::make_shared<Impl>();
    int64_t lb = 0, ub = 0; bool any = false;
    for (int i = 0; i < count; ++i) {
        const int64_t base = i * stride_bytes;
        append_block(*p, o, base, blocklen);
        if (blocklen > 0) {
            const int64_t l = base + o.lb, u = base + o.lb + o.extent * blocklen;
            lb = any ? std::min(lb, l) : l; ub = any ? std::max(ub, u) : u; any = true;
        }
    }
    finish_bounds(*p, lb, ub, any);
    p->name = "hvector(" + std::to_string(count) + "," + std::to_string(blocklen) + "," +
              std::to_string(stride_bytes) + "," + o.name + ")";
    return Datatype(p);
}

Datatype Datatype::vector(int count, int blocklen, int stride, const Datatype& old) {
    return hvector(count, blocklen, (int64_t)stride * need(old.impl_).extent, old);
}

Datatype Datatype::hindexed(const std::vector<int>& blocklens,
                            const std::vector<int64_t>& byte_disps, const Datatype& old) {
    const Impl& o = need(old.impl_);
    B200_REQUIRE(blocklens.size() == byte_disps.size(), ERR_ARG,
                 "blocklens and displacements differ in length");
    auto p = std::make_shared<Impl>();
    int64_t lb = 0, ub = 0; bool any = false;
    for (size_t i = 0; i < blocklens.size(); ++i) {
        B200_REQUIRE(blocklens[i] >= 0, ERR_COUNT, "negative block length");
        append_block(*p, o, byte_disps[i], blocklens[i]);
        if (blocklens[i] > 0) {
            const int64_t l = byte_disps[i] + o.lb, u = l + o.extent * blocklens[i];
            lb = any ? std::min(lb, l) : l; ub = any ? std::max(ub, u) : u; any = true;
        }
    }
    finish_bounds(*p, lb, ub, any);
    p->name = "hindexed(" + std::to_string(blocklens.size()) + "," + o.name + ")";
    return Datatype(p);
}

Datatype Datatype::indexed(const std::vector<int>& blocklens, const std::vector<int>& disps,
                           const Datatype& old) {
    const Impl& o = need(old.impl_);
    std::vector<int64_t> bd(disps.size());
    for (size_t i = 0; i < disps.size(); ++i) bd[i] = (int64_t)disps[i] * o.extent;
    Datatype t = hindexed(blocklens, bd, old);
    t.impl_->name = "indexed(" + std::to_string(blocklens.size()) + "," + o.name + ")";
    return t;
}

Datatype Datatype::create_struct(const std::vector<int>& blocklens,
                                 const std::vector<int64_t>& byte_disps,
                                 const std::vector<Datatype>& types) {
    B200_REQUIRE(blocklens.size() == byte_disps.size() && blocklens.size() == types.size(),
                 ERR_ARG, "struct argument arrays differ in length");
    auto p = std::make_shared<Impl>();
    int64_t lb = 0, ub = 0; bool any = false;
    for (size_t i = 0; i < blocklens.size(); ++i) {
        const Impl& o = need(types[i].impl_);
        B200_REQUIRE(blocklens[i] >= 0, ERR_COUNT, "negative block length");
        append_block(*p, o, byte_disps[i], blocklens[i]);
        if (blocklens[i] > 0) {
            const int64_t l = byte_disps[i] + o.lb, u = l + o.extent * blocklens[i];
            lb = any ? std::min(lb, l) : l; ub = any ? std::max(ub, u) : u; any = true;
        }
    }
    finish_bounds(*p, lb, ub, any);
    p->name = "struct(" + std::to_string(blocklens.size()) + ")";
    return Datatype(p);
}

Datatype Datatype::subarray(const std::vector<int>& sizes, const std::vector<int>& subsizes,
                            const std::vector<int>& starts, Order order, const Datatype& old) {
    const Impl& o = need(old.impl_);
    const size_t nd = sizes.size();
    B200_REQUIRE(nd >= 1 && subsizes.size() == nd && starts.size() == nd, ERR_ARG,
                 "subarray argument arrays differ in length");
    for (size_t d = 0; d < nd; ++d)
        B200_REQUIRE(sizes[d] >= 1 && subsizes[d] >= 0 && starts[d] >= 0 &&
                         starts[d] + subsizes[d] <= sizes[d],
                     ERR_ARG, "subarray does not fit inside the array");
    // normalise to C order: dimension 0 slowest
    std::vector<int64_t> sz(nd), sub(nd), st(nd);
    for (size_t d = 0; d < nd; ++d) {
        const size_t s = order == Order::C ? d : nd - 1 - d;
        sz[d] = sizes[s]; sub[d] = subsizes[s]; st[d] = starts[s];
    }
    std::vector<int64_t> stride(nd);   // in elements of `old`
    stride[nd - 1] = 1;
    for (size_t d = nd - 1; d-- > 0;) stride[d] = stride[d + 1] * sz[d + 1];
    auto p = std::make_shared<Impl>();
    int64_t nrows = 1;
    for (size_t d = 0; d + 1 < nd; ++d) nrows *= sub[d];
    if (sub[nd - 1] == 0) nrows = 0;
    std::vector<int64_t> idx(nd, 0);
    for (int64_t r = 0; r < nrows; ++r) {
        int64_t off = st[nd - 1];
        for (size_t d = 0; d + 1 < nd; ++d) off += (st[d] + idx[d]) * stride[d];
        append_block(*p, o, off * o.extent, sub[nd - 1]);
        for (size_t d = nd - 1; d-- > 0;) {   // odometer over the outer dimensions
            if (++idx[d] < sub[d]) break;
            idx[d] = 0;
        }
    }
    // MPI: the extent of a subarray type is the extent of the FULL array
    int64_t total = 1;
    for (size_t d = 0; d < nd; ++d) total *= sz[d];
    p->lb = 0;
    p->extent = total * o.extent;
    p->name = "subarray(" + std::to_string(nd) + "d," + o.name + ")";
    return Datatype(p);
}

Datatype Datatype::resized(const Datatype& old, int64_t lb, int64_t extent) {
    const Impl& o = need(old.impl_);
    auto p = std::make_shared<Impl>(o);
    p->lb = lb; p->extent = extent; p->committed = false; p->basic = false;
    p->name = "resized(" + o.name + ")";
    return Datatype(p);
}

Datatype& Datatype::commit() {
    B200_REQUIRE((bool)impl_, ERR_TYPE, "null datatype");
    impl_->committed = true;   // runs are merged while they are appended
    return *this;
}
void Datatype::free() { impl_.reset(); }
bool Datatype::committed() const { return impl_ && impl_->committed; }
bool Datatype::is_basic() const { return impl_ && impl_->basic; }
int64_t Datatype::size() const { return need(impl_).size; }
int64_t Datatype::extent() const { return need(impl_).extent; }
int64_t Datatype::lb() const { return need(impl_).lb; }
const std::vector<Run>& Datatype::runs() const { return need(impl_).runs; }
bool Datatype::contiguous_layout() const {
    const Impl& t = need(impl_);
    return t.runs.size() == 1 && t.runs[0].disp == 0 && t.runs[0].len == t.extent;
}
bool Datatype::homogeneous(Basic* kind, int64_t* n) const {
    const Impl& t = need(impl_);
    if (t.sig.size() != 1) return false;
    if (kind) *kind = (Basic)t.sig[0].kind;
    if (n) *n = t.sig[0].rep;
    return true;
}
uint64_t Datatype::signature_hash() const {
    const Impl& t = need(impl_);
    // hash of the size-class sequence (types of equal size are interchangeable
    // on the wire; BYTE/CHAR/INT8/UINT8 all hash alike)
    uint64_t h = 1469598103934665603ull;
    int prev = -1; int64_t rep = 0;
    auto flush = [&]() {
        if (prev < 0) return;
        h = (h ^ (uint64_t)prev) * 1099511628211ull;
        h = (h ^ (uint64_t)rep) * 1099511628211ull;
    };
    for (const SigItem& s : t.sig) {
        const int cls = (int)basic_size((Basic)s.kind) * 16 +
                        ((Basic)s.kind == Basic::FLOAT || (Basic)s.kind == Basic::DOUBLE ? 1 : 0);
        if (cls == prev) rep += s.rep;
        else { flush(); prev = cls; rep = s.rep; }
    }
    flush();
    return h;
}
std::string Datatype::describe() const {
    const Impl& t = need(impl_);
    std::ostringstream os;
    os << t.name << " size=" << t.size << " extent=" << t.extent << " lb=" << t.lb << " runs=[";
    for (size_t i = 0; i < t.runs.size() && i < 8; ++i)
        os << (i ? "," : "") << "(" << t.runs[i].disp << "," << t.runs[i].len << ")";
    if (t.runs.size() > 8) os << ",...x" << t.runs.size();
    os << "]";
    return os.str();
}

std::vector<Run> Datatype::flatten(int64_t count) const {
    const Impl& t = need(impl_);
    B200_REQUIRE(count >= 0, ERR_COUNT, "negative count");
    std::vector<Run> out;
    if (t.runs.size() == 1 && t.runs[0].len == t.extent) {
        if (count > 0) out.push_back({t.runs[0].disp, t.extent * count});
        return out;
    }
    out.reserve((size_t)count * t.runs.size());
    for (int64_t k = 0; k < count; ++k)
        for (const Run& r : t.runs) push_run(out, k * t.extent + r.disp, r.len);
    return out;
}

bool Datatype::uniform(int64_t count, int64_t* n, int64_t* len, int64_t* stride,
                       int64_t* disp0) const {
    std::vector<Run> r = flatten(count);
    if (r.empty()) return false;
    const int64_t l = r[0].len;
    int64_t s = r.size() > 1 ? r[1].disp - r[0].disp : l;
    for (size_t i = 0; i < r.size(); ++i) {
        if (r[i].len != l) return false;
        if (i > 0 && r[i].disp - r[i - 1].disp != s) return false;
    }
    *n = (int64_t)r.size(); *len = l; *stride = s; *disp0 = r[0].disp;
    return true;
}

void Datatype::pack(const void* buf, int64_t count, void* packed) const {
    const Impl& t = need(impl_);
    B200_REQUIRE(t.committed, ERR_TYPE, "datatype not committed");
    const char* src = static_cast<const char*>(buf);
    char* dst = static_cast<char*>(packed);
    for (int64_t k = 0; k < count; ++k)
        for (const Run& r : t.runs) {
            std::memcpy(dst, src + k * t.extent + r.disp, (size_t)r.len);
            dst += r.len;
        }
}

void Datatype::unpack(const void* packed, int64_t packed_bytes, void* buf, int64_t count) const {
    const Impl& t = need(impl_);
    B200_REQUIRE(t.committed, ERR_TYPE, "datatype not committed");
    const char* src = static_cast<const char*>(packed);
    char* dst = static_cast<char*>(buf);
    int64_t left = packed_bytes;
    for (int64_t k = 0; k < count && left > 0; ++k)
        for (const Run& r : t.runs) {
            const int64_t n = std::min(left, r.len);
            std::memcpy(dst + k * t.extent + r.disp, src, (size_t)n);
            src += n; left -= n;
            if (left <= 0) break;
        }
}

std::vector<Segment> merge_runs(const std::vector<Run>& src, const std::vector<Run>& dst) {
    std::vector<Segment> out;
    size_t i = 0, j = 0;
    int64_t so = 0, dof = 0;   // consumed bytes inside the current runs
    while (i < src.size() && j < dst.size()) {
        const int64_t n = std::min(src[i].len - so, dst[j].len - dof);
        if (n > 0) {
            const int64_t s = src[i].disp + so, d = dst[j].disp + dof;
            if (!out.empty() && out.back().src_off + out.back().len == s &&
                out.back().dst_off + out.back().len == d)
                out.back().len += n;
            else
                out.push_back({s, d, n});
        }
        so += n; dof += n;
        if (so == src[i].len) { ++i; so = 0; }
        if (dof == dst[j].len) { ++j; dof = 0; }
    }
    return out;
}

}  // namespace b200
