// Communicators, groups, Cartesian topology, host collectives, environment.
// See include/b200comm/comm.hpp for the reference call sites covered.
#include "core/trace.hpp"
#include "b200comm/comm.hpp"

#include <unistd.h>

#include <algorithm>
#include <chrono>
#include <cstdlib>
#include <cstring>
#include <map>
#include <numeric>

#include "comm/engine.hpp"

namespace b200 {
namespace comm {

struct CommImpl {
    std::shared_ptr<Engine> eng;
    uint32_t ctx = 0;                       // p2p context; ctx + 1 is the collective context
    std::vector<int> ranks;                 // communicator rank -> world rank
    std::shared_ptr<std::vector<int>> inv;  // world rank -> communicator rank (UNDEFINED if none)
    int myrank = 0;
    bool cart = false;
    std::vector<int> dims, periods;
};

static std::shared_ptr<CommImpl> make_impl(std::shared_ptr<Engine> eng, uint32_t ctx,
                                           std::vector<int> ranks) {
    auto p = std::make_shared<CommImpl>();
    p->eng = std::move(eng);
    p->ctx = ctx;
    p->ranks = std::move(ranks);
    p->inv = std::make_shared<std::vector<int>>((size_t)p->eng->size(), UNDEFINED);
    p->myrank = UNDEFINED;
    for (size_t i = 0; i < p->ranks.size(); ++i) {
        (*p->inv)[(size_t)p->ranks[i]] = (int)i;
        if (p->ranks[i] == p->eng->rank()) p->myrank = (int)i;
    }
    return p;
}

static const CommImpl& need(const std::shared_ptr<CommImpl>& p) {
    B200_REQUIRE((bool)p, ERR_COMM, "null communicator");
    return *p;
}

// ------------------------------------------------------------------ requests
void Request::wait(Status* st) {
    if (!s_) return;
    trace::Scope ts("wait", "p2p", -1, s_->kind == ReqState::RECV ? s_->src_world : s_->dest_world,
                    s_->kind == ReqState::RECV ? s_->tag : s_->hdr.tag);
    if (!s_->done.load(std::memory_order_acquire)) s_->eng->wait(s_);
    ts.set_bytes(s_->kind == ReqState::RECV ? s_->status.bytes : (int64_t)s_->total);
    if (st) *st = s_->status;
    const int err = s_->status.error;
    if (err) B200_FAIL(err, s_->kind == ReqState::RECV ? "receive failed" : "send failed");
}
bool Request::test(Status* st) {
    if (!s_) return true;
    const bool d = s_->eng ? s_->eng->test(s_) : s_->done.load();
    if (d && st) *st = s_->status;
    return d;
}
void waitall(std::vector<Request>& reqs, std::vector<Status>* sts) {
    if (sts) sts->resize(reqs.size());
    for (size_t i = 0; i < reqs.size(); ++i) reqs[i].wait(sts ? &(*sts)[i] : nullptr);
}

// ------------------------------------------------------------------ groups
int Group::rank_of_world(int w) const {
    for (size_t i = 0; i < ranks_.size(); ++i)
        if (ranks_[i] == w) return (int)i;
    return UNDEFINED;
}
Group Group::incl(const std::vector<int>& r) const {
    std::vector<int> out;
    for (int i : r) {
        B200_REQUIRE(i >= 0 && i < size(), ERR_RANK, "group rank out of range");
        out.push_back(ranks_[(size_t)i]);
    }
    return Group(out);
}
Group Group::excl(const std::vector<int>& r) const {
    std::vector<int> out;
    for (int i = 0; i < size(); ++i)
        if (std::find(r.begin(), r.end(), i) == r.end()) out.push_back(ranks_[(size_t)i]);
    return Group(out);
}

// ------------------------------------------------------------------ comm basics
int Comm::rank() const { return need(p_).myrank; }
int Comm::size() const { return (int)need(p_).ranks.size(); }
Engine& Comm::engine() const { return *need(p_).eng; }
uint32_t Comm::context() const { return need(p_).ctx; }
int Comm::world_rank_of(int r) const {
    if (r == PROC_NULL || r == ANY_SOURCE) return r;
    const CommImpl& c = need(p_);
    B200_REQUIRE(r >= 0 && r < (int)c.ranks.size(), ERR_RANK,
                 "rank " + std::to_string(r) + " outside communicator of size " +
                     std::to_string(c.ranks.size()));
    return c.ranks[(size_t)r];
}

Request Comm::isend(const void* buf, int64_t count, const Datatype& t, int dest, int tag) const {
    const CommImpl& c = need(p_);
    trace::Scope ts("isend", "p2p", c.eng->rank(), dest, tag, t.valid() ? t.size() * count : -1);
    return Request(c.eng->isend(buf, count, t, world_rank_of(dest), tag, c.ctx));
}
Request Comm::irecv(void* buf, int64_t count, const Datatype& t, int src, int tag) const {
    const CommImpl& c = need(p_);
    trace::Scope ts("irecv", "p2p", c.eng->rank(), src, tag, t.valid() ? t.size() * count : -1);
    return Request(c.eng->irecv(buf, count, t, world_rank_of(src), tag, c.ctx, c.inv));
}
void Comm::send(const void* buf, int64_t count, const Datatype& t, int dest, int tag) const {
    trace::Scope ts("send", "p2p", need(p_).eng->rank(), dest, tag, t.valid() ? t.size() * count : -1);
    isend(buf, count, t, dest, tag).wait();
}
void Comm::recv(void* buf, int64_t count, const Datatype& t, int src, int tag, Status* st) const {
    trace::Scope ts("recv", "p2p", need(p_).eng->rank(), src, tag, t.valid() ? t.size() * count : -1);
    irecv(buf, count, t, src, tag).wait(st);
}
Status Comm::probe(int src, int tag) const {
    const CommImpl& c = need(p_);
    trace::Scope ts("probe", "p2p", c.eng->rank(), src, tag);
    return c.eng->probe(world_rank_of(src), tag, c.ctx, c.inv);
}
bool Comm::iprobe(int src, int tag, Status* st) const {
    const CommImpl& c = need(p_);
    return c.eng->iprobe(world_rank_of(src), tag, c.ctx, st, c.inv);
}
void Comm::sendrecv(const void* sbuf, int64_t scount, const Datatype& st, int dest, int stag,
                    void* rbuf, int64_t rcount, const Datatype& rt, int src, int rtag,
                    Status* status) const {
    trace::Scope ts("sendrecv", "p2p", need(p_).eng->rank(), dest, stag, st.valid() ? st.size() * scount : -1);
    Request r = irecv(rbuf, rcount, rt, src, rtag);
    Request s = isend(sbuf, scount, st, dest, stag);
    r.wait(status);
    s.wait();
}

// ------------------------------------------------------------------ collectives
namespace {

// p2p on the collective context
struct Coll {
    const CommImpl& c;
    Request isend(const void* b, int64_t n, const Datatype& t, int dest, int tag) const {
        return Request(c.eng->isend(b, n, t, c.ranks[(size_t)dest], tag, c.ctx + 1));
    }
    Request irecv(void* b, int64_t n, const Datatype& t, int src, int tag) const {
        return Request(c.eng->irecv(b, n, t, c.ranks[(size_t)src], tag, c.ctx + 1, c.inv));
    }
    void send(const void* b, int64_t n, const Datatype& t, int dest, int tag) const {
        isend(b, n, t, dest, tag).wait();
    }
    void recv(void* b, int64_t n, const Datatype& t, int src, int tag) const {
        irecv(b, n, t, src, tag).wait();
    }
};

enum CollTag { T_BARRIER = 1, T_BCAST = 2, T_REDUCE = 3, T_GATHER = 4, T_SCATTER = 5 };

template <typename T>
void apply_op(T* acc, const T* in, int64_t n, Op op) {
    for (int64_t i = 0; i < n; ++i) {
        switch (op) {
            case Op::SUM: acc[i] = acc[i] + in[i]; break;
            case Op::PROD: acc[i] = acc[i] * in[i]; break;
            case Op::MAX: acc[i] = std::max(acc[i], in[i]); break;
            case Op::MIN: acc[i] = std::min(acc[i], in[i]); break;
            case Op::LAND: acc[i] = (T)((acc[i] != (T)0) && (in[i] != (T)0)); break;
            case Op::LOR: acc[i] = (T)((acc[i] != (T)0) || (in[i] != (T)0)); break;
            default: B200_FAIL(ERR_ARG, "operation not defined for this type");
        }
    }
}
template <typename T>
void apply_op_int(T* acc, const T* in, int64_t n, Op op) {
    if (op == Op::BAND) { for (int64_t i = 0; i < n; ++i) acc[i] &= in[i]; return; }
    if (op == Op::BOR) { for (int64_t i = 0; i < n; ++i) acc[i] |= in[i]; return; }
    apply_op(acc, in, n, op);
}

// The element kind of a reduce: the datatype must be dense and of one basic kind.
Basic reduce_kind(const Datatype& t, int64_t* per_elem) {
    Basic k;
    B200_REQUIRE(t.contiguous_layout() && t.homogeneous(&k, per_elem), ERR_TYPE,
                 "reductions need a dense datatype of a single basic kind");
    return k;
}

void reduce_into(void* acc, const void* in, int64_t n, Basic k, Op op) {
    switch (k) {
        case Basic::DOUBLE: apply_op((double*)acc, (const double*)in, n, op); break;
        case Basic::FLOAT: apply_op((float*)acc, (const float*)in, n, op); break;
        case Basic::INT32: apply_op_int((int32_t*)acc, (const int32_t*)in, n, op); break;
        case Basic::UINT32: apply_op_int((uint32_t*)acc, (const uint32_t*)in, n, op); break;
        case Basic::INT64: apply_op_int((int64_t*)acc, (const int64_t*)in, n, op); break;
        case Basic::UINT64: apply_op_int((uint64_t*)acc, (const uint64_t*)in, n, op); break;
        case Basic::INT16: apply_op_int((int16_t*)acc, (const int16_t*)in, n, op); break;
        case Basic::INT8: apply_op_int((int8_t*)acc, (const int8_t*)in, n, op); break;
        case Basic::CHAR: apply_op_int((char*)acc, (const char*)in, n, op); break;
        default: apply_op_int((uint8_t*)acc, (const uint8_t*)in, n, op); break;
    }
}

}  // namespace

void Comm::barrier() const {
    trace::Scope ts("barrier", "coll", need(p_).eng->rank());
    const CommImpl& c = need(p_);
    Coll k{c};
    const int n = size(), me = rank();
    Datatype byte = Datatype::basic(Basic::BYTE);
    // dissemination barrier: ceil(log2 n) rounds of zero-byte messages
    for (int d = 1; d < n; d <<= 1) {
        Request r = k.irecv(nullptr, 0, byte, (me - d + n) % n, T_BARRIER);
        Request s = k.isend(nullptr, 0, byte, (me + d) % n, T_BARRIER);
        r.wait();
        s.wait();
    }
}

void Comm::bcast(void* buf, int64_t count, const Datatype& t, int root) const {
    trace::Scope ts("bcast", "coll", need(p_).eng->rank());
    const CommImpl& c = need(p_);
    Coll k{c};
    const int n = size(), me = rank();
    B200_REQUIRE(root >= 0 && root < n, ERR_RANK, "invalid root");
    const int rel = (me - root + n) % n;
    // binomial tree
    int mask = 1;
    while (mask < n) {
        if (rel & mask) {
            k.recv(buf, count, t, (rel - mask + root) % n, T_BCAST);
            break;
        }
        mask <<= 1;
    }
    mask >>= 1;
    while (mask > 0) {
        if (rel + mask < n) k.send(buf, count, t, (rel + mask + root) % n, T_BCAST);
        mask >>= 1;
    }
}

void Comm::reduce(const void* sbuf, void* rbuf, int64_t count, const Datatype& t, Op op,
                  int root) const {
    trace::Scope ts("reduce", "coll", need(p_).eng->rank());
    const CommImpl& c = need(p_);
    Coll k{c};
    const int n = size(), me = rank();
    B200_REQUIRE(root >= 0 && root < n, ERR_RANK, "invalid root");
    int64_t per = 1;
    const Basic kind = reduce_kind(t, &per);
    const int64_t bytes = t.size() * count, elems = per * count;
    std::vector<char> acc((size_t)bytes), tmp((size_t)bytes);
    if (bytes) std::memcpy(acc.data(), sbuf, (size_t)bytes);
    const int rel = (me - root + n) % n;
    // binomial tree towards relative rank 0; children are combined in rank order
    // so that the result does not depend on arrival order
    int mask = 1;
    while (mask < n) {
        if (rel & mask) {
            k.send(acc.data(), count, t, (rel - mask + root) % n, T_REDUCE);
            break;
        }
        if (rel + mask < n) {
            k.recv(tmp.data(), count, t, (rel + mask + root) % n, T_REDUCE);
            reduce_into(acc.data(), tmp.data(), elems, kind, op);
        }
        mask <<= 1;
    }
    if (me == root && bytes) std::memcpy(rbuf, acc.data(), (size_t)bytes);
}

void Comm::allreduce(const void* sbuf, void* rbuf, int64_t count, const Datatype& t, Op op) const {
    trace::Scope ts("allreduce", "coll", need(p_).eng->rank());
    std::vector<char> tmp((size_t)(t.size() * count));
    reduce(sbuf, rank() == 0 ? rbuf : tmp.data(), count, t, op, 0);
    bcast(rbuf, count, t, 0);
}

void Comm::gather(const void* sbuf, int64_t scount, const Datatype& st, void* rbuf,
                  int64_t rcount, const Datatype& rt, int root) const {
    trace::Scope ts("gather", "coll", need(p_).eng->rank());
    const CommImpl& c = need(p_);
    Coll k{c};
    const int n = size(), me = rank();
    B200_REQUIRE(root >= 0 && root < n, ERR_RANK, "invalid root");
    if (me != root) {   // non-root receive arguments are ignored (mpi6.cpp:96-100)
        k.send(sbuf, scount, st, root, T_GATHER);
        return;
    }
    std::vector<Request> reqs;
    char* out = static_cast<char*>(rbuf);
    for (int r = 0; r < n; ++r) {
        char* slot = out + (int64_t)r * rcount * rt.extent();
        if (r == me) {
            std::vector<char> packed((size_t)(st.size() * scount));
            st.pack(sbuf, scount, packed.data());
            rt.unpack(packed.data(), (int64_t)packed.size(), slot, rcount);
        } else {
            reqs.push_back(k.irecv(slot, rcount, rt, r, T_GATHER));
        }
    }
    waitall(reqs);
}

void Comm::scatter(const void* sbuf, int64_t scount, const Datatype& st, void* rbuf,
                   int64_t rcount, const Datatype& rt, int root) const {
    trace::Scope ts("scatter", "coll", need(p_).eng->rank());
    const CommImpl& c = need(p_);
    Coll k{c};
    const int n = size(), me = rank();
    B200_REQUIRE(root >= 0 && root < n, ERR_RANK, "invalid root");
    if (me != root) {
        k.recv(rbuf, rcount, rt, root, T_SCATTER);
        return;
    }
    std::vector<Request> reqs;
    const char* in = static_cast<const char*>(sbuf);
    for (int r = 0; r < n; ++r) {
        const char* slot = in + (int64_t)r * scount * st.extent();
        if (r == me) {
            std::vector<char> packed((size_t)(st.size() * scount));
            st.pack(slot, scount, packed.data());
            rt.unpack(packed.data(), (int64_t)packed.size(), rbuf, rcount);
        } else {
            reqs.push_back(k.isend(slot, scount, st, r, T_SCATTER));
        }
    }
    waitall(reqs);
}

void Comm::allgather(const void* sbuf, int64_t scount, const Datatype& st, void* rbuf,
                     int64_t rcount, const Datatype& rt) const {
    trace::Scope ts("allgather", "coll", need(p_).eng->rank());
    gather(sbuf, scount, st, rbuf, rcount, rt, 0);
    bcast(rbuf, rcount * size(), rt, 0);
}

// ------------------------------------------------------------------ groups / sub-comms
Group Comm::group() const { return Group(need(p_).ranks); }

Comm Comm::create(const Group& g) const {
    const CommImpl& c = need(p_);
    // rank 0 of the parent draws a globally unique context pair and broadcasts it
    uint32_t ctx = 0;
    if (rank() == 0) ctx = c.eng->alloc_contexts(1);
    bcast(&ctx, 1, Datatype::basic(Basic::UINT32), 0);
    for (int w : g.world_ranks())
        B200_REQUIRE((*c.inv)[(size_t)w] != UNDEFINED, ERR_RANK,
                     "group member is not part of the parent communicator");
    if (g.rank_of_world(c.eng->rank()) == UNDEFINED) return Comm();   // MPI_COMM_NULL
    return Comm(make_impl(c.eng, ctx, g.world_ranks()));
}

Comm Comm::dup() const { return create(group()); }

Comm Comm::split(int color, int key) const {
    const CommImpl& c = need(p_);
    const int n = size();
    std::vector<int> mine = {color, key}, all((size_t)2 * n);
    Datatype i32 = Datatype::basic(Basic::INT32);
    allgather(mine.data(), 2, i32, all.data(), 2, i32);
    std::vector<int> colors;
    for (int r = 0; r < n; ++r)
        if (all[2 * r] != UNDEFINED) colors.push_back(all[2 * r]);
    std::sort(colors.begin(), colors.end());
    colors.erase(std::unique(colors.begin(), colors.end()), colors.end());
    uint32_t base = 0;
    if (rank() == 0) base = c.eng->alloc_contexts((uint32_t)std::max<size_t>(colors.size(), 1));
    bcast(&base, 1, Datatype::basic(Basic::UINT32), 0);
    if (color == UNDEFINED) return Comm();
    const size_t idx = std::lower_bound(colors.begin(), colors.end(), color) - colors.begin();
    std::vector<std::pair<std::pair<int, int>, int>> members;   // ((key, parent rank), world)
    for (int r = 0; r < n; ++r)
        if (all[2 * r] == color) members.push_back({{all[2 * r + 1], r}, c.ranks[(size_t)r]});
    std::sort(members.begin(), members.end());
    std::vector<int> ranks;
    for (auto& m : members) ranks.push_back(m.second);
    return Comm(make_impl(c.eng, base + 2 * (uint32_t)idx, ranks));
}

void Comm::free() { p_.reset(); }

// ------------------------------------------------------------------ Cartesian topology
std::vector<int> dims_create(int nranks, int ndims) {
    B200_REQUIRE(nranks >= 1 && ndims >= 1, ERR_ARG, "nranks and ndims must be positive");
    std::vector<int> dims((size_t)ndims, 1), primes;
    int n = nranks;
    for (int p = 2; p * p <= n; ++p)
        while (n % p == 0) { primes.push_back(p); n /= p; }
    if (n > 1) primes.push_back(n);
    std::sort(primes.rbegin(), primes.rend());
    for (int f : primes) *std::min_element(dims.begin(), dims.end()) *= f;
    std::sort(dims.begin(), dims.end());
    return dims;
}

Comm Comm::cart_create(const std::vector<int>& dims, const std::vector<int>& periods,
                       bool reorder) const {
    (void)reorder;   // ranks are never reordered: every GPU pair is equidistant over NVSwitch
    const CommImpl& c = need(p_);
    B200_REQUIRE(!dims.empty() && dims.size() == periods.size(), ERR_TOPOLOGY,
                 "dims and periods must have the same, non-zero length");
    int64_t total = 1;
    for (int d : dims) {
        B200_REQUIRE(d >= 1, ERR_TOPOLOGY, "dimension sizes must be positive");
        total *= d;
    }
    B200_REQUIRE(total <= size(), ERR_TOPOLOGY,
                 "Cartesian grid (" + std::to_string(total) + " ranks) larger than the communicator (" +
                     std::to_string(size()) + ")");
    std::vector<int> members((size_t)total);
    std::iota(members.begin(), members.end(), 0);
    Comm sub = create(group().incl(members));
    if (sub.is_null()) return sub;
    sub.p_->cart = true;
    sub.p_->dims = dims;
    sub.p_->periods = periods;
    (void)c;
    return sub;
}

bool Comm::is_cart() const { return need(p_).cart; }
std::vector<int> Comm::cart_dims() const { return need(p_).dims; }
std::vector<int> Comm::cart_periods() const { return need(p_).periods; }

std::vector<int> Comm::cart_coords(int r) const {
    const CommImpl& c = need(p_);
    B200_REQUIRE(c.cart, ERR_TOPOLOGY, "communicator has no Cartesian topology");
    B200_REQUIRE(r >= 0 && r < size(), ERR_RANK, "rank out of range");
    std::vector<int> out(c.dims.size());
    for (size_t d = c.dims.size(); d-- > 0;) {
        out[d] = r % c.dims[d];
        r /= c.dims[d];
    }
    return out;
}

int Comm::cart_rank(const std::vector<int>& coords) const {
    const CommImpl& c = need(p_);
    B200_REQUIRE(c.cart, ERR_TOPOLOGY, "communicator has no Cartesian topology");
    B200_REQUIRE(coords.size() == c.dims.size(), ERR_TOPOLOGY, "wrong number of coordinates");
    int r = 0;
    for (size_t d = 0; d < c.dims.size(); ++d) {
        int x = coords[d];
        if (c.periods[d]) x = ((x % c.dims[d]) + c.dims[d]) % c.dims[d];
        else if (x < 0 || x >= c.dims[d]) return PROC_NULL;
        r = r * c.dims[d] + x;
    }
    return r;
}

void Comm::cart_shift(int dim, int disp, int* src, int* dst) const {
    const CommImpl& c = need(p_);
    B200_REQUIRE(c.cart, ERR_TOPOLOGY, "communicator has no Cartesian topology");
    B200_REQUIRE(dim >= 0 && dim < (int)c.dims.size(), ERR_TOPOLOGY, "dimension out of range");
    std::vector<int> me = cart_coords(rank()), up = me, dn = me;
    up[(size_t)dim] += disp;
    dn[(size_t)dim] -= disp;
    if (src) *src = cart_rank(dn);
    if (dst) *dst = cart_rank(up);
}

// ------------------------------------------------------------------ environment
std::vector<Comm> init_local(int nranks, size_t ring_bytes) {
    std::shared_ptr<Fabric> f(Fabric::create_local(nranks, ring_bytes));
    std::vector<int> all((size_t)nranks);
    std::iota(all.begin(), all.end(), 0);
    std::vector<Comm> out;
    for (int r = 0; r < nranks; ++r)
        out.emplace_back(make_impl(std::make_shared<Engine>(f, r), 2, all));
    return out;
}

Comm init_process(const std::string& job, int rank, int nranks, size_t ring_bytes,
                  double timeout_s) {
    std::string name = job;
    if (name.empty() || name[0] != '/') name = "/" + name;
    std::shared_ptr<Fabric> f(Fabric::attach_shm(name, rank, nranks, ring_bytes, timeout_s));
    f->startup_barrier(rank);
    std::vector<int> all((size_t)nranks);
    std::iota(all.begin(), all.end(), 0);
    return Comm(make_impl(std::make_shared<Engine>(f, rank), 2, all));
}

static int env_i(const char* n, int dflt) {
    const char* v = std::getenv(n);
    return (v && *v) ? std::atoi(v) : dflt;
}

Comm init_from_env() {
    int rank = env_i("B200_RANK", -1), size = env_i("B200_WORLD_SIZE", -1);
    if (rank < 0) { rank = env_i("RANK", -1); size = env_i("WORLD_SIZE", -1); }
    if (rank < 0) { rank = env_i("OMPI_COMM_WORLD_RANK", -1); size = env_i("OMPI_COMM_WORLD_SIZE", -1); }
    if (rank < 0) { rank = env_i("PMI_RANK", -1); size = env_i("PMI_SIZE", -1); }
    if (rank < 0) { rank = env_i("SLURM_PROCID", -1); size = env_i("SLURM_NTASKS", -1); }
    if (rank < 0 || size < 1) { rank = 0; size = 1; }
    std::string job;
    if (const char* j = std::getenv("B200_JOB")) job = j;
    else if (const char* p = std::getenv("MASTER_PORT")) job = std::string("b200comm-") + p;
    else job = "b200comm-" + std::to_string((long)getppid());
    job += "-" + std::to_string((long)getuid());
    return init_process(job, rank, size);
}

void finalize(Comm& world) {
    if (world.is_null()) return;
    world.barrier();
    world.free();
}

[[noreturn]] void abort(const Comm& world, int code) {
    if (!world.is_null()) world.engine().fabric().header()->aborted.store(1);
    std::_Exit(code ? code : 1);
}

double wtime() {
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

std::string processor_name() {
    char buf[256] = {0};
    gethostname(buf, sizeof(buf) - 1);
    return buf;
}

}  // namespace comm
}  // namespace b200
