// C ABI of the comm + datatype layers (declared in include/b200comm.h).
#include <cstring>
#include <string>
#include <vector>

#include "b200comm.h"
#include "b200comm/comm.hpp"
#include "b200comm/datatype.hpp"
#include "comm/engine.hpp"
#include "core/capi_util.hpp"
#include "core/handles.hpp"

using namespace b200;
using namespace b200::comm;


static const Datatype& T(const b2c_type* t) {
    B200_REQUIRE(t != nullptr, ERR_TYPE, "null datatype handle");
    return t->t;
}
static const Comm& Cm(const b2c_comm* c) {
    B200_REQUIRE(c != nullptr && !c->c.is_null(), ERR_COMM, "null communicator handle");
    return c->c;
}
static void put_status(b2c_status* out, const Status& s) {
    if (!out) return;
    out->source = s.source; out->tag = s.tag; out->error = s.error; out->bytes = s.bytes;
}
static b2c_type* wrap(Datatype t) { return new b2c_type{std::move(t)}; }
static b2c_comm* wrap(Comm c) { return c.is_null() ? nullptr : new b2c_comm{std::move(c)}; }

extern "C" {

// ------------------------------------------------------------------ datatypes
B2C_EXPORT int b2c_type_basic(int kind, b2c_type** out) {
    B2C_TRY *out = wrap(Datatype::basic((Basic)kind)); B2C_CATCH
}
B2C_EXPORT int b2c_type_contiguous(int count, b2c_type* old, b2c_type** out) {
    B2C_TRY *out = wrap(Datatype::contiguous(count, T(old))); B2C_CATCH
}
B2C_EXPORT int b2c_type_vector(int count, int blocklen, int stride, b2c_type* old, b2c_type** out) {
    B2C_TRY *out = wrap(Datatype::vector(count, blocklen, stride, T(old))); B2C_CATCH
}
B2C_EXPORT int b2c_type_hvector(int count, int blocklen, int64_t stride, b2c_type* old, b2c_type** out) {
    B2C_TRY *out = wrap(Datatype::hvector(count, blocklen, stride, T(old))); B2C_CATCH
}
B2C_EXPORT int b2c_type_indexed(int n, const int* blocklens, const int* disps, b2c_type* old,
                                b2c_type** out) {
    B2C_TRY
    *out = wrap(Datatype::indexed(std::vector<int>(blocklens, blocklens + n),
                                  std::vector<int>(disps, disps + n), T(old)));
    B2C_CATCH
}
B2C_EXPORT int b2c_type_hindexed(int n, const int* blocklens, const int64_t* disps, b2c_type* old,
                                 b2c_type** out) {
    B2C_TRY
    *out = wrap(Datatype::hindexed(std::vector<int>(blocklens, blocklens + n),
                                   std::vector<int64_t>(disps, disps + n), T(old)));
    B2C_CATCH
}
B2C_EXPORT int b2c_type_struct(int n, const int* blocklens, const int64_t* disps,
                               b2c_type* const* types, b2c_type** out) {
    B2C_TRY
    std::vector<Datatype> ts;
    for (int i = 0; i < n; ++i) ts.push_back(T(types[i]));
    *out = wrap(Datatype::create_struct(std::vector<int>(blocklens, blocklens + n),
                                        std::vector<int64_t>(disps, disps + n), ts));
    B2C_CATCH
}
B2C_EXPORT int b2c_type_subarray(int ndims, const int* sizes, const int* subsizes, const int* starts,
                                 int order, b2c_type* old, b2c_type** out) {
    B2C_TRY
    *out = wrap(Datatype::subarray(std::vector<int>(sizes, sizes + ndims),
                                   std::vector<int>(subsizes, subsizes + ndims),
                                   std::vector<int>(starts, starts + ndims), (Order)order, T(old)));
    B2C_CATCH
}
B2C_EXPORT int b2c_type_resized(b2c_type* old, int64_t lb, int64_t extent, b2c_type** out) {
    B2C_TRY *out = wrap(Datatype::resized(T(old), lb, extent)); B2C_CATCH
}
B2C_EXPORT int b2c_type_commit(b2c_type* t) {
    B2C_TRY B200_REQUIRE(t, ERR_TYPE, "null datatype handle"); t->t.commit(); B2C_CATCH
}
B2C_EXPORT int b2c_type_free(b2c_type* t) {
    B2C_TRY delete t; B2C_CATCH
}
B2C_EXPORT int b2c_type_size(b2c_type* t, int64_t* size) {
    B2C_TRY *size = T(t).size(); B2C_CATCH
}
B2C_EXPORT int b2c_type_extent(b2c_type* t, int64_t* lb, int64_t* extent) {
    B2C_TRY if (lb) *lb = T(t).lb(); if (extent) *extent = T(t).extent(); B2C_CATCH
}
// Flattened runs of `count` elements as (disp, len) pairs; *n is the run count.
B2C_EXPORT int b2c_type_runs(b2c_type* t, int64_t count, int64_t* pairs, int64_t cap, int64_t* n) {
    B2C_TRY
    std::vector<Run> r = T(t).flatten(count);
    *n = (int64_t)r.size();
    for (int64_t i = 0; i < *n && i < cap; ++i) { pairs[2 * i] = r[i].disp; pairs[2 * i + 1] = r[i].len; }
    B2C_CATCH
}
B2C_EXPORT int b2c_type_pack(b2c_type* t, const void* buf, int64_t count, void* packed) {
    B2C_TRY T(t).pack(buf, count, packed); B2C_CATCH
}
B2C_EXPORT int b2c_type_unpack(b2c_type* t, const void* packed, int64_t bytes, void* buf, int64_t count) {
    B2C_TRY T(t).unpack(packed, bytes, buf, count); B2C_CATCH
}
B2C_EXPORT int b2c_type_describe(b2c_type* t, char* buf, int cap) {
    B2C_TRY
    std::string s = T(t).describe();
    std::strncpy(buf, s.c_str(), (size_t)cap - 1);
    buf[cap - 1] = 0;
    B2C_CATCH
}

// ------------------------------------------------------------------ environment
B2C_EXPORT int b2c_init_local(int nranks, size_t ring_bytes, b2c_comm** out) {
    B2C_TRY
    std::vector<Comm> cs = init_local(nranks, ring_bytes ? ring_bytes : (4u << 20));
    for (int r = 0; r < nranks; ++r) out[r] = wrap(cs[(size_t)r]);
    B2C_CATCH
}
B2C_EXPORT int b2c_init_process(const char* job, int rank, int nranks, size_t ring_bytes,
                                double timeout_s, b2c_comm** out) {
    B2C_TRY
    *out = wrap(init_process(job ? job : "", rank, nranks, ring_bytes ? ring_bytes : (8u << 20),
                             timeout_s > 0 ? timeout_s : 60.0));
    B2C_CATCH
}
B2C_EXPORT int b2c_init_env(b2c_comm** out) {
    B2C_TRY *out = wrap(init_from_env()); B2C_CATCH
}
B2C_EXPORT int b2c_comm_finalize(b2c_comm* c) {
    B2C_TRY if (c) { finalize(c->c); delete c; } B2C_CATCH
}
B2C_EXPORT int b2c_comm_free(b2c_comm* c) {
    B2C_TRY delete c; B2C_CATCH
}
B2C_EXPORT int b2c_comm_rank(b2c_comm* c, int* rank) { B2C_TRY *rank = Cm(c).rank(); B2C_CATCH }
B2C_EXPORT int b2c_comm_size(b2c_comm* c, int* size) { B2C_TRY *size = Cm(c).size(); B2C_CATCH }
B2C_EXPORT int b2c_comm_set_timeout(b2c_comm* c, double seconds) {
    B2C_TRY Cm(c).engine().set_timeout(seconds); B2C_CATCH
}
B2C_EXPORT int b2c_abort(b2c_comm* c, int code) {
    comm::abort(c ? c->c : Comm(), code);
}
B2C_EXPORT double b2c_wtime(void) { return wtime(); }
B2C_EXPORT int b2c_processor_name(char* buf, int cap) {
    B2C_TRY
    std::string s = processor_name();
    std::strncpy(buf, s.c_str(), (size_t)cap - 1);
    buf[cap - 1] = 0;
    B2C_CATCH
}

// ------------------------------------------------------------------ point to point
B2C_EXPORT int b2c_send(b2c_comm* c, const void* buf, int64_t count, b2c_type* t, int dest, int tag) {
    B2C_TRY Cm(c).send(buf, count, T(t), dest, tag); B2C_CATCH
}
B2C_EXPORT int b2c_recv(b2c_comm* c, void* buf, int64_t count, b2c_type* t, int src, int tag,
                        b2c_status* st) {
    B2C_TRY
    Status s;
    Request r = Cm(c).irecv(buf, count, T(t), src, tag);
    try { r.wait(&s); } catch (...) { put_status(st, r.state()->status); throw; }
    put_status(st, s);
    B2C_CATCH
}
B2C_EXPORT int b2c_isend(b2c_comm* c, const void* buf, int64_t count, b2c_type* t, int dest, int tag,
                         b2c_request** out) {
    B2C_TRY *out = new b2c_request{Cm(c).isend(buf, count, T(t), dest, tag)}; B2C_CATCH
}
B2C_EXPORT int b2c_irecv(b2c_comm* c, void* buf, int64_t count, b2c_type* t, int src, int tag,
                         b2c_request** out) {
    B2C_TRY *out = new b2c_request{Cm(c).irecv(buf, count, T(t), src, tag)}; B2C_CATCH
}
// Completes and releases the request.
B2C_EXPORT int b2c_wait(b2c_request* r, b2c_status* st) {
    B2C_TRY
    if (!r) return 0;
    std::unique_ptr<b2c_request> own(r);
    Status s;
    try { r->r.wait(&s); } catch (...) { put_status(st, r->r.state()->status); throw; }
    put_status(st, s);
    B2C_CATCH
}
// *flag = 1 and the request is released if it completed.
B2C_EXPORT int b2c_test(b2c_request* r, int* flag, b2c_status* st) {
    B2C_TRY
    if (!r) { *flag = 1; return 0; }
    Status s;
    *flag = r->r.test(&s) ? 1 : 0;
    if (*flag) { put_status(st, s); delete r; }
    B2C_CATCH
}
B2C_EXPORT int b2c_waitall(int n, b2c_request** reqs, b2c_status* sts) {
    B2C_TRY
    int first_err = 0;
    std::string msg;
    for (int i = 0; i < n; ++i) {
        if (!reqs[i]) continue;
        std::unique_ptr<b2c_request> own(reqs[i]);
        reqs[i] = nullptr;
        Status s;
        try { own->r.wait(&s); }
        catch (const Error& e) { if (!first_err) { first_err = e.code(); msg = e.what(); } s = own->r.state()->status; }
        if (sts) put_status(&sts[i], s);
    }
    if (first_err) { set_last_error(msg); return first_err; }
    B2C_CATCH
}
B2C_EXPORT int b2c_probe(b2c_comm* c, int src, int tag, b2c_status* st) {
    B2C_TRY put_status(st, Cm(c).probe(src, tag)); B2C_CATCH
}
B2C_EXPORT int b2c_iprobe(b2c_comm* c, int src, int tag, int* flag, b2c_status* st) {
    B2C_TRY
    Status s;
    *flag = Cm(c).iprobe(src, tag, &s) ? 1 : 0;
    if (*flag) put_status(st, s);
    B2C_CATCH
}
B2C_EXPORT int b2c_get_count(const b2c_status* st, b2c_type* t, int64_t* count) {
    B2C_TRY
    Status s; s.bytes = st->bytes;
    *count = s.count(T(t));
    B2C_CATCH
}
B2C_EXPORT int b2c_sendrecv(b2c_comm* c, const void* sbuf, int64_t scount, b2c_type* st, int dest,
                            int stag, void* rbuf, int64_t rcount, b2c_type* rt, int src, int rtag,
                            b2c_status* status) {
    B2C_TRY
    Status s;
    Cm(c).sendrecv(sbuf, scount, T(st), dest, stag, rbuf, rcount, T(rt), src, rtag, &s);
    put_status(status, s);
    B2C_CATCH
}

// ------------------------------------------------------------------ collectives
B2C_EXPORT int b2c_barrier(b2c_comm* c) { B2C_TRY Cm(c).barrier(); B2C_CATCH }
B2C_EXPORT int b2c_bcast(b2c_comm* c, void* buf, int64_t count, b2c_type* t, int root) {
    B2C_TRY Cm(c).bcast(buf, count, T(t), root); B2C_CATCH
}
B2C_EXPORT int b2c_reduce(b2c_comm* c, const void* sbuf, void* rbuf, int64_t count, b2c_type* t,
                          int op, int root) {
    B2C_TRY Cm(c).reduce(sbuf, rbuf, count, T(t), (Op)op, root); B2C_CATCH
}
B2C_EXPORT int b2c_allreduce(b2c_comm* c, const void* sbuf, void* rbuf, int64_t count, b2c_type* t,
                             int op) {
    B2C_TRY Cm(c).allreduce(sbuf, rbuf, count, T(t), (Op)op); B2C_CATCH
}
B2C_EXPORT int b2c_gather(b2c_comm* c, const void* sbuf, int64_t scount, b2c_type* st, void* rbuf,
                          int64_t rcount, b2c_type* rt, int root) {
    B2C_TRY
    // non-root ranks may pass null receive arguments (mpi6.cpp:96-100)
    Cm(c).gather(sbuf, scount, T(st), rbuf, rcount, rt ? rt->t : T(st), root);
    B2C_CATCH
}
B2C_EXPORT int b2c_scatter(b2c_comm* c, const void* sbuf, int64_t scount, b2c_type* st, void* rbuf,
                           int64_t rcount, b2c_type* rt, int root) {
    B2C_TRY Cm(c).scatter(sbuf, scount, st ? st->t : T(rt), rbuf, rcount, T(rt), root); B2C_CATCH
}
B2C_EXPORT int b2c_allgather(b2c_comm* c, const void* sbuf, int64_t scount, b2c_type* st, void* rbuf,
                             int64_t rcount, b2c_type* rt) {
    B2C_TRY Cm(c).allgather(sbuf, scount, T(st), rbuf, rcount, T(rt)); B2C_CATCH
}

// ------------------------------------------------------------------ groups / topology
B2C_EXPORT int b2c_comm_group(b2c_comm* c, int* world_ranks, int cap, int* n) {
    B2C_TRY
    Group g = Cm(c).group();
    *n = g.size();
    for (int i = 0; i < g.size() && i < cap; ++i) world_ranks[i] = g.world_ranks()[(size_t)i];
    B2C_CATCH
}
// MPI_Comm_group + MPI_Group_incl + MPI_Comm_create in one call: `ranks` are
// ranks of the parent communicator.  *out is NULL on non-members.
B2C_EXPORT int b2c_comm_create(b2c_comm* c, int n, const int* ranks, b2c_comm** out) {
    B2C_TRY
    *out = wrap(Cm(c).create(Cm(c).group().incl(std::vector<int>(ranks, ranks + n))));
    B2C_CATCH
}
B2C_EXPORT int b2c_comm_split(b2c_comm* c, int color, int key, b2c_comm** out) {
    B2C_TRY *out = wrap(Cm(c).split(color, key)); B2C_CATCH
}
B2C_EXPORT int b2c_comm_dup(b2c_comm* c, b2c_comm** out) {
    B2C_TRY *out = wrap(Cm(c).dup()); B2C_CATCH
}
B2C_EXPORT int b2c_cart_create(b2c_comm* c, int ndims, const int* dims, const int* periods,
                               int reorder, b2c_comm** out) {
    B2C_TRY
    *out = wrap(Cm(c).cart_create(std::vector<int>(dims, dims + ndims),
                                  std::vector<int>(periods, periods + ndims), reorder != 0));
    B2C_CATCH
}
B2C_EXPORT int b2c_cart_coords(b2c_comm* c, int rank, int maxdims, int* coords) {
    B2C_TRY
    std::vector<int> v = Cm(c).cart_coords(rank);
    for (int i = 0; i < (int)v.size() && i < maxdims; ++i) coords[i] = v[(size_t)i];
    B2C_CATCH
}
B2C_EXPORT int b2c_cart_rank(b2c_comm* c, int ndims, const int* coords, int* rank) {
    B2C_TRY *rank = Cm(c).cart_rank(std::vector<int>(coords, coords + ndims)); B2C_CATCH
}
B2C_EXPORT int b2c_cart_shift(b2c_comm* c, int dim, int disp, int* src, int* dst) {
    B2C_TRY Cm(c).cart_shift(dim, disp, src, dst); B2C_CATCH
}
B2C_EXPORT int b2c_cart_get(b2c_comm* c, int maxdims, int* ndims, int* dims, int* periods) {
    B2C_TRY
    std::vector<int> d = Cm(c).cart_dims(), p = Cm(c).cart_periods();
    *ndims = (int)d.size();
    for (int i = 0; i < *ndims && i < maxdims; ++i) { dims[i] = d[(size_t)i]; periods[i] = p[(size_t)i]; }
    B2C_CATCH
}
B2C_EXPORT int b2c_dims_create(int nranks, int ndims, int* dims) {
    B2C_TRY
    std::vector<int> d = dims_create(nranks, ndims);
    for (int i = 0; i < ndims; ++i) dims[i] = d[(size_t)i];
    B2C_CATCH
}

}  // extern "C"
