// mpi.h -- a SHIM, not an MPI implementation.
//
// Implements the MPI surface that ugovaretto-accel/cuda-mpi-scratch exercises (SURVEY.md 2.3,
// ~45 functions) plus the close relatives a program written in that style reaches for next
// (Sendrecv, Test, Waitany, Iprobe, Allgather, Comm_split/dup, Dims_create, Cart_get, hvector,
// resized, Get_address, ...) on top of the b200comm C ABI, so that the reference's
// tutorial sources compile UNCHANGED (`-I include/b200comm/shim`) and run over this
// library's host transport, launched with `python -m cuda_mpi_scratch_b200.launch -n N`.
// Purpose: a conformance suite (tests/test_reference_conformance.py builds the sources
// straight from the mounted reference tree; nothing of the reference is copied here) and
// a "reference source over shim transport" data point.  Numbers obtained this way must
// never be labelled "CUDA-aware MPI".
//
// C++17, header only.  Device pointers work in every point-to-point call (the receiver
// pulls with a kernel), which is what the reference's CUDA drivers rely on.
#ifndef B200COMM_SHIM_MPI_H_
#define B200COMM_SHIM_MPI_H_

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "b200comm.h"

typedef b2c_comm* MPI_Comm;
typedef b2c_type* MPI_Datatype;
typedef b2c_request* MPI_Request;
typedef long long MPI_Aint;
typedef int MPI_Op;
typedef int MPI_Errhandler;
struct MPI_Status {
    int MPI_SOURCE, MPI_TAG, MPI_ERROR;
    long long b2c_bytes;
};
struct b2c_shim_group {
    MPI_Comm parent;
    std::vector<int> ranks;   // ranks of the parent communicator, in group order
};
typedef b2c_shim_group* MPI_Group;

#define MPI_SUCCESS 0
#define MPI_ERR_OTHER 15
#define MPI_MAX_PROCESSOR_NAME 256
#define MPI_MAX_ERROR_STRING 512
#define MPI_ANY_SOURCE B2C_ANY_SOURCE
#define MPI_ANY_TAG B2C_ANY_TAG
#define MPI_PROC_NULL B2C_PROC_NULL
#define MPI_UNDEFINED B2C_UNDEFINED
#define MPI_ORDER_C 0
#define MPI_ORDER_FORTRAN 1
#define MPI_ERRORS_RETURN 1
#define MPI_ERRORS_ARE_FATAL 0
#define MPI_COMM_NULL ((MPI_Comm)0)
#define MPI_GROUP_NULL ((MPI_Group)0)
#define MPI_REQUEST_NULL ((MPI_Request)0)
#define MPI_STATUS_IGNORE ((MPI_Status*)0)
#define MPI_STATUSES_IGNORE ((MPI_Status*)0)
#define MPI_SUM 0
#define MPI_PROD 1
#define MPI_MAX 2
#define MPI_MIN 3
#define MPI_LAND 4
#define MPI_LOR 5
#define MPI_BAND 6
#define MPI_BOR 7
#define MPI_VERSION 3
#define MPI_SUBVERSION 1

namespace b2c_shim {
inline MPI_Comm world = nullptr;
inline bool finalized = false;
inline MPI_Datatype basic(int kind) {
    static MPI_Datatype cache[16] = {};
    if (!cache[kind]) b2c_type_basic(kind, &cache[kind]);
    return cache[kind];
}
inline void to_status(MPI_Status* out, const b2c_status& s) {
    if (!out) return;
    out->MPI_SOURCE = s.source; out->MPI_TAG = s.tag; out->MPI_ERROR = s.error; out->b2c_bytes = s.bytes;
}
}  // namespace b2c_shim

#define MPI_COMM_WORLD (b2c_shim::world)
#define MPI_BYTE (b2c_shim::basic(0))
#define MPI_CHAR (b2c_shim::basic(1))
#define MPI_UNSIGNED_CHAR (b2c_shim::basic(3))
#define MPI_SHORT (b2c_shim::basic(4))
#define MPI_INT (b2c_shim::basic(5))
#define MPI_LONG_LONG (b2c_shim::basic(6))
#define MPI_UNSIGNED_LONG_LONG (b2c_shim::basic(7))
#define MPI_FLOAT (b2c_shim::basic(8))
#define MPI_DOUBLE (b2c_shim::basic(9))
#define MPI_DOUBLE_PRECISION (b2c_shim::basic(9))   /* Fortran handle used from C by the reference */
#define MPI_UNSIGNED (b2c_shim::basic(10))
#define MPI_SIGNED_CHAR (b2c_shim::basic(2))
#define MPI_INT8_T (b2c_shim::basic(2))
#define MPI_UINT8_T (b2c_shim::basic(3))
#define MPI_INT16_T (b2c_shim::basic(4))
#define MPI_INT32_T (b2c_shim::basic(5))
#define MPI_INT64_T (b2c_shim::basic(6))
#define MPI_UINT64_T (b2c_shim::basic(7))
#define MPI_UINT32_T (b2c_shim::basic(10))
#define MPI_LONG (b2c_shim::basic(6))                /* LP64 */
#define MPI_UNSIGNED_LONG (b2c_shim::basic(7))
#define MPI_AINT (b2c_shim::basic(6))

// ---- environment ------------------------------------------------------------------
inline int MPI_Init(int*, char***) { return b2c_init_env(&b2c_shim::world); }
inline int MPI_Finalize() {
    int rc = b2c_comm_finalize(b2c_shim::world);
    b2c_shim::world = nullptr;
    b2c_shim::finalized = true;
    return rc;
}
inline int MPI_Initialized(int* flag) { *flag = b2c_shim::world != nullptr; return MPI_SUCCESS; }
inline int MPI_Finalized(int* flag) { *flag = b2c_shim::finalized; return MPI_SUCCESS; }
inline int MPI_Get_version(int* v, int* sv) { *v = MPI_VERSION; *sv = MPI_SUBVERSION; return MPI_SUCCESS; }
inline int MPI_Abort(MPI_Comm c, int code) { return b2c_abort(c, code); }
inline double MPI_Wtime() { return b2c_wtime(); }
inline double MPI_Wtick() { return 1e-9; }
inline int MPI_Get_address(const void* p, MPI_Aint* a) { *a = (MPI_Aint)(intptr_t)p; return MPI_SUCCESS; }
inline int MPI_Get_processor_name(char* name, int* len) {
    int rc = b2c_processor_name(name, MPI_MAX_PROCESSOR_NAME);
    if (len) *len = (int)std::strlen(name);
    return rc;
}
// Errors are always returned as codes here (the MPI_ERRORS_RETURN behaviour).
inline int MPI_Errhandler_set(MPI_Comm, MPI_Errhandler) { return MPI_SUCCESS; }
inline int MPI_Comm_set_errhandler(MPI_Comm, MPI_Errhandler) { return MPI_SUCCESS; }
inline int MPI_Error_class(int code, int* cls) { *cls = code; return MPI_SUCCESS; }
inline int MPI_Error_string(int code, char* s, int* len) {
    const char* last = b2c_last_error();
    const char* txt = (last && *last) ? last : b2c_error_class_string(code);
    std::snprintf(s, MPI_MAX_ERROR_STRING, "%s", txt);
    if (len) *len = (int)std::strlen(s);
    return MPI_SUCCESS;
}

// ---- communicators, groups --------------------------------------------------------
inline int MPI_Comm_rank(MPI_Comm c, int* r) { return b2c_comm_rank(c, r); }
inline int MPI_Comm_size(MPI_Comm c, int* s) { return b2c_comm_size(c, s); }
inline int MPI_Comm_free(MPI_Comm* c) { int rc = b2c_comm_free(*c); *c = MPI_COMM_NULL; return rc; }
inline int MPI_Comm_group(MPI_Comm c, MPI_Group* g) {
    int n = 0;
    int rc = b2c_comm_size(c, &n);
    if (rc) return rc;
    *g = new b2c_shim_group{c, {}};
    for (int i = 0; i < n; ++i) (*g)->ranks.push_back(i);
    return MPI_SUCCESS;
}
inline int MPI_Group_incl(MPI_Group g, int n, const int* ranks, MPI_Group* out) {
    *out = new b2c_shim_group{g->parent, {}};
    for (int i = 0; i < n; ++i) {
        if (ranks[i] < 0 || ranks[i] >= (int)g->ranks.size()) return 7;
        (*out)->ranks.push_back(g->ranks[(size_t)ranks[i]]);
    }
    return MPI_SUCCESS;
}
inline int MPI_Group_excl(MPI_Group g, int n, const int* ranks, MPI_Group* out) {
    *out = new b2c_shim_group{g->parent, {}};
    for (size_t i = 0; i < g->ranks.size(); ++i) {
        bool drop = false;
        for (int k = 0; k < n; ++k) drop = drop || ranks[k] == (int)i;
        if (!drop) (*out)->ranks.push_back(g->ranks[i]);
    }
    return MPI_SUCCESS;
}
inline int MPI_Group_rank(MPI_Group g, int* rank) {
    int me = 0;
    int rc = b2c_comm_rank(g->parent, &me);
    *rank = MPI_UNDEFINED;
    for (size_t i = 0; i < g->ranks.size(); ++i)
        if (g->ranks[i] == me) *rank = (int)i;
    return rc;
}
inline int MPI_Group_size(MPI_Group g, int* n) { *n = (int)g->ranks.size(); return MPI_SUCCESS; }
inline int MPI_Group_free(MPI_Group* g) { delete *g; *g = MPI_GROUP_NULL; return MPI_SUCCESS; }
inline int MPI_Comm_create(MPI_Comm c, MPI_Group g, MPI_Comm* out) {
    return b2c_comm_create(c, (int)g->ranks.size(), g->ranks.data(), out);
}

inline int MPI_Comm_split(MPI_Comm c, int color, int key, MPI_Comm* out) {
    return b2c_comm_split(c, color, key, out);     // MPI_UNDEFINED == B2C_UNDEFINED: *out = MPI_COMM_NULL
}
inline int MPI_Comm_dup(MPI_Comm c, MPI_Comm* out) { return b2c_comm_dup(c, out); }

// ---- Cartesian topology -------------------------------------------------------------
inline int MPI_Dims_create(int nnodes, int ndims, int* dims) {
    // the C ABI returns a balanced factorisation; honour entries the caller fixed (non-zero)
    int fixed = 1, free_dims = 0;
    for (int i = 0; i < ndims; ++i) {
        if (dims[i] > 0) fixed *= dims[i];
        else ++free_dims;
    }
    if (free_dims == 0) return fixed == nnodes ? MPI_SUCCESS : 13;
    if (nnodes % fixed) return 13;
    std::vector<int> f((size_t)free_dims);
    int rc = b2c_dims_create(nnodes / fixed, free_dims, f.data());
    if (rc) return rc;
    for (int i = 0, k = 0; i < ndims; ++i)
        if (dims[i] <= 0) dims[i] = f[(size_t)k++];
    return MPI_SUCCESS;
}
inline int MPI_Cart_get(MPI_Comm c, int maxdims, int* dims, int* periods, int* coords) {
    int nd = 0, me = 0;
    int rc = b2c_cart_get(c, maxdims, &nd, dims, periods);
    if (rc) return rc;
    rc = b2c_comm_rank(c, &me);
    if (rc) return rc;
    return b2c_cart_coords(c, me, maxdims, coords);
}
inline int MPI_Cartdim_get(MPI_Comm c, int* ndims) {
    int d[8], p[8];
    return b2c_cart_get(c, 8, ndims, d, p);
}
inline int MPI_Cart_create(MPI_Comm c, int nd, const int* dims, const int* periods, int reorder,
                           MPI_Comm* out) {
    return b2c_cart_create(c, nd, dims, periods, reorder, out);
}
inline int MPI_Cart_coords(MPI_Comm c, int rank, int maxdims, int* coords) {
    return b2c_cart_coords(c, rank, maxdims, coords);
}
inline int MPI_Cart_rank(MPI_Comm c, const int* coords, int* rank) {
    int nd = 0, d[8], p[8];
    int rc = b2c_cart_get(c, 8, &nd, d, p);
    if (rc) return rc;
    return b2c_cart_rank(c, nd, coords, rank);
}
// MPI order of the outputs: (source, dest)
inline int MPI_Cart_shift(MPI_Comm c, int dim, int disp, int* src, int* dst) {
    return b2c_cart_shift(c, dim, disp, src, dst);
}

// ---- point to point -------------------------------------------------------------------
inline int MPI_Send(const void* b, int n, MPI_Datatype t, int dest, int tag, MPI_Comm c) {
    return b2c_send(c, b, n, t, dest, tag);
}
inline int MPI_Recv(void* b, int n, MPI_Datatype t, int src, int tag, MPI_Comm c, MPI_Status* st) {
    b2c_status s{};
    int rc = b2c_recv(c, b, n, t, src, tag, &s);
    b2c_shim::to_status(st, s);
    return rc;
}
inline int MPI_Isend(const void* b, int n, MPI_Datatype t, int dest, int tag, MPI_Comm c, MPI_Request* r) {
    return b2c_isend(c, b, n, t, dest, tag, r);
}
inline int MPI_Irecv(void* b, int n, MPI_Datatype t, int src, int tag, MPI_Comm c, MPI_Request* r) {
    return b2c_irecv(c, b, n, t, src, tag, r);
}
inline int MPI_Wait(MPI_Request* r, MPI_Status* st) {
    b2c_status s{};
    int rc = b2c_wait(*r, &s);
    *r = MPI_REQUEST_NULL;
    b2c_shim::to_status(st, s);
    return rc;
}
inline int MPI_Waitall(int n, MPI_Request* reqs, MPI_Status* sts) {
    std::vector<b2c_status> s((size_t)(n > 0 ? n : 1));
    int rc = b2c_waitall(n, reqs, s.data());
    for (int i = 0; i < n; ++i) {
        reqs[i] = MPI_REQUEST_NULL;
        if (sts) b2c_shim::to_status(&sts[i], s[(size_t)i]);
    }
    return rc;
}
inline int MPI_Test(MPI_Request* r, int* flag, MPI_Status* st) {
    if (*r == MPI_REQUEST_NULL) { *flag = 1; return MPI_SUCCESS; }
    b2c_status s{};
    int rc = b2c_test(*r, flag, &s);
    if (*flag) {
        *r = MPI_REQUEST_NULL;
        b2c_shim::to_status(st, s);
    }
    return rc;
}
// Completes exactly one active request (polling, like the progress engine underneath).
inline int MPI_Waitany(int n, MPI_Request* reqs, int* index, MPI_Status* st) {
    bool any = false;
    for (int i = 0; i < n; ++i) any = any || reqs[i] != MPI_REQUEST_NULL;
    *index = MPI_UNDEFINED;
    if (!any) return MPI_SUCCESS;
    for (;;) {
        for (int i = 0; i < n; ++i) {
            if (reqs[i] == MPI_REQUEST_NULL) continue;
            int flag = 0;
            int rc = MPI_Test(&reqs[i], &flag, st);
            if (rc) return rc;
            if (flag) { *index = i; return MPI_SUCCESS; }
        }
    }
}
inline int MPI_Testall(int n, MPI_Request* reqs, int* flag, MPI_Status* sts) {
    // completes the requests that are done; *flag only when ALL are
    *flag = 1;
    for (int i = 0; i < n; ++i) {
        int f = 0;
        int rc = MPI_Test(&reqs[i], &f, sts ? &sts[i] : MPI_STATUS_IGNORE);
        if (rc) return rc;
        if (!f) *flag = 0;
    }
    return MPI_SUCCESS;
}
inline int MPI_Sendrecv(const void* sb, int sn, MPI_Datatype st, int dest, int stag, void* rb, int rn,
                        MPI_Datatype rt, int src, int rtag, MPI_Comm c, MPI_Status* status) {
    b2c_status s{};
    int rc = b2c_sendrecv(c, sb, sn, st, dest, stag, rb, rn, rt, src, rtag, &s);
    b2c_shim::to_status(status, s);
    return rc;
}
inline int MPI_Iprobe(int src, int tag, MPI_Comm c, int* flag, MPI_Status* st) {
    b2c_status s{};
    int rc = b2c_iprobe(c, src, tag, flag, &s);
    if (*flag) b2c_shim::to_status(st, s);
    return rc;
}
inline int MPI_Probe(int src, int tag, MPI_Comm c, MPI_Status* st) {
    b2c_status s{};
    int rc = b2c_probe(c, src, tag, &s);
    b2c_shim::to_status(st, s);
    return rc;
}
inline int MPI_Get_count(const MPI_Status* st, MPI_Datatype t, int* count) {
    b2c_status s{st->MPI_SOURCE, st->MPI_TAG, st->MPI_ERROR, st->b2c_bytes};
    int64_t n = 0;
    int rc = b2c_get_count(&s, t, &n);
    *count = (int)n;
    return rc;
}

// ---- collectives ----------------------------------------------------------------------
inline int MPI_Barrier(MPI_Comm c) { return b2c_barrier(c); }
inline int MPI_Bcast(void* b, int n, MPI_Datatype t, int root, MPI_Comm c) { return b2c_bcast(c, b, n, t, root); }
inline int MPI_Reduce(const void* s, void* r, int n, MPI_Datatype t, MPI_Op op, int root, MPI_Comm c) {
    return b2c_reduce(c, s, r, n, t, op, root);
}
inline int MPI_Allreduce(const void* s, void* r, int n, MPI_Datatype t, MPI_Op op, MPI_Comm c) {
    return b2c_allreduce(c, s, r, n, t, op);
}
inline int MPI_Gather(const void* s, int sn, MPI_Datatype st, void* r, int rn, MPI_Datatype rt, int root,
                      MPI_Comm c) {
    return b2c_gather(c, s, sn, st, r, rn, rt, root);
}
inline int MPI_Scatter(const void* s, int sn, MPI_Datatype st, void* r, int rn, MPI_Datatype rt, int root,
                       MPI_Comm c) {
    return b2c_scatter(c, s, sn, st, r, rn, rt, root);
}

inline int MPI_Allgather(const void* s, int sn, MPI_Datatype st, void* r, int rn, MPI_Datatype rt, MPI_Comm c) {
    return b2c_allgather(c, s, sn, st, r, rn, rt);
}

// ---- datatypes ------------------------------------------------------------------------
inline int MPI_Type_create_hvector(int n, int bl, MPI_Aint stride, MPI_Datatype old, MPI_Datatype* out) {
    return b2c_type_hvector(n, bl, stride, old, out);
}
inline int MPI_Type_create_indexed_block(int n, int bl, const int* disp, MPI_Datatype old, MPI_Datatype* out) {
    std::vector<int> bls((size_t)(n > 0 ? n : 1), bl);
    return b2c_type_indexed(n, bls.data(), disp, old, out);
}
// MPI_Pack / MPI_Unpack on host buffers: the packed stream is the concatenation of the
// typemap's bytes, `*position` advances by the packed size.
inline int MPI_Pack_size(int count, MPI_Datatype t, MPI_Comm, int* size) {
    int64_t s = 0;
    int rc = b2c_type_size(t, &s);
    *size = (int)(s * count);
    return rc;
}
inline int MPI_Pack(const void* inbuf, int count, MPI_Datatype t, void* outbuf, int outsize, int* position,
                    MPI_Comm) {
    int64_t s = 0;
    int rc = b2c_type_size(t, &s);
    if (rc) return rc;
    if (*position + s * count > outsize) return 14;   /* truncation */
    rc = b2c_type_pack(t, inbuf, count, static_cast<char*>(outbuf) + *position);
    if (!rc) *position += (int)(s * count);
    return rc;
}
inline int MPI_Unpack(const void* inbuf, int insize, int* position, void* outbuf, int count, MPI_Datatype t,
                      MPI_Comm) {
    int64_t s = 0;
    int rc = b2c_type_size(t, &s);
    if (rc) return rc;
    if (*position + s * count > insize) return 14;
    rc = b2c_type_unpack(t, static_cast<const char*>(inbuf) + *position, s * count, outbuf, count);
    if (!rc) *position += (int)(s * count);
    return rc;
}
#define MPI_PACKED (b2c_shim::basic(0))
inline int MPI_Type_create_resized(MPI_Datatype old, MPI_Aint lb, MPI_Aint extent, MPI_Datatype* out) {
    return b2c_type_resized(old, lb, extent, out);
}
inline int MPI_Type_indexed(int n, const int* bl, const int* disp, MPI_Datatype old, MPI_Datatype* out) {
    return b2c_type_indexed(n, bl, disp, old, out);
}
inline int MPI_Type_create_hindexed(int n, const int* bl, const MPI_Aint* disp, MPI_Datatype old,
                                    MPI_Datatype* out) {
    std::vector<int64_t> d(disp, disp + n);
    return b2c_type_hindexed(n, bl, d.data(), old, out);
}
inline int MPI_Type_create_struct(int n, const int* bl, const MPI_Aint* disp, const MPI_Datatype* types,
                                  MPI_Datatype* out) {
    std::vector<int64_t> d(disp, disp + n);
    return b2c_type_struct(n, bl, d.data(), const_cast<b2c_type* const*>(types), out);
}
inline int MPI_Type_create_subarray(int nd, const int* sizes, const int* sub, const int* starts, int order,
                                    MPI_Datatype old, MPI_Datatype* out) {
    return b2c_type_subarray(nd, sizes, sub, starts, order, old, out);
}
inline int MPI_Type_contiguous(int n, MPI_Datatype old, MPI_Datatype* out) {
    return b2c_type_contiguous(n, old, out);
}
inline int MPI_Type_vector(int n, int bl, int stride, MPI_Datatype old, MPI_Datatype* out) {
    return b2c_type_vector(n, bl, stride, old, out);
}
inline int MPI_Type_extent(MPI_Datatype t, MPI_Aint* extent) {   // MPI-1 name used by mpi8.cpp
    int64_t lb = 0, ex = 0;
    int rc = b2c_type_extent(t, &lb, &ex);
    *extent = ex;
    return rc;
}
inline int MPI_Type_get_extent(MPI_Datatype t, MPI_Aint* lb, MPI_Aint* extent) {
    int64_t l = 0, e = 0;
    int rc = b2c_type_extent(t, &l, &e);
    *lb = l; *extent = e;
    return rc;
}
inline int MPI_Type_size(MPI_Datatype t, int* size) {
    int64_t s = 0;
    int rc = b2c_type_size(t, &s);
    *size = (int)s;
    return rc;
}
inline int MPI_Type_commit(MPI_Datatype* t) { return b2c_type_commit(*t); }
inline int MPI_Type_free(MPI_Datatype* t) { int rc = b2c_type_free(*t); *t = nullptr; return rc; }

#endif  // B200COMM_SHIM_MPI_H_
