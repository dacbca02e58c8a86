// MPI-like communication API of b200comm (C++ face; the C ABI and the Python
// bindings mirror it).
//
// Capability parity with the MPI surface the reference exercises (SURVEY.md
// 2.3): init/finalize/rank/size/processor name/wtime/abort (mpi1.cpp,
// mpi2.cpp); blocking and non-blocking point-to-point with tags, probe,
// get_count, self-send and PROC_NULL (mpi3-6.cpp, mpi10.cpp); gather / bcast /
// reduce / allreduce / scatter (mpi6.cpp, mpi9.cpp, mpicuda2-4.cu); groups and
// sub-communicators (mpi9.cpp); Cartesian topologies (mpi10.cpp,
// stencil2d/stencil2D.h:232-244); derived datatypes in every p2p call
// (mpi7.cpp, mpi8.cpp, mpi-complex-types.cpp).  Host buffers travel through
// the host transport (csrc/comm/fabric.hpp); device buffers are moved by
// kernels over peer-mapped memory and only their descriptors use the fabric.
#pragma once
#include <atomic>
#include <cstdint>
#include <memory>
#include <string>
#include <vector>

#include "b200comm/datatype.hpp"
#include "b200comm/error.hpp"

#pragma GCC visibility push(default)   // public C++ API: exported from the library
namespace b200 {
namespace comm {

constexpr int ANY_SOURCE = -1;
constexpr int ANY_TAG = -1;
constexpr int PROC_NULL = -2;
constexpr int UNDEFINED = -32766;
constexpr int TAG_UB = (1 << 30) - 1;

enum class Op : int { SUM = 0, PROD = 1, MAX = 2, MIN = 3, LAND = 4, LOR = 5, BAND = 6, BOR = 7 };

struct Status {
    int source = ANY_SOURCE;   // rank in the communicator of the call
    int tag = ANY_TAG;
    int error = 0;
    int64_t bytes = 0;         // payload bytes received
    // MPI_Get_count (mpi3.cpp:28-31): number of whole `t` elements, or UNDEFINED
    int64_t count(const Datatype& t) const;
};

struct ReqState;
class Engine;
class Fabric;

class Request {
public:
    Request() = default;
    explicit Request(std::shared_ptr<ReqState> s) : s_(std::move(s)) {}
    bool valid() const { return (bool)s_; }
    void wait(Status* st = nullptr);           // MPI_Wait
    bool test(Status* st = nullptr);           // MPI_Test
    std::shared_ptr<ReqState>& state() { return s_; }
private:
    std::shared_ptr<ReqState> s_;
};
void waitall(std::vector<Request>& reqs, std::vector<Status>* sts = nullptr);  // MPI_Waitall

class Group {
public:
    Group() = default;
    explicit Group(std::vector<int> world_ranks) : ranks_(std::move(world_ranks)) {}
    int size() const { return (int)ranks_.size(); }
    int rank_of_world(int world_rank) const;          // UNDEFINED if not a member
    Group incl(const std::vector<int>& ranks) const;  // MPI_Group_incl (mpi9.cpp:33-40)
    Group excl(const std::vector<int>& ranks) const;
    const std::vector<int>& world_ranks() const { return ranks_; }
private:
    std::vector<int> ranks_;
};

struct CommImpl;

class Comm {
public:
    Comm() = default;                      // COMM_NULL
    explicit Comm(std::shared_ptr<CommImpl> p) : p_(std::move(p)) {}
    bool is_null() const { return !p_; }
    int rank() const;
    int size() const;
    Engine& engine() const;
    uint32_t context() const;

    // ---- point to point (host or device buffers; count elements of `t`) ----
    void send(const void* buf, int64_t count, const Datatype& t, int dest, int tag) const;
    void recv(void* buf, int64_t count, const Datatype& t, int src, int tag, Status* st = nullptr) const;
    Request isend(const void* buf, int64_t count, const Datatype& t, int dest, int tag) const;
    Request irecv(void* buf, int64_t count, const Datatype& t, int src, int tag) const;
    Status probe(int src, int tag) const;                    // MPI_Probe (mpi3.cpp:28)
    bool iprobe(int src, int tag, Status* st) const;
    void sendrecv(const void* sbuf, int64_t scount, const Datatype& st, int dest, int stag,
                  void* rbuf, int64_t rcount, const Datatype& rt, int src, int rtag,
                  Status* status = nullptr) const;

    // ---- collectives (host buffers) ----
    void barrier() const;
    void bcast(void* buf, int64_t count, const Datatype& t, int root) const;
    void reduce(const void* sbuf, void* rbuf, int64_t count, const Datatype& t, Op op, int root) const;
    void allreduce(const void* sbuf, void* rbuf, int64_t count, const Datatype& t, Op op) const;
    void gather(const void* sbuf, int64_t scount, const Datatype& st, void* rbuf, int64_t rcount,
                const Datatype& rt, int root) const;
    void scatter(const void* sbuf, int64_t scount, const Datatype& st, void* rbuf, int64_t rcount,
                 const Datatype& rt, int root) const;
    void allgather(const void* sbuf, int64_t scount, const Datatype& st, void* rbuf,
                   int64_t rcount, const Datatype& rt) const;

    // ---- groups / sub-communicators (mpi9.cpp:26-54) ----
    Group group() const;                    // MPI_Comm_group
    Comm create(const Group& g) const;      // MPI_Comm_create (collective over *this)
    Comm split(int color, int key) const;   // MPI_Comm_split
    Comm dup() const;
    void free();                            // MPI_Comm_free

    // ---- Cartesian topology (mpi10.cpp:22-54, stencil2D.h:232-244) ----
    Comm cart_create(const std::vector<int>& dims, const std::vector<int>& periods,
                     bool reorder = false) const;
    bool is_cart() const;
    std::vector<int> cart_dims() const;
    std::vector<int> cart_periods() const;
    std::vector<int> cart_coords(int rank) const;
    int cart_rank(const std::vector<int>& coords) const;       // wraps / PROC_NULL
    void cart_shift(int dim, int disp, int* src, int* dst) const;

    int world_rank_of(int rank) const;
private:
    std::shared_ptr<CommImpl> p_;
};

// Balanced factorisation (MPI_Dims_create-like).
std::vector<int> dims_create(int nranks, int ndims);

// ---- environment ------------------------------------------------------------------
// Ranks as threads of this process: returns one world communicator per rank;
// hand comm[r] to thread r.
std::vector<Comm> init_local(int nranks, size_t ring_bytes = 4u << 20);
// Ranks as processes of this node attached to the POSIX shared-memory fabric
// `job` (e.g. "b200comm-<MASTER_PORT>").
Comm init_process(const std::string& job, int rank, int nranks, size_t ring_bytes = 8u << 20,
                  double timeout_s = 60.0);
// Reads RANK / WORLD_SIZE (torchrun), OMPI_COMM_WORLD_*, PMI_*, or B200_* from
// the environment.
Comm init_from_env();
void finalize(Comm& world);
[[noreturn]] void abort(const Comm& world, int code);
double wtime();
std::string processor_name();

}  // namespace comm
}  // namespace b200
#pragma GCC visibility pop
