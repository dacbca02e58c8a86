// Per-rank progress engine: tag matching, unexpected-message queue, chunked
// eager sends over the Fabric, rendezvous descriptors for device buffers.
//
// This is the piece an MPI library provides to the reference (matching by
// communicator + source + tag, mpi5.cpp:47-52 comment; probe, mpi3.cpp:28-31;
// self-send, mpi7.cpp:48-54).  Single-threaded progress: every blocking call
// drives progress() until its request completes.
#pragma once
#include <atomic>
#include <deque>
#include <functional>
#include <list>
#include <map>
#include <memory>
#include <mutex>
#include <vector>

#include "b200comm/comm.hpp"
#include "comm/fabric.hpp"

namespace b200 {
namespace comm {

// Descriptor of a device-resident send buffer, carried by a REC_DEV_RTS record.
struct DevDesc {
    unsigned char ipc_handle[64];
    int64_t pid;            // owner process
    int32_t device;         // owner CUDA device
    int32_t same_process;   // filled by the receiver
    uint64_t base_ptr;      // owner-side base address of the allocation
    uint64_t offset;        // buffer address - base_ptr
    uint64_t alloc_bytes;   // size of the allocation
    int64_t count;
    int64_t nruns;          // flattened runs of `count` elements follow the descriptor
    int64_t total_bytes;    // payload bytes
    uint64_t buffer_id;     // CU_POINTER_ATTRIBUTE_BUFFER_ID of the allocation: a cudaFree + cudaMalloc
                            // that returns the same address is a DIFFERENT allocation
};

// Payload of a REC_DEV_EAGER record: where the (contiguous, packed) message sits in the
// receiver's device mailbox.
struct DevEagerMsg {
    uint32_t slot;          // slot of the (sender -> receiver) ring
    uint32_t flags;         // bit0: the put kernel had already finished when the record was pushed
    uint64_t seq;           // the slot's flag reaches this value when the payload has landed
    uint64_t bytes;
};

struct Incoming {
    RecordHeader first{};
    std::vector<char> data;      // assembled payload when not received in place
    uint64_t received = 0;
    std::shared_ptr<ReqState> req;
    bool in_place = false;
    bool device = false;         // REC_DEV_RTS: rendezvous, the receiver pulls
    bool eager_dev = false;      // REC_DEV_EAGER: payload already in my device mailbox
    bool complete = false;
};

struct ReqState {
    enum Kind { SEND, RECV } kind = SEND;
    std::atomic<bool> done{false};
    Status status;
    Engine* eng = nullptr;
    // world -> communicator rank translation for Status.source
    std::shared_ptr<const std::vector<int>> inv;
    // ---- send side ----
    int dest_world = 0;
    RecordHeader hdr{};
    const char* data = nullptr;      // contiguous user buffer or `owned`
    std::vector<char> owned;
    uint64_t total = 0, sent = 0;
    bool wait_fin = false;           // device send: completes on REC_DEV_FIN
    bool dev_eager = false;          // small device send through the receiver's mailbox: completes
                                     // when the put kernel has read the buffer AND the record is out
    bool hdr_pushed = false;
    uint64_t dev_seq = 0;            // sequence number of that put (per destination)
    bool dev_gated = false;          // the record may only leave once the put kernel has finished
    // ---- recv side ----
    void* buf = nullptr;
    int64_t count = 0;
    Datatype type;
    uint32_t context = 0;
    int src_world = ANY_SOURCE;
    int tag = ANY_TAG;
    bool buf_is_device = false;
    int buf_device = -1;             // CUDA device of `buf` (known from the classification at irecv)
    std::shared_ptr<Incoming> in;
};

// Hooks implemented by the device layer (csrc/comm/device_p2p.cpp).  Kept as
// function pointers so the host-only engine has no CUDA dependency.
struct DeviceHooks {
    // true if `p` is device memory; fills base/size of its allocation
    bool (*classify)(const void* p, int* device, uint64_t* base, uint64_t* bytes) = nullptr;
    // export the allocation containing `desc->base_ptr`
    void (*export_handle)(DevDesc* desc) = nullptr;
    // receiver: move the described source into (buf,count,type); blocks until done
    void (*pull)(const DevDesc& desc, const Run* runs, void* buf, int64_t count,
                 const Datatype& type, bool dst_is_device, int64_t* moved) = nullptr;
    // host message -> device buffer
    void (*unpack_to_device)(const void* packed, int64_t bytes, void* buf, int64_t count,
                             const Datatype& type) = nullptr;
    // ---- eager path for small contiguous device messages (kernel-initiated, no host sync) ----
    // `state` is per Engine, created on first use.  All calls come from the engine's owner thread.
    // publish my mailbox (first device send / receive of this rank)
    void (*eager_prepare)(void** state, Fabric* f, int me, int device) = nullptr;
    // sender: put `bytes` at `buf` into a free slot of dest's mailbox.  false -> use the rendezvous.
    // *gated: sender and receiver share one physical GPU, so the receiver must not be told before
    // the put kernel has FINISHED (kernels of two ranks on one GPU must never wait for each other).
    // `type`/`count` describe the layout at `buf`; a derived type is GATHERED by the put (the slot
    // always holds the packed bytes), so strided data crosses NVLink as contiguous bursts.
    bool (*eager_send)(void** state, Fabric* f, int me, int dest, const void* buf, int device,
                       uint64_t bytes, const Datatype& type, int64_t count, DevEagerMsg* m,
                       bool* gated) = nullptr;
    bool (*eager_sent)(void* state, int dest, uint64_t seq) = nullptr;    // put kernel done?
    void (*eager_ack)(void* state, int dest, uint32_t slot) = nullptr;    // receiver freed the slot
    // receiver: start moving message `m` from `src` out of my mailbox into (buf, count, type);
    // returns a token for eager_recv_done.  Never blocks.
    uint64_t (*eager_recv_start)(void** state, Fabric* f, int me, int src, const DevEagerMsg& m,
                                 void* buf, int64_t count, const Datatype& type, int dst_device /* -1: host */,
                                 int64_t* moved) = nullptr;
    bool (*eager_recv_done)(void* state, uint64_t token) = nullptr;
    void (*eager_destroy)(void* state) = nullptr;
};
DeviceHooks& device_hooks();

class Engine {
public:
    Engine(std::shared_ptr<Fabric> f, int world_rank);
    ~Engine();
    int rank() const { return rank_; }
    int size() const { return fabric_->nranks(); }
    Fabric& fabric() { return *fabric_; }
    std::shared_ptr<Fabric> fabric_ptr() { return fabric_; }

    std::shared_ptr<ReqState> isend(const void* buf, int64_t count, const Datatype& t,
                                    int dest_world, int tag, uint32_t context);
    std::shared_ptr<ReqState> irecv(void* buf, int64_t count, const Datatype& t, int src_world,
                                    int tag, uint32_t context,
                                    std::shared_ptr<const std::vector<int>> inv);
    void wait(const std::shared_ptr<ReqState>& r);
    bool test(const std::shared_ptr<ReqState>& r);
    bool iprobe(int src_world, int tag, uint32_t context, Status* st,
                const std::shared_ptr<const std::vector<int>>& inv);
    Status probe(int src_world, int tag, uint32_t context,
                 const std::shared_ptr<const std::vector<int>>& inv);
    bool progress();   // true if anything moved
    void set_timeout(double seconds) { timeout_s_ = seconds; }
    double timeout() const { return timeout_s_; }
    uint32_t alloc_contexts(uint32_t n);   // globally unique context ids

private:
    bool push_sends();
    bool flush_ctrl();   // FIN/ERR records that could not be pushed yet; never blocks, never drains
    bool progress_device();   // completions of eager device sends / receives
    bool drain_inbox();
    void deliver_first(const RecordHeader& h, const char* payload);
    void deliver_chunk(Incoming& in, const RecordHeader& h, const char* payload);
    void finish_incoming(const std::shared_ptr<Incoming>& in);
    void attach(const std::shared_ptr<ReqState>& r, const std::shared_ptr<Incoming>& in);
    static bool matches(const ReqState& r, const RecordHeader& h);
    void idle_backoff(int& spins, double t0);

    std::shared_ptr<Fabric> fabric_;
    int rank_;
    std::recursive_mutex mu_;
    uint64_t next_msg_id_ = 1;
    std::vector<std::deque<std::shared_ptr<ReqState>>> sendq_;
    std::list<std::shared_ptr<ReqState>> posted_;
    std::list<std::shared_ptr<Incoming>> unexpected_;
    std::map<std::pair<int, uint64_t>, std::shared_ptr<Incoming>> inflight_;
    std::map<uint64_t, std::shared_ptr<ReqState>> awaiting_fin_;
    std::deque<std::pair<int, RecordHeader>> ctrlq_;   // control records waiting for ring space
    // eager device path
    void* dev_state_ = nullptr;
    struct DevRecv { std::shared_ptr<Incoming> in; uint64_t token; uint32_t slot; };
    std::deque<std::shared_ptr<ReqState>> dev_sends_;   // put launched, not yet complete
    std::deque<DevRecv> dev_recvs_;                     // copy-out launched, not yet complete
    double timeout_s_ = 120.0;
};

}  // namespace comm
}  // namespace b200
