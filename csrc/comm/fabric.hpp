// Host transport: one shared region with a byte ring ("inbox") per rank.
//
// The reference has no transport of its own -- it links an external MPI
// (MVAPICH2 / Cray MPI, README:3-8).  This image has no MPI at all, so the
// MPI-like semantics (tags, probe, self-send, derived datatypes, groups) run
// over this transport instead.  The same code serves two deployments:
//   * ranks = threads of one process  (region on the heap)      -> CPU tests
//   * ranks = processes of one node   (region in POSIX shm)     -> torchrun / b200run
// Bulk device data never goes through here: device buffers are peer-mapped
// and moved by kernels; the fabric only carries the rendezvous descriptors.
#pragma once
#include <atomic>
#include <cstddef>
#include <cstdint>
#include <string>

namespace b200 {
namespace comm {

constexpr uint32_t kFabricMagic = 0xB200C0DEu;
constexpr int kMaxRanks = 64;

enum RecordKind : uint32_t {
    REC_PAD = 0,       // filler up to the end of the ring
    REC_EAGER = 1,     // chunk of a host message
    REC_DEV_RTS = 2,   // device send: descriptor of the source buffer
    REC_DEV_FIN = 3,   // device send completed by the receiver (control)
    REC_DEV_ERR = 4,   // receiver could not complete the device send (control)
    REC_DEV_EAGER = 5, // small device send: the payload sits in a slot of the receiver's device mailbox
    REC_DEV_ACK = 6,   // receiver has emptied a mailbox slot (control; slot index in `tag`)
};

struct alignas(16) RecordHeader {
    uint32_t kind;
    uint32_t rec_bytes;     // total bytes of this record incl. header, 16-aligned
    int32_t src;            // world rank of the sender
    uint32_t context;       // communicator context id
    int32_t tag;
    uint32_t flags;
    uint64_t msg_id;        // sender-unique message id
    uint64_t total_bytes;   // payload bytes of the whole message
    uint64_t offset;        // offset of this chunk inside the message
    uint32_t chunk_bytes;
    uint32_t reserved;
    uint64_t sig_hash;      // datatype signature hash (0: unknown)
};
static_assert(sizeof(RecordHeader) == 64, "record header is 64 bytes");

struct alignas(64) Inbox {
    std::atomic<uint32_t> lock;       // producers serialise on this
    char pad0[60];
    std::atomic<uint64_t> head;       // consumer position (bytes, monotonic)
    char pad1[56];
    std::atomic<uint64_t> tail;       // producer position (bytes, monotonic)
    char pad2[56];
};

struct FabricHeader {
    std::atomic<uint32_t> magic;
    uint32_t nranks;
    uint64_t ring_bytes;
    uint64_t region_bytes;
    std::atomic<uint32_t> attached;
    std::atomic<uint32_t> detached;
    std::atomic<uint32_t> aborted;    // a rank called abort(): everybody bails out
    std::atomic<uint32_t> next_context;
    // sense-reversing barrier for set-up / tear-down only
    std::atomic<uint32_t> bar_count;
    std::atomic<uint32_t> bar_sense;
    char pad[64];
};

// What a rank publishes about its DEVICE mailbox (csrc/comm/device_p2p.cu) so that peers can
// map it and put small device messages straight into it.  Lives in the shared region, all zero
// until the rank touches a device buffer for the first time.
struct alignas(64) DevInfo {
    std::atomic<uint32_t> ready;      // 1 once every other field is valid
    int32_t device;                   // owner's CUDA device index
    int64_t pid;                      // owner process
    uint64_t base;                    // owner-side address of the mailbox
    uint64_t bytes;
    unsigned char gpu_uuid[16];       // physical GPU (same-GPU peers must not wait for each other's kernels)
    unsigned char ipc_handle[64];
    char pad[16];
};
static_assert(sizeof(DevInfo) == 128, "DevInfo is two cache lines");

class Fabric {
public:
    // Heap-backed region for `nranks` threads of this process.
    static Fabric* create_local(int nranks, size_t ring_bytes);
    // POSIX shared-memory region `name`; rank 0 creates, the others attach.
    static Fabric* attach_shm(const std::string& name, int rank, int nranks, size_t ring_bytes,
                              double timeout_s);
    ~Fabric();

    int nranks() const { return (int)hdr_->nranks; }
    size_t ring_bytes() const { return hdr_->ring_bytes; }
    size_t max_chunk() const { return hdr_->ring_bytes / 4; }
    bool is_shm() const { return shm_; }
    FabricHeader* header() { return hdr_; }

    // Try to append one record (header + payload) to `dest`'s inbox.  Returns
    // false if the ring is full right now.
    bool try_push(int dest, RecordHeader h, const void* payload, size_t payload_bytes);

    // Consume the next record of `rank`'s inbox, if any: fills *h and returns a
    // pointer to its payload inside the ring (valid until release()).
    bool peek(int rank, RecordHeader* h, const char** payload);
    void release(int rank, const RecordHeader& h);

    void startup_barrier(int rank);
    DevInfo* dev_info(int r) const;   // rank r's entry of the device-mailbox table

private:
    Fabric() = default;
    Inbox* inbox(int r) const;
    char* ring(int r) const;
    FabricHeader* hdr_ = nullptr;
    char* base_ = nullptr;
    size_t bytes_ = 0;
    bool shm_ = false;
    bool owner_ = false;
    std::string name_;
};

}  // namespace comm
}  // namespace b200
