// Host transport implementation -- see fabric.hpp.
#include "comm/fabric.hpp"

#include <fcntl.h>
#include <sched.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <chrono>
#include <cstdlib>
#include <cstring>
#include <new>
#include <thread>

#include "b200comm/error.hpp"

namespace b200 {
namespace comm {

// header | nranks x (Inbox + ring) | 4 KiB slack | nranks x DevInfo (zero = not published)
static size_t rings_end(int nranks, size_t ring_bytes) {
    return sizeof(FabricHeader) + (size_t)nranks * (sizeof(Inbox) + ring_bytes) + 4096;
}
static size_t region_size(int nranks, size_t ring_bytes) {
    return rings_end(nranks, ring_bytes) + (size_t)nranks * sizeof(DevInfo) + 64;
}

static size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

static void init_region(char* base, int nranks, size_t ring_bytes, size_t bytes) {
    auto* h = new (base) FabricHeader();
    h->nranks = (uint32_t)nranks;
    h->ring_bytes = ring_bytes;
    h->region_bytes = bytes;
    h->attached.store(0);
    h->detached.store(0);
    h->aborted.store(0);
    h->next_context.store(16);
    h->bar_count.store(0);
    h->bar_sense.store(0);
    char* p = base + align_up(sizeof(FabricHeader), 64);
    for (int r = 0; r < nranks; ++r) {
        auto* ib = new (p) Inbox();
        ib->lock.store(0);
        ib->head.store(0);
        ib->tail.store(0);
        p += sizeof(Inbox) + ring_bytes;
    }
    // device-mailbox table: all zero = nobody has published yet
    char* tab = base + align_up(rings_end(nranks, ring_bytes), 64);
    std::memset(tab, 0, (size_t)(base + bytes - tab));
    h->magic.store(kFabricMagic, std::memory_order_release);
}

Inbox* Fabric::inbox(int r) const {
    char* p = base_ + align_up(sizeof(FabricHeader), 64) + (size_t)r * (sizeof(Inbox) + hdr_->ring_bytes);
    return reinterpret_cast<Inbox*>(p);
}
char* Fabric::ring(int r) const { return reinterpret_cast<char*>(inbox(r)) + sizeof(Inbox); }
DevInfo* Fabric::dev_info(int r) const {
    char* p = base_ + align_up(rings_end((int)hdr_->nranks, hdr_->ring_bytes), 64);
    return reinterpret_cast<DevInfo*>(p) + r;
}

Fabric* Fabric::create_local(int nranks, size_t ring_bytes) {
    B200_REQUIRE(nranks >= 1 && nranks <= kMaxRanks, ERR_ARG, "rank count out of range");
    ring_bytes = align_up(ring_bytes < 65536 ? 65536 : ring_bytes, 4096);
    auto* f = new Fabric();
    f->bytes_ = region_size(nranks, ring_bytes);
    void* mem = nullptr;
    if (posix_memalign(&mem, 4096, f->bytes_) != 0) {
        delete f;
        B200_FAIL(ERR_NOMEM, "cannot allocate the local fabric");
    }
    f->base_ = static_cast<char*>(mem);
    init_region(f->base_, nranks, ring_bytes, f->bytes_);
    f->hdr_ = reinterpret_cast<FabricHeader*>(f->base_);
    f->shm_ = false;
    f->owner_ = true;
    return f;
}

Fabric* Fabric::attach_shm(const std::string& name, int rank, int nranks, size_t ring_bytes,
                           double timeout_s) {
    B200_REQUIRE(nranks >= 1 && nranks <= kMaxRanks, ERR_ARG, "rank count out of range");
    B200_REQUIRE(rank >= 0 && rank < nranks, ERR_RANK, "rank out of range");
    ring_bytes = align_up(ring_bytes < 65536 ? 65536 : ring_bytes, 4096);
    const size_t bytes = region_size(nranks, ring_bytes);
    auto* f = new Fabric();
    f->name_ = name;
    f->bytes_ = bytes;
    f->shm_ = true;
    int fd = -1;
    const auto t0 = std::chrono::steady_clock::now();
    if (rank == 0) {
        shm_unlink(name.c_str());  // stale segment from a crashed job
        fd = shm_open(name.c_str(), O_CREAT | O_EXCL | O_RDWR, 0600);
        if (fd < 0 || ftruncate(fd, (off_t)bytes) != 0) {
            if (fd >= 0) close(fd);
            delete f;
            B200_FAIL(ERR_NOMEM, "cannot create shared-memory fabric " + name);
        }
        f->owner_ = true;
    } else {
        for (;;) {
            fd = shm_open(name.c_str(), O_RDWR, 0600);
            if (fd >= 0) {
                struct stat sb;
                if (fstat(fd, &sb) == 0 && (size_t)sb.st_size >= bytes) break;
                close(fd);
                fd = -1;
            }
            const double el = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
            if (el > timeout_s) {
                delete f;
                B200_FAIL(ERR_TIMEOUT, "timed out attaching to shared-memory fabric " + name);
            }
            std::this_thread::sleep_for(std::chrono::milliseconds(2));
        }
    }
    void* mem = mmap(nullptr, bytes, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
    close(fd);
    if (mem == MAP_FAILED) {
        delete f;
        B200_FAIL(ERR_NOMEM, "mmap of the fabric failed");
    }
    f->base_ = static_cast<char*>(mem);
    f->hdr_ = reinterpret_cast<FabricHeader*>(f->base_);
    if (rank == 0) {
        init_region(f->base_, nranks, ring_bytes, bytes);
    } else {
        while (f->hdr_->magic.load(std::memory_order_acquire) != kFabricMagic) {
            const double el = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
            if (el > timeout_s) {
                munmap(mem, bytes);
                f->base_ = nullptr;
                delete f;
                B200_FAIL(ERR_TIMEOUT, "fabric " + name + " never became ready");
            }
            std::this_thread::sleep_for(std::chrono::milliseconds(1));
        }
    }
    const uint32_t n = f->hdr_->attached.fetch_add(1) + 1;
    // once everybody is mapped the name can go: the memory lives until the last unmap
    if (n == (uint32_t)nranks) shm_unlink(name.c_str());
    return f;
}

Fabric::~Fabric() {
    if (!base_) return;
    if (shm_) {
        hdr_->detached.fetch_add(1);
        munmap(base_, bytes_);
        if (owner_) shm_unlink(name_.c_str());
    } else {
        std::free(base_);
    }
}

void Fabric::startup_barrier(int rank) {
    (void)rank;
    const uint32_t n = hdr_->nranks;
    const uint32_t sense = hdr_->bar_sense.load(std::memory_order_acquire);
    if (hdr_->bar_count.fetch_add(1, std::memory_order_acq_rel) + 1 == n) {
        hdr_->bar_count.store(0, std::memory_order_relaxed);
        hdr_->bar_sense.store(sense + 1, std::memory_order_release);
    } else {
        int spins = 0;
        while (hdr_->bar_sense.load(std::memory_order_acquire) == sense) {
            if (hdr_->aborted.load(std::memory_order_relaxed))
                B200_FAIL(ERR_PEER, "a peer aborted during start-up");
            if (++spins > 64) { sched_yield(); }
        }
    }
}

bool Fabric::try_push(int dest, RecordHeader h, const void* payload, size_t payload_bytes) {
    Inbox* ib = inbox(dest);
    char* rg = ring(dest);
    const uint64_t cap = hdr_->ring_bytes;
    const uint64_t rec = align_up(sizeof(RecordHeader) + payload_bytes, 16);
    if (rec > cap / 2) B200_FAIL(ERR_INTERN, "record larger than half the ring");
    // spin lock among producers (short critical section: one memcpy)
    uint32_t exp = 0;
    int spins = 0;
    while (!ib->lock.compare_exchange_weak(exp, 1, std::memory_order_acquire)) {
        exp = 0;
        if (++spins > 128) { sched_yield(); spins = 0; }
    }
    const uint64_t head = ib->head.load(std::memory_order_acquire);
    uint64_t tail = ib->tail.load(std::memory_order_relaxed);
    uint64_t pos = tail % cap;
    uint64_t need = rec;
    uint64_t padlen = 0;
    if (pos + rec > cap) {  // does not fit before the end: pad and wrap
        padlen = cap - pos;
        need += padlen;
    }
    if (cap - (tail - head) < need) {
        ib->lock.store(0, std::memory_order_release);
        return false;
    }
    if (padlen) {
        if (padlen >= sizeof(RecordHeader)) {
            RecordHeader ph{};
            ph.kind = REC_PAD;
            ph.rec_bytes = (uint32_t)padlen;
            std::memcpy(rg + pos, &ph, sizeof(ph));
        } else {
            // cannot even hold a header: consumer recognises "less than a header left"
        }
        tail += padlen;
        pos = 0;
    }
    h.rec_bytes = (uint32_t)rec;
    h.chunk_bytes = (uint32_t)payload_bytes;
    std::memcpy(rg + pos, &h, sizeof(h));
    if (payload_bytes) std::memcpy(rg + pos + sizeof(RecordHeader), payload, payload_bytes);
    ib->tail.store(tail + rec, std::memory_order_release);
    ib->lock.store(0, std::memory_order_release);
    return true;
}

bool Fabric::peek(int rank, RecordHeader* h, const char** payload) {
    Inbox* ib = inbox(rank);
    char* rg = ring(rank);
    const uint64_t cap = hdr_->ring_bytes;
    for (;;) {
        uint64_t head = ib->head.load(std::memory_order_relaxed);
        const uint64_t tail = ib->tail.load(std::memory_order_acquire);
        if (head == tail) return false;
        const uint64_t pos = head % cap;
        if (cap - pos < sizeof(RecordHeader)) {  // implicit pad
            ib->head.store(head + (cap - pos), std::memory_order_release);
            continue;
        }
        std::memcpy(h, rg + pos, sizeof(RecordHeader));
        if (h->kind == REC_PAD) {
            ib->head.store(head + h->rec_bytes, std::memory_order_release);
            continue;
        }
        *payload = rg + pos + sizeof(RecordHeader);
        return true;
    }
}

void Fabric::release(int rank, const RecordHeader& h) {
    Inbox* ib = inbox(rank);
    ib->head.store(ib->head.load(std::memory_order_relaxed) + h.rec_bytes, std::memory_order_release);
}

}  // namespace comm
}  // namespace b200
