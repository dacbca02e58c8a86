// Progress engine implementation -- see engine.hpp.
#include "comm/engine.hpp"

#include <sched.h>

#include <algorithm>
#include <chrono>
#include <cstring>
#include <thread>

namespace b200 {
namespace comm {

DeviceHooks& device_hooks() {
    static DeviceHooks h;
    return h;
}

static double now_s() {
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

int64_t Status::count(const Datatype& t) const {
    const int64_t s = t.size();
    if (s == 0) return bytes == 0 ? 0 : UNDEFINED;
    return bytes % s == 0 ? bytes / s : UNDEFINED;
}

Engine::Engine(std::shared_ptr<Fabric> f, int world_rank)
    : fabric_(std::move(f)), rank_(world_rank), sendq_((size_t)fabric_->nranks()) {}

Engine::~Engine() {
    DeviceHooks& dh = device_hooks();
    if (dev_state_ && dh.eager_destroy) dh.eager_destroy(dev_state_);
    dev_state_ = nullptr;
}

uint32_t Engine::alloc_contexts(uint32_t n) {
    return fabric_->header()->next_context.fetch_add(2 * n) ;
}

bool Engine::matches(const ReqState& r, const RecordHeader& h) {
    return r.context == h.context && (r.src_world == ANY_SOURCE || r.src_world == h.src) &&
           (r.tag == ANY_TAG || r.tag == h.tag);
}

std::shared_ptr<ReqState> Engine::isend(const void* buf, int64_t count, const Datatype& t,
                                        int dest_world, int tag, uint32_t context) {
    B200_REQUIRE(t.valid() && t.committed(), ERR_TYPE, "send datatype is not committed");
    B200_REQUIRE(count >= 0, ERR_COUNT, "negative send count");
    auto r = std::make_shared<ReqState>();
    r->kind = ReqState::SEND;
    r->eng = this;
    if (dest_world == PROC_NULL) {
        r->done.store(true);
        return r;
    }
    B200_REQUIRE(dest_world >= 0 && dest_world < size(), ERR_RANK, "destination rank out of range");
    B200_REQUIRE(tag >= 0 && tag <= TAG_UB, ERR_TAG, "tag out of range");
    B200_REQUIRE(buf != nullptr || count == 0 || t.size() == 0, ERR_BUFFER, "null send buffer");
    std::lock_guard<std::recursive_mutex> lk(mu_);
    r->dest_world = dest_world;
    r->hdr.kind = REC_EAGER;
    r->hdr.src = rank_;
    r->hdr.context = context;
    r->hdr.tag = tag;
    r->hdr.msg_id = next_msg_id_++;
    r->hdr.sig_hash = t.signature_hash();
    r->hdr.flags = 0;
    const int64_t bytes = t.size() * count;
    DeviceHooks& dh = device_hooks();
    int dev = -1;
    uint64_t base = 0, alloc = 0;
    DevEagerMsg em{};
    bool gated = false;
    const bool is_dev = bytes > 0 && dh.classify && dh.classify(buf, &dev, &base, &alloc);
    if (is_dev && dh.eager_send &&
        dh.eager_send(&dev_state_, fabric_.get(), rank_, dest_world, buf, dev, (uint64_t)bytes, t, count,
                      &em, &gated)) {
        // small device message: a put kernel is already copying (gathering) it into a slot of the
        // receiver's device mailbox; the record only says where (no IPC handle, no run list,
        // no host synchronisation on either side)
        r->owned.resize(sizeof(em));
        std::memcpy(r->owned.data(), &em, sizeof(em));
        r->data = r->owned.data();
        r->total = r->owned.size();
        r->hdr.kind = REC_DEV_EAGER;
        r->dev_eager = true;
        r->dev_seq = em.seq;
        r->dev_gated = gated;
        dev_sends_.push_back(r);
    } else if (is_dev) {
        // device buffer: ship a descriptor, the receiver pulls with a kernel
        std::vector<Run> runs = t.flatten(count);
        DevDesc d{};
        d.device = dev;
        d.base_ptr = base;
        d.offset = (uint64_t)reinterpret_cast<uintptr_t>(buf) - base;
        d.alloc_bytes = alloc;
        d.count = count;
        d.nruns = (int64_t)runs.size();
        d.total_bytes = bytes;
        dh.export_handle(&d);
        r->owned.resize(sizeof(DevDesc) + runs.size() * sizeof(Run));
        std::memcpy(r->owned.data(), &d, sizeof(d));
        std::memcpy(r->owned.data() + sizeof(d), runs.data(), runs.size() * sizeof(Run));
        r->data = r->owned.data();
        r->total = r->owned.size();
        r->hdr.kind = REC_DEV_RTS;
        r->wait_fin = true;
        awaiting_fin_[r->hdr.msg_id] = r;
    } else if (t.contiguous_layout() || bytes == 0) {
        r->data = static_cast<const char*>(buf);
        r->total = (uint64_t)bytes;
    } else {
        r->owned.resize((size_t)bytes);
        t.pack(buf, count, r->owned.data());
        r->data = r->owned.data();
        r->total = (uint64_t)bytes;
    }
    r->hdr.total_bytes = r->total;
    sendq_[(size_t)dest_world].push_back(r);
    progress();
    return r;
}

std::shared_ptr<ReqState> Engine::irecv(void* buf, int64_t count, const Datatype& t,
                                        int src_world, int tag, uint32_t context,
                                        std::shared_ptr<const std::vector<int>> inv) {
    B200_REQUIRE(t.valid() && t.committed(), ERR_TYPE, "receive datatype is not committed");
    B200_REQUIRE(count >= 0, ERR_COUNT, "negative receive count");
    auto r = std::make_shared<ReqState>();
    r->kind = ReqState::RECV;
    r->eng = this;
    r->inv = std::move(inv);
    if (src_world == PROC_NULL) {
        r->status.source = PROC_NULL;
        r->status.tag = ANY_TAG;
        r->status.bytes = 0;
        r->done.store(true);
        return r;
    }
    B200_REQUIRE(src_world == ANY_SOURCE || (src_world >= 0 && src_world < size()), ERR_RANK,
                 "source rank out of range");
    B200_REQUIRE(tag == ANY_TAG || (tag >= 0 && tag <= TAG_UB), ERR_TAG, "tag out of range");
    B200_REQUIRE(buf != nullptr || count == 0 || t.size() == 0, ERR_BUFFER, "null receive buffer");
    r->buf = buf;
    r->count = count;
    r->type = t;
    r->context = context;
    r->src_world = src_world;
    r->tag = tag;
    DeviceHooks& dh = device_hooks();
    int dev = -1;
    uint64_t b, n;
    r->buf_is_device = buf && dh.classify && dh.classify(buf, &dev, &b, &n);
    r->buf_device = r->buf_is_device ? dev : -1;
    std::lock_guard<std::recursive_mutex> lk(mu_);
    if (r->buf_is_device && dh.eager_prepare) dh.eager_prepare(&dev_state_, fabric_.get(), rank_, dev);
    for (auto it = unexpected_.begin(); it != unexpected_.end(); ++it) {
        if (matches(*r, (*it)->first)) {
            std::shared_ptr<Incoming> in = *it;
            unexpected_.erase(it);
            in->req = r;
            r->in = in;
            if (in->complete) finish_incoming(in);
            return r;
        }
    }
    posted_.push_back(r);
    progress();
    return r;
}

void Engine::attach(const std::shared_ptr<ReqState>& r, const std::shared_ptr<Incoming>& in) {
    in->req = r;
    r->in = in;
    const uint64_t cap = (uint64_t)(r->type.size() * r->count);
    in->in_place = !in->device && !in->eager_dev && !r->buf_is_device && r->type.contiguous_layout() &&
                   in->first.total_bytes <= cap;
    if (!in->in_place) in->data.resize((size_t)in->first.total_bytes);
}

void Engine::deliver_first(const RecordHeader& h, const char* payload) {
    auto in = std::make_shared<Incoming>();
    in->first = h;
    in->device = h.kind == REC_DEV_RTS;
    in->eager_dev = h.kind == REC_DEV_EAGER;
    bool matched = false;
    for (auto it = posted_.begin(); it != posted_.end(); ++it) {
        if (matches(**it, h)) {
            std::shared_ptr<ReqState> r = *it;
            posted_.erase(it);
            attach(r, in);
            matched = true;
            break;
        }
    }
    if (!matched) {
        in->data.resize((size_t)h.total_bytes);
        unexpected_.push_back(in);
    }
    if (h.chunk_bytes < h.total_bytes) inflight_[{h.src, h.msg_id}] = in;
    deliver_chunk(*in, h, payload);
    if (in->complete && in->req) finish_incoming(in);
}

void Engine::deliver_chunk(Incoming& in, const RecordHeader& h, const char* payload) {
    if (h.chunk_bytes) {
        char* dst = in.in_place ? static_cast<char*>(in.req->buf) : in.data.data();
        std::memcpy(dst + h.offset, payload, h.chunk_bytes);
    }
    in.received += h.chunk_bytes;
    if (in.received >= in.first.total_bytes) in.complete = true;
}

void Engine::finish_incoming(const std::shared_ptr<Incoming>& in) {
    std::shared_ptr<ReqState> r = in->req;
    const RecordHeader& h = in->first;
    const int64_t cap = r->type.size() * r->count;
    r->status.tag = h.tag;
    r->status.source = (r->inv && h.src >= 0 && h.src < (int)r->inv->size()) ? (*r->inv)[(size_t)h.src]
                                                                              : h.src;
    DeviceHooks& dh = device_hooks();
    if (in->eager_dev) {
        // the payload is (or will shortly be) in my device mailbox: start the copy-out and let
        // progress_device() complete the request -- nothing here waits for the GPU
        DevEagerMsg m;
        std::memcpy(&m, in->data.data(), sizeof(m));
        int64_t moved = 0;
        r->status.error = (int64_t)m.bytes > cap ? ERR_TRUNCATE : 0;
        try {
            B200_REQUIRE(dh.eager_recv_start != nullptr, ERR_STATE, "device message but no device layer");
            const uint64_t token = dh.eager_recv_start(&dev_state_, fabric_.get(), rank_, h.src, m, r->buf,
                                                       r->count, r->type, r->buf_device, &moved);
            r->status.bytes = moved;
            dev_recvs_.push_back(DevRecv{in, token, m.slot});
        } catch (const Error& e) {
            r->status.error = e.code();
            r->status.bytes = 0;
            r->in.reset();
            in->req.reset();
            r->done.store(true, std::memory_order_release);
        }
        inflight_.erase({h.src, h.msg_id});
        return;
    }
    if (in->device) {
        DevDesc d;
        std::memcpy(&d, in->data.data(), sizeof(d));
        const Run* runs = reinterpret_cast<const Run*>(in->data.data() + sizeof(DevDesc));
        int64_t moved = 0;
        int err = 0;
        std::string what;
        if (d.total_bytes > cap) err = ERR_TRUNCATE;
        try {
            B200_REQUIRE(dh.pull != nullptr, ERR_STATE, "device message but no device layer");
            dh.pull(d, runs, r->buf, r->count, r->type, r->buf_is_device, &moved);
        } catch (const Error& e) {
            err = e.code();
            what = e.what();
        }
        r->status.bytes = moved;
        r->status.error = err;
        RecordHeader fin{};
        fin.kind = err && err != ERR_TRUNCATE ? REC_DEV_ERR : REC_DEV_FIN;
        fin.src = rank_;
        fin.msg_id = h.msg_id;
        fin.tag = err;
        // Never wait here: this function runs from inside drain_inbox() with the current record
        // peeked but not yet released, so draining again would deliver that record twice.  The
        // control record goes out now if the sender's ring has room, else with the next
        // progress() of this rank (every wait, test, send, receive, barrier and finalize).
        ctrlq_.emplace_back(h.src, fin);
        flush_ctrl();
    } else {
        int64_t n = (int64_t)h.total_bytes;
        if (n > cap) {
            r->status.error = ERR_TRUNCATE;
            n = cap;
        }
        // MPI leaves signature mismatches undefined; we move the bytes and only
        // diagnose truncation.  (sig_hash travels in the header for debugging.)
        if (!in->in_place && n > 0) {
            if (r->buf_is_device) {
                if (dh.unpack_to_device) dh.unpack_to_device(in->data.data(), n, r->buf, r->count, r->type);
                else r->status.error = ERR_STATE;
            } else {
                r->type.unpack(in->data.data(), n, r->buf, r->count);
            }
        }
        r->status.bytes = n;
    }
    inflight_.erase({h.src, h.msg_id});
    r->in.reset();
    in->req.reset();
    in->data.clear();
    in->data.shrink_to_fit();
    r->done.store(true, std::memory_order_release);
}

bool Engine::drain_inbox() {
    bool any = false;
    RecordHeader h;
    const char* payload = nullptr;
    while (fabric_->peek(rank_, &h, &payload)) {
        any = true;
        if (h.kind == REC_DEV_ACK) {
            DeviceHooks& dh = device_hooks();
            if (dh.eager_ack && dev_state_) dh.eager_ack(dev_state_, h.src, (uint32_t)h.tag);
        } else if (h.kind == REC_DEV_FIN || h.kind == REC_DEV_ERR) {
            auto it = awaiting_fin_.find(h.msg_id);
            if (it != awaiting_fin_.end()) {
                std::shared_ptr<ReqState> r = it->second;
                awaiting_fin_.erase(it);
                if (h.kind == REC_DEV_ERR) r->status.error = h.tag ? h.tag : ERR_PEER;
                r->done.store(true, std::memory_order_release);
            }
        } else if (h.offset == 0) {
            deliver_first(h, payload);
        } else {
            auto it = inflight_.find({h.src, h.msg_id});
            if (it != inflight_.end()) {
                std::shared_ptr<Incoming> in = it->second;
                deliver_chunk(*in, h, payload);
                if (in->complete) {
                    if (in->req) finish_incoming(in);
                    else inflight_.erase(it);
                }
            }
        }
        fabric_->release(rank_, h);
    }
    return any;
}

bool Engine::flush_ctrl() {
    bool any = false;
    while (!ctrlq_.empty()) {
        if (!fabric_->try_push(ctrlq_.front().first, ctrlq_.front().second, nullptr, 0)) break;
        ctrlq_.pop_front();
        any = true;
    }
    return any;
}

bool Engine::push_sends() {
    bool any = flush_ctrl();
    const uint64_t max_chunk = fabric_->max_chunk() - sizeof(RecordHeader);
    for (size_t d = 0; d < sendq_.size(); ++d) {
        auto& q = sendq_[d];
        while (!q.empty()) {
            std::shared_ptr<ReqState>& r = q.front();
            bool blocked = false;
            if (r->dev_gated && r->sent == 0) {
                // same physical GPU on both ends: announce the message only when its put kernel
                // has finished (later messages to this rank queue up behind it: order is kept)
                DeviceHooks& dh = device_hooks();
                if (!dh.eager_sent(dev_state_, r->dest_world, r->dev_seq)) break;
            }
            do {
                const uint64_t n = std::min<uint64_t>(max_chunk, r->total - r->sent);
                RecordHeader h = r->hdr;
                h.offset = r->sent;
                if (!fabric_->try_push((int)d, h, r->data + r->sent, (size_t)n)) {
                    blocked = true;
                    break;
                }
                r->sent += n;
                any = true;
            } while (r->sent < r->total);
            if (blocked) break;
            if (r->dev_eager) {
                r->hdr_pushed = true;          // completes in progress_device()
            } else if (!r->wait_fin) {
                r->owned.clear();
                r->owned.shrink_to_fit();
                r->done.store(true, std::memory_order_release);
            }
            q.pop_front();
        }
    }
    return any;
}

bool Engine::progress_device() {
    if (dev_sends_.empty() && dev_recvs_.empty()) return false;
    DeviceHooks& dh = device_hooks();
    bool any = false;
    // copy-outs complete in launch order (one stream)
    while (!dev_recvs_.empty() && dh.eager_recv_done(dev_state_, dev_recvs_.front().token)) {
        DevRecv d = dev_recvs_.front();
        dev_recvs_.pop_front();
        std::shared_ptr<ReqState> r = d.in->req;
        RecordHeader ack{};
        ack.kind = REC_DEV_ACK;
        ack.src = rank_;
        ack.tag = (int32_t)d.slot;
        ack.msg_id = d.in->first.msg_id;
        ctrlq_.emplace_back(d.in->first.src, ack);      // the sender may reuse the slot
        if (r) {
            r->in.reset();
            d.in->req.reset();
            r->done.store(true, std::memory_order_release);
        }
        any = true;
    }
    for (auto it = dev_sends_.begin(); it != dev_sends_.end();) {
        ReqState& r = **it;
        if (r.hdr_pushed && dh.eager_sent(dev_state_, r.dest_world, r.dev_seq)) {
            r.owned.clear();
            r.done.store(true, std::memory_order_release);
            it = dev_sends_.erase(it);
            any = true;
        } else {
            ++it;
        }
    }
    if (any) flush_ctrl();
    return any;
}

bool Engine::progress() {
    std::lock_guard<std::recursive_mutex> lk(mu_);
    bool a = push_sends();
    bool b = drain_inbox();
    bool c = b ? push_sends() : false;
    bool d = progress_device();
    if (d && !sendq_.empty()) c = push_sends() || c;   // a gated record may be free to go now
    return a || b || c || d;
}

void Engine::idle_backoff(int& spins, double t0) {
    if (fabric_->header()->aborted.load(std::memory_order_relaxed))
        B200_FAIL(ERR_PEER, "a peer rank aborted");
    ++spins;
    if (spins < 200) return;
    if (spins < 2000) { sched_yield(); return; }
    std::this_thread::sleep_for(std::chrono::microseconds(50));
    if ((spins & 1023) == 0 && timeout_s_ > 0 && now_s() - t0 > timeout_s_)
        B200_FAIL(ERR_TIMEOUT, "communication wait exceeded " + std::to_string(timeout_s_) +
                               " s (deadlock or dead peer?)");
}

void Engine::wait(const std::shared_ptr<ReqState>& r) {
    int spins = 0;
    const double t0 = now_s();
    while (!r->done.load(std::memory_order_acquire)) {
        if (progress()) spins = 0;
        else idle_backoff(spins, t0);
    }
}

bool Engine::test(const std::shared_ptr<ReqState>& r) {
    if (r->done.load(std::memory_order_acquire)) return true;
    progress();
    return r->done.load(std::memory_order_acquire);
}

bool Engine::iprobe(int src_world, int tag, uint32_t context, Status* st,
                    const std::shared_ptr<const std::vector<int>>& inv) {
    if (src_world == PROC_NULL) {
        if (st) { st->source = PROC_NULL; st->tag = ANY_TAG; st->bytes = 0; st->error = 0; }
        return true;
    }
    std::lock_guard<std::recursive_mutex> lk(mu_);
    progress();
    ReqState probe;
    probe.context = context;
    probe.src_world = src_world;
    probe.tag = tag;
    for (auto& in : unexpected_) {
        if (!matches(probe, in->first)) continue;
        int64_t bytes = (int64_t)in->first.total_bytes;
        if (in->device) {
            if (in->received < sizeof(DevDesc)) return false;
            DevDesc d;
            std::memcpy(&d, in->data.data(), sizeof(d));
            bytes = d.total_bytes;
        } else if (in->eager_dev) {
            if (in->received < sizeof(DevEagerMsg)) return false;
            DevEagerMsg m;
            std::memcpy(&m, in->data.data(), sizeof(m));
            bytes = (int64_t)m.bytes;
        }
        if (st) {
            st->source = (inv && in->first.src < (int)inv->size()) ? (*inv)[(size_t)in->first.src]
                                                                   : in->first.src;
            st->tag = in->first.tag;
            st->bytes = bytes;
            st->error = 0;
        }
        return true;
    }
    return false;
}

Status Engine::probe(int src_world, int tag, uint32_t context,
                     const std::shared_ptr<const std::vector<int>>& inv) {
    Status st;
    int spins = 0;
    const double t0 = now_s();
    while (!iprobe(src_world, tag, context, &st, inv)) idle_backoff(spins, t0);
    return st;
}

}  // namespace comm
}  // namespace b200
