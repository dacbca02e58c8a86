// Host-side call tracing -- see trace.hpp.
#include "core/trace.hpp"

#include <nvtx3/nvToolsExt.h>
#include <unistd.h>

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <mutex>
#include <string>
#include <vector>

#include "b200comm.h"
#include "core/capi_util.hpp"

namespace b200 {
namespace trace {

std::atomic<int> g_enabled{0};

namespace {

struct Event {
    const char* name;
    const char* cat;
    uint64_t t0_ns, dur_ns;
    int64_t bytes;
    int rank, peer, tag;
    uint32_t tid;
};

std::mutex g_mu;
std::vector<Event> g_events;
std::string g_exit_prefix;
thread_local int tl_rank = -1;
thread_local uint32_t tl_tid = 0;
std::atomic<uint32_t> g_next_tid{1};

uint64_t now_ns() {
    return (uint64_t)std::chrono::duration_cast<std::chrono::nanoseconds>(
               std::chrono::steady_clock::now().time_since_epoch())
        .count();
}

void dump_at_exit() {
    if (g_exit_prefix.empty()) return;
    const std::string path = g_exit_prefix + "." + std::to_string((long long)getpid()) + ".json";
    dump(path.c_str());
}

// Reads B200COMM_TRACE / B200COMM_NVTX once, when the library is loaded.
struct EnvInit {
    EnvInit() {
        const char* t = std::getenv("B200COMM_TRACE");
        const char* n = std::getenv("B200COMM_NVTX");
        const bool rec = t && *t, nv = n && *n && *n != '0';
        if (rec) {
            g_exit_prefix = t;
            std::atexit(dump_at_exit);
        }
        if (rec || nv) enable(rec, nv);
    }
} g_env_init;

}  // namespace

void enable(bool record, bool nvtx) {
    g_enabled.store((record ? 1 : 0) | (nvtx ? 2 : 0), std::memory_order_relaxed);
}

void clear() {
    std::lock_guard<std::mutex> lk(g_mu);
    g_events.clear();
}

int64_t count() {
    std::lock_guard<std::mutex> lk(g_mu);
    return (int64_t)g_events.size();
}

void set_thread_rank(int rank) { tl_rank = rank; }

uint64_t begin_span(const char* name) {
    if (g_enabled.load(std::memory_order_relaxed) & 2) nvtxRangePushA(name);
    return now_ns();
}

void end_span(const char* name, const char* cat, uint64_t t0, int rank, int peer, int tag, int64_t bytes) {
    const uint64_t t1 = now_ns();
    const int mode = g_enabled.load(std::memory_order_relaxed);
    if (mode & 2) nvtxRangePop();
    if (!(mode & 1)) return;
    if (rank >= 0) tl_rank = rank;
    if (!tl_tid) tl_tid = g_next_tid.fetch_add(1);
    Event e{name, cat, t0, t1 - t0, bytes, tl_rank, peer, tag, tl_tid};
    std::lock_guard<std::mutex> lk(g_mu);
    g_events.push_back(e);
}

int64_t dump(const char* path) {
    std::vector<Event> ev;
    {
        std::lock_guard<std::mutex> lk(g_mu);
        ev = g_events;
    }
    FILE* f = std::fopen(path, "w");
    if (!f) B200_FAIL(ERR_ARG, std::string("cannot open trace file ") + path);
    std::fprintf(f, "{\"displayTimeUnit\":\"ns\",\"traceEvents\":[\n");
    for (size_t i = 0; i < ev.size(); ++i) {
        const Event& e = ev[i];
        std::fprintf(f, "{\"name\":\"%s\",\"cat\":\"%s\",\"ph\":\"X\",\"ts\":%.3f,\"dur\":%.3f,\"pid\":%d,\"tid\":%u,"
                        "\"args\":{\"peer\":%d,\"tag\":%d,\"bytes\":%lld}}%s\n",
                     e.name, e.cat, e.t0_ns / 1e3, e.dur_ns / 1e3, e.rank, e.tid, e.peer, e.tag,
                     (long long)e.bytes, i + 1 < ev.size() ? "," : "");
    }
    std::fprintf(f, "]}\n");
    std::fclose(f);
    return (int64_t)ev.size();
}

}  // namespace trace
}  // namespace b200

extern "C" {

// mode bit 0: record spans, bit 1: NVTX ranges; 0 switches tracing off.
B2C_EXPORT int b2c_trace_enable(int mode) {
    B2C_TRY b200::trace::enable((mode & 1) != 0, (mode & 2) != 0); B2C_CATCH
}
B2C_EXPORT int b2c_trace_clear(void) { B2C_TRY b200::trace::clear(); B2C_CATCH }
B2C_EXPORT int b2c_trace_count(int64_t* n) {
    B2C_TRY
    B200_REQUIRE(n, b200::ERR_ARG, "null output");
    *n = b200::trace::count();
    B2C_CATCH
}
B2C_EXPORT int b2c_trace_dump(const char* path, int64_t* n) {
    B2C_TRY
    B200_REQUIRE(path, b200::ERR_ARG, "null path");
    const int64_t c = b200::trace::dump(path);
    if (n) *n = c;
    B2C_CATCH
}
// User spans from the binding layer: begin returns the start time stamp to hand back to end.
// Names are interned (copied once), so they need not outlive the call.
static const char* intern_name(const char* name) {
    static std::mutex mu;
    static std::vector<std::string*> names;    // kept for the life of the process
    std::lock_guard<std::mutex> lk(mu);
    for (auto* s : names)
        if (*s == name) return s->c_str();
    names.push_back(new std::string(name));
    return names.back()->c_str();
}
B2C_EXPORT int b2c_trace_begin(const char* name, uint64_t* t0) {
    B2C_TRY
    B200_REQUIRE(name && t0, b200::ERR_ARG, "null argument");
    *t0 = b200::trace::on() ? b200::trace::begin_span(intern_name(name)) : 0;
    B2C_CATCH
}
B2C_EXPORT int b2c_trace_end(const char* name, uint64_t t0, int rank, int64_t bytes) {
    B2C_TRY
    B200_REQUIRE(name, b200::ERR_ARG, "null name");
    if (t0 != 0) b200::trace::end_span(intern_name(name), "user", t0, rank, -3, -1, bytes);
    B2C_CATCH
}
}
