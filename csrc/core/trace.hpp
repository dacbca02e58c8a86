// Host-side call tracing: every communication call (point-to-point, waits, probes,
// collectives, halo exchanges) and every enqueue of a device operation (stencil steps, dot,
// typed moves, puts) can be recorded as a timed span with rank / peer / tag / bytes and written
// as a Chrome trace-event file (chrome://tracing, Perfetto).  Off by default; when off a span
// costs one relaxed atomic load.
//
// The reference's tracing is MPI_Wtime brackets and clock() deltas gathered by hand
// (test-benchmark/mpi-pingpong-gpu.cpp:51-57, mpicuda3.cu:176-179,315-325) and a `time` wrapper
// in the job script (mpi_cuda_pbs_ref.sh:18); this is the library-level form of it.
//
//   B200COMM_TRACE=/tmp/run   -> enabled at load; written to /tmp/run.<pid>.json at exit
//   B200COMM_NVTX=1           -> every span is also an NVTX range (nsys / ncu timelines)
#pragma once
#include <atomic>
#include <cstdint>

namespace b200 {
namespace trace {

extern std::atomic<int> g_enabled;    // bit 0: record spans, bit 1: NVTX ranges
inline bool on() { return g_enabled.load(std::memory_order_relaxed) != 0; }

void enable(bool record, bool nvtx);
void clear();
// Number of recorded spans.
int64_t count();
// Writes all spans recorded so far as Chrome trace-event JSON; returns the span count.
int64_t dump(const char* path);
// Rank shown as "pid" for spans of this thread (thread ranks share a process).
void set_thread_rank(int rank);

uint64_t begin_span(const char* name);
void end_span(const char* name, const char* cat, uint64_t t0, int rank, int peer, int tag, int64_t bytes);

// RAII span; `name` and `cat` must be string literals (stored by pointer).
class Scope {
public:
    Scope(const char* name, const char* cat, int rank = -1, int peer = -3, int tag = -1, int64_t bytes = -1)
        : name_(name), cat_(cat), rank_(rank), peer_(peer), tag_(tag), bytes_(bytes), live_(on()) {
        if (live_) t0_ = begin_span(name);
    }
    ~Scope() {
        if (live_) end_span(name_, cat_, t0_, rank_, peer_, tag_, bytes_);
    }
    void set_bytes(int64_t b) { bytes_ = b; }
    Scope(const Scope&) = delete;
    Scope& operator=(const Scope&) = delete;

private:
    const char* name_;
    const char* cat_;
    int rank_, peer_, tag_;
    int64_t bytes_;
    bool live_;
    uint64_t t0_ = 0;
};

}  // namespace trace
}  // namespace b200
