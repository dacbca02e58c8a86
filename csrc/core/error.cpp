// Error policy implementation -- see include/b200comm/error.hpp.
#include "b200comm/error.hpp"

#include <cuda_runtime.h>

#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <sstream>

namespace b200 {

static std::atomic<int> g_mode{(int)ErrorMode::Throw};
static thread_local std::string t_last_error;

const char* error_class_string(int code) noexcept {
    switch (code) {
        case OK: return "no error";
        case ERR_ARG: return "invalid argument";
        case ERR_BUFFER: return "invalid buffer";
        case ERR_COUNT: return "invalid count";
        case ERR_TYPE: return "invalid datatype";
        case ERR_TAG: return "invalid tag";
        case ERR_COMM: return "invalid communicator";
        case ERR_RANK: return "invalid rank";
        case ERR_TRUNCATE: return "message truncated";
        case ERR_TOPOLOGY: return "invalid topology";
        case ERR_CUDA: return "CUDA failure";
        case ERR_TIMEOUT: return "timeout";
        case ERR_STATE: return "invalid state";
        case ERR_NOMEM: return "out of memory";
        case ERR_PEER: return "peer failure";
        case ERR_INTERN: return "internal error";
        default: return "unknown error";
    }
}

void set_error_mode(ErrorMode m) noexcept { g_mode.store((int)m); }
ErrorMode error_mode() noexcept { return (ErrorMode)g_mode.load(); }

Error::Error(int code, const std::string& what, const char* file, int line)
    : std::runtime_error(what), code_(code), file_(file), line_(line) {}

void set_last_error(const std::string& s) noexcept {
    try { t_last_error = s; } catch (...) {}
}
const char* last_error() noexcept { return t_last_error.c_str(); }

static std::string format(int code, const std::string& msg, const char* file, int line) {
    std::ostringstream os;
    os << "Error " << code << ":\n  error message: " << msg
       << "\n  error class message: " << error_class_string(code)
       << "\n  at " << file << ":" << line;
    return os.str();
}

[[noreturn]] void fail_comm(int code, const std::string& msg, const char* file, int line) {
    std::string text = format(code, msg, file, line);
    if (error_mode() == ErrorMode::Abort) {
        std::fprintf(stderr, "%s\n", text.c_str());
        std::fflush(stderr);
        std::_Exit(code == 0 ? 1 : code);
    }
    throw CommError(code, text, file, line);
}

[[noreturn]] void fail_cuda(int cuda_status, const char* expr, const char* file, int line,
                            const char* extra) {
    std::ostringstream m;
    if (extra) m << extra << " ";
    m << expr << " -> " << cudaGetErrorName((cudaError_t)cuda_status) << ": "
      << cudaGetErrorString((cudaError_t)cuda_status);
    std::string text = format(ERR_CUDA, m.str(), file, line);
    if (error_mode() == ErrorMode::Abort) {
        std::fprintf(stderr, "%s\n", text.c_str());
        std::fflush(stderr);
        std::_Exit(1);
    }
    throw CudaError(ERR_CUDA, text, file, line);
}

}  // namespace b200
