// Error policy of the b200comm runtime.
//
// Capability parity (behaviour, not code) with the reference's two error
// headers: mpierr.h:15-52 (return code -> "message + error-class message",
// then either throw std::runtime_error or print + MPI_Abort, chosen at build
// time by MPI_ERR_USE_EXCEPTIONS) and cuda_error_handler.h:47-86 (throwing and
// exit(1) variants for API status and for post-launch cudaGetLastError()).
//
// Here the choice is a run-time switch (set_error_mode), every error carries
// file:line, and one hierarchy covers communication and CUDA failures.
#pragma once
#include <stdexcept>
#include <string>

#pragma GCC visibility push(default)   // public C++ API: exported from the library
namespace b200 {

// Error classes (the analogue of MPI error classes, mpierr.h:23-25).
enum ErrorCode : int {
    OK = 0,
    ERR_ARG = 1,       // invalid argument
    ERR_BUFFER = 2,    // null / misaligned / too-small buffer
    ERR_COUNT = 3,     // negative or inconsistent count
    ERR_TYPE = 4,      // datatype not committed / signature mismatch
    ERR_TAG = 5,       // tag out of range
    ERR_COMM = 6,      // invalid or freed communicator
    ERR_RANK = 7,      // rank out of range for the communicator
    ERR_TRUNCATE = 8,  // message longer than the receive buffer
    ERR_TOPOLOGY = 9,  // invalid Cartesian request
    ERR_CUDA = 10,     // CUDA runtime / driver failure
    ERR_TIMEOUT = 11,  // device- or host-side wait exceeded its deadline
    ERR_STATE = 12,    // call not legal in the current state (e.g. before init)
    ERR_NOMEM = 13,    // allocation failure
    ERR_PEER = 14,     // a peer failed / no peer access
    ERR_INTERN = 15,   // internal invariant violated
    ERR_LASTCODE = 16
};

const char* error_class_string(int code) noexcept;

enum class ErrorMode : int { Throw = 0, Abort = 1 };
void set_error_mode(ErrorMode m) noexcept;
ErrorMode error_mode() noexcept;

class Error : public std::runtime_error {
public:
    Error(int code, const std::string& what, const char* file, int line);
    int code() const noexcept { return code_; }
    const char* file() const noexcept { return file_; }
    int line() const noexcept { return line_; }
private:
    int code_;
    const char* file_;
    int line_;
};
class CommError : public Error { using Error::Error; };
class CudaError : public Error { using Error::Error; };

// Formats "Error <code>:\n  error message: ...\n  error class message: ..."
// (the layout of format_mpi_err_msg, mpierr.h:15-28) and then throws or
// aborts according to error_mode().
[[noreturn]] void fail_comm(int code, const std::string& msg, const char* file, int line);
[[noreturn]] void fail_cuda(int cuda_status, const char* expr, const char* file, int line,
                            const char* extra = nullptr);

// Per-thread text of the last error that crossed the C ABI.
void set_last_error(const std::string& s) noexcept;
const char* last_error() noexcept;

}  // namespace b200
#pragma GCC visibility pop

#define B200_FAIL(code, msg) ::b200::fail_comm((code), (msg), __FILE__, __LINE__)
#define B200_REQUIRE(cond, code, msg)                                   \
    do {                                                                \
        if (!(cond)) ::b200::fail_comm((code), (msg), __FILE__, __LINE__); \
    } while (0)

// CUDA status check (cuda_error_handler.h:73-75 HANDLE_/DIE_ON_CUDA_ERROR).
#define B200_CHECK_CUDA(expr)                                                   \
    do {                                                                        \
        cudaError_t b200_st__ = (expr);                                         \
        if (b200_st__ != cudaSuccess)                                           \
            ::b200::fail_cuda((int)b200_st__, #expr, __FILE__, __LINE__);       \
    } while (0)

// Post-launch check (cuda_error_handler.h:80-86 LAUNCH_CUDA_KERNEL /
// DIE_ON_FAILED_KERNEL_LAUNCH): catches launch-configuration failures only;
// execution errors surface at the next synchronising call.
#define B200_CHECK_LAUNCH()                                                     \
    do {                                                                        \
        cudaError_t b200_st__ = cudaGetLastError();                             \
        if (b200_st__ != cudaSuccess)                                           \
            ::b200::fail_cuda((int)b200_st__, "(kernel launch)", __FILE__, __LINE__); \
    } while (0)
