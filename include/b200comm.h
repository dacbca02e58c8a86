/* b200comm -- C ABI of the Blackwell-native communication + stencil library.
 *
 * One shared object (cuda_mpi_scratch_b200/lib/libb200comm.so, sm_100a only)
 * exports everything below.  Python binds it with ctypes; C/C++ programs
 * include this header and link the library.  Every function returns 0 on
 * success or an error class from b200comm/error.hpp; b2c_last_error() gives
 * the message of the calling thread's last failure.
 *
 * "Reference" in the comments is ugovaretto-accel/cuda-mpi-scratch; citations
 * are path:line in that tree and name the CAPABILITY a function covers, not
 * code it was derived from.
 */
#ifndef B200COMM_H_
#define B200COMM_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ------------------------------------------------------------------ errors */
/* Reference: mpierr.h:15-52, cuda_error_handler.h:47-86. */
const char* b2c_last_error(void);
const char* b2c_error_class_string(int code);
int b2c_set_error_mode(int mode); /* 0 = return codes/exceptions, 1 = print + abort */
int b2c_version(void);

/* ------------------------------------------------------- device + memory */
/* Rank -> GPU binding.  Reference: mpicuda2.cu:198-202 (task % device_count,
 * round-robin over nodes) and stencil2d/mpi-2d-stencil-subarray-cuda.cu:40-73
 * (local rank from the launcher's environment, NUM_GPU_DEVICES override).
 * policy 0: local rank (LOCAL_RANK / OMPI_COMM_WORLD_LOCAL_RANK /
 *           MV2_COMM_WORLD_LOCAL_RANK / SLURM_LOCALID, else `rank`) % devices
 * policy 1: rank % devices
 * policy 2: (rank / node_count) % devices                                     */
int b2c_device_count(int* n);
int b2c_bind_device(int policy, int rank, int node_count, int* device_out);
int b2c_set_device(int device);
int b2c_get_device(int* device);
int b2c_device_synchronize(void);
int b2c_stream_synchronize(void* stream);
int b2c_device_sm_count(int device, int* sms);

int b2c_malloc(void** ptr, size_t bytes);          /* device memory, zero-filled */
int b2c_free(void* ptr);
int b2c_memset_async(void* ptr, int value, size_t bytes, void* stream);
int b2c_memcpy_async(void* dst, const void* src, size_t bytes, int kind, void* stream);
/* Pinned host memory.  Reference: host_allocator.h:58-93 (the C++ allocator
 * itself is include/b200comm/pinned_allocator.hpp). */
int b2c_pinned_alloc(void** ptr, size_t bytes);
int b2c_pinned_free(void* ptr);

/* Peer mapping of device allocations between processes (CUDA IPC) or between
 * devices of one process.  This replaces "hand the device pointer to a
 * CUDA-aware MPI" (reference: stencil2d/stencil2D.h:369-373). */
#define B2C_IPC_HANDLE_BYTES 64
int b2c_ipc_export(void* ptr, unsigned char handle[B2C_IPC_HANDLE_BYTES]);
int b2c_ipc_open(const unsigned char handle[B2C_IPC_HANDLE_BYTES], void** ptr);
int b2c_ipc_close(void* ptr);
int b2c_enable_peer_access(int device, int peer_device);
int b2c_can_access_peer(int device, int peer_device, int* yes);

/* ---------------------------------------------------------------- stencil */
/* Fused 2-D stencil + halo exchange.  Reference capability: the exchange/
 * compute/terminate loop of stencil2d/mpi-2d-stencil-subarray-cuda.cu:169-172
 * over the plan of stencil2d/stencil2D.h:381-437 (Compute() there is an empty
 * stub, so the update itself is new functionality).                          */
typedef struct b2c_stencil b2c_stencil;

typedef struct {
    int W, H;            /* core cells                                   */
    int radius, ghost;   /* ghost width == radius                        */
    int lead;            /* column of the first core cell in a padded row */
    int pitch, rows;     /* padded row length (elements) and row count   */
    size_t pad_bytes;    /* control pad at the start of the symmetric block */
    size_t tile_bytes;   /* one padded buffer                            */
    size_t off_buf0, off_buf1, off_red_slots;
    size_t total_bytes;  /* size of the per-rank symmetric block         */
} b2c_stencil_layout_t;

typedef struct {
    int cur;                 /* which buffer holds the current state */
    unsigned long long steps_done;
    int grid, threads, nstages, tiles_x, tiles_y, ncoef, launches_per_step;
    size_t smem_bytes;
    void* buf[2];
    double* resid;           /* device address of the last local residual */
} b2c_stencil_info_t;

/* Neighbour directions: 0 NW, 1 N, 2 NE, 3 W, 4 E, 5 SW, 6 S, 7 SE
 * (the order of MPIGridCellID, stencil2d/stencil2D.h:86-88). */
int b2c_stencil_layout(int W, int H, int radius, b2c_stencil_layout_t* out);
int b2c_stencil_create(b2c_stencil** out, int W, int H, int radius, int kind /*0 star,1 box*/,
                       void* local_block, void* const neighbour_block[8], int push_corners);
int b2c_stencil_destroy(b2c_stencil* s);
int b2c_stencil_set_coef(b2c_stencil* s, const double* host_coef, int n, void* stream);
int b2c_stencil_set_timeout_ms(b2c_stencil* s, unsigned long long ms);
int b2c_stencil_info(b2c_stencil* s, b2c_stencil_info_t* out);
/* flags: bit0 residual of the last step, bit1 drain the last step's halo */
int b2c_stencil_step(b2c_stencil* s, int nsteps, int flags, void* stream);
int b2c_stencil_drain(b2c_stencil* s, void* stream);
int b2c_stencil_exchange_push(b2c_stencil* s, void* stream);
int b2c_stencil_exchange_wait(b2c_stencil* s, void* stream);
int b2c_stencil_fill(b2c_stencil* s, int which, int region, double value, void* stream);
int b2c_stencil_status(b2c_stencil* s, void* stream, unsigned int* status);

#ifdef __cplusplus
}
#endif
#endif /* B200COMM_H_ */
