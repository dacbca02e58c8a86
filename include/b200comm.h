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

/* ---------------------------------------------------------------- tracing */
/* Host-side spans of every communication call and every enqueue of a device operation,
 * written as Chrome trace-event JSON.  Reference: MPI_Wtime / clock() brackets by hand
 * (test-benchmark/mpi-pingpong-gpu.cpp:51-57, mpicuda3.cu:176-179,315-325).
 * mode bit 0: record spans, bit 1: NVTX ranges.  Environment: B200COMM_TRACE=<prefix> records
 * from load time and writes <prefix>.<pid>.json at exit; B200COMM_NVTX=1 emits NVTX ranges. */
int b2c_trace_enable(int mode);
int b2c_trace_clear(void);
int b2c_trace_count(int64_t* n);
int b2c_trace_dump(const char* path, int64_t* n);
int b2c_trace_begin(const char* name, uint64_t* t0);
int b2c_trace_end(const char* name, uint64_t t0, int rank, int64_t bytes);

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

/* Symmetric heap building blocks (csrc/core/vmm.cpp): physical allocations that are exported as
 * POSIX file descriptors, imported and mapped by the peers (no cudaIpc*, one mapping per arena),
 * and NVLink multicast objects bound to them, so that one store reaches every rank's memory and
 * one multimem.ld_reduce returns the in-switch sum.  The Python SymmetricHeap drives these calls
 * and moves the descriptors between processes (SCM_RIGHTS).  Reference counterpart: none -- the
 * reference hands device pointers to a CUDA-aware MPI (stencil2d/stencil2D.h:369-373). */
int b2c_vmm_support(int device, int* vmm, int* multicast);
int b2c_vmm_granularity(int device, int ndevices, size_t* alloc_gran, size_t* mc_gran);
int b2c_vmm_create(size_t bytes, int device, unsigned long long* handle, int* fd);
int b2c_vmm_import(int fd, unsigned long long* handle);
int b2c_vmm_map(unsigned long long handle, size_t bytes, size_t align, int device, void** ptr);
int b2c_vmm_unmap(void* ptr, size_t bytes);
int b2c_vmm_release(unsigned long long handle);
int b2c_mc_create(int ndevices, size_t bytes, unsigned long long* handle, int* fd);
int b2c_mc_add_device(unsigned long long mc, int device);
int b2c_mc_bind(unsigned long long mc, unsigned long long mem, size_t bytes);
int b2c_close_fd(int fd);

/* -------------------------------------------------------------- datatypes */
/* Derived datatypes.  Reference: MPI_Type_indexed (mpi7.cpp:32-41),
 * MPI_Type_create_struct + MPI_Type_extent (mpi8.cpp:38-54),
 * MPI_Type_create_subarray (mpi-complex-types.cpp:33-36, stencil2d/stencil2D.h:210-228),
 * MPI_Type_create_hindexed (mpi-complex-types.cpp:38-50), commit/free.
 * basic kinds: 0 BYTE 1 CHAR 2 INT8 3 UINT8 4 INT16 5 INT32 6 INT64 7 UINT64
 *              8 FLOAT 9 DOUBLE 10 UINT32                                    */
typedef struct b2c_type b2c_type;
int b2c_type_basic(int kind, b2c_type** out);
int b2c_type_contiguous(int count, b2c_type* old, b2c_type** out);
int b2c_type_vector(int count, int blocklen, int stride, b2c_type* old, b2c_type** out);
int b2c_type_hvector(int count, int blocklen, int64_t stride_bytes, b2c_type* old, b2c_type** out);
int b2c_type_indexed(int n, const int* blocklens, const int* disps, b2c_type* old, b2c_type** out);
int b2c_type_hindexed(int n, const int* blocklens, const int64_t* byte_disps, b2c_type* old,
                      b2c_type** out);
int b2c_type_struct(int n, const int* blocklens, const int64_t* byte_disps, b2c_type* const* types,
                    b2c_type** out);
int b2c_type_subarray(int ndims, const int* sizes, const int* subsizes, const int* starts,
                      int order /*0 C, 1 Fortran*/, b2c_type* old, b2c_type** out);
int b2c_type_resized(b2c_type* old, int64_t lb, int64_t extent, b2c_type** out);
int b2c_type_commit(b2c_type* t);
int b2c_type_free(b2c_type* t);
int b2c_type_size(b2c_type* t, int64_t* size);
int b2c_type_extent(b2c_type* t, int64_t* lb, int64_t* extent);
int b2c_type_runs(b2c_type* t, int64_t count, int64_t* disp_len_pairs, int64_t cap, int64_t* n);
int b2c_type_pack(b2c_type* t, const void* buf, int64_t count, void* packed);
int b2c_type_unpack(b2c_type* t, const void* packed, int64_t bytes, void* buf, int64_t count);
int b2c_type_describe(b2c_type* t, char* buf, int cap);

/* ------------------------------------------------------------ communication */
/* MPI-like communicators over the built-in host transport (threads of one
 * process, or processes attached to a POSIX shared-memory segment).
 * Reference: MPI_Init/Finalize/Comm_rank/Comm_size/Get_processor_name/Wtime/
 * Abort (mpi1.cpp:11-16, mpierr.h:41), Send/Recv/Probe/Get_count (mpi3.cpp:26-46),
 * Isend/Irecv/Wait/Waitall (mpi5.cpp:34-75), Gather/Bcast/Reduce/Allreduce
 * (mpi6.cpp:89-100, mpi9.cpp:51-54, mpicuda2.cu:154,293), groups (mpi9.cpp:26-54),
 * Cartesian topology (mpi10.cpp:22-54).                                       */
#define B2C_ANY_SOURCE (-1)
#define B2C_ANY_TAG (-1)
#define B2C_PROC_NULL (-2)
#define B2C_UNDEFINED (-32766)
typedef struct b2c_comm b2c_comm;
typedef struct b2c_request b2c_request;
typedef struct { int source, tag, error; int64_t bytes; } b2c_status;
/* reduction ops: 0 SUM 1 PROD 2 MAX 3 MIN 4 LAND 5 LOR 6 BAND 7 BOR */

int b2c_init_local(int nranks, size_t ring_bytes, b2c_comm** out /* nranks handles */);
int b2c_init_process(const char* job, int rank, int nranks, size_t ring_bytes, double timeout_s,
                     b2c_comm** out);
int b2c_init_env(b2c_comm** out);
int b2c_comm_finalize(b2c_comm* c);
int b2c_comm_free(b2c_comm* c);
int b2c_comm_rank(b2c_comm* c, int* rank);
int b2c_comm_size(b2c_comm* c, int* size);
int b2c_comm_set_timeout(b2c_comm* c, double seconds);
int b2c_abort(b2c_comm* c, int code);
double b2c_wtime(void);
int b2c_processor_name(char* buf, int cap);

int b2c_send(b2c_comm* c, const void* buf, int64_t count, b2c_type* t, int dest, int tag);
int b2c_recv(b2c_comm* c, void* buf, int64_t count, b2c_type* t, int src, int tag, b2c_status* st);
int b2c_isend(b2c_comm* c, const void* buf, int64_t count, b2c_type* t, int dest, int tag,
              b2c_request** out);
int b2c_irecv(b2c_comm* c, void* buf, int64_t count, b2c_type* t, int src, int tag,
              b2c_request** out);
int b2c_wait(b2c_request* r, b2c_status* st);
int b2c_test(b2c_request* r, int* flag, b2c_status* st);
int b2c_waitall(int n, b2c_request** reqs, b2c_status* sts);
int b2c_probe(b2c_comm* c, int src, int tag, b2c_status* st);
int b2c_iprobe(b2c_comm* c, int src, int tag, int* flag, b2c_status* st);
int b2c_get_count(const b2c_status* st, b2c_type* t, int64_t* count);
int b2c_sendrecv(b2c_comm* c, const void* sbuf, int64_t scount, b2c_type* st, int dest, int stag,
                 void* rbuf, int64_t rcount, b2c_type* rt, int src, int rtag, b2c_status* status);

int b2c_barrier(b2c_comm* c);
int b2c_bcast(b2c_comm* c, void* buf, int64_t count, b2c_type* t, int root);
int b2c_reduce(b2c_comm* c, const void* sbuf, void* rbuf, int64_t count, b2c_type* t, int op, int root);
int b2c_allreduce(b2c_comm* c, const void* sbuf, void* rbuf, int64_t count, b2c_type* t, int op);
int b2c_gather(b2c_comm* c, const void* sbuf, int64_t scount, b2c_type* st, void* rbuf,
               int64_t rcount, b2c_type* rt, int root);
int b2c_scatter(b2c_comm* c, const void* sbuf, int64_t scount, b2c_type* st, void* rbuf,
                int64_t rcount, b2c_type* rt, int root);
int b2c_allgather(b2c_comm* c, const void* sbuf, int64_t scount, b2c_type* st, void* rbuf,
                  int64_t rcount, b2c_type* rt);

int b2c_comm_group(b2c_comm* c, int* world_ranks, int cap, int* n);
int b2c_comm_create(b2c_comm* c, int n, const int* ranks_in_parent, b2c_comm** out);
int b2c_comm_split(b2c_comm* c, int color, int key, b2c_comm** out);
int b2c_comm_dup(b2c_comm* c, b2c_comm** out);
int b2c_cart_create(b2c_comm* c, int ndims, const int* dims, const int* periods, int reorder,
                    b2c_comm** out);
int b2c_cart_coords(b2c_comm* c, int rank, int maxdims, int* coords);
int b2c_cart_rank(b2c_comm* c, int ndims, const int* coords, int* rank);
int b2c_cart_shift(b2c_comm* c, int dim, int disp, int* src, int* dst);
int b2c_cart_get(b2c_comm* c, int maxdims, int* ndims, int* dims, int* periods);
int b2c_dims_create(int nranks, int ndims, int* dims);

/* ---------------------------------------------------------- halo exchange */
/* Layout descriptors, tile regions and the 8-neighbour halo plan with its
 * UNFUSED exchange over the comm layer (any datatype, host or device tiles).
 * Reference: stencil2d/stencil2D.h -- Array2D :30-42, RegionID :79-82,
 * SubArrayRegion :107-201, CreateSendRecvArrays :381-437, ExchangeData :363-377,
 * TestSubRegionExtraction :441-510.  C++ face: include/b200comm/stencil2d.hpp.
 * regions: 0 TOP_LEFT 1 TOP_CENTER 2 TOP_RIGHT 3 CENTER_LEFT 4 CENTER
 *          5 CENTER_RIGHT 6 BOTTOM_LEFT 7 BOTTOM_CENTER 8 BOTTOM_RIGHT
 *          9 TOP 10 LEFT 11 BOTTOM 12 RIGHT                                  */
typedef struct { int width, height, x_offset, y_offset, row_stride; } b2c_array2d;
typedef struct b2c_halo_plan b2c_halo_plan;
int b2c_sub_array_region(const b2c_array2d* g, int stencil_w, int stencil_h, int region,
                         b2c_array2d* out);
int b2c_halo_plan_create(b2c_comm* cart, const b2c_array2d* tile, int stencil_w, int stencil_h,
                         b2c_type* elem, b2c_halo_plan** out);
int b2c_halo_plan_destroy(b2c_halo_plan* p);
int b2c_halo_exchange(b2c_halo_plan* p, void* tile_data);
int b2c_halo_plan_entry(b2c_halo_plan* p, int which /*0 recv,1 send*/, int i, int* peer, int* tag,
                        int* region, b2c_array2d* layout, int64_t* bytes);
int b2c_region_report(int w, int h, int stencil_w, int stencil_h, char* buf, int cap);

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
    unsigned long long launches; /* kernels launched through this plan so far */
} b2c_stencil_info_t;

/* Neighbour directions: 0 NW, 1 N, 2 NE, 3 W, 4 E, 5 SW, 6 S, 7 SE
 * (the order of MPIGridCellID, stencil2d/stencil2D.h:86-88). */
int b2c_stencil_layout(int W, int H, int radius, b2c_stencil_layout_t* out);
int b2c_stencil_create(b2c_stencil** out, int W, int H, int radius,
                       int kind /*0 star, 1 box, 2 radius-1 star x `radius` time levels per sweep*/,
                       void* local_block, void* const neighbour_block[8], int push_corners);
int b2c_stencil_destroy(b2c_stencil* s);
int b2c_stencil_set_coef(b2c_stencil* s, const double* host_coef, int n, void* stream);
int b2c_stencil_set_timeout_ms(b2c_stencil* s, unsigned long long ms);
int b2c_stencil_info(b2c_stencil* s, b2c_stencil_info_t* out);
/* flags: bit0 residual of the last step, bit1 drain the last step's halo */
int b2c_stencil_step(b2c_stencil* s, int nsteps, int flags, void* stream);
/* kind 2 (temporal blocking: radius-1 star, `radius` time levels per sweep): one sweep that
 * computes only `levels` (1..radius) levels -- the tail of a run that is not a multiple */
int b2c_stencil_step_levels(b2c_stencil* s, int levels, int flags, void* stream);
/* One call of a pipelined loop: `nsteps` time steps (any count) with this call's coefficients
 * taken from host memory (NULL: unchanged) and its residual (flags bit0) returned to host
 * memory.  Both copies run on copy streams of the plan, next to the kernels of the neighbouring
 * calls; `stream` carries the kernels only.  *ticket names the call; b2c_stencil_io_wait blocks
 * until that call has finished on the device and hands out its residual (local sum of squared
 * updates).  The results of the last two calls are kept. */
int b2c_stencil_step_io(b2c_stencil* s, int nsteps, int flags, const double* host_coef, int n,
                        void* stream, unsigned long long* ticket);
int b2c_stencil_io_wait(b2c_stencil* s, unsigned long long ticket, double* resid);
/* after replaying a captured CUDA graph of `nsteps` steps: advance the host-side step count */
int b2c_stencil_advance(b2c_stencil* s, long long nsteps);
int b2c_stencil_drain(b2c_stencil* s, void* stream);
int b2c_stencil_exchange_push(b2c_stencil* s, void* stream);
int b2c_stencil_exchange_wait(b2c_stencil* s, void* stream);
int b2c_stencil_fill(b2c_stencil* s, int which, int region, double value, void* stream);
int b2c_stencil_status(b2c_stencil* s, void* stream, unsigned int* status);
/* Per-phase device timing of the last sweep (temporal strip kernel, plan created with
 * B2C_STENCIL_DEBUG=32): boundary phase, CTA lifetime, halo-flag waits, grid span, in microseconds
 * (mean / max over CTAs).  Reference counterpart: the clock() brackets of mpicuda3.cu:176-179. */
int b2c_stencil_phase_stats(b2c_stencil* s, void* stream, double out[7]);

/* ------------------------------------------------- P2P copy engine / flags */
/* Kernel-initiated put into a peer-mapped buffer + device-side epoch flags, and a
 * device-resident ping-pong loop.  Reference capability: the GPU ping-pong
 * benchmarks test-benchmark/mpi-pingpong-gpu.cpp:24-83 (blocking) and
 * mpi-pingpong-gpu-async.cpp:43-106 (Isend/Irecv/Wait, host-staged variants). */
int b2c_p2p_put(void* dst_peer, const void* src, size_t bytes, void* peer_flag,
                unsigned long long flag_value, void* counter, int mode /*0 LD/ST, 1 TMA*/,
                int max_ctas, void* stream);
int b2c_flag_wait(const void* flag, unsigned long long value, void* status,
                  unsigned long long timeout_ms, void* stream);
int b2c_flag_set(void* flag, unsigned long long value, void* stream);
int b2c_pingpong_device(int role, int iters, size_t bytes, const void* send_local, void* recv_peer,
                        const void* my_flag, void* peer_flag, unsigned long long base_epoch,
                        double* out_ns, void* status, unsigned long long timeout_ms, void* stream);

/* Device sends through the MPI-like API: counts of the paths taken so far in this process.
 * out[0] eager (put kernel into the receiver's device mailbox, <= 64 KiB contiguous);
 * rendezvous because: out[1] too large / eager disabled (B200COMM_NO_EAGER=1), out[2] no mapped
 * mailbox of the receiver yet, out[3] all slots to that receiver in flight, out[4] buffer on a
 * different GPU than this rank's mailbox.  Reference counterpart: none (a CUDA-aware MPI hides
 * its eager/rendezvous switch, test-benchmark/mpi-pingpong-gpu.cpp:52-53). */
int b2c_dev_eager_stats(unsigned long long out[5]);

/* ------------------------------------------------ datatype-driven device moves */
/* Pack + move + unpack in one kernel; src and dst may be peer mapped.  Reference
 * capability: derived datatypes on device pointers handed to a CUDA-aware MPI
 * (stencil2d/stencil2D.h:369-373); see also b2c_send/b2c_recv, which accept
 * device buffers and use these kernels on the receiver side.                   */
int b2c_typed_move(const void* src, b2c_type* stype, int64_t scount, void* dst, b2c_type* dtype,
                   int64_t dcount, void* stream, int64_t* moved_bytes);
int b2c_move_uniform(const void* src, void* dst, int64_t nruns, int64_t len, int64_t sstride,
                     int64_t dstride, void* stream);
int b2c_move_segments(const void* src, void* dst, const void* dev_segments, int64_t nseg,
                      int64_t max_len, int64_t items_64k, void* stream);

/* ------------------------------------------------------ device collectives */
/* Collectives on DEVICE tensors through peer-mapped staging + epoch flags (no host,
 * no NCCL).  Reference capability: MPI_Reduce / MPI_Allreduce / MPI_Gather /
 * MPI_Bcast / MPI_Scatter (mpicuda2.cu:154,293, mpi9.cpp:51-54, mpi6.cpp:89-100),
 * which the reference only ever applies to host scalars after a D2H copy.
 * dtype 5 int32, 6 int64, 8 float32, 9 float64; op 0 SUM 1 PROD 2 MAX 3 MIN;
 * algo 0 one-shot, 1 two-shot; kind 0 bcast 1 gather 2 allgather 3 scatter.     */
typedef struct b2c_coll b2c_coll;
int b2c_coll_workspace_bytes(size_t max_bytes, size_t* bytes);
int b2c_coll_create(b2c_coll** out, int rank, int nranks, void* const* workspaces, size_t max_bytes);
int b2c_coll_destroy(b2c_coll* c);
/* Multicast mapping of the workspace (NULL: none).  Enables algo 2 of b2c_coll_allreduce -- the
 * sum is formed inside the NVSwitch (multimem.ld_reduce) and fanned out by it (multimem.st) -- and,
 * with mc_bcast != 0, bcast through one multicast store per vector.  Reference capability:
 * MPI_Allreduce mpi9.cpp:51-54, MPI_Bcast mpicuda2.cu:154 (host scalars there). */
int b2c_coll_set_multicast(b2c_coll* c, void* mc_base, int mc_bcast);
/* In-place SUM allreduce of a tensor that lives in a multicast arena of the symmetric heap
 * (`mc_data` = its multicast address): no staging copies, reduction and fan-out by the switch. */
int b2c_coll_allreduce_symmetric(b2c_coll* c, void* mc_data, int64_t count, int dtype, void* stream);
/* bcast / allgather in place on such a tensor: this rank's byte range [lo, hi) of it (`local_data`
 * = this rank's own address of the tensor) is stored once to the multicast address and lands in
 * every rank's copy.  bcast: [0, bytes) on the root, lo == hi elsewhere; allgather: block `rank`. */
int b2c_coll_spread_symmetric(b2c_coll* c, void* mc_data, const void* local_data, size_t lo, size_t hi,
                              void* stream);
int b2c_coll_allreduce(b2c_coll* c, const void* in, void* out, int64_t count, int dtype, int op,
                       int root /* <0: allreduce */, int algo, int phases, void* stream);
int b2c_coll_move(b2c_coll* c, int kind, const void* sendbuf, void* recvbuf, size_t block_bytes,
                  int root, int phases, void* stream);
int b2c_coll_barrier(b2c_coll* c, int phases, void* stream);
int b2c_coll_check(b2c_coll* c, void* stream);

/* -------------------------------------------------------- distributed dot */
/* Dot product whose local reduction is fused with a one-shot P2P allreduce
 * (or reduce-to-root) in ONE kernel per rank.  Reference capability:
 * mpicuda2.cu:65-100,259-293, mpicuda3.cu, mpicuda4.cu:97-185,347-368
 * (kernel + D2H + MPI_Reduce on the host).                                   */
typedef struct b2c_dot b2c_dot;
int b2c_dot_symmetric_bytes(int nranks, size_t* bytes);
int b2c_dot_create(b2c_dot** out, int rank, int nranks, void* const* slot_tables);
int b2c_dot_destroy(b2c_dot* d);
/* dtype 8 = float32, 9 = float64; root < 0 = allreduce; phases 3 = one launch */
int b2c_dot_run(b2c_dot* d, const void* x, const void* y, int64_t n, int dtype, int root,
                int phases, void* stream);
int b2c_dot_result_ptr(b2c_dot* d, double** device_scalar);
int b2c_dot_result(b2c_dot* d, void* stream, double* value);
int b2c_dot_set_timeout_ms(b2c_dot* d, unsigned long long ms);

/* Local dot product that allocates NOTHING (no counterpart in the reference).
 * For callers that may not call cudaMalloc where they need the result: code
 * inside another runtime's callback or a latency-critical loop (cudaMalloc
 * synchronises the device), or a process whose GPU memory already belongs to a
 * framework allocator / fixed arena.  The kernel's scratch is a fixed array
 * inside the library image (8448 bytes per GPU, held until exit), so:
 *   - at most 1024 thread blocks (b2c_dot_local_noalloc_limits); more is an error;
 *   - all calls share that scratch: calls that overlap on one GPU race -- keep
 *     them on one stream;
 *   - the first call on a device also resolves the scratch address; make that
 *     call before starting any stream capture.
 * result is a device pointer to one double; synchronise `stream` before reading. */
int b2c_dot_local_noalloc_limits(int* max_ctas, size_t* scratch_bytes);
int b2c_dot_local_noalloc(const void* x, const void* y, int64_t n, int dtype /*8 f32, 9 f64*/,
                          int ctas /*0 = 4 per SM*/, double* result, void* stream);

/* Reference-STYLE local dot kernels, baselines only (csrc/baseline/):
 * variant 0 atomicAdd per block (mpicuda2.cu:65-81), 1 per-block partials for a
 * host finish (mpicuda2.cu:84-100), 2 last-block finish (mpicuda4.cu:157-185). */
int b2c_baseline_dot_blocks(int64_t n);
int b2c_baseline_dot(int variant, const void* a, const void* b, int64_t n, int dtype, void* out,
                     void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200COMM_H_ */
