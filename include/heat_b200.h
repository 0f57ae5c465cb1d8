/* heat_b200.h — C ABI of libheat.so: the FTCS 5-point heat-equation time-march for NVIDIA B200.
 *
 * Drop-in boundary for the one hot path of cssrikanth/CUDA-HIP-MPI-Heat-equation-test.
 * Plain C: pointers, sizes and PODs only; no C++/torch types.  Load with dlopen / ctypes.CDLL.
 * "ref:" comments name the reference interface (file:line, relative to the reference root)
 * that each entry point replaces.
 *
 * Field layout: row-major [ny][nx], x contiguous — byte-identical to the reference's Fortran
 * T(i,j) (i contiguous) and to NumPy u[y, x].  Boundary cells (row 0, row ny-1, column 0,
 * column nx-1) are Dirichlet: they are never updated and keep the values passed to heat_init.
 *
 * Conventions: every function returns HEAT_OK (0) or a negative error code and never throws or
 * aborts; heat_last_error() returns the message of the calling thread's last failure.  A handle
 * is not thread-safe; different handles are independent.  Work is asynchronous on the handle's
 * stream unless stated; heat_gather / heat_sync / heat_get_timing synchronise.
 */
#ifndef HEAT_B200_H
#define HEAT_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define HEAT_OK               0
#define HEAT_ERR_INVALID     -1   /* bad argument / configuration */
#define HEAT_ERR_CUDA        -2   /* a CUDA runtime or driver call failed (no CUDA device included) */
#define HEAT_ERR_NCCL        -3   /* NCCL unavailable or a NCCL call failed */
#define HEAT_ERR_NOMEM       -4
#define HEAT_ERR_UNSUPPORTED -5   /* valid request this build / device cannot serve */

/* dtype — the arithmetic type of the march.  ref: fortran/hip/heat_kernel.cpp:5-9 (real). */
#define HEAT_F64 0
#define HEAT_F32 1

/* op_order — which of the reference's three floating-point evaluation orders to reproduce
 * (E/W = +-1 in x, the contiguous index; N/S = +-1 in y):
 *   FORTRAN   C + r*((((E+N)+W)+S) - 4*C)          ref: fortran/serial/heat.f90:66,
 *                                                       fortran/hip/heat_kernel.cpp:43
 *   PYCUDA    C + r*((((W+E)+N)+S) - 4*C)          ref: python/cuda/cuda.py:78
 *   PYSERIAL  (C + rx*((E-2C)+W)) + ry*((N-2C)+S)  ref: python/serial/heat.py:50-54        */
#define HEAT_ORDER_FORTRAN  0
#define HEAT_ORDER_PYCUDA   1
#define HEAT_ORDER_PYSERIAL 2

/* kernel — which hand-written sm_100a kernel advances the field (AUTO picks per shape). */
#define HEAT_KERNEL_AUTO    0
#define HEAT_KERNEL_SIMPLE  1   /* one thread per cell; any pitch/alignment                      */
#define HEAT_KERNEL_STREAM  2   /* 128-bit row loads, register row window, shuffle neighbours    */
#define HEAT_KERNEL_TMA     3   /* TMA 2-D tile loads (1-cell halo) into shared memory, mbarrier */
#define HEAT_KERNEL_RESIDENT 4  /* whole field in one CTA's shared memory, all steps in 1 launch */
#define HEAT_KERNEL_WTMA    5   /* one private TMA pipeline per WARP, rows rotate through registers
                                   (the temporal-blocking kernel, also used for single steps)     */
#define HEAT_KERNEL_PERSIST 6   /* field resident in the shared memory of ALL SMs (one row slab per CTA), all
                                   steps of a heat_step call in ONE cooperative launch, neighbour CTAs trade
                                   ghost rows through L2 (calls shorter than 8 steps use the TMA kernel)     */

/* flags */
#define HEAT_FLAG_FMAD          1u  /* FMA contraction.  Default (flag clear) is every add/mul individually rounded:
                                       bit-exact with the reference's CPU programs (gfortran -O0, NumPy) and with its
                                       GPU kernel text read literally (goldens built with -ffp-contract=off).
                                       With the flag, FORTRAN and PYCUDA orders evaluate fma(r, fma(-4,C,t), C) —
                                       bit-exact with what the reference's GPU toolchains make of that text by default
                                       (nvcc -fmad=true under PyCUDA, hipcc -ffp-contract=fast; goldens *_fma, built
                                       with gcc -mfma -ffp-contract=fast — a run of the nvcc/hipcc binaries themselves
                                       is not available here, so that pairing stays unpinned) and <= 1e-12 relative
                                       off the uncontracted result; PYSERIAL (NumPy never contracts) has no
                                       reference counterpart under this flag: <= 1e-12 only */
#define HEAT_FLAG_OVERLAP       2u  /* nranks>1: exchange halos on a second stream behind the interior update */
#define HEAT_FLAG_SKIP_EXCHANGE 4u  /* DIAGNOSTIC: no halo traffic (wrong results) — to measure halo time share */
#define HEAT_FLAG_PEER_HALO     8u  /* nranks>1: neighbours' rows are written by peer stores (CUDA IPC) instead of NCCL */

typedef struct heat_handle heat_handle;

/* ref: the parameters every reference program derives after reading input.dat
 * (fortran/serial/heat.f90:13-17,59; python/cuda/cuda.py:31-39) plus the launch choices the
 * reference hard-codes (block sizes cuda.py:99, heat_kernel.cpp:20-21; ng=1 mpi+cuda/heat.F90:41). */
typedef struct heat_config {
    uint32_t struct_size;  /* = sizeof(heat_config); guards against ABI drift                        */
    int32_t  nx, ny;       /* GLOBAL grid size including the boundary cells                           */
    int32_t  dtype;        /* HEAT_F64 | HEAT_F32                                                    */
    int32_t  op_order;     /* HEAT_ORDER_*                                                           */
    int32_t  kernel;       /* HEAT_KERNEL_*                                                          */
    int32_t  tblock;       /* k >= 1 time steps fused per HBM round trip (temporal blocking); 0 = 1  */
    uint32_t flags;        /* HEAT_FLAG_*                                                            */
    double   rx, ry;       /* nu*dt/dx^2, nu*dt/dy^2 — derived by the HOST in fp64 exactly as the
                              reference does; FORTRAN and PYCUDA orders use rx only                  */
    int32_t  device;       /* CUDA device ordinal                                                    */
    int32_t  rank, nranks; /* row-slab decomposition: this process owns heat_slab_partition() rows  */
    int32_t  ghost;        /* ghost depth g >= 1 (rows exchanged every g steps); 0 = default         */
    const void* nccl_id;   /* nranks>1: 128-byte ncclUniqueId from heat_nccl_unique_id on rank 0     */
    void*    ext_buf[2];   /* optional caller-owned DEVICE buffers (e.g. torch data_ptr()), each
                              [(rows + 2g)][ext_pitch]; caller keeps them alive until heat_destroy   */
    int64_t  ext_pitch;    /* row pitch of ext_buf in ELEMENTS (>= nx)                               */
    uint64_t stream;       /* optional cudaStream_t to run on (0: the library creates its own)       */
    int32_t  ranks_x;      /* 2-D decomposition: ranks along x (0 or 1 = row slabs, the default and the fast
                              path); nranks must be a multiple; rank = ry * ranks_x + rx owns the tile
                              heat_tile_partition() gives it.  ref: ndims / nblocks / MPI_Dims_create,
                              fortran/mpi+cuda/heat.F90:28-29,87-97                                  */
    int32_t  reserved0;
} heat_config;

typedef struct heat_info {
    int32_t nx, ny;            /* global                                                             */
    int32_t row0, nrows;       /* this rank's rows [row0, row0+nrows) of the global field            */
    int32_t ghost;             /* g; local buffers hold nrows + 2g rows, own rows start at local row g */
    int32_t kernel, tblock;    /* what AUTO resolved to                                              */
    int64_t pitch;             /* elements                                                           */
    void*   buf[2];            /* device pointers of the two ping-pong buffers                       */
    int32_t current;           /* which of buf[] holds the current field                             */
    int32_t device;
    int32_t col0, ncols;       /* this rank's columns [col0, col0+ncols) (the whole width for row slabs)             */
    int32_t ghost_left;        /* ghost columns held left of the own ones; local rows are local_nx wide              */
    int32_t local_nx;
} heat_info;

typedef struct heat_timing {
    double   last_step_ms;     /* device time of the last heat_step call (CUDA events on its stream) */
    double   last_halo_ms;     /* device time of the last halo exchange (its own stream's events)    */
    uint64_t steps_total;      /* time steps advanced since heat_init                                */
    uint64_t kernel_launches;  /* stencil kernel launches since heat_init                            */
    uint64_t halo_exchanges;   /* exchanges since heat_init                                          */
} heat_timing;

/* Create a solver.  `field` (may be NULL: then call heat_set_field) points to THIS RANK's rows
 * [row0, row0+nrows) x nx, dense, HOST memory, of cfg->dtype.  Both ping-pong buffers receive it,
 * so the Dirichlet ring is present in both (ref: the ring survives in the reference only because
 * T_old = T is a full copy, fortran/serial/heat.f90:63).  With nranks>1 the call is collective
 * (NCCL communicator creation + first ghost fill).
 * ref: allocation + H2D of python/cuda/cuda.py:99 (drv.InOut/In), fortran/mpi+cuda/heat.F90:107-121,256. */
int heat_init(heat_handle** out, const heat_config* cfg, const void* field);

/* Replace the current field (this rank's rows). src_is_device: 0 host, 1 device pointer; src_pitch in elements. */
int heat_set_field(heat_handle* h, const void* src, int64_t src_pitch, int src_is_device);

/* Advance nt time steps (asynchronous).  Includes the halo exchanges when nranks>1.
 * ref: the time loops python/cuda/cuda.py:96-99, python/serial/heat.py:48-58,
 * fortran/serial/heat.f90:61-69, fortran/mpi+cuda/heat.F90:206-219 (callers apply the
 * reference's own count: nt+1 for the Python scripts, ntime for Fortran). */
int heat_step(heat_handle* h, int nt);

/* Refresh this rank's ghost rows from its neighbours now (collective).
 * ref: swap(), fortran/mpi+cuda/heat.F90:145-193; launch_send/recv1/recv2 heat_kernel.cpp:63-150. */
int heat_halo_exchange(heat_handle* h);

/* Peer-memory halo (HEAT_FLAG_PEER_HALO, one process per GPU on one NVLink node, row slabs only): instead of
 * NCCL, the edge-strip stencil kernels store their result rows into the own buffer AND, as peer stores over
 * NVLink, into the neighbours' ghost rows, which are mapped here through CUDA IPC (kernels without that second
 * destination and the first ghost fill use the copy engines); ordering uses one flag word per direction, signalled
 * by a 1-thread kernel and awaited with cuStreamWaitValue32 on the exchange stream (HEAT_PEER_WAIT=kernel: one
 * kernel that signals and spins).
 * After heat_init every rank calls
 * heat_peer_export(blob) (HEAT_PEER_BLOB_BYTES bytes), the host layer hands each rank the blobs of rank-1 and
 * rank+1 (NULL at the ends), and heat_peer_attach opens them (no cross-GPU wait: the ghosts are filled by the
 * first heat_step / heat_halo_exchange, which all ranks must then call).
 * cfg.nccl_id may then be NULL (heat_gather with global_root >= 0 still needs it).
 * ref: replaces swap() — pack, 2 x MPI_Sendrecv, unpack — of fortran/mpi+cuda/heat.F90:145-193 entirely. */
#define HEAT_PEER_BLOB_BYTES 256
int heat_peer_export(heat_handle* h, void* blob);
int heat_peer_attach(heat_handle* h, const void* up_blob, const void* down_blob);

/* Synchronise and copy the current field out: this rank's rows, dense [nrows][nx], to `dst`
 * (dst_is_device 0/1).  With global_root >= 0 the call is collective and rank `global_root`
 * receives the whole [ny][nx] field instead (other ranks pass dst = NULL).
 * ref: D2H of python/cuda/cuda.py:99,102; T = Td fortran/mpi+cuda/heat.F90:261. */
int heat_gather(heat_handle* h, void* dst, int dst_is_device, int global_root);

/* Job pipeline: heat_submit = heat_set_field(host_src) + heat_step(nt) + heat_gather(host_dst) for this rank's
 * rows, fully asynchronous.  Jobs submitted back to back overlap on the GPU: the next job's rows cross PCIe
 * (host -> device) and the previous job's result leaves (device -> host) on two copy streams while the current
 * job marches, through two device staging buffers the first call allocates (2 x this rank's rows).  Use PINNED
 * host memory (pageable memory makes the copies synchronous and nothing overlaps).  host_src must stay unchanged
 * and host_dst must not be read until heat_drain — which waits for every submitted job — has returned; the
 * same host_dst may be given to several jobs (their downloads are ordered).  With nranks>1 every rank submits the
 * same sequence of jobs (each includes the collective ghost fill).
 * ref: the per-launch host<->device traffic of python/cuda/cuda.py:99 (drv.InOut / drv.In) — here once per
 * job instead of three times per time step, and hidden behind the previous job's march. */
int heat_submit(heat_handle* h, const void* host_src, int64_t src_pitch, int nt, void* host_dst);
int heat_drain(heat_handle* h);

/* Sum of this rank's own cells, deterministic (fixed reduction tree), as fp64.
 * ref: the commented-out debug check fortran/mpi+cuda/heat.F90:266-273. */
int heat_checksum(heat_handle* h, double* sum);

/* No counterpart in the reference.  The checksum for callers that must not allocate device memory:
 * programs that, like the reference's Fortran drivers, own every device buffer themselves and only
 * pass raw pointers to launchers; processes whose device heap another allocator has already
 * reserved; multi-stream time loops that cannot afford the device-wide synchronisation of
 * cudaMalloc/cudaFree.  Sums the caller's DEVICE array field_dev ([ny][pitch] elements of `dtype`,
 * nx cells used per row) into *sum_dev (DEVICE double), asynchronously on `stream` (a cudaStream_t,
 * 0 = default stream), on the CURRENT device.  The per-block partial sums live in an 8 KiB array built
 * into the library (held by every GPU that loads it until the process exits) instead of a
 * cudaMalloc'ed buffer; the call performs no cudaMalloc/cudaFree and uses no allocating routine.
 * nblocks = number of partial sums: 0 selects heat_checksum's own reduction tree (same bits);
 * more than 1024 (the array's capacity) is refused with HEAT_ERR_UNSUPPORTED.  The first call on each
 * device looks up the array's device address; make that call outside any stream capture.  All
 * calls share the one array (whatever the previous call left in it is overwritten before it is read),
 * so two calls that may overlap on the same GPU — different streams or threads — race: serialise them.
 * The looked-up address is remembered per device for the life of the process; it does not survive
 * cudaDeviceReset() (do not call this entry point on a device after resetting it). */
int heat_checksum_noalloc(const void* field_dev, int dtype, int nx, int ny, int64_t pitch,
                          int nblocks, double* sum_dev, uint64_t stream);

int heat_sync(heat_handle* h);
int heat_get_info(heat_handle* h, heat_info* info);
int heat_get_timing(heat_handle* h, heat_timing* t);
int heat_destroy(heat_handle* h);

/* Message for the calling thread's most recent error ("" if none).  Never NULL. */
const char* heat_last_error(void);
const char* heat_version(void);

/* Pure host helpers (usable without a GPU). */

/* Rows [row0, row0+nrows) of a ny-row field owned by `rank` of `nranks`; the remainder ny % nranks is
 * spread over the first ranks.  ref: nx = n/nblocks(1), fortran/mpi+cuda/heat.F90:92 (which drops it). */
int heat_slab_partition(int ny, int nranks, int rank, int* row0, int* nrows);

/* The tile of `rank` = ry * ranks_x + rx in a ranks_x x ranks_y decomposition: rows as heat_slab_partition(ny,
 * ranks_y, ry), columns as heat_slab_partition(nx, ranks_x, rx).  With ranks_x > 1 every `field` / `src` / `dst`
 * of heat_init, heat_set_field, heat_submit and heat_gather (global_root < 0) is this tile, dense [nrows][ncols].
 * ref: MPI_Dims_create / mpi_cart_create, fortran/mpi+cuda/heat.F90:87-97 (the reference reserves the second
 * dimension in ndims / nblocks, :28-29, and splits only one). */
int heat_tile_partition(int nx, int ny, int ranks_x, int ranks_y, int rank, int* row0, int* nrows, int* col0, int* ncols);

/* delta, dt, r with the reference's exact operation sequence.  ref: fortran/serial/heat.f90:15,17,59. */
int heat_derive_fortran(int n, double sigma, double nu, double dom_len, double* delta, double* dt, double* r);

/* Fill out[128] with a fresh ncclUniqueId (rank 0 calls it, the host layer broadcasts it).
 * ref: MPI_Init / MPI_Cart_create, fortran/mpi+cuda/heat.F90:60-62,97. */
int heat_nccl_unique_id(void* out128);

/* ---- One-shot, caller-owned buffers -------------------------------------------------------- */

/* Legacy launchers with the reference's exact signatures on ghosted, i-contiguous DEVICE arrays
 * ((nx+2ng) x (ny+2ng) reals, element (i,j) at (i-1+ng) + (nx+2ng)*(j-1+ng)); asynchronous on the
 * default stream, return nothing, caller synchronises — as the originals.
 * ref: fortran/hip/heat_kernel.cpp:48-61, 137-150, 78-90, 106-118; bind(c) fortran/hip/heat.F90:48-102.
 * `real` is double, or float when the caller compiles with -DSINGLE_PRECISION, as in the reference
 * (heat_kernel.cpp:5-9): the float flavour is exported as launch_*_sp and the macros below give it
 * the reference's names (Fortran: bind(c, name="launch_heat_eqn_sp")). */
void launch_heat_eqn(double* Td, double* Td_old, double r, int nx, int ny, int ng);
void launch_send(double* Td, double* Tds1, double* Tds2, int nx, int ny, int ng);
void launch_recv1(double* Td, double* Tdr, int nx, int ny, int ng);
void launch_recv2(double* Td, double* Tdr, int nx, int ny, int ng);
void launch_heat_eqn_sp(float* Td, float* Td_old, float r, int nx, int ny, int ng);
void launch_send_sp(float* Td, float* Tds1, float* Tds2, int nx, int ny, int ng);
void launch_recv1_sp(float* Td, float* Tdr, int nx, int ny, int ng);
void launch_recv2_sp(float* Td, float* Tdr, int nx, int ny, int ng);
#if defined(SINGLE_PRECISION) && !defined(HEAT_B200_NO_LEGACY_MACROS)
#define launch_heat_eqn launch_heat_eqn_sp
#define launch_send     launch_send_sp
#define launch_recv1    launch_recv1_sp
#define launch_recv2    launch_recv2_sp
#endif

/* Process-wide switch for launch_heat_eqn[_sp]: 0 (default) rounds every add/mul separately (the
 * kernel text read literally); 1 fuses c + r*(t - 4c) the way hipcc compiles heat_kernel.cpp:43 by
 * default (-ffp-contract=fast).  Returns the previous setting.  ref: fortran/hip/Makefile (hipcc flags). */
int heat_legacy_set_fmad(int on);

#ifdef __cplusplus
}
#endif
#endif /* HEAT_B200_H */
