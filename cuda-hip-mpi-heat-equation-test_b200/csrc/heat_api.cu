// heat_api.cu — the C ABI of libheat.so (include/heat_b200.h): handle, buffers, streams, kernel
// dispatch, row-slab halo exchange.  B200 (sm_100a) only; there is no CPU fallback anywhere: without
// a CUDA device heat_init fails with HEAT_ERR_CUDA.
#include "../../include/heat_b200.h"

#include <cuda_runtime.h>

#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <pthread.h>
#include <string>

#include "kernels_basic.cuh"
#include "kernels_tma.cuh"
#include "kernels_temporal.cuh"
#include "kernels_persist.cuh"
#include "nccl_dyn.h"

using namespace heat;

// ------------------------------------------------------------------------------------ errors
static thread_local std::string g_err;

static int fail(int code, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_err = buf;
    return code;
}

#define CUDA_TRY(expr)                                                                          \
    do {                                                                                        \
        cudaError_t e_ = (expr);                                                                \
        if (e_ != cudaSuccess)                                                                  \
            return fail(HEAT_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e_),  \
                        __FILE__, __LINE__);                                                    \
    } while (0)

#define NCCL_TRY(expr)                                                                          \
    do {                                                                                        \
        ncclResult_t r_ = (expr);                                                               \
        if (r_ != 0)                                                                            \
            return fail(HEAT_ERR_NCCL, "%s failed: %s (%s:%d)", #expr,                          \
                        nccl_api().GetErrorString(r_), __FILE__, __LINE__);                     \
    } while (0)

// ------------------------------------------------------------------------------------ handle
struct heat_handle {
    heat_config cfg{};
    int dev = 0;
    int nx = 0;                    // LOCAL width: what the kernels see (own columns + ghost columns); == gnx for row slabs
    int gnx = 0, ny = 0;           // global
    int row0 = 0, nrows = 0;       // owned global rows
    // 2-D (px x py) decomposition: rank = ry * px + rx.  Row slabs are px == 1.
    int px = 1, py = 1, rxi = 0, ryi = 0;
    int col0 = 0, ncols = 0;       // owned global columns
    int gl = 0, gr = 0;            // ghost columns held on the left / right (g next to a neighbour, 0 at the domain's edge)
    void* colbuf[4] = {nullptr, nullptr, nullptr, nullptr};   // packed column halos: send left, send right, recv left, recv right
    int g = 0;                     // ghost depth (0 when nranks == 1)
    int lrows = 0;                 // nrows + 2g
    long pitch = 0;                // elements
    size_t esize = 8;
    void* buf[2] = {nullptr, nullptr};
    bool owns = false;
    int cur = 0;
    int kernel = HEAT_KERNEL_SIMPLE;
    int tblock = 1;
    bool fma = false;
    bool vec_ok = false;           // buffers 16-byte aligned and pitch a multiple of the vector width
    int valid_ghost = 0;           // ghost rows of buf[cur] that hold current-time values
    bool comm_pending = false;     // an exchange is in flight on s_comm
    cudaStream_t s_compute = nullptr, s_comm = nullptr;
    bool owns_stream = false;
    cudaEvent_t ev_t0 = nullptr, ev_t1 = nullptr, ev_boundary = nullptr, ev_comm = nullptr,
                ev_h0 = nullptr, ev_h1 = nullptr;
    bool timed = false, halo_timed = false;
    ncclComm_t comm = nullptr;
    double* d_partials = nullptr;  // checksum scratch (lazily allocated)
    double* d_sum = nullptr;
    void* d_stage = nullptr;       // global-gather staging on the root (lazily allocated)
    size_t stage_bytes = 0;
    // job pipeline (heat_submit / heat_drain): device staging for the next job's input and the previous job's
    // result, moved over PCIe on their own streams while the current job marches
    void* pipe_in = nullptr;
    void* pipe_out = nullptr;
    cudaStream_t s_h2d = nullptr, s_d2h = nullptr;
    cudaEvent_t ev_up = nullptr, ev_in_free = nullptr, ev_res = nullptr, ev_out_free = nullptr;
    uint64_t pipe_jobs = 0;
    TmaState tma;                  // tensor maps for the TMA kernels (kernels_tma.cuh)
    TemporalState temporal;        // tensor maps for the fused K-step kernel (kernels_temporal.cuh)
    // peer-memory halo (HEAT_FLAG_PEER_HALO): neighbours' buffers and flag words mapped through CUDA IPC
    uint32_t* d_flags = nullptr;   // own: [0] ready_from_up [1] ready_from_down [2] done_from_up [3] done_from_down
    void* up_buf[2] = {nullptr, nullptr};
    void* down_buf[2] = {nullptr, nullptr};
    uint32_t* up_flags = nullptr;
    uint32_t* down_flags = nullptr;
    int up_nrows = 0;
    bool peer_ready = false, peer_pending_fill = false;
    uint32_t epoch = 0;
    int sm_count = 148;
    int stream_rows = 0, stream_threads = 0;   // tuning overrides (env)
    uint64_t steps_total = 0, launches = 0, exchanges = 0;
    PersistPlan persist;           // on-chip persistent stepping (kernels_persist.cuh): slabs, ghost depth, smem
    void* p_mailbox = nullptr;     // its neighbour mailbox (global scratch) and per-slab epoch flags
    unsigned* p_flags = nullptr;
    unsigned p_epoch = 0;
    int persist_min_nt = 8;        // shorter heat_step calls are not worth loading the field on chip
    void* mirror_base = nullptr;   // transient: second destination (peer ghost rows) of the next update launch
    int mirror_shift = 0;
    int kflags = 0;                // KF_* of the next launches: sweep direction (alternates per advance) + L2 policy
};

static const int kSumBlocks = 592;   // 148 SMs x 4; fixed so the checksum's reduction tree is fixed

static inline size_t row_bytes(const heat_handle* h) { return (size_t)h->pitch * h->esize; }
static inline char* row_ptr(const heat_handle* h, int which, int lrow) {
    return (char*)h->buf[which] + (size_t)lrow * row_bytes(h);
}
// local row index of a global row
static inline int lrow(const heat_handle* h, int grow) { return grow - h->row0 + h->g; }

// ------------------------------------------------------------------------------------ dispatch
template <typename T, int ORDER, bool FMA>
static cudaError_t launch_rows_t(heat_handle* h, const void* src, void* dst, int yb, int ye,
                                 cudaStream_t s) {
    if (yb >= ye || h->nx < 3) return cudaSuccess;
    const T rx = (T)h->cfg.rx, ry = (T)h->cfg.ry;
    const T* S = (const T*)src;
    T* D = (T*)dst;
    const int rows = ye - yb;
    h->launches++;
    if (h->kernel == HEAT_KERNEL_STREAM) {
        constexpr int V = Pack<T>::V;
        const int threads = h->stream_threads ? h->stream_threads : 256;
        const long nvec = h->vec_ok ? h->pitch / V : (long)h->nx - 2;      // lanes needed per row
        const int gx = (int)((nvec + threads - 1) / threads);
        int rpb = h->stream_rows;
        if (rpb <= 0) {
            const int target_blocks = h->sm_count * 16;
            const int ychunks = std::max(1, target_blocks / gx);
            rpb = std::max(32, (rows + ychunks - 1) / ychunks);
        }
        rpb = std::min(rpb, rows);
        const int gy = (rows + rpb - 1) / rpb;
        if (h->vec_ok)
            heat_stream_kernel<T, ORDER, FMA, 4><<<dim3(gx, gy), threads, 0, s>>>(
                S, D, h->nx, h->pitch, yb, ye, rpb, rx, ry);
        else        // pitch or base not 16-byte friendly: one cell per lane
            heat_stream1_kernel<T, ORDER, FMA, 4><<<dim3(gx, gy), threads, 0, s>>>(
                S, D, 1, h->nx - 1, h->pitch, yb, ye, rpb, rx, ry);
    } else if (h->kernel == HEAT_KERNEL_WTMA) {
        const int ybc0 = h->g - h->row0, ybc1 = h->ny - 1 - h->row0 + h->g;
        return launch_temporal_rows<T, ORDER, FMA, 1>(h->temporal, src == h->buf[0] ? 0 : 1, D, h->nx, h->pitch, yb, ye,
                                                     ybc0, ybc1, rx, ry, h->sm_count, s, h->kflags, (T*)h->mirror_base, h->mirror_shift);
    } else if (h->kernel == HEAT_KERNEL_TMA || h->kernel == HEAT_KERNEL_PERSIST) {
        return launch_tma_rows<T, ORDER, FMA>(h->tma, src == h->buf[0] ? 0 : 1, D, h->nx, h->pitch,
                                              yb, ye, rx, ry, h->sm_count, s, h->kflags, (T*)h->mirror_base, h->mirror_shift);
    } else {
        dim3 block(32, 8);
        dim3 grid((h->nx - 2 + 31) / 32, (rows + 7) / 8);
        heat_simple_kernel<T, ORDER, FMA><<<grid, block, 0, s>>>(S, D, h->nx, h->pitch, yb, ye, rx, ry);
    }
    return cudaGetLastError();
}

template <typename T, int ORDER, bool FMA>
static cudaError_t launch_resident_t(heat_handle* h, const void* src, void* dst, int nt, cudaStream_t s) {
    const size_t smem = 2 * (size_t)h->nx * h->ny * sizeof(T);
    auto k = heat_resident_kernel<T, ORDER, FMA>;
    cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    const int cells = h->nx * h->ny;
    const int threads = std::min(1024, std::max(32, ((cells + 31) / 32) * 32));
    h->launches++;
    k<<<1, threads, smem, s>>>((const T*)src, (T*)dst, h->nx, h->ny, h->pitch, nt, (T)h->cfg.rx, (T)h->cfg.ry);
    return cudaGetLastError();
}

template <typename T, int ORDER, bool FMA>
static cudaError_t launch_persist_t(heat_handle* h, const void* src, void* dst, int nt, cudaStream_t s) {
    auto k = heat_persist_kernel<T, ORDER, FMA>;
    // the attribute is set once per kernel and device, so ask for the most any plan can need
    int smem_cap = 0;
    cudaDeviceGetAttribute(&smem_cap, cudaDevAttrMaxSharedMemoryPerBlockOptin, h->dev);
    cudaError_t e = ensure_dynamic_smem<heat_persist_kernel<T, ORDER, FMA>>((size_t)smem_cap);
    if (e != cudaSuccess) return e;
    const T* S = (const T*)src;
    T* D = (T*)dst;
    int nx = h->nx, ny = h->ny, G = h->persist.G, spitch = h->persist.spitch, lmax = h->persist.lmax;
    long pitch = h->pitch;
    T rx = (T)h->cfg.rx, ry = (T)h->cfg.ry;
    T* mailbox = (T*)h->p_mailbox;
    unsigned* flags = h->p_flags;
    unsigned epoch0 = h->p_epoch;
    void* args[] = {&S, &D, &nx, &ny, &pitch, &nt, &G, &spitch, &lmax, &rx, &ry, &mailbox, &flags, &epoch0};
    h->launches++;
    h->p_epoch += (unsigned)((nt + G - 1) / G);      // one epoch per neighbour exchange, an upper bound is enough
    return cudaLaunchCooperativeKernel((const void*)k, dim3(h->persist.P), dim3(kPersistThreads), args, h->persist.smem, s);
}

template <typename T, int ORDER, bool FMA>
static cudaError_t launch_fused_t(heat_handle* h, const void* src, void* dst, int yb, int ye, cudaStream_t s) {
    if (yb >= ye || h->nx < 3) return cudaSuccess;
    const T rx = (T)h->cfg.rx, ry = (T)h->cfg.ry;
    const int which = src == h->buf[0] ? 0 : 1;
    const int ybc0 = h->g - h->row0, ybc1 = h->ny - 1 - h->row0 + h->g;   // local rows of global rows 0, ny-1
    h->launches++;
    switch (h->tblock) {
        case 2: return launch_temporal_rows<T, ORDER, FMA, 2>(h->temporal, which, (T*)dst, h->nx, h->pitch, yb, ye, ybc0, ybc1, rx, ry, h->sm_count, s, h->kflags, (T*)h->mirror_base, h->mirror_shift);
        case 4: return launch_temporal_rows<T, ORDER, FMA, 4>(h->temporal, which, (T*)dst, h->nx, h->pitch, yb, ye, ybc0, ybc1, rx, ry, h->sm_count, s, h->kflags, (T*)h->mirror_base, h->mirror_shift);
    }
    return cudaErrorInvalidValue;
}

#define HEAT_DISPATCH(h, FN, ...)                                                                   \
    [&]() -> cudaError_t {                                                                          \
        const int key_ = ((h)->cfg.dtype == HEAT_F32 ? 6 : 0) + (h)->cfg.op_order * 2 + ((h)->fma ? 1 : 0); \
        switch (key_) {                                                                             \
            case 0: return FN<double, 0, false>(__VA_ARGS__);                                       \
            case 1: return FN<double, 0, true>(__VA_ARGS__);                                        \
            case 2: return FN<double, 1, false>(__VA_ARGS__);                                       \
            case 3: return FN<double, 1, true>(__VA_ARGS__);                                        \
            case 4: return FN<double, 2, false>(__VA_ARGS__);                                       \
            case 5: return FN<double, 2, true>(__VA_ARGS__);                                        \
            case 6: return FN<float, 0, false>(__VA_ARGS__);                                        \
            case 7: return FN<float, 0, true>(__VA_ARGS__);                                         \
            case 8: return FN<float, 1, false>(__VA_ARGS__);                                        \
            case 9: return FN<float, 1, true>(__VA_ARGS__);                                         \
            case 10: return FN<float, 2, false>(__VA_ARGS__);                                       \
            case 11: return FN<float, 2, true>(__VA_ARGS__);                                        \
        }                                                                                           \
        return cudaErrorInvalidValue;                                                               \
    }()

// update global rows [glo, ghi] (inclusive) of buf[cur] -> buf[cur^1]
static cudaError_t update_rows(heat_handle* h, int glo, int ghi, cudaStream_t s) {
    glo = std::max(glo, 1);
    ghi = std::min(ghi, h->ny - 2);
    if (glo > ghi) return cudaSuccess;
    const void* src = h->buf[h->cur];
    void* dst = h->buf[h->cur ^ 1];
    const int yb = lrow(h, glo), ye = lrow(h, ghi) + 1;
    return HEAT_DISPATCH(h, launch_rows_t, h, src, dst, yb, ye, s);
}

// K fused steps on global rows [glo, ghi] (inclusive): buf[cur] -> buf[cur^1] holds time level +K there
static cudaError_t update_rows_fused(heat_handle* h, int glo, int ghi, cudaStream_t s) {
    glo = std::max(glo, 1);
    ghi = std::min(ghi, h->ny - 2);
    if (glo > ghi) return cudaSuccess;
    const void* src = h->buf[h->cur];
    void* dst = h->buf[h->cur ^ 1];
    const int yb = lrow(h, glo), ye = lrow(h, ghi) + 1;
    return HEAT_DISPATCH(h, launch_fused_t, h, src, dst, yb, ye, s);
}

// ------------------------------------------------------------------------------------ halo exchange
// Row slabs along the slow index make every halo a contiguous block of g rows, so NCCL reads the
// field rows and writes the neighbour's ghost rows directly: no pack/unpack kernels and no staging
// (contrast fortran/mpi+cuda/heat.F90:153-159,163-168,175-190).
//
// HEAT_FLAG_PEER_HALO replaces NCCL by the GPU's copy engines writing straight into the neighbour's ghost
// rows (its buffers are mapped here through CUDA IPC), so no SM runs communication code at all.  Ordering
// between the two processes uses one 32-bit flag word per direction in each GPU's memory and only
// stream-ordered operations — a 1-thread kernel stores the epoch into the neighbour's flag, the receiving
// stream blocks in cuStreamWaitValue32 on its own flag — never a kernel that spins on another GPU:
//   ready: "every earlier kernel of mine that touches these ghost rows is done — you may overwrite them"
//   done : "my rows have landed in your ghost rows"
__global__ void heat_signal_kernel(uint32_t* a, uint32_t* b, uint32_t v) {
    if (a) *reinterpret_cast<volatile uint32_t*>(a) = v;
    if (b) *reinterpret_cast<volatile uint32_t*>(b) = v;
    __threadfence_system();
}

// Signal and wait in ONE 1-thread kernel: store the epoch into the neighbours' flags, then spin (ld.acquire.sys) until
// both neighbours have stored it into ours.  It reacts within a memory round trip where a stream-ordered
// cuStreamWaitValue32 is polled by the front end; it occupies one CTA slot while it waits, and it cannot deadlock: what
// it waits for is stored by a kernel on ANOTHER GPU, whose progress never depends on this GPU's SMs.
__global__ void heat_signal_wait_kernel(uint32_t* sig_a, uint32_t* sig_b, const uint32_t* wait_a, const uint32_t* wait_b, uint32_t v) {
    if (sig_a) *reinterpret_cast<volatile uint32_t*>(sig_a) = v;
    if (sig_b) *reinterpret_cast<volatile uint32_t*>(sig_b) = v;
    __threadfence_system();
    for (const uint32_t* w : {wait_a, wait_b}) {
        if (!w) continue;
        uint32_t got;
        do {
            asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(got) : "l"(w) : "memory");
        } while ((int32_t)(got - v) < 0);
    }
}

// Opt-in (HEAT_PEER_WAIT=kernel).  Measured on 2 GPUs only (cfg2-weak 870.1 vs 865.0 GLUPS, nothing on the 32768^2
// strong-scaling run); the default stays the stream-ordered wait, which the 8-GPU parity and bench runs were taken with.
static bool peer_wait_in_kernel() {
    static const bool on = [] { const char* v = getenv("HEAT_PEER_WAIT"); return v && v[0] == 'k'; }();
    return on;
}

typedef CUresult (*PFN_streamWaitValue32)(CUstream, CUdeviceptr, cuuint32_t, unsigned int);
static PFN_streamWaitValue32 get_wait_value32() {
    static PFN_streamWaitValue32 fn = [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuStreamWaitValue32", &p, cudaEnableDefault, &q) != cudaSuccess ||
            q != cudaDriverEntryPointSuccess)
            p = nullptr;
        return (PFN_streamWaitValue32)p;
    }();
    return fn;
}

// phase 0: "ready" flags, phase 1: "done" flags.  Signal the neighbours, then block the stream until both have signalled us.
static int peer_handshake(heat_handle* h, int phase, uint32_t epoch, cudaStream_t s) {
    PFN_streamWaitValue32 wait32 = get_wait_value32();
    if (!wait32) return fail(HEAT_ERR_UNSUPPORTED, "peer halo: cuStreamWaitValue32 is not available from this driver");
    const bool up = h->cfg.rank > 0, down = h->cfg.rank + 1 < h->cfg.nranks;
    if (peer_wait_in_kernel()) {
        heat_signal_wait_kernel<<<1, 1, 0, s>>>(up ? h->up_flags + 2 * phase + 1 : nullptr, down ? h->down_flags + 2 * phase : nullptr,
                                                up ? h->d_flags + 2 * phase : nullptr, down ? h->d_flags + 2 * phase + 1 : nullptr, epoch);
        CUDA_TRY(cudaGetLastError());
        return HEAT_OK;
    }
    // I am the up neighbour's "down" and the down neighbour's "up"
    heat_signal_kernel<<<1, 1, 0, s>>>(up ? h->up_flags + 2 * phase + 1 : nullptr, down ? h->down_flags + 2 * phase : nullptr, epoch);
    CUDA_TRY(cudaGetLastError());
    for (int side = 0; side < 2; ++side) {
        if (!(side == 0 ? up : down)) continue;
        CUresult r = wait32((CUstream)s, (CUdeviceptr)(uintptr_t)(h->d_flags + 2 * phase + side), epoch, CU_STREAM_WAIT_VALUE_GEQ);
        if (r != CUDA_SUCCESS) return fail(HEAT_ERR_CUDA, "cuStreamWaitValue32 failed with CUresult %d", (int)r);
    }
    return HEAT_OK;
}

static int exchange_peer(heat_handle* h, int which, cudaStream_t s) {
    const uint32_t epoch = ++h->epoch;
    const bool up = h->cfg.rank > 0, down = h->cfg.rank + 1 < h->cfg.nranks;
    const size_t bytes = (size_t)h->g * row_bytes(h);
    CUDA_TRY(cudaEventRecord(h->ev_h0, s));
    int rc = peer_handshake(h, 0, epoch, s);
    if (rc) return rc;
    if (up)     // my first g rows -> the bottom ghost rows of the rank above
        CUDA_TRY(cudaMemcpyAsync((char*)h->up_buf[which] + (size_t)(h->g + h->up_nrows) * row_bytes(h),
                                 row_ptr(h, which, h->g), bytes, cudaMemcpyDeviceToDevice, s));
    if (down)   // my last g rows -> the top ghost rows of the rank below
        CUDA_TRY(cudaMemcpyAsync((char*)h->down_buf[which], row_ptr(h, which, h->nrows), bytes,
                                 cudaMemcpyDeviceToDevice, s));
    rc = peer_handshake(h, 1, epoch, s);
    if (rc) return rc;
    CUDA_TRY(cudaEventRecord(h->ev_h1, s));
    h->halo_timed = true;
    return HEAT_OK;
}

// Column halos of the 2-D decomposition are strided (g cells out of every row), so they travel packed:
// one kernel gathers both edge blocks into [rows][g] buffers (pack = 1) or scatters the received blocks into the
// ghost columns (pack = 0).  ref: the reference's own strided direction, pack -> sendrecv -> unpack,
// fortran/mpi+cuda/heat.F90:153-159,175-190.
template <typename T>
__global__ void __launch_bounds__(256)
heat_cols_kernel(T* __restrict__ field, long pitch, int rows, int g, int c_left, int c_right, T* __restrict__ bleft,
                 T* __restrict__ bright, int pack) {
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long)rows * g) return;
    const long r = i / g;
    const int j = (int)(i % g);
    T* row = field + r * pitch;
    if (pack) {
        if (bleft) bleft[i] = row[c_left + j];
        if (bright) bright[i] = row[c_right + j];
    } else {
        if (bleft) row[c_left + j] = bleft[i];
        if (bright) row[c_right + j] = bright[i];
    }
}

template <typename T>
static cudaError_t launch_cols(heat_handle* h, int which, int c_left, int c_right, void* bl, void* br, int pack, cudaStream_t s) {
    const long n = (long)h->lrows * h->g;
    heat_cols_kernel<T><<<(unsigned)((n + 255) / 256), 256, 0, s>>>((T*)h->buf[which], h->pitch, h->lrows, h->g, c_left, c_right,
                                                                    (T*)bl, (T*)br, pack);
    return cudaGetLastError();
}

static int exchange(heat_handle* h, int which, cudaStream_t s) {
    if (h->cfg.nranks <= 1) return HEAT_OK;
    h->exchanges++;
    if (h->cfg.flags & HEAT_FLAG_SKIP_EXCHANGE) return HEAT_OK;
    if (h->cfg.flags & HEAT_FLAG_PEER_HALO) {
        if (!h->peer_ready) return fail(HEAT_ERR_INVALID, "peer halo: call heat_peer_attach on every rank before stepping");
        return exchange_peer(h, which, s);
    }
    const NcclApi& N = nccl_api();
    const size_t count = (size_t)h->g * h->pitch;
    const int dt = h->cfg.dtype == HEAT_F32 ? ncclFloat32 : ncclFloat64;
    // rank = ry * px + rx: the rank above / below is px away (row slabs: px == 1)
    const int up = h->ryi > 0 ? h->cfg.rank - h->px : -1, down = h->ryi + 1 < h->py ? h->cfg.rank + h->px : -1;
    CUDA_TRY(cudaEventRecord(h->ev_h0, s));
    ncclResult_t gr = 0;                       // first failure inside a group; the group is closed regardless
    auto in_group = [&](ncclResult_t r) { if (gr == 0) gr = r; };
    if (h->py > 1) {
        // phase 1: g full local rows up and down (with px > 1 they include the ghost columns, whose corner blocks
        // are put right by phase 2)
        NCCL_TRY(N.GroupStart());
        if (up >= 0) {
            in_group(N.Send(row_ptr(h, which, h->g), count, dt, up, h->comm, s));
            in_group(N.Recv(row_ptr(h, which, 0), count, dt, up, h->comm, s));
        }
        if (down >= 0) {
            in_group(N.Send(row_ptr(h, which, h->nrows), count, dt, down, h->comm, s));
            in_group(N.Recv(row_ptr(h, which, h->g + h->nrows), count, dt, down, h->comm, s));
        }
        const ncclResult_t ge = N.GroupEnd();
        if (gr != 0) return fail(HEAT_ERR_NCCL, "halo exchange: ncclSend/ncclRecv failed: %s", N.GetErrorString(gr));
        if (ge != 0) return fail(HEAT_ERR_NCCL, "halo exchange: ncclGroupEnd failed: %s", N.GetErrorString(ge));
    }
    if (h->px > 1) {
        // phase 2: g own columns left and right over ALL local rows — ghost rows included, which is what carries the
        // diagonal neighbours' cells into the corner blocks (needed as soon as g > 1)
        const int left = h->rxi > 0 ? h->cfg.rank - 1 : -1, right = h->rxi + 1 < h->px ? h->cfg.rank + 1 : -1;
        const size_t ccount = (size_t)h->lrows * h->g;
        void* sl = left >= 0 ? h->colbuf[0] : nullptr;
        void* sr = right >= 0 ? h->colbuf[1] : nullptr;
        void* rl = left >= 0 ? h->colbuf[2] : nullptr;
        void* rr = right >= 0 ? h->colbuf[3] : nullptr;
        const int own_first = h->gl, own_last = h->gl + h->ncols - h->g;        // first / last g own columns
        CUDA_TRY(h->cfg.dtype == HEAT_F32 ? launch_cols<float>(h, which, own_first, own_last, sl, sr, 1, s)
                                          : launch_cols<double>(h, which, own_first, own_last, sl, sr, 1, s));
        NCCL_TRY(N.GroupStart());
        if (left >= 0) {
            in_group(N.Send(sl, ccount, dt, left, h->comm, s));
            in_group(N.Recv(rl, ccount, dt, left, h->comm, s));
        }
        if (right >= 0) {
            in_group(N.Send(sr, ccount, dt, right, h->comm, s));
            in_group(N.Recv(rr, ccount, dt, right, h->comm, s));
        }
        const ncclResult_t ge = N.GroupEnd();
        if (gr != 0) return fail(HEAT_ERR_NCCL, "halo exchange: ncclSend/ncclRecv (columns) failed: %s", N.GetErrorString(gr));
        if (ge != 0) return fail(HEAT_ERR_NCCL, "halo exchange: ncclGroupEnd (columns) failed: %s", N.GetErrorString(ge));
        CUDA_TRY(h->cfg.dtype == HEAT_F32 ? launch_cols<float>(h, which, 0, h->gl + h->ncols, rl, rr, 0, s)
                                          : launch_cols<double>(h, which, 0, h->gl + h->ncols, rl, rr, 0, s));
    }
    CUDA_TRY(cudaEventRecord(h->ev_h1, s));
    h->halo_timed = true;
    return HEAT_OK;
}

static int wait_comm(heat_handle* h) {
    if (h->comm_pending) {
        CUDA_TRY(cudaStreamWaitEvent(h->s_compute, h->ev_comm, 0));
        h->comm_pending = false;
    }
    return HEAT_OK;
}

// ------------------------------------------------------------------------------------ helpers
// rows x width elements between a pitched device buffer and a (possibly pitched) other side: ONE flat
// cudaMemcpyAsync when both sides are dense (pitch == width — every power-of-two benchmark size), the 2-D
// form only when there is padding to skip
static cudaError_t copy_rows(void* dst, size_t dpitch_bytes, const void* src, size_t spitch_bytes, size_t width_bytes,
                             size_t rows, cudaMemcpyKind kind, cudaStream_t s) {
    if (rows == 0 || width_bytes == 0) return cudaSuccess;
    if (dpitch_bytes == width_bytes && spitch_bytes == width_bytes)
        return cudaMemcpyAsync(dst, src, width_bytes * rows, kind, s);
    return cudaMemcpy2DAsync(dst, dpitch_bytes, src, spitch_bytes, width_bytes, rows, kind, s);
}

static int resolve_kernel(heat_handle* h) {
    const heat_config& c = h->cfg;
    const size_t V = 16 / h->esize;
    const bool aligned = (h->pitch % (long)V == 0) && ((uintptr_t)h->buf[0] % 16 == 0) &&
                         ((uintptr_t)h->buf[1] % 16 == 0);
    const bool fits_resident = c.nranks == 1 && h->nx >= 3 && h->ny >= 3 &&
                               2 * (size_t)h->nx * h->ny * h->esize <= 200 * 1024;
    h->vec_ok = aligned;
    int smem_cap = 0;
    cudaDeviceGetAttribute(&smem_cap, cudaDevAttrMaxSharedMemoryPerBlockOptin, h->dev);
    if (c.nranks == 1 && aligned && h->tblock == 1)
        h->persist = persist_plan(h->nx, h->ny, h->esize, h->sm_count, (size_t)std::max(0, smem_cap - 1024));
    const bool fits_persist = h->persist.ok;
    int k = c.kernel;
    if (k == HEAT_KERNEL_AUTO) {
        if (fits_resident) k = HEAT_KERNEL_RESIDENT;
        // the whole field fits in the SMs' shared memory: one launch for all steps.  Measured (tools/bench_small.py)
        // it beats per-step launches while a ghost depth >= 3 fits (fp64 up to 1024^2), not beyond (1100^2: G = 2)
        else if (fits_persist && h->persist.G >= 3) k = HEAT_KERNEL_PERSIST;
        else if (aligned) k = HEAT_KERNEL_TMA;       // measured fastest for fp64 and fp32 at 4096^2 and 16384^2 (profiles/)
        else k = HEAT_KERNEL_STREAM;                 // its one-cell-per-lane variant needs no alignment
    }
    if (k == HEAT_KERNEL_RESIDENT && !fits_resident)
        return fail(HEAT_ERR_UNSUPPORTED, "RESIDENT kernel needs nranks == 1 and 2*nx*ny*sizeof(T) <= 200 KiB");
    if (k == HEAT_KERNEL_PERSIST && !fits_persist)
        return fail(HEAT_ERR_UNSUPPORTED, "PERSIST kernel needs nranks == 1, tblock == 1, 16-byte friendly buffers and a field whose row slabs "
                                          "(two copies, with ghost rows) fit in the SMs' shared memory");
    if ((k == HEAT_KERNEL_TMA || k == HEAT_KERNEL_WTMA) && !aligned)
        return fail(HEAT_ERR_UNSUPPORTED, "TMA kernels need 16-byte aligned buffers and a pitch that is a multiple of %d elements", (int)V);
    if (k < HEAT_KERNEL_SIMPLE || k > HEAT_KERNEL_PERSIST) return fail(HEAT_ERR_INVALID, "unknown kernel %d", k);
    h->kernel = k;
    return HEAT_OK;
}

static int env_int(const char* name, int dflt) {
    const char* v = getenv(name);
    return (v && *v) ? atoi(v) : dflt;
}

// ------------------------------------------------------------------------------------ C ABI
extern "C" {

const char* heat_last_error(void) { return g_err.c_str(); }
const char* heat_version(void) { return "heat_b200 0.2 (sm_100a)"; }

int heat_slab_partition(int ny, int nranks, int rank, int* row0, int* nrows) {
    if (ny < 1 || nranks < 1 || rank < 0 || rank >= nranks || !row0 || !nrows)
        return fail(HEAT_ERR_INVALID, "heat_slab_partition: bad arguments (ny=%d nranks=%d rank=%d)", ny, nranks, rank);
    const int base = ny / nranks, rem = ny % nranks;
    *nrows = base + (rank < rem ? 1 : 0);
    *row0 = rank * base + std::min(rank, rem);
    return HEAT_OK;
}

int heat_tile_partition(int nx, int ny, int ranks_x, int ranks_y, int rank, int* row0, int* nrows, int* col0, int* ncols) {
    if (ranks_x < 1 || ranks_y < 1 || rank < 0 || rank >= ranks_x * ranks_y || !row0 || !nrows || !col0 || !ncols)
        return fail(HEAT_ERR_INVALID, "heat_tile_partition: bad arguments (%d x %d ranks, rank %d)", ranks_x, ranks_y, rank);
    int rc = heat_slab_partition(ny, ranks_y, rank / ranks_x, row0, nrows);
    if (rc) return rc;
    return heat_slab_partition(nx, ranks_x, rank % ranks_x, col0, ncols);
}

int heat_derive_fortran(int n, double sigma, double nu, double dom_len, double* delta, double* dt, double* r) {
    if (n < 2 || !delta || !dt || !r) return fail(HEAT_ERR_INVALID, "heat_derive_fortran: bad arguments");
    // ref: fortran/serial/heat.f90:15,17,59 — real(n-1) is a default (fp32) real; delta**2 = delta*delta
    volatile double d = dom_len / (double)(float)(n - 1);
    volatile double d2 = d * d;
    volatile double sd = sigma * d2;
    volatile double t = sd / nu;
    volatile double nt = nu * t;
    *delta = d;
    *dt = t;
    *r = nt / d2;
    return HEAT_OK;
}

int heat_nccl_unique_id(void* out128) {
    if (!out128) return fail(HEAT_ERR_INVALID, "heat_nccl_unique_id: null output");
    const NcclApi& N = nccl_api();
    if (!N.ok) {
        const char* why = dlerror();             // reading it clears it: read once
        return fail(HEAT_ERR_NCCL, "NCCL not available: %s", why ? why : "libnccl.so.2 not found");
    }
    ncclUniqueId id;
    NCCL_TRY(N.GetUniqueId(&id));
    memcpy(out128, &id, sizeof id);
    return HEAT_OK;
}

int heat_destroy(heat_handle* h) {
    if (!h) return HEAT_OK;
    cudaSetDevice(h->dev);
    if (h->s_compute) cudaStreamSynchronize(h->s_compute);
    if (h->s_comm) cudaStreamSynchronize(h->s_comm);
    if (h->comm) nccl_api().CommDestroy(h->comm);
    for (void* p : {h->up_buf[0], h->up_buf[1], h->down_buf[0], h->down_buf[1], (void*)h->up_flags, (void*)h->down_flags})
        if (p) cudaIpcCloseMemHandle(p);
    cudaFree(h->d_flags);
    cudaFree(h->p_mailbox);
    cudaFree(h->p_flags);
    for (void* p : h->colbuf) cudaFree(p);
    if (h->owns) { cudaFree(h->buf[0]); cudaFree(h->buf[1]); }
    cudaFree(h->d_partials);
    cudaFree(h->d_sum);
    cudaFree(h->d_stage);
    for (cudaStream_t st : {h->s_h2d, h->s_d2h})
        if (st) { cudaStreamSynchronize(st); cudaStreamDestroy(st); }
    cudaFree(h->pipe_in);
    cudaFree(h->pipe_out);
    for (cudaEvent_t e : {h->ev_t0, h->ev_t1, h->ev_boundary, h->ev_comm, h->ev_h0, h->ev_h1, h->ev_up, h->ev_in_free,
                          h->ev_res, h->ev_out_free})
        if (e) cudaEventDestroy(e);
    if (h->s_comm) cudaStreamDestroy(h->s_comm);
    if (h->owns_stream && h->s_compute) cudaStreamDestroy(h->s_compute);
    delete h;
    return HEAT_OK;
}

// stream-ordered on s_compute, no host synchronisation: `src` becomes this rank's rows of both buffers
static int install_field(heat_handle* h, const void* src, int64_t src_pitch, int src_is_device) {
    int rc = wait_comm(h);
    if (rc) return rc;
    const cudaMemcpyKind kind = src_is_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice;
    // one trip over PCIe into buffer 0, ONE ghost exchange there (collective), then a device-side copy of
    // own rows and ghost rows into buffer 1 (both buffers need the Dirichlet ring; the ghosts come for free)
    CUDA_TRY(copy_rows(row_ptr(h, 0, h->g) + (size_t)h->gl * h->esize, row_bytes(h), src, (size_t)src_pitch * h->esize,
                       (size_t)h->ncols * h->esize, h->nrows, kind, h->s_compute));
    h->cur = 0;
    h->valid_ghost = 0;
    if (h->cfg.nranks > 1 && (h->cfg.flags & HEAT_FLAG_PEER_HALO) && !h->peer_ready) {
        h->peer_pending_fill = true;               // ghosts are filled by the first heat_step after heat_peer_attach
    } else if (h->cfg.nranks > 1) {
        rc = exchange(h, 0, h->s_compute);
        if (rc) return rc;
        h->valid_ghost = h->g;
    }
    CUDA_TRY(cudaMemcpyAsync(h->buf[1], h->buf[0], (size_t)h->lrows * row_bytes(h), cudaMemcpyDeviceToDevice, h->s_compute));
    return HEAT_OK;
}

int heat_set_field(heat_handle* h, const void* src, int64_t src_pitch, int src_is_device) {
    if (!h || !src) return fail(HEAT_ERR_INVALID, "heat_set_field: null argument");
    if (src_pitch <= 0) src_pitch = h->ncols;
    if (src_pitch < h->ncols) return fail(HEAT_ERR_INVALID, "heat_set_field: src_pitch %lld < own columns %d", (long long)src_pitch, h->ncols);
    CUDA_TRY(cudaSetDevice(h->dev));
    int rc = install_field(h, src, src_pitch, src_is_device);
    if (rc) return rc;
    CUDA_TRY(cudaStreamSynchronize(h->s_compute));
    return HEAT_OK;
}

int heat_init(heat_handle** out, const heat_config* cfg, const void* field) {
    if (!out || !cfg) return fail(HEAT_ERR_INVALID, "heat_init: null argument");
    *out = nullptr;
    if (cfg->struct_size != sizeof(heat_config))
        return fail(HEAT_ERR_INVALID, "heat_init: heat_config.struct_size is %u, this library expects %zu",
                    cfg->struct_size, sizeof(heat_config));
    if (cfg->nx < 1 || cfg->ny < 1) return fail(HEAT_ERR_INVALID, "heat_init: nx, ny must be >= 1 (got %d x %d)", cfg->nx, cfg->ny);
    if (cfg->dtype != HEAT_F64 && cfg->dtype != HEAT_F32) return fail(HEAT_ERR_INVALID, "heat_init: bad dtype %d", cfg->dtype);
    if (cfg->op_order < 0 || cfg->op_order > 2) return fail(HEAT_ERR_INVALID, "heat_init: bad op_order %d", cfg->op_order);
    if (cfg->nranks < 1 || cfg->rank < 0 || cfg->rank >= cfg->nranks)
        return fail(HEAT_ERR_INVALID, "heat_init: bad rank %d of %d", cfg->rank, cfg->nranks);
    if (cfg->tblock != 0 && cfg->tblock != 1 && cfg->tblock != 2 && cfg->tblock != 4)
        return fail(HEAT_ERR_INVALID, "heat_init: tblock must be 0 (= 1), 1, 2 or 4 (got %d)", cfg->tblock);
    if (cfg->nranks > 1 && !cfg->nccl_id && !(cfg->flags & (HEAT_FLAG_SKIP_EXCHANGE | HEAT_FLAG_PEER_HALO)))
        return fail(HEAT_ERR_INVALID, "heat_init: nranks > 1 needs cfg.nccl_id (or HEAT_FLAG_PEER_HALO)");
    if ((cfg->ext_buf[0] == nullptr) != (cfg->ext_buf[1] == nullptr))
        return fail(HEAT_ERR_INVALID, "heat_init: supply both ext_buf pointers or neither");

    int ndev = 0;
    cudaError_t ce = cudaGetDeviceCount(&ndev);
    if (ce != cudaSuccess || ndev == 0)
        return fail(HEAT_ERR_CUDA, "heat_init: no CUDA device (%s); libheat has no CPU path", cudaGetErrorString(ce));
    if (cfg->device < 0 || cfg->device >= ndev) return fail(HEAT_ERR_INVALID, "heat_init: device %d out of range (0..%d)", cfg->device, ndev - 1);

    heat_handle* h = new (std::nothrow) heat_handle;
    if (!h) return fail(HEAT_ERR_NOMEM, "heat_init: out of host memory");
    h->cfg = *cfg;
    h->dev = cfg->device;
    h->gnx = cfg->nx;
    h->ny = cfg->ny;
    h->esize = cfg->dtype == HEAT_F32 ? 4 : 8;
    h->fma = (cfg->flags & HEAT_FLAG_FMAD) != 0;
    h->tblock = cfg->tblock > 0 ? cfg->tblock : 1;
    h->kflags = sweep_policy_flags();
    h->stream_rows = env_int("HEAT_STREAM_ROWS", 0);
    h->stream_threads = env_int("HEAT_STREAM_THREADS", 0);
    h->px = cfg->ranks_x > 0 ? cfg->ranks_x : 1;
    if (cfg->nranks % h->px != 0) {
        const int px = h->px;
        delete h;
        return fail(HEAT_ERR_INVALID, "heat_init: ranks_x %d does not divide nranks %d", px, cfg->nranks);
    }
    h->py = cfg->nranks / h->px;
    h->rxi = cfg->rank % h->px;
    h->ryi = cfg->rank / h->px;
    int rc = heat_tile_partition(h->gnx, h->ny, h->px, h->py, cfg->rank, &h->row0, &h->nrows, &h->col0, &h->ncols);
    if (rc) { delete h; return rc; }
    h->g = cfg->nranks > 1 ? (cfg->ghost > 0 ? cfg->ghost : 1) : 0;
    // judged on the THINNEST tile of the partition, not on this rank's: every rank then takes the same
    // decision and none is left waiting inside the collective calls below
    if ((h->py > 1 && h->ny / h->py < h->g) || (h->px > 1 && h->gnx / h->px < h->g)) {
        const int px = h->px, py = h->py, g = h->g;
        delete h;
        return fail(HEAT_ERR_INVALID, "heat_init: the thinnest tile (%d rows x %d columns: %d x %d over %d x %d ranks) is thinner than ghost depth %d",
                    cfg->ny / py, cfg->nx / px, cfg->ny, cfg->nx, py, px, g);
    }
    if (h->px > 1 && (cfg->flags & HEAT_FLAG_PEER_HALO)) {
        delete h;
        return fail(HEAT_ERR_UNSUPPORTED, "heat_init: HEAT_FLAG_PEER_HALO serves row slabs only (ranks_x <= 1); the 2-D decomposition exchanges through NCCL");
    }
    if (h->px > 1 && cfg->ext_buf[0]) {
        delete h;
        return fail(HEAT_ERR_UNSUPPORTED, "heat_init: caller-owned buffers are not supported with ranks_x > 1");
    }
    h->gl = (h->px > 1 && h->rxi > 0) ? h->g : 0;
    h->gr = (h->px > 1 && h->rxi + 1 < h->px) ? h->g : 0;
    h->nx = h->ncols + h->gl + h->gr;          // (px x 1 keeps its g ghost rows allocated; they are simply never exchanged)
    if (h->g % h->tblock != 0 && cfg->nranks > 1) {
        const int g = h->g, k = h->tblock;
        delete h;
        return fail(HEAT_ERR_INVALID, "heat_init: ghost depth %d must be a multiple of tblock %d", g, k);
    }
    h->lrows = h->nrows + 2 * h->g;

#define INIT_TRY(expr)                                                                          \
    do {                                                                                        \
        cudaError_t e_ = (expr);                                                                \
        if (e_ != cudaSuccess) {                                                                \
            int c_ = fail(e_ == cudaErrorMemoryAllocation ? HEAT_ERR_NOMEM : HEAT_ERR_CUDA,     \
                          "%s failed: %s", #expr, cudaGetErrorString(e_));                      \
            std::string keep = g_err; heat_destroy(h); g_err = keep;                            \
            return c_;                                                                          \
        }                                                                                       \
    } while (0)

    INIT_TRY(cudaSetDevice(h->dev));
    INIT_TRY(cudaDeviceGetAttribute(&h->sm_count, cudaDevAttrMultiProcessorCount, h->dev));
    if (cfg->ext_buf[0]) {
        h->buf[0] = cfg->ext_buf[0];
        h->buf[1] = cfg->ext_buf[1];
        h->pitch = cfg->ext_pitch > 0 ? cfg->ext_pitch : h->nx;
        if (h->pitch < h->nx) { heat_destroy(h); return fail(HEAT_ERR_INVALID, "heat_init: ext_pitch < nx"); }
        h->owns = false;
    } else {
        h->pitch = ((long)h->nx + 15) / 16 * 16;     // rows start on 128-byte (f64) / 64-byte (f32) boundaries
        const size_t bytes = (size_t)h->lrows * row_bytes(h);
        h->owns = true;
        INIT_TRY(cudaMalloc(&h->buf[0], bytes));
        INIT_TRY(cudaMalloc(&h->buf[1], bytes));
    }
    if (cfg->stream) {
        h->s_compute = (cudaStream_t)(uintptr_t)cfg->stream;
    } else {
        INIT_TRY(cudaStreamCreateWithFlags(&h->s_compute, cudaStreamNonBlocking));
        h->owns_stream = true;
    }
    int prio_lo = 0, prio_hi = 0;
    INIT_TRY(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
    INIT_TRY(cudaStreamCreateWithPriority(&h->s_comm, cudaStreamNonBlocking, prio_hi));
    for (cudaEvent_t* e : {&h->ev_t0, &h->ev_t1, &h->ev_h0, &h->ev_h1}) INIT_TRY(cudaEventCreate(e));
    for (cudaEvent_t* e : {&h->ev_boundary, &h->ev_comm}) INIT_TRY(cudaEventCreateWithFlags(e, cudaEventDisableTiming));
    if (h->owns) {
        // zero the pitch padding and ghost rows ON THE COMPUTE STREAM: s_compute is non-blocking, so a
        // cudaMemset on the legacy default stream would not be ordered before the first field upload
        const size_t bytes = (size_t)h->lrows * row_bytes(h);
        INIT_TRY(cudaMemsetAsync(h->buf[0], 0, bytes, h->s_compute));
        INIT_TRY(cudaMemsetAsync(h->buf[1], 0, bytes, h->s_compute));
    }

    if (h->px > 1)
        for (void*& p : h->colbuf) INIT_TRY(cudaMalloc(&p, (size_t)h->lrows * h->g * h->esize));
    rc = resolve_kernel(h);
    if (rc) { std::string keep = g_err; heat_destroy(h); g_err = keep; return rc; }
    if (h->kernel == HEAT_KERNEL_PERSIST) {
        const size_t mb = persist_mailbox_bytes(h->nx, h->esize, h->sm_count);
        INIT_TRY(cudaMalloc(&h->p_mailbox, mb));
        INIT_TRY(cudaMalloc((void**)&h->p_flags, (size_t)h->sm_count * sizeof(unsigned)));
        INIT_TRY(cudaMemsetAsync(h->p_flags, 0, (size_t)h->sm_count * sizeof(unsigned), h->s_compute));
        h->persist_min_nt = env_int("HEAT_PERSIST_MIN_NT", 8);
    }
    if (h->kernel == HEAT_KERNEL_TMA || h->kernel == HEAT_KERNEL_PERSIST) {
        rc = tma_setup(h->tma, h->buf, cfg->dtype == HEAT_F32, h->nx, h->lrows, h->pitch, g_err);
        if (rc) { std::string keep = g_err; heat_destroy(h); g_err = keep; return rc; }
    }

    if (h->tblock > 1 || h->kernel == HEAT_KERNEL_WTMA) {
        if (h->kernel == HEAT_KERNEL_SIMPLE || h->kernel == HEAT_KERNEL_RESIDENT || !h->vec_ok) {
            h->tblock = 1;             // unaligned caller buffers / tiny fields: fall back to single steps
        } else {
            rc = temporal_setup(h->temporal, h->buf, cfg->dtype == HEAT_F32, h->lrows, h->pitch, h->tblock, g_err);
            if (rc) { std::string keep = g_err; heat_destroy(h); g_err = keep; return rc; }
        }
    }

    if (cfg->nranks > 1 && (cfg->flags & HEAT_FLAG_PEER_HALO)) {
        if (!h->owns) { heat_destroy(h); return fail(HEAT_ERR_UNSUPPORTED, "heat_init: peer halo needs library-owned buffers (CUDA IPC handles)"); }
        INIT_TRY(cudaMalloc(&h->d_flags, 4 * sizeof(uint32_t)));
        INIT_TRY(cudaMemsetAsync(h->d_flags, 0, 4 * sizeof(uint32_t), h->s_compute));
        INIT_TRY(cudaStreamSynchronize(h->s_compute));
    }
    if (cfg->nranks > 1 && !(cfg->flags & HEAT_FLAG_SKIP_EXCHANGE) && cfg->nccl_id) {
        const NcclApi& N = nccl_api();
        if (!N.ok) { heat_destroy(h); return fail(HEAT_ERR_NCCL, "heat_init: NCCL not available (libnccl.so.2 could not be loaded)"); }
        ncclUniqueId id;
        memcpy(&id, cfg->nccl_id, sizeof id);
        ncclResult_t r = N.CommInitRank(&h->comm, cfg->nranks, id, cfg->rank);
        if (r != 0) {
            int c = fail(HEAT_ERR_NCCL, "ncclCommInitRank failed: %s", N.GetErrorString(r));
            std::string keep = g_err; h->comm = nullptr; heat_destroy(h); g_err = keep;
            return c;
        }
    }
    if (field) {
        rc = heat_set_field(h, field, h->ncols, 0);
        if (rc) { std::string keep = g_err; heat_destroy(h); g_err = keep; return rc; }
    }
    *out = h;
    return HEAT_OK;
#undef INIT_TRY
}

struct PeerBlob {                      // what one rank tells its neighbours (fits HEAT_PEER_BLOB_BYTES)
    cudaIpcMemHandle_t buf[2];
    cudaIpcMemHandle_t flags;
    int32_t rank, nrows, ghost, dtype;
    int64_t pitch;
};
static_assert(sizeof(PeerBlob) <= 256, "PeerBlob must fit HEAT_PEER_BLOB_BYTES");

int heat_peer_export(heat_handle* h, void* blob) {
    if (!h || !blob) return fail(HEAT_ERR_INVALID, "heat_peer_export: null argument");
    if (!(h->cfg.flags & HEAT_FLAG_PEER_HALO) || !h->d_flags)
        return fail(HEAT_ERR_INVALID, "heat_peer_export: handle was not created with HEAT_FLAG_PEER_HALO and nranks > 1");
    CUDA_TRY(cudaSetDevice(h->dev));
    PeerBlob b;
    memset(&b, 0, sizeof b);
    CUDA_TRY(cudaIpcGetMemHandle(&b.buf[0], h->buf[0]));
    CUDA_TRY(cudaIpcGetMemHandle(&b.buf[1], h->buf[1]));
    CUDA_TRY(cudaIpcGetMemHandle(&b.flags, h->d_flags));
    b.rank = h->cfg.rank; b.nrows = h->nrows; b.ghost = h->g; b.dtype = h->cfg.dtype; b.pitch = h->pitch;
    memset(blob, 0, 256);
    memcpy(blob, &b, sizeof b);
    return HEAT_OK;
}

int heat_peer_attach(heat_handle* h, const void* up_blob, const void* down_blob) {
    if (!h) return fail(HEAT_ERR_INVALID, "heat_peer_attach: null handle");
    if (!(h->cfg.flags & HEAT_FLAG_PEER_HALO) || !h->d_flags)
        return fail(HEAT_ERR_INVALID, "heat_peer_attach: handle was not created with HEAT_FLAG_PEER_HALO and nranks > 1");
    const bool up = h->cfg.rank > 0, down = h->cfg.rank + 1 < h->cfg.nranks;
    if ((up && !up_blob) || (down && !down_blob)) return fail(HEAT_ERR_INVALID, "heat_peer_attach: missing neighbour blob");
    CUDA_TRY(cudaSetDevice(h->dev));
    auto open_one = [&](const void* blob, int want_rank, void* bufs[2], uint32_t** flags, int* nrows) -> int {
        PeerBlob b;
        memcpy(&b, blob, sizeof b);
        if (b.rank != want_rank || b.ghost != h->g || b.pitch != h->pitch || b.dtype != h->cfg.dtype)
            return fail(HEAT_ERR_INVALID, "heat_peer_attach: blob of rank %d (ghost %d, pitch %lld) does not match neighbour %d (ghost %d, pitch %ld)",
                        b.rank, b.ghost, (long long)b.pitch, want_rank, h->g, h->pitch);
        for (int i = 0; i < 2; ++i) CUDA_TRY(cudaIpcOpenMemHandle(&bufs[i], b.buf[i], cudaIpcMemLazyEnablePeerAccess));
        void* f = nullptr;
        CUDA_TRY(cudaIpcOpenMemHandle(&f, b.flags, cudaIpcMemLazyEnablePeerAccess));
        *flags = (uint32_t*)f;
        if (nrows) *nrows = b.nrows;
        return HEAT_OK;
    };
    int rc;
    if (up && (rc = open_one(up_blob, h->cfg.rank - 1, h->up_buf, &h->up_flags, &h->up_nrows))) return rc;
    if (down && (rc = open_one(down_blob, h->cfg.rank + 1, h->down_buf, &h->down_flags, nullptr))) return rc;
    h->peer_ready = true;
    // No cross-GPU wait happens here, so a rank whose attach failed cannot leave its neighbours blocked: the
    // ghosts are filled by the first heat_step / heat_halo_exchange (valid_ghost == 0 forces an exchange).
    h->valid_ghost = 0;
    h->peer_pending_fill = false;
    return HEAT_OK;
}

int heat_halo_exchange(heat_handle* h) {
    if (!h) return fail(HEAT_ERR_INVALID, "heat_halo_exchange: null handle");
    if (h->cfg.nranks <= 1) return HEAT_OK;
    CUDA_TRY(cudaSetDevice(h->dev));
    int rc = wait_comm(h);
    if (rc) return rc;
    rc = exchange(h, h->cur, h->s_compute);
    if (rc) return rc;
    h->valid_ghost = h->g;
    return HEAT_OK;
}

int heat_step(heat_handle* h, int nt) {
    if (!h) return fail(HEAT_ERR_INVALID, "heat_step: null handle");
    if (nt < 0) return fail(HEAT_ERR_INVALID, "heat_step: nt must be >= 0 (got %d)", nt);
    CUDA_TRY(cudaSetDevice(h->dev));
    CUDA_TRY(cudaEventRecord(h->ev_t0, h->s_compute));
    const bool multi = h->cfg.nranks > 1;
    if (nt > 0 && h->nx >= 3 && h->ny >= 3) {
        if (h->kernel == HEAT_KERNEL_RESIDENT) {
            CUDA_TRY(HEAT_DISPATCH(h, launch_resident_t, h, h->buf[h->cur], h->buf[h->cur ^ 1], nt, h->s_compute));
            h->cur ^= 1;
        } else if (h->kernel == HEAT_KERNEL_PERSIST && nt >= h->persist_min_nt) {
            CUDA_TRY(HEAT_DISPATCH(h, launch_persist_t, h, h->buf[h->cur], h->buf[h->cur ^ 1], nt, h->s_compute));
            h->cur ^= 1;
        } else if (!multi) {
            int t = 0;
            for (; h->tblock > 1 && t + h->tblock <= nt; t += h->tblock) {
                CUDA_TRY(update_rows_fused(h, 1, h->ny - 2, h->s_compute));
                h->cur ^= 1;
                if (serpentine_enabled()) h->kflags ^= KF_REVERSE;
            }
            for (; t < nt; ++t) {
                CUDA_TRY(update_rows(h, 1, h->ny - 2, h->s_compute));
                h->cur ^= 1;
                if (serpentine_enabled()) h->kflags ^= KF_REVERSE;
            }
        } else {
            const bool overlap = (h->cfg.flags & HEAT_FLAG_OVERLAP) && h->nrows >= 2 * h->g && h->px == 1;
            const int own_lo = h->row0, own_hi = h->row0 + h->nrows - 1;
            for (int t = 0; t < nt;) {
                int rc = wait_comm(h);
                if (rc) return rc;
                if (h->valid_ghost == 0) {      // only after SKIP_EXCHANGE or an external reset
                    rc = exchange(h, h->cur, h->s_compute);
                    if (rc) return rc;
                    h->valid_ghost = h->g;
                }
                // advance K fused steps while the ghost rows allow it, single steps otherwise
                const int adv = (h->tblock > 1 && h->valid_ghost >= h->tblock && nt - t >= h->tblock) ? h->tblock : 1;
                auto upd = [&](int lo, int hi) { return adv > 1 ? update_rows_fused(h, lo, hi, h->s_compute) : update_rows(h, lo, hi, h->s_compute); };
                const int e = h->valid_ghost - adv;   // ghost rows still valid after this advance
                if (e == 0 && overlap) {
                    // The two g-row edge strips and their exchange run on the high-priority stream s_comm,
                    // concurrently with the interior update on s_compute: all three read buf[cur] and write
                    // disjoint rows of buf[cur^1].  s_comm first waits for everything already queued on
                    // s_compute (the previous steps, which wrote buf[cur]).
                    auto upd_on = [&](int lo, int hi, cudaStream_t st) { return adv > 1 ? update_rows_fused(h, lo, hi, st) : update_rows(h, lo, hi, st); };
                    CUDA_TRY(cudaEventRecord(h->ev_boundary, h->s_compute));
                    CUDA_TRY(cudaStreamWaitEvent(h->s_comm, h->ev_boundary, 0));
                    const bool fused_peer = (h->cfg.flags & HEAT_FLAG_PEER_HALO) && h->peer_ready &&
                                            !(h->cfg.flags & HEAT_FLAG_SKIP_EXCHANGE) &&
                                            (adv > 1 || h->kernel == HEAT_KERNEL_TMA || h->kernel == HEAT_KERNEL_WTMA);
                    if (fused_peer) {
                        // ONE kernel computes an edge strip and pushes it over NVLink: its stores go to the own
                        // buffer and, as peer stores, to the neighbour's ghost rows.  Handshakes around it:
                        // "ready" before (the neighbour is done with those ghost rows), "done" after.
                        const uint32_t epoch = ++h->epoch;
                        h->exchanges++;
                        CUDA_TRY(cudaEventRecord(h->ev_h0, h->s_comm));
                        rc = peer_handshake(h, 0, epoch, h->s_comm);
                        if (rc) return rc;
                        const int dstb = h->cur ^ 1;
                        if (h->cfg.rank > 0) { h->mirror_base = h->up_buf[dstb]; h->mirror_shift = h->up_nrows; }
                        cudaError_t ce1 = upd_on(own_lo, own_lo + h->g - 1, h->s_comm);
                        h->mirror_base = nullptr;
                        CUDA_TRY(ce1);
                        if (h->cfg.rank + 1 < h->cfg.nranks) { h->mirror_base = h->down_buf[dstb]; h->mirror_shift = -h->nrows; }
                        cudaError_t ce2 = upd_on(own_hi - h->g + 1, own_hi, h->s_comm);
                        h->mirror_base = nullptr;
                        CUDA_TRY(ce2);
                        rc = peer_handshake(h, 1, epoch, h->s_comm);
                        if (rc) return rc;
                        CUDA_TRY(cudaEventRecord(h->ev_h1, h->s_comm));
                        h->halo_timed = true;
                    } else {
                        CUDA_TRY(upd_on(own_lo, own_lo + h->g - 1, h->s_comm));
                        CUDA_TRY(upd_on(own_hi - h->g + 1, own_hi, h->s_comm));
                        rc = exchange(h, h->cur ^ 1, h->s_comm);
                        if (rc) return rc;
                    }
                    CUDA_TRY(cudaEventRecord(h->ev_comm, h->s_comm));
                    h->comm_pending = true;
                    CUDA_TRY(upd(own_lo + h->g, own_hi - h->g));
                    h->valid_ghost = h->g;
                } else {
                    CUDA_TRY(upd(own_lo - e, own_hi + e));
                    h->valid_ghost = e;
                    if (e == 0) {
                        rc = exchange(h, h->cur ^ 1, h->s_compute);
                        if (rc) return rc;
                        h->valid_ghost = h->g;
                    }
                }
                h->cur ^= 1;
                if (serpentine_enabled()) h->kflags ^= KF_REVERSE;
                t += adv;
            }
            int rc = wait_comm(h);   // the timed region of this call ends after its last exchange
            if (rc) return rc;
        }
        h->steps_total += (uint64_t)nt;
    }
    CUDA_TRY(cudaEventRecord(h->ev_t1, h->s_compute));
    h->timed = true;
    return HEAT_OK;
}

// Job pipeline.  A job = upload this rank's rows, march nt steps, download the result.  Jobs submitted back to
// back overlap: while job i marches on s_compute, job i+1's input crosses PCIe into `pipe_in` on s_h2d and job
// i-1's result leaves from `pipe_out` on s_d2h (PCIe is full duplex).  The two staging buffers cost two device-
// side copies per job (about 3 launch-times of the stencil kernel in total) and leave the stepping code untouched.
int heat_submit(heat_handle* h, const void* host_src, int64_t src_pitch, int nt, void* host_dst) {
    if (!h || !host_src || !host_dst) return fail(HEAT_ERR_INVALID, "heat_submit: null argument");
    if (nt < 0) return fail(HEAT_ERR_INVALID, "heat_submit: nt must be >= 0 (got %d)", nt);
    if (src_pitch <= 0) src_pitch = h->ncols;
    if (src_pitch < h->ncols) return fail(HEAT_ERR_INVALID, "heat_submit: src_pitch %lld < own columns %d", (long long)src_pitch, h->ncols);
    CUDA_TRY(cudaSetDevice(h->dev));
    const size_t own_bytes = (size_t)h->nrows * row_bytes(h), dense = (size_t)h->ncols * h->esize;
    const size_t goff = (size_t)h->gl * h->esize;            // own columns start here inside a local row
    if (!h->pipe_in) {                       // first job: staging buffers, copy streams, events
        CUDA_TRY(cudaMalloc(&h->pipe_in, own_bytes));
        CUDA_TRY(cudaMalloc(&h->pipe_out, own_bytes));
        CUDA_TRY(cudaStreamCreateWithFlags(&h->s_h2d, cudaStreamNonBlocking));
        CUDA_TRY(cudaStreamCreateWithFlags(&h->s_d2h, cudaStreamNonBlocking));
        for (cudaEvent_t* e : {&h->ev_up, &h->ev_in_free, &h->ev_res, &h->ev_out_free})
            CUDA_TRY(cudaEventCreateWithFlags(e, cudaEventDisableTiming));
        CUDA_TRY(cudaMemsetAsync(h->pipe_in, 0, own_bytes, h->s_h2d));      // the pitch padding stays zero for good
    }
    // upload, once the previous job has taken its input out of pipe_in
    if (h->pipe_jobs) CUDA_TRY(cudaStreamWaitEvent(h->s_h2d, h->ev_in_free, 0));
    CUDA_TRY(copy_rows((char*)h->pipe_in + goff, row_bytes(h), host_src, (size_t)src_pitch * h->esize, dense, h->nrows,
                       cudaMemcpyHostToDevice, h->s_h2d));
    CUDA_TRY(cudaEventRecord(h->ev_up, h->s_h2d));
    // install + march on the compute stream
    CUDA_TRY(cudaStreamWaitEvent(h->s_compute, h->ev_up, 0));
    int rc = install_field(h, (char*)h->pipe_in + goff, h->pitch, 1);
    if (rc) return rc;
    CUDA_TRY(cudaEventRecord(h->ev_in_free, h->s_compute));
    rc = heat_step(h, nt);
    if (rc) return rc;
    // result -> pipe_out once the previous job's download has drained it, then over PCIe on s_d2h
    if (h->pipe_jobs) CUDA_TRY(cudaStreamWaitEvent(h->s_compute, h->ev_out_free, 0));
    CUDA_TRY(cudaMemcpyAsync(h->pipe_out, row_ptr(h, h->cur, h->g), own_bytes, cudaMemcpyDeviceToDevice, h->s_compute));
    CUDA_TRY(cudaEventRecord(h->ev_res, h->s_compute));
    CUDA_TRY(cudaStreamWaitEvent(h->s_d2h, h->ev_res, 0));
    CUDA_TRY(copy_rows(host_dst, dense, (char*)h->pipe_out + goff, row_bytes(h), dense, h->nrows, cudaMemcpyDeviceToHost, h->s_d2h));
    CUDA_TRY(cudaEventRecord(h->ev_out_free, h->s_d2h));
    h->pipe_jobs++;
    return HEAT_OK;
}

int heat_drain(heat_handle* h) {
    if (!h) return fail(HEAT_ERR_INVALID, "heat_drain: null handle");
    CUDA_TRY(cudaSetDevice(h->dev));
    if (h->s_h2d) CUDA_TRY(cudaStreamSynchronize(h->s_h2d));
    CUDA_TRY(cudaStreamSynchronize(h->s_comm));
    CUDA_TRY(cudaStreamSynchronize(h->s_compute));
    if (h->s_d2h) CUDA_TRY(cudaStreamSynchronize(h->s_d2h));
    return HEAT_OK;
}

int heat_sync(heat_handle* h) {
    if (!h) return fail(HEAT_ERR_INVALID, "heat_sync: null handle");
    CUDA_TRY(cudaSetDevice(h->dev));
    CUDA_TRY(cudaStreamSynchronize(h->s_comm));
    CUDA_TRY(cudaStreamSynchronize(h->s_compute));
    return HEAT_OK;
}

int heat_gather(heat_handle* h, void* dst, int dst_is_device, int global_root) {
    if (!h) return fail(HEAT_ERR_INVALID, "heat_gather: null handle");
    CUDA_TRY(cudaSetDevice(h->dev));
    int rc = wait_comm(h);
    if (rc) return rc;
    const cudaMemcpyKind kind = dst_is_device ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost;
    const size_t dense = (size_t)h->ncols * h->esize;        // this rank's own tile, dense
    const size_t goff = (size_t)h->gl * h->esize;
    if (global_root < 0 || h->cfg.nranks == 1) {
        if (!dst) return fail(HEAT_ERR_INVALID, "heat_gather: null destination");
        CUDA_TRY(copy_rows(dst, dense, row_ptr(h, h->cur, h->g) + goff, row_bytes(h), dense, h->nrows, kind, h->s_compute));
        CUDA_TRY(cudaStreamSynchronize(h->s_compute));
        return HEAT_OK;
    }
    if (global_root >= h->cfg.nranks) return fail(HEAT_ERR_INVALID, "heat_gather: root %d out of range", global_root);
    if (!h->comm) return fail(HEAT_ERR_NCCL, "heat_gather: no communicator (SKIP_EXCHANGE build of the handle?)");
    const NcclApi& N = nccl_api();
    const int dt = h->cfg.dtype == HEAT_F32 ? ncclFloat32 : ncclFloat64;
    if (h->cfg.rank != global_root) {
        NCCL_TRY(N.Send(row_ptr(h, h->cur, h->g), (size_t)h->nrows * h->pitch, dt, global_root, h->comm, h->s_compute));
        CUDA_TRY(cudaStreamSynchronize(h->s_compute));
        return HEAT_OK;
    }
    if (!dst) return fail(HEAT_ERR_INVALID, "heat_gather: root needs a destination");
    const size_t gdense = (size_t)h->gnx * h->esize;         // a row of the global field
    for (int r = 0; r < h->cfg.nranks; ++r) {
        int r0, nr, c0, nc;
        heat_tile_partition(h->gnx, h->ny, h->px, h->py, r, &r0, &nr, &c0, &nc);
        char* out = (char*)dst + (size_t)r0 * gdense + (size_t)c0 * h->esize;
        if (r == h->cfg.rank) {
            CUDA_TRY(copy_rows(out, gdense, row_ptr(h, h->cur, h->g) + goff, row_bytes(h), dense, nr, kind, h->s_compute));
            continue;
        }
        // the sender's local geometry (library-owned buffers: pitch = its local width rounded up to 16 elements)
        const int rgl = (h->px > 1 && r % h->px > 0) ? h->g : 0, rgr = (h->px > 1 && r % h->px + 1 < h->px) ? h->g : 0;
        const long rpitch = h->px > 1 ? ((long)(nc + rgl + rgr) + 15) / 16 * 16 : h->pitch;
        const size_t need = (size_t)nr * rpitch * h->esize;
        if (need > h->stage_bytes) {
            CUDA_TRY(cudaStreamSynchronize(h->s_compute));
            cudaFree(h->d_stage);
            h->d_stage = nullptr;
            h->stage_bytes = 0;
            CUDA_TRY(cudaMalloc(&h->d_stage, need));
            h->stage_bytes = need;
        }
        NCCL_TRY(N.Recv(h->d_stage, (size_t)nr * rpitch, dt, r, h->comm, h->s_compute));
        CUDA_TRY(copy_rows(out, gdense, (char*)h->d_stage + (size_t)rgl * h->esize, (size_t)rpitch * h->esize,
                           (size_t)nc * h->esize, nr, kind, h->s_compute));
        CUDA_TRY(cudaStreamSynchronize(h->s_compute));
    }
    CUDA_TRY(cudaStreamSynchronize(h->s_compute));
    return HEAT_OK;
}

int heat_checksum(heat_handle* h, double* sum) {
    if (!h || !sum) return fail(HEAT_ERR_INVALID, "heat_checksum: null argument");
    CUDA_TRY(cudaSetDevice(h->dev));
    int rc = wait_comm(h);
    if (rc) return rc;
    if (!h->d_partials) CUDA_TRY(cudaMalloc(&h->d_partials, kSumBlocks * sizeof(double)));
    if (!h->d_sum) CUDA_TRY(cudaMalloc(&h->d_sum, sizeof(double)));
    const int yb = h->g, ye = h->g + h->nrows;
    if (h->cfg.dtype == HEAT_F32)
        heat_sum_partial_kernel<float><<<kSumBlocks, 256, 0, h->s_compute>>>((const float*)h->buf[h->cur] + h->gl, h->ncols, h->pitch, yb, ye, h->d_partials);
    else
        heat_sum_partial_kernel<double><<<kSumBlocks, 256, 0, h->s_compute>>>((const double*)h->buf[h->cur] + h->gl, h->ncols, h->pitch, yb, ye, h->d_partials);
    CUDA_TRY(cudaGetLastError());
    heat_sum_final_kernel<<<1, 32, 0, h->s_compute>>>(h->d_partials, kSumBlocks, h->d_sum);
    CUDA_TRY(cudaGetLastError());
    CUDA_TRY(cudaMemcpyAsync(sum, h->d_sum, sizeof(double), cudaMemcpyDeviceToHost, h->s_compute));
    CUDA_TRY(cudaStreamSynchronize(h->s_compute));
    return HEAT_OK;
}

int heat_get_info(heat_handle* h, heat_info* info) {
    if (!h || !info) return fail(HEAT_ERR_INVALID, "heat_get_info: null argument");
    info->nx = h->gnx; info->ny = h->ny;
    info->col0 = h->col0; info->ncols = h->ncols;
    info->ghost_left = h->gl; info->local_nx = h->nx;
    info->row0 = h->row0; info->nrows = h->nrows;
    info->ghost = h->g;
    info->kernel = h->kernel; info->tblock = h->tblock;
    info->pitch = h->pitch;
    info->buf[0] = h->buf[0]; info->buf[1] = h->buf[1];
    info->current = h->cur;
    info->device = h->dev;
    return HEAT_OK;
}

int heat_get_timing(heat_handle* h, heat_timing* t) {
    if (!h || !t) return fail(HEAT_ERR_INVALID, "heat_get_timing: null argument");
    CUDA_TRY(cudaSetDevice(h->dev));
    float ms = 0.f, hms = 0.f;
    if (h->timed) {
        CUDA_TRY(cudaEventSynchronize(h->ev_t1));
        CUDA_TRY(cudaEventElapsedTime(&ms, h->ev_t0, h->ev_t1));
    }
    if (h->halo_timed) {
        CUDA_TRY(cudaEventSynchronize(h->ev_h1));
        CUDA_TRY(cudaEventElapsedTime(&hms, h->ev_h0, h->ev_h1));
    }
    t->last_step_ms = ms;
    t->last_halo_ms = hms;
    t->steps_total = h->steps_total;
    t->kernel_launches = h->launches;
    t->halo_exchanges = h->exchanges;
    return HEAT_OK;
}

}  // extern "C"

// ------------------------------------------------------------------------------------ no-allocation checksum
// For callers that must not allocate device memory: a host program in the style of the reference's
// Fortran drivers, which owns every device buffer itself and hands raw pointers to launchers; a
// process whose device heap is already fully reserved by another allocator (a framework's caching
// pool, an OpenACC/CUDA-Fortran runtime); or a time loop on several streams where cudaMalloc and
// cudaFree would stall every stream because they synchronise the whole device.
//
// heat_checksum keeps its per-block partial sums in a buffer it cudaMallocs on first use.  This
// variant runs the same two kernels with the partials in the fixed array below instead, so it never
// calls cudaMalloc, cudaFree or anything else that allocates.  The price: the array is part of the
// library image, so every GPU that loads libheat.so carries these 8 KiB until the process exits,
// and the number of partial sums is capped at HEAT_NOALLOC_MAX_PARTIALS.
#define HEAT_NOALLOC_MAX_PARTIALS 1024
#define HEAT_NOALLOC_MAX_DEVICES 64

static __device__ __align__(256) double g_noalloc_partials[HEAT_NOALLOC_MAX_PARTIALS];

// In host code `g_noalloc_partials` is only a handle for the symbol, not a device address, so the
// address each device gives the array is looked up once per device with cudaGetSymbolAddress.
static double* g_noalloc_addr[HEAT_NOALLOC_MAX_DEVICES];
static bool g_noalloc_ready[HEAT_NOALLOC_MAX_DEVICES];
static pthread_mutex_t g_noalloc_lock = PTHREAD_MUTEX_INITIALIZER;

extern "C" int heat_checksum_noalloc(const void* field_dev, int dtype, int nx, int ny, int64_t pitch,
                                     int nblocks, double* sum_dev, uint64_t stream) {
    if (!field_dev || !sum_dev) return fail(HEAT_ERR_INVALID, "heat_checksum_noalloc: null pointer");
    if (dtype != HEAT_F64 && dtype != HEAT_F32) return fail(HEAT_ERR_INVALID, "heat_checksum_noalloc: bad dtype %d", dtype);
    if (nx < 1 || ny < 1 || pitch < nx)
        return fail(HEAT_ERR_INVALID, "heat_checksum_noalloc: need nx, ny >= 1 and pitch >= nx (got %d x %d, pitch %lld)", nx, ny, (long long)pitch);
    if (nblocks == 0) nblocks = kSumBlocks;      // the tree heat_checksum uses, hence the same bits
    if (nblocks < 0 || nblocks > HEAT_NOALLOC_MAX_PARTIALS)
        return fail(HEAT_ERR_UNSUPPORTED,
                    "heat_checksum_noalloc: %d partial sums requested but the built-in scratch holds at most %d; "
                    "use heat_checksum (which allocates) or fewer blocks", nblocks, HEAT_NOALLOC_MAX_PARTIALS);
    int dev = -1;
    CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= HEAT_NOALLOC_MAX_DEVICES)
        return fail(HEAT_ERR_UNSUPPORTED, "heat_checksum_noalloc: device ordinal %d is beyond the %d this build tracks", dev, HEAT_NOALLOC_MAX_DEVICES);
    pthread_mutex_lock(&g_noalloc_lock);
    if (!g_noalloc_ready[dev]) {                 // set-up: first call on this device, with it current
        double* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress((void**)&p, g_noalloc_partials);
        if (e != cudaSuccess || !p) {
            pthread_mutex_unlock(&g_noalloc_lock);
            return fail(HEAT_ERR_CUDA, "heat_checksum_noalloc: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
        }
        g_noalloc_addr[dev] = p;
        g_noalloc_ready[dev] = true;
    }
    double* partials = g_noalloc_addr[dev];
    pthread_mutex_unlock(&g_noalloc_lock);
    cudaStream_t s = (cudaStream_t)(uintptr_t)stream;
    // Every slot the second kernel reads is written by the first one in this same call, so nothing
    // left in the array by an earlier call matters and nothing has to be zeroed.
    if (dtype == HEAT_F32)
        heat_sum_partial_kernel<float><<<nblocks, 256, 0, s>>>((const float*)field_dev, nx, (long)pitch, 0, ny, partials);
    else
        heat_sum_partial_kernel<double><<<nblocks, 256, 0, s>>>((const double*)field_dev, nx, (long)pitch, 0, ny, partials);
    CUDA_TRY(cudaGetLastError());
    heat_sum_final_kernel<<<1, 32, 0, s>>>(partials, nblocks, sum_dev);
    CUDA_TRY(cudaGetLastError());
    return HEAT_OK;
}
