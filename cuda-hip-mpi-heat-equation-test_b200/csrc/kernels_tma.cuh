// kernels_tma.cuh — TMA-fed variant of the single-step update (sm_100a).
//
// A CTA owns a column strip BX cells wide and a segment of rows.  One elected thread streams the
// strip down through a ring of S shared-memory stages with cp.async.bulk.tensor.2d (SASS: UTMALDG):
// each stage is a (BX + 2V) x BY box whose first column is x0 - V, i.e. the strip plus one 16-byte
// vector of halo on either side, so that a lane's own vector sits 16-byte aligned in shared memory
// and the x-1 / x+1 cells of the strip's edge lanes are already there.  Out-of-range box columns
// and rows are zero-filled by the TMA unit and only ever feed cells that are not updated.
// Completion is signalled on an mbarrier per stage (complete_tx::bytes); the 256 consumer threads
// keep a rolling 3-row register window per row-subgroup, take x-neighbours by warp shuffle, and
// store results straight from registers with 128-bit coalesced stores.  S-1 boxes (S-1 x ~17 KB)
// are in flight per CTA, 3 CTAs per SM — enough bytes in flight to cover HBM latency without
// depending on occupancy or register-staged loads.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>

#include <algorithm>
#include <string>
#include <type_traits>

#include "stencil_ops.cuh"

namespace heat {

struct TmaState {
    CUtensorMap map[2];
    bool ready = false;
    int box_w = 0, box_h = 0;
};

// Stage shape (compile-time knobs).  Measured at 16384^2 fp64 with 46-row segments: 16 rows x 4 stages 432.6 GLUPS,
// 8 x 8 432.8 (fp32 852 vs 858), 8 x 4 425.8, 32 x 2 281.1 — the kernel sits at its DRAM ceiling, the default stays.
#ifndef HEAT_TMA_BY
#define HEAT_TMA_BY 16
#endif
#ifndef HEAT_TMA_STAGES
#define HEAT_TMA_STAGES 4
#endif
constexpr int kTmaBX = 128;      // strip width in cells
constexpr int kTmaBY = HEAT_TMA_BY;       // rows per stage
constexpr int kTmaStages = HEAT_TMA_STAGES;
constexpr int kTmaThreads = 256;

__device__ __forceinline__ unsigned smem_u32(const void* p) {
    return (unsigned)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, unsigned parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void tma_load_2d(void* smem_dst, const CUtensorMap* map, int cx, int cy, uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
        ::"r"(smem_u32(smem_dst)), "l"(map), "r"(cx), "r"(cy), "r"(smem_u32(bar)) : "memory");
}

// Programmatic dependent launch: consecutive time steps are separate launches whose only true
// dependency is the field itself.  A step's CTAs may be scheduled while the previous step's last
// wave drains (launch_dependents is signalled at CTA start); they set up their barriers, then block
// in griddepcontrol.wait until the previous grid has completed and flushed before touching memory.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

inline bool pdl_enabled() {
    static const bool on = [] { const char* v = getenv("HEAT_PDL"); return !(v && v[0] == '0'); }();
    return on;
}

// L2 residency experiment (OFF unless HEAT_L2_PERSIST_MB is set): pin the bytes a step writes last — the
// bytes the next, opposite-direction step reads first — in the persisting part of L2 with a per-launch
// access-policy window.  Measured on B200 it LOSES (4096^2: 407 -> 335-401 GLUPS for 32-128 MB windows;
// 16384^2: 385 -> 227): the carve-out costs more ordinary L2 capacity than the pinned lines give back.
// Plain serpentine sweeping (below) keeps the gain without it.
struct L2Window {
    void* base = nullptr;
    size_t bytes = 0;
};

inline size_t l2_persist_bytes() {       // how much L2 this process may pin (set up once per process/device)
    static const size_t n = [] {
        int dev = 0, max_persist = 0, max_window = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&max_persist, cudaDevAttrMaxPersistingL2CacheSize, dev);
        cudaDeviceGetAttribute(&max_window, cudaDevAttrMaxAccessPolicyWindowSize, dev);
        const char* v = getenv("HEAT_L2_PERSIST_MB");
        if (!v) return (size_t)0;
        const size_t want = std::min((size_t)std::min(max_persist, max_window), (size_t)atol(v) << 20);
        if (want == 0) return (size_t)0;
        if (cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, want) != cudaSuccess) { cudaGetLastError(); return (size_t)0; }
        return want;
    }();
    return n;
}

template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, L2Window win,
                              Args... args) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = s;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    if (win.base && win.bytes) {
        attr[1].id = cudaLaunchAttributeAccessPolicyWindow;
        attr[1].val.accessPolicyWindow.base_ptr = win.base;
        attr[1].val.accessPolicyWindow.num_bytes = win.bytes;
        attr[1].val.accessPolicyWindow.hitRatio = 1.0f;
        attr[1].val.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
        attr[1].val.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
        cfg.numAttrs = 2;
    }
    return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}

// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) once per kernel instantiation and device, not per launch
// (keyed by the kernel itself — a non-type template argument — because instantiations that differ only in
// ORDER/FMA share one function-pointer type)
template <auto Kernel>
inline cudaError_t ensure_dynamic_smem(size_t smem) {
    static bool done[64] = {};
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev >= 0 && dev < 64 && done[dev]) return cudaSuccess;
    e = cudaFuncSetAttribute(Kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e == cudaSuccess && dev >= 0 && dev < 64) done[dev] = true;
    return e;
}

// the tail of a sweep over rows [yb, ye) of `dst`: its last rows when sweeping forward, its first when reversed
template <typename T>
inline L2Window sweep_tail_window(T* dst, long pitch, int yb, int ye, int kflags) {
    L2Window w;
    const size_t cap = l2_persist_bytes();
    if (cap == 0) return w;
    const size_t row = (size_t)pitch * sizeof(T);
    const size_t span = (size_t)(ye - yb) * row;
    w.bytes = std::min(cap, span);
    w.base = (kflags & 1) ? (char*)dst + (size_t)yb * row : (char*)dst + (size_t)ye * row - w.bytes;
    return w;
}

// L2 cache-policy variants (tuning knobs; measured in profiles/README.md).  The field is swept in opposite
// directions on alternate steps (serpentine), so the rows written last — still in the 126 MB L2 — are the
// rows read first by the next step.
__device__ __forceinline__ uint64_t policy_evict_first() {
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
    return p;
}
__device__ __forceinline__ uint64_t policy_evict_last() {
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
    return p;
}
__device__ __forceinline__ void tma_load_2d_hint(void* smem_dst, const CUtensorMap* map, int cx, int cy, uint64_t* bar,
                                                 uint64_t policy) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1, {%2, %3}], [%4], %5;"
        ::"r"(smem_u32(smem_dst)), "l"(map), "r"(cx), "r"(cy), "r"(smem_u32(bar)), "l"(policy) : "memory");
}
__device__ __forceinline__ void st_hint(double* p, const Pack<double>& v, uint64_t policy) {
    asm volatile("st.global.L2::cache_hint.v2.f64 [%0], {%1, %2}, %3;" ::"l"(p), "d"(v.v[0]), "d"(v.v[1]), "l"(policy) : "memory");
}
__device__ __forceinline__ void st_hint(float* p, const Pack<float>& v, uint64_t policy) {
    asm volatile("st.global.L2::cache_hint.v4.f32 [%0], {%1, %2, %3, %4}, %5;" ::"l"(p), "f"(v.v[0]), "f"(v.v[1]),
                 "f"(v.v[2]), "f"(v.v[3]), "l"(policy) : "memory");
}

// kflags: bit 0 = sweep the grid in reverse block order; bit 1 = evict-first loads; bit 2 = evict-last stores
enum : int { KF_REVERSE = 1, KF_LOAD_EVICT_FIRST = 2, KF_STORE_EVICT_LAST = 4 };

inline int sweep_policy_flags() {
    static const int f = [] {
        const char* v = getenv("HEAT_CACHE_HINT");       // tuning knob: 0 none, 1 loads, 2 loads+stores, 3 stores
        const int m = v ? atoi(v) : 0;
        return (m == 1 || m == 2 ? KF_LOAD_EVICT_FIRST : 0) | (m == 2 || m == 3 ? KF_STORE_EVICT_LAST : 0);
    }();
    return f;
}
inline bool serpentine_enabled() {
    static const bool on = [] { const char* v = getenv("HEAT_SERPENTINE"); return !(v && v[0] == '0'); }();
    return on;
}

template <typename T, int ORDER, bool FMA>
__global__ void __launch_bounds__(kTmaThreads)
heat_tma_kernel(const __grid_constant__ CUtensorMap tmap, T* __restrict__ dst, int nx, long pitch,
                int y_begin, int y_end, int seg_rows, T rx, T ry, int kflags, T* __restrict__ dst2, int shift2) {
    // dst2 (optional): the same rows are ALSO stored, shifted by shift2 rows, into a second buffer — the ghost
    // rows of the neighbouring GPU, mapped through CUDA IPC — so an edge strip's results cross NVLink from
    // inside the stencil kernel, as peer stores issued next to the local ones (fused compute + halo push).
    constexpr int V = Pack<T>::V;
    using P = Pack<T>;
    constexpr int BX = kTmaBX, BY = kTmaBY, S = kTmaStages;
    constexpr int BW = BX + 2 * V;                 // box width (cells)
    constexpr int LPR = BX / V;                    // lanes per row
    constexpr int NSUB = kTmaThreads / LPR;        // row subgroups
    constexpr int RPS = BY / NSUB;                 // rows per subgroup per stage
    constexpr int RING = S * BY;                   // ring depth in rows
    static_assert(BY % NSUB == 0 && LPR % 32 == 0, "tile shape");
    constexpr unsigned kStageBytes = BY * BW * sizeof(T);
    static_assert(kStageBytes % 128 == 0, "stage must keep 128-byte alignment");

    extern __shared__ __align__(128) unsigned char smem_raw[];
    T* ring = reinterpret_cast<T*>(smem_raw);
    uint64_t* full = reinterpret_cast<uint64_t*>(smem_raw + (size_t)S * kStageBytes);

    const int tid = threadIdx.x, lane = tid & 31;
    const int col = tid % LPR, sub = tid / LPR;
    const bool rev = kflags & KF_REVERSE;
    const int bx = rev ? gridDim.x - 1 - blockIdx.x : blockIdx.x;
    const int by = rev ? gridDim.y - 1 - blockIdx.y : blockIdx.y;
    const int x0 = bx * BX;
    const int ys = y_begin + by * seg_rows;
    const int ye = min(ys + seg_rows, y_end);
    if (ys >= ye) return;
    const bool ld_hint = kflags & KF_LOAD_EVICT_FIRST, st_hinted = kflags & KF_STORE_EVICT_LAST;
    const uint64_t pol_ld = policy_evict_first(), pol_st = policy_evict_last();
    const int np = ye - ys;                         // centres p = 1..np  (global row ys-1+p)
    const int nstages = (np + 2 + BY - 1) / BY;     // input rows p = 0..np+1

    pdl_launch_dependents();
    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < S; ++s) mbar_init(&full[s], 1);
        mbar_fence_init();
    }
    __syncthreads();
    pdl_wait();                                     // previous step's field is complete and visible

    auto issue = [&](int j) {
        const int slot = j % S;
        mbar_expect_tx(&full[slot], kStageBytes);
        if (ld_hint) tma_load_2d_hint(ring + (size_t)slot * BY * BW, &tmap, x0 - V, ys - 1 + j * BY, &full[slot], pol_ld);
        else tma_load_2d(ring + (size_t)slot * BY * BW, &tmap, x0 - V, ys - 1 + j * BY, &full[slot]);
    };
    if (tid == 0) {
        for (int j = 0; j < S - 1 && j < nstages; ++j) issue(j);
    }

    const long xg = (long)x0 + (long)col * V;       // this lane's first global column
    const bool active = xg < pitch;
    const int sx = (col + 1) * V;                   // its column inside the box
    // lanes at a warp's edge read the adjacent cell from shared memory instead of shuffling
    const int sxe = (lane == 0) ? sx - 1 : sx + V;

    // A stage is PLAIN when none of its rows needs a guard: every row it reads has been loaded, every row it
    // produces lies inside this CTA's segment, the strip touches no Dirichlet column and no pitch padding, and
    // no cache-policy or mirror store is requested.  Plain stages — almost all of them — run a copy of the body
    // without the selects, predicates and signed index arithmetic (fewer instructions, less power).
    const bool plain_cta = x0 >= V && (long)x0 + BX <= (long)nx - 1 && (long)x0 + BX <= pitch && !st_hinted && !dst2;
    static_assert((RING & (RING - 1)) == 0, "ring depth must be a power of two");

    auto stage_body = [&](auto plain_tag, int j) {
        constexpr bool PLAIN = decltype(plain_tag)::value;
        const int p0 = j * BY - 1 + sub * RPS;      // first centre row of this subgroup
        P win[RPS + 2];
        T edg[RPS];
#pragma unroll
        for (int i = 0; i < RPS + 2; ++i) {
            const int p = p0 - 1 + i;
            if (PLAIN || p >= 0) win[i] = *reinterpret_cast<const P*>(ring + (size_t)((unsigned)p & (RING - 1)) * BW + sx);
            else {
#pragma unroll
                for (int v = 0; v < V; ++v) win[i].v[v] = T(0);
            }
        }
        if (lane == 0 || lane == 31) {
#pragma unroll
            for (int i = 0; i < RPS; ++i) {
                const int p = p0 + i;
                edg[i] = (PLAIN || p >= 0) ? ring[(size_t)((unsigned)p & (RING - 1)) * BW + sxe] : T(0);
            }
        }
        T* q = dst + (long)(ys - 1 + p0) * pitch + xg;
#pragma unroll
        for (int i = 0; i < RPS; ++i, q += pitch) {
            const int p = p0 + i;
            const P& s = win[i];
            const P& c = win[i + 1];
            const P& n = win[i + 2];
            T left = __shfl_up_sync(0xffffffffu, c.v[V - 1], 1);
            T right = __shfl_down_sync(0xffffffffu, c.v[0], 1);
            if (lane == 0) left = edg[i];
            if (lane == 31) right = edg[i];
            P o;
#pragma unroll
            for (int v = 0; v < V; ++v) {
                const T w = (v == 0) ? left : c.v[v - 1];
                const T e = (v == V - 1) ? right : c.v[v + 1];
                const T r = stencil<T, ORDER, FMA>(c.v[v], w, e, s.v[v], n.v[v], rx, ry);
                if constexpr (PLAIN) o.v[v] = r;
                else {
                    const long x = xg + v;
                    o.v[v] = (x >= 1 && x <= nx - 2) ? r : c.v[v];
                }
            }
            if constexpr (PLAIN) {
                *reinterpret_cast<P*>(q) = o;
            } else if (active && p >= 1 && p <= np) {
                if (st_hinted) st_hint(q, o, pol_st);
                else *reinterpret_cast<P*>(q) = o;
                if (dst2) *reinterpret_cast<P*>(dst2 + (long)(ys - 1 + p + shift2) * pitch + xg) = o;
            }
        }
    };

    for (int j = 0; j < nstages; ++j) {
        mbar_wait(&full[j % S], (unsigned)((j / S) & 1));
        // rows produced by stage j: p = j*BY-1 .. (j+1)*BY-2, all inside [1, np] and all inputs loaded?
        const bool plain = plain_cta && j >= 1 && (j + 1) * BY - 2 <= np;
        if (plain) stage_body(std::true_type{}, j);
        else stage_body(std::false_type{}, j);
        __syncthreads();                            // everyone is done with stage j-1 (and j's reads)
        if (tid == 0 && j + S - 1 < nstages) issue(j + S - 1);
    }
}

// ------------------------------------------------------------------------------------ host side
typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                    CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                    CUtensorMapFloatOOBfill);

inline PFN_encodeTiled get_encode_tiled() {
    static PFN_encodeTiled fn = [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
            q != cudaDriverEntryPointSuccess)
            p = nullptr;
        return (PFN_encodeTiled)p;
    }();
    return fn;
}

// one 2-D map per ping-pong buffer: dims {pitch, rows}, box {box_w, box_h}, zero fill out of range
inline int tma_make_map(CUtensorMap* map, void* base, bool f32, long pitch, int rows, int box_w, int box_h,
                        std::string& err) {
    PFN_encodeTiled enc = get_encode_tiled();
    if (!enc) { err = "cuTensorMapEncodeTiled is not available from this driver"; return -5; }
    const size_t es = f32 ? 4 : 8;
    cuuint64_t dims[2] = {(cuuint64_t)pitch, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)pitch * es};
    cuuint32_t box[2] = {(cuuint32_t)box_w, (cuuint32_t)box_h};
    cuuint32_t estr[2] = {1, 1};
    // L2 promotion: measured (profiles/README.md) — 256 B helps fields of about L2 size (4096^2), 64 B is best
    // once the field is many times larger than the 126 MB L2 (16384^2: 388.7 vs 380.7 GLUPS)
    const size_t field_bytes = (size_t)pitch * rows * es;
    CUtensorMapL2promotion promo = field_bytes > (size_t)512 << 20 ? CU_TENSOR_MAP_L2_PROMOTION_L2_64B
                                                                   : CU_TENSOR_MAP_L2_PROMOTION_L2_256B;
    if (const char* v = getenv("HEAT_TMA_L2PROMO")) {          // tuning knob: 0 none, 1 64 B, 2 128 B, 3 256 B
        const int k = atoi(v);
        promo = k == 0 ? CU_TENSOR_MAP_L2_PROMOTION_NONE : k == 1 ? CU_TENSOR_MAP_L2_PROMOTION_L2_64B
              : k == 2 ? CU_TENSOR_MAP_L2_PROMOTION_L2_128B : CU_TENSOR_MAP_L2_PROMOTION_L2_256B;
    }
    CUresult r = enc(map, f32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 2, base, dims,
                     strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                     promo, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        err = "cuTensorMapEncodeTiled failed with CUresult " + std::to_string((int)r);
        return -2;
    }
    return 0;
}

inline int tma_setup(TmaState& st, void* const buf[2], bool f32, int nx, int rows, long pitch, std::string& err) {
    (void)nx;
    const int V = f32 ? 4 : 2;
    st.box_w = kTmaBX + 2 * V;
    st.box_h = kTmaBY;
    for (int b = 0; b < 2; ++b) {
        int rc = tma_make_map(&st.map[b], buf[b], f32, pitch, rows, st.box_w, st.box_h, err);
        if (rc) return rc;
    }
    st.ready = true;
    return 0;
}

template <typename T, int ORDER, bool FMA>
inline cudaError_t launch_tma_rows(const TmaState& st, int which, T* dst, int nx, long pitch, int yb, int ye,
                                   T rx, T ry, int sm_count, cudaStream_t s, int kflags, T* dst2 = nullptr,
                                   int shift2 = 0) {
    constexpr int V = Pack<T>::V;
    constexpr size_t smem = (size_t)kTmaStages * kTmaBY * (kTmaBX + 2 * V) * sizeof(T) + kTmaStages * sizeof(uint64_t);
    auto k = heat_tma_kernel<T, ORDER, FMA>;
    cudaError_t e = ensure_dynamic_smem<heat_tma_kernel<T, ORDER, FMA>>(smem);
    if (e != cudaSuccess) return e;
    const int rows = ye - yb;
    const int gx = (int)((pitch + kTmaBX - 1) / kTmaBX);
    // Segment length.  SHORT segments, issued strip-fastest, make the resident CTAs a thin wavefront that sweeps the
    // field row band by row band: DRAM sees two compact streams (one read, one write) instead of ~13 scattered ones,
    // the two halo rows a segment shares with its neighbours are L2 hits, and the tail of the grid is short.  Measured
    // on B200 at 16384^2 fp64 (profiles/README.md, round 2): 1264-row segments 387 GLUPS, 94 -> 423, 46 -> 433,
    // 30 -> 431, 16 -> 292 (per-CTA prologue dominates).  46 + 2 halo rows = 48 input rows = a whole number of stages.
    int seg = 0;
    if (const char* v = getenv("HEAT_TMA_SEG")) seg = atoi(v);
    if (seg <= 0) seg = 48 - 2;
    (void)sm_count;
    seg = std::min(seg, rows);
    const int gy = (rows + seg - 1) / seg;
    return launch_pdl(k, dim3(gx, gy), dim3(kTmaThreads), smem, s, sweep_tail_window(dst, pitch, yb, ye, kflags), st.map[which], dst, nx, pitch, yb, ye, seg, rx, ry, kflags, dst2, shift2);
}

}  // namespace heat
