// kernels_temporal.cuh — temporal blocking: K time steps per HBM round trip (sm_100a).
//
// Unit of work = one WARP marching down a column strip.  The warp owns a private TMA pipeline:
// lane 0 streams (32*M*V) x RS boxes of the time-level-0 field into the warp's own ring of S
// shared-memory stages (cp.async.bulk.tensor.2d + one mbarrier per stage); nothing in the kernel
// synchronises wider than a warp.  Each lane owns M*V contiguous cells of the strip and keeps, for
// every time level 0..K-1, the two previous rows in registers.  When level-0 row y arrives the
// warp produces level-1 row y-1, level-2 row y-2, ... level-K row y-K (a wavefront skewed along
// y), exchanging only the two lane-edge cells per level by warp shuffle, and stores level K.
// The field therefore crosses HBM once per K steps: 16/K B per lattice update (fp64) plus the
// strip overlap (2*HX of 32*M*V columns, HX = K rounded up to the vector width) and 2K rows per
// row segment that neighbouring warps re-read (mostly from L2).
//
// Exactness: every cell at every level is computed with exactly the same operation sequence as
// K separate single steps — only the schedule changes — so results are bit-identical to K
// launches of the single-step kernels (tests/test_gpu_parity.py checks this).
// Dirichlet cells (global boundary rows/columns) are passed through unchanged at every level;
// out-of-range box cells (zero-filled by TMA) and the strip's outer HX columns only ever feed
// cells outside the region that is stored.
#pragma once
#include "kernels_tma.cuh"

namespace heat {

constexpr int kTempWarps = 4;     // warps per CTA (independent workers)
constexpr int kTempRS = 4;        // rows per TMA stage
constexpr int kTempStages = 4;    // stages in each warp's ring

template <typename T, int K> struct TemporalGeom {
    static constexpr int V = Pack<T>::V;
    static constexpr int HX = (K + V - 1) / V * V;       // halo columns each side (vector aligned)
};

template <typename T, int ORDER, bool FMA, int K, int M>
__global__ void __launch_bounds__(kTempWarps * 32)
heat_temporal_kernel(const __grid_constant__ CUtensorMap tmap, T* __restrict__ dst, int nx, long pitch,
                     int y_begin, int y_end, int seg_rows, int nstrips, int nitems, int ybc0, int ybc1,
                     T rx, T ry) {
    constexpr int V = Pack<T>::V;
    using P = Pack<T>;
    constexpr int MV = M * V;                     // cells per lane
    constexpr int WIN = 32 * MV;                  // strip width loaded (cells)
    constexpr int HX = TemporalGeom<T, K>::HX;
    constexpr int WOUT = WIN - 2 * HX;            // strip width stored
    constexpr int RS = kTempRS, S = kTempStages;
    constexpr unsigned kStageBytes = RS * WIN * sizeof(T);
    static_assert(kStageBytes % 128 == 0, "stage alignment");

    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int item = blockIdx.x * kTempWarps + warp;
    if (item >= nitems) return;                   // whole warps leave; nothing below syncs across warps
    const int strip = item % nstrips, seg = item / nstrips;
    const int xs = strip * WOUT - HX;             // first column of the loaded strip
    const int ys = y_begin + seg * seg_rows;
    const int ye = min(ys + seg_rows, y_end);
    const int nin = (ye - ys) + 2 * K;            // input rows ys-K .. ye+K-1
    const int nst = (nin + RS - 1) / RS;

    T* ring = reinterpret_cast<T*>(smem_raw + (size_t)warp * S * kStageBytes);
    uint64_t* full = reinterpret_cast<uint64_t*>(smem_raw + (size_t)kTempWarps * S * kStageBytes) + warp * S;

    auto issue = [&](int j) {
        const int slot = j % S;
        mbar_expect_tx(&full[slot], kStageBytes);
        tma_load_2d(ring + (size_t)slot * RS * WIN, &tmap, xs, ys - K + j * RS, &full[slot]);
    };
    if (lane == 0) {
#pragma unroll
        for (int s = 0; s < S; ++s) mbar_init(&full[s], 1);
        mbar_fence_init();
        for (int j = 0; j < S && j < nst; ++j) issue(j);
    }
    __syncwarp();

    const long xl = (long)xs + (long)lane * MV;   // this lane's first global column
    bool colfix[MV];                               // Dirichlet (or outside the field): pass through
    bool any_fix = false;
#pragma unroll
    for (int i = 0; i < MV; ++i) {
        colfix[i] = (xl + i <= 0) || (xl + i >= nx - 1);
        any_fix |= colfix[i];
    }
    const bool warp_fix = __any_sync(0xffffffffu, any_fix);

    T prev1[K][MV], prev2[K][MV];                  // levels 0..K-1: rows y-1 and y-2 of the wavefront
#pragma unroll
    for (int t = 0; t < K; ++t)
#pragma unroll
        for (int i = 0; i < MV; ++i) prev1[t][i] = prev2[t][i] = T(0);

    for (int p = 0; p < nin; ++p) {
        const int j = p / RS, q = p % RS;
        if (q == 0) mbar_wait(&full[j % S], (unsigned)((j / S) & 1));
        T newest[MV];
        {
            const T* row = ring + (size_t)(j % S) * RS * WIN + (size_t)q * WIN + lane * MV;
#pragma unroll
            for (int m = 0; m < M; ++m) {
                const P v = *reinterpret_cast<const P*>(row + m * V);
#pragma unroll
                for (int i = 0; i < V; ++i) newest[m * V + i] = v.v[i];
            }
        }
        const int gy = ys - K + p;                 // level-0 row that just arrived
#pragma unroll
        for (int t = 1; t <= K; ++t) {
            const int r = gy - t;                  // level-t row produced now
            T (&c)[MV] = prev1[t - 1];
            T (&s)[MV] = prev2[t - 1];
            const T left = __shfl_up_sync(0xffffffffu, c[MV - 1], 1);
            const T right = __shfl_down_sync(0xffffffffu, c[0], 1);
            const bool rowfix = (r == ybc0) || (r == ybc1);
            T nw[MV];
#pragma unroll
            for (int i = 0; i < MV; ++i) {
                const T w = (i == 0) ? left : c[i - 1];
                const T e = (i == MV - 1) ? right : c[i + 1];
                nw[i] = stencil<T, ORDER, FMA>(c[i], w, e, s[i], newest[i], rx, ry);
            }
            if (rowfix) {
#pragma unroll
                for (int i = 0; i < MV; ++i) nw[i] = c[i];
            } else if (warp_fix) {
#pragma unroll
                for (int i = 0; i < MV; ++i) nw[i] = colfix[i] ? c[i] : nw[i];
            }
#pragma unroll
            for (int i = 0; i < MV; ++i) {
                s[i] = c[i];
                c[i] = newest[i];
                newest[i] = nw[i];
            }
        }
        const int o = gy - K;                      // level-K row, complete
        if (o >= ys && o < ye) {
#pragma unroll
            for (int m = 0; m < M; ++m) {
                const int rel = lane * MV + m * V;
                if (rel >= HX && rel < HX + WOUT && xl + m * V < pitch) {
                    P out;
#pragma unroll
                    for (int i = 0; i < V; ++i) out.v[i] = newest[m * V + i];
                    *reinterpret_cast<P*>(dst + (long)o * pitch + xl + m * V) = out;
                }
            }
        }
        if (q == RS - 1) {                         // stage j fully consumed (its values are in registers)
            __syncwarp();
            if (lane == 0 && j + S < nst) issue(j + S);
        }
    }
}

// ------------------------------------------------------------------------------------ host side
struct TemporalState {
    CUtensorMap map[2];
    int k = 0, m = 0;
    bool ready = false;
};

template <typename T> constexpr int temporal_m() { return 2; }

inline int temporal_setup(TemporalState& st, void* const buf[2], bool f32, int rows, long pitch, int k,
                          std::string& err) {
    const int V = f32 ? 4 : 2;
    const int M = 2;
    for (int b = 0; b < 2; ++b) {
        int rc = tma_make_map(&st.map[b], buf[b], f32, pitch, rows, 32 * M * V, kTempRS, err);
        if (rc) return rc;
    }
    st.k = k;
    st.m = M;
    st.ready = true;
    return 0;
}

template <typename T, int ORDER, bool FMA, int K>
inline cudaError_t launch_temporal_rows(const TemporalState& st, int which, T* dst, int nx, long pitch, int yb,
                                        int ye, int ybc0, int ybc1, T rx, T ry, int sm_count, cudaStream_t s) {
    constexpr int V = Pack<T>::V;
    constexpr int M = 2;
    constexpr int WIN = 32 * M * V, HX = TemporalGeom<T, K>::HX, WOUT = WIN - 2 * HX;
    constexpr size_t smem = (size_t)kTempWarps * kTempStages * kTempRS * WIN * sizeof(T) +
                            (size_t)kTempWarps * kTempStages * sizeof(uint64_t);
    auto k = heat_temporal_kernel<T, ORDER, FMA, K, M>;
    cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    const int rows = ye - yb;
    const int nstrips = (int)((pitch + WOUT - 1) / WOUT);
    int seg = 0;
    if (const char* v = getenv("HEAT_TEMPORAL_SEG")) seg = atoi(v);
    if (seg <= 0) {
        const int target = sm_count * 12 * 3;                 // ~3 waves of 12 resident warps per SM
        const int nseg = std::max(1, target / nstrips);
        seg = std::max(64 * K, (rows + nseg - 1) / nseg);
    }
    seg = std::min(seg, rows);
    const int nsegs = (rows + seg - 1) / seg;
    const int nitems = nstrips * nsegs;
    const int grid = (nitems + kTempWarps - 1) / kTempWarps;
    k<<<grid, kTempWarps * 32, smem, s>>>(st.map[which], dst, nx, pitch, yb, ye, seg, nstrips, nitems, ybc0, ybc1,
                                          rx, ry);
    return cudaGetLastError();
}

}  // namespace heat
