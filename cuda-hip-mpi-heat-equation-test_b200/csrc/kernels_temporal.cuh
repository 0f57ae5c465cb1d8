// kernels_temporal.cuh — temporal blocking: K time steps per HBM round trip (sm_100a).
//
// Unit of work = one WARP marching down a column strip.  The warp owns a private TMA pipeline:
// lane 0 streams (32*M*V) x RS boxes of the time-level-0 field into the warp's own ring of S
// shared-memory stages (cp.async.bulk.tensor.2d + one mbarrier per stage); nothing in the kernel
// synchronises wider than a warp.  Each lane owns M*V contiguous cells of the strip and keeps, for
// every time level 0..K-1, three consecutive rows in registers (their roles rotate, nothing is copied).  When level-0 row y arrives the
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
#include <type_traits>

#include "kernels_tma.cuh"

namespace heat {

constexpr int kTempWarps = 4;     // warps per CTA (independent workers)
constexpr int kTempRS = 3;        // rows per TMA stage == period of the register-row rotation
constexpr int kTempStages = 4;    // stages in each warp's ring

template <typename T, int K> struct TemporalGeom {
    static constexpr int V = Pack<T>::V;
    static constexpr int HX = (K + V - 1) / V * V;       // halo columns each side (vector aligned)
};

#ifndef HEAT_TEMP_MINB
#define HEAT_TEMP_MINB 1
#endif
template <typename T, int ORDER, bool FMA, int K, int M>
__global__ void __launch_bounds__(kTempWarps * 32, HEAT_TEMP_MINB)
heat_temporal_kernel(const __grid_constant__ CUtensorMap tmap, T* __restrict__ dst, int nx, long pitch,
                     int y_begin, int y_end, int seg_rows, int nstrips, int nitems, int ybc0, int ybc1,
                     T rx, T ry, int pdl_early, int kflags, T* __restrict__ dst2, int shift2) {
    constexpr int V = Pack<T>::V;
    using P = Pack<T>;
    constexpr int MV = M * V;                     // cells per lane
    constexpr int WIN = 32 * MV;                  // strip width loaded (cells)
    constexpr int HX = TemporalGeom<T, K>::HX;
    constexpr int WOUT = WIN - 2 * HX;            // strip width stored
    constexpr int RS = kTempRS, S = kTempStages;
    static_assert(RS == 3, "the row rotation below has period 3");
    constexpr unsigned kStageBytes = RS * WIN * sizeof(T);
    static_assert(kStageBytes % 128 == 0, "stage alignment");

    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int item_fwd = blockIdx.x * kTempWarps + warp;
    if (pdl_early) pdl_launch_dependents();
    if (item_fwd >= nitems) return;               // whole warps leave; nothing below syncs across warps
    const int item = (kflags & KF_REVERSE) ? nitems - 1 - item_fwd : item_fwd;
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
    }
    pdl_wait();                                   // previous launch's field is complete and visible
    if (lane == 0) {
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
    bool st_ok[M];
#pragma unroll
    for (int m = 0; m < M; ++m) {
        const int rel = lane * MV + m * V;
        st_ok[m] = rel >= HX && rel < HX + WOUT && xl + m * V < pitch;
    }
    T* const out_base = dst + xl;

    // R[t][slot]: three consecutive rows of time level t (t = 0..K-1).  The roles south/centre/north
    // rotate through the slots with period 3 == RS, so with the row loop unrolled by 3 every index is
    // a compile-time constant and no register is ever copied.
    T R[K][3][MV];
#pragma unroll
    for (int t = 0; t < K; ++t)
#pragma unroll
        for (int b = 0; b < 3; ++b)
#pragma unroll
            for (int i = 0; i < MV; ++i) R[t][b][i] = T(0);

    // One stage = RS(=3) rows through all K levels.  FIX = this stage may touch Dirichlet cells (boundary
    // columns in this warp's strip, or a boundary row within reach); only then are the pass-through
    // selects compiled in.  Interior stages — almost all of them — run the select-free copy.
    auto stage_rows = [&](auto fix_tag, const T* stage, int gy0) {
        constexpr bool FIX = decltype(fix_tag)::value;
#pragma unroll
        for (int ph = 0; ph < 3; ++ph) {
            const int so = ph % 3, ce = (ph + 1) % 3, no = (ph + 2) % 3;   // slots: south, centre, north
            {
                const T* row = stage + ph * WIN;
#pragma unroll
                for (int m = 0; m < M; ++m) {
                    const P v = *reinterpret_cast<const P*>(row + m * V);
#pragma unroll
                    for (int i = 0; i < V; ++i) R[0][no][m * V + i] = v.v[i];
                }
            }
            const int gy = gy0 + ph;               // level-0 row that just arrived
            T outrow[MV];
#pragma unroll
            for (int t = 1; t <= K; ++t) {
                const T* c = R[t - 1][ce];
                const T* s = R[t - 1][so];
                const T* n = R[t - 1][no];
                T* nw = (t < K) ? R[t < K ? t : 0][no] : outrow;
                const T left = __shfl_up_sync(0xffffffffu, c[MV - 1], 1);
                const T right = __shfl_down_sync(0xffffffffu, c[0], 1);
#pragma unroll
                for (int i = 0; i < MV; ++i) {
                    const T w = (i == 0) ? left : c[i - 1];
                    const T e = (i == MV - 1) ? right : c[i + 1];
                    nw[i] = stencil<T, ORDER, FMA>(c[i], w, e, s[i], n[i], rx, ry);
                }
                if constexpr (FIX) {
                    const int r = gy - t;          // level-t row produced now
                    const bool rowfix = (r == ybc0) || (r == ybc1);
#pragma unroll
                    for (int i = 0; i < MV; ++i) nw[i] = (rowfix || colfix[i]) ? c[i] : nw[i];
                }
            }
            const int o = gy - K;                  // level-K row, complete
            if (o >= ys && o < ye) {
                T* orow = out_base + (long)o * pitch;
#pragma unroll
                for (int m = 0; m < M; ++m) {
                    if (st_ok[m]) {
                        P out;
#pragma unroll
                        for (int i = 0; i < V; ++i) out.v[i] = outrow[m * V + i];
                        *reinterpret_cast<P*>(orow + m * V) = out;
                        if (dst2)   // peer store into the neighbour GPU's ghost row (see kernels_tma.cuh)
                            *reinterpret_cast<P*>(dst2 + (long)(o + shift2) * pitch + xl + m * V) = out;
                    }
                }
            }
        }
    };

    for (int j = 0; j < nst; ++j) {
        mbar_wait(&full[j % S], (unsigned)((j / S) & 1));
        const T* stage = ring + (size_t)(j % S) * RS * WIN + lane * MV;
        const int gy0 = ys - K + j * RS;
        // rows produced by this stage span gy0-K .. gy0+RS-2
        const bool fix = warp_fix || (ybc0 >= gy0 - K && ybc0 <= gy0 + RS) || (ybc1 >= gy0 - K && ybc1 <= gy0 + RS);
        if (fix) stage_rows(std::true_type{}, stage, gy0);
        else stage_rows(std::false_type{}, stage, gy0);
        __syncwarp();                              // stage j fully consumed (its values are in registers)
        if (lane == 0 && j + S < nst) issue(j + S);
    }
    if (!pdl_early) pdl_launch_dependents();
}

// ------------------------------------------------------------------------------------ host side
struct TemporalState {
    CUtensorMap map[2];
    int k = 0, m = 0;
    bool ready = false;
};

inline int temporal_setup(TemporalState& st, void* const buf[2], bool f32, int rows, long pitch, int k,
                          std::string& err) {
    const int V = f32 ? 4 : 2;
    // cells per lane = M * V
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

template <typename T, int ORDER, bool FMA, int K, int M>
inline cudaError_t launch_temporal_rows_m(const TemporalState& st, int which, T* dst, int nx, long pitch, int yb,
                                          int ye, int ybc0, int ybc1, T rx, T ry, int sm_count, cudaStream_t s, int kflags,
                                          T* dst2, int shift2) {
    constexpr int V = Pack<T>::V;
    constexpr int WIN = 32 * M * V, HX = TemporalGeom<T, K>::HX, WOUT = WIN - 2 * HX;
    constexpr size_t smem = (size_t)kTempWarps * kTempStages * kTempRS * WIN * sizeof(T) +
                            (size_t)kTempWarps * kTempStages * sizeof(uint64_t);
    auto k = heat_temporal_kernel<T, ORDER, FMA, K, M>;
    cudaError_t e = ensure_dynamic_smem<heat_temporal_kernel<T, ORDER, FMA, K, M>>(smem);
    if (e != cudaSuccess) return e;
    const int rows = ye - yb;
    const int nstrips = (int)((pitch + WOUT - 1) / WOUT);
    int seg = 0;
    if (const char* v = getenv("HEAT_TEMPORAL_SEG")) seg = atoi(v);
    if (seg <= 0) {
        // Work items are warps; about 12 are resident per SM (3 CTAs, shared-memory bound).  As in the CTA-level
        // kernel, short segments issued strip-fastest sweep the field as a thin wavefront, which DRAM likes (16384^2
        // fp64 k=2: 1024-row segments 641 GLUPS, 44 -> 790, 14 -> 856, 11 -> 869, 8 -> 796; profiles/README.md, round 2) — but every
        // segment recomputes 2K rows of pipeline fill, so the best length grows with K: k=4 fp64 peaks at ~160 rows
        // (1492 GLUPS), fp32 at ~88.  That needs a grid of several waves; smaller grids keep whole waves instead.
        const int resident = sm_count * (M == 1 ? 20 : 12);
        const int pref = K == 1 ? 14 : K == 2 ? (sizeof(T) == 8 ? 11 : 14) : (sizeof(T) == 8 ? 160 : 88);
        if ((long)nstrips * ((rows + pref - 1) / pref) >= 3L * resident) seg = pref;
        // Otherwise prefer a whole number of waves (3, 2 or 1) with segments long enough that the 2K re-read rows
        // stay cheap; a grid too small for that gets exactly one wave, whatever the segment length.
        for (int w = 3; w >= 1 && seg <= 0; --w) {
            const int nseg = std::max(1, w * resident / nstrips);
            const int cand = (rows + nseg - 1) / nseg;
            if (cand >= 48 * K) seg = cand;
        }
        if (seg <= 0) {
            const int nseg = std::max(1, resident / nstrips);
            seg = std::max(8 * K, (rows + nseg - 1) / nseg);
        }
    }
    seg = std::min(seg, rows);
    const int nsegs = (rows + seg - 1) / seg;
    const int nitems = nstrips * nsegs;
    const int grid = (nitems + kTempWarps - 1) / kTempWarps;
    // signal dependents at CTA start only when the grid spans several waves: early-placed CTAs of the next
    // launch would otherwise skew the balance of a grid that does not even fill the machine once
    const int pdl_early = grid >= 2 * sm_count * 4 ? 1 : 0;
    return launch_pdl(k, dim3(grid), dim3(kTempWarps * 32), smem, s, sweep_tail_window(dst, pitch, yb, ye, kflags), st.map[which], dst, nx, pitch, yb, ye, seg,
                      nstrips, nitems, ybc0, ybc1, rx, ry, pdl_early, kflags, dst2, shift2);
}

template <typename T, int ORDER, bool FMA, int K>
inline cudaError_t launch_temporal_rows(const TemporalState& st, int which, T* dst, int nx, long pitch, int yb,
                                        int ye, int ybc0, int ybc1, T rx, T ry, int sm_count, cudaStream_t s, int kflags,
                                        T* dst2 = nullptr, int shift2 = 0) {
    // M = 1 (half the registers per level, twice the warps, narrower strips) was measured and rejected: 16384^2
    // fp64 k=4 1326 vs 1492 GLUPS, fp32 k=4 2414 vs 2520, k=2 796 vs 869 (profiles/r02_sweep_temporal_m.txt)
    return launch_temporal_rows_m<T, ORDER, FMA, K, 2>(st, which, dst, nx, pitch, yb, ye, ybc0, ybc1, rx, ry, sm_count, s, kflags, dst2, shift2);
}

}  // namespace heat
