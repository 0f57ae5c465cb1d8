// kernels_basic.cuh — SIMPLE, STREAM and RESIDENT kernels of the time-march (sm_100a).
//
// All kernels read `src` and write `dst` (ping-pong; the reference instead copies T_old = T every
// step, fortran/cuda_kernel/heat.F90:32, and moves 32 B/LUP).  Rows [y_begin, y_end) of the local
// buffer are updated; columns 1..nx-2 get the stencil.  Row-range arguments let the same kernels
// serve whole fields, multi-GPU slabs with shrinking ghost regions, and boundary strips.
#pragma once
#include "stencil_ops.cuh"
#include "kernels_stream1.cuh"

namespace heat {

// ------------------------------------------------------------------------------------ SIMPLE
// One thread per cell, 32x8 blocks — the reference's own mapping (python/cuda/cuda.py:99,
// fortran/cuda_kernel/heat.F90:23) minus its out-of-bounds reads.  Works for any pitch/alignment;
// it is the fallback for caller-owned buffers that are not 16-byte friendly.
template <typename T, int ORDER, bool FMA>
__global__ void __launch_bounds__(256)
heat_simple_kernel(const T* __restrict__ src, T* __restrict__ dst, int nx, long pitch,
                   int y_begin, int y_end, T rx, T ry) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x + 1;
    const int y = y_begin + blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= nx - 1 || y >= y_end) return;
    const T* p = src + (long)y * pitch + x;
    dst[(long)y * pitch + x] = stencil<T, ORDER, FMA>(p[0], p[-1], p[1], p[-pitch], p[pitch], rx, ry);
}

// ------------------------------------------------------------------------------------ STREAM
// Each lane owns one 16-byte column group (2 doubles / 4 floats) and marches down `rows_per_block`
// rows keeping a 3-row window in registers, so every row is loaded from L2/HBM once per block
// (+2 halo rows per chunk).  x-neighbours come from the adjacent lanes by warp shuffle; only the
// two lanes at the warp's edges issue one extra scalar load per row.  U rows are loaded back to
// back before they are consumed, to keep U x 512 B per warp in flight.
//
// Requirements: src/dst 16-byte aligned, pitch % V == 0.  Boundary columns 0 and nx-1 (and any
// pitch padding) are written back with their own old value, which is what they hold in both
// ping-pong buffers anyway, so full vector stores are safe.
template <typename T, int ORDER, bool FMA, int U>
__global__ void __launch_bounds__(256)
heat_stream_kernel(const T* __restrict__ src, T* __restrict__ dst, int nx, long pitch,
                   int y_begin, int y_end, int rows_per_block, T rx, T ry) {
    constexpr int V = Pack<T>::V;
    using P = Pack<T>;
    const int lane = threadIdx.x & 31;
    const long x0 = ((long)blockIdx.x * blockDim.x + threadIdx.x) * V;
    const bool active = x0 < pitch;
    const int yb = y_begin + blockIdx.y * rows_per_block;
    const int ye = min(yb + rows_per_block, y_end);
    if (yb >= ye) return;

    // lane 0 fetches the cell left of the warp's span, lane 31 the cell right of it
    const long xe = (lane == 0) ? x0 - 1 : x0 + V;
    const bool edge_lane = active && ((lane == 0 && x0 > 0) || (lane == 31 && xe < nx));

    P win[U + 2];
    T edg[U + 2];
    auto load_row = [&](int slot, int y) {
        const T* row = src + (long)y * pitch;
        if (active) win[slot] = *reinterpret_cast<const P*>(row + x0);
        else {
#pragma unroll
            for (int i = 0; i < V; ++i) win[slot].v[i] = T(0);
        }
        edg[slot] = edge_lane ? row[xe] : T(0);
    };

    load_row(0, yb - 1);
    load_row(1, yb);
    for (int y = yb; y < ye; y += U) {
#pragma unroll
        for (int u = 0; u < U; ++u)
            if (y + 1 + u <= ye) load_row(2 + u, y + 1 + u);
#pragma unroll
        for (int u = 0; u < U; ++u) {
            if (y + u < ye) {
                const P& s = win[u];
                const P& c = win[u + 1];
                const P& n = win[u + 2];
                T left = __shfl_up_sync(0xffffffffu, c.v[V - 1], 1);
                T right = __shfl_down_sync(0xffffffffu, c.v[0], 1);
                if (lane == 0) left = edg[u + 1];
                if (lane == 31) right = edg[u + 1];
                P o;
#pragma unroll
                for (int i = 0; i < V; ++i) {
                    const T w = (i == 0) ? left : c.v[i - 1];
                    const T e = (i == V - 1) ? right : c.v[i + 1];
                    const T r = stencil<T, ORDER, FMA>(c.v[i], w, e, s.v[i], n.v[i], rx, ry);
                    const long x = x0 + i;
                    o.v[i] = (x >= 1 && x <= nx - 2) ? r : c.v[i];
                }
                if (active) *reinterpret_cast<P*>(dst + (long)(y + u) * pitch + x0) = o;
            }
        }
        win[0] = win[U];
        win[1] = win[U + 1];
        edg[0] = edg[U];
        edg[1] = edg[U + 1];
    }
}

// ------------------------------------------------------------------------------------ RESIDENT
// Small fields (the reference's shipped 31x31 and 100x100 cases) are launch-bound: 1000 steps of
// a 100x100 grid are 1000 x ~2 us of launch overhead for ~80 KB of data.  Here ONE CTA keeps both
// ping-pong copies in shared memory and runs all nt steps in one launch with a block barrier
// between steps; HBM is touched once on the way in and once on the way out.
// Dynamic shared memory: 2 * nx * ny * sizeof(T) bytes (<= 227 KB).
template <typename T, int ORDER, bool FMA>
__global__ void __launch_bounds__(1024)
heat_resident_kernel(const T* __restrict__ src, T* __restrict__ dst, int nx, int ny, long pitch,
                     int nt, T rx, T ry) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    T* a = reinterpret_cast<T*>(smem_raw);
    T* b = a + (long)nx * ny;
    const int cells = nx * ny;
    for (int i = threadIdx.x; i < cells; i += blockDim.x) {
        const T v = src[(long)(i / nx) * pitch + (i % nx)];
        a[i] = v;
        b[i] = v;
    }
    __syncthreads();
    const int ix = nx - 2, inner = ix * (ny - 2);
    for (int t = 0; t < nt; ++t) {
        for (int i = threadIdx.x; i < inner; i += blockDim.x) {
            const int p = (i / ix + 1) * nx + (i % ix) + 1;
            b[p] = stencil<T, ORDER, FMA>(a[p], a[p - 1], a[p + 1], a[p - nx], a[p + nx], rx, ry);
        }
        __syncthreads();
        T* tmp = a; a = b; b = tmp;
    }
    for (int i = threadIdx.x; i < cells; i += blockDim.x)
        dst[(long)(i / nx) * pitch + (i % nx)] = a[i];
}

// ------------------------------------------------------------------------------------ checksum
// Deterministic sum: fixed per-block partials (fixed grid), fixed tree inside a block, partials
// added in index order by a second single-block launch.  partials[] is scratch of gridDim.x doubles.
template <typename T>
__global__ void __launch_bounds__(256)
heat_sum_partial_kernel(const T* __restrict__ src, int nx, long pitch, int y_begin, int y_end,
                        double* __restrict__ partials) {
    __shared__ double sh[256];
    double acc = 0.0;
    const long total = (long)(y_end - y_begin) * nx;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        const long y = y_begin + i / nx, x = i % nx;
        acc += (double)src[y * pitch + x];
    }
    sh[threadIdx.x] = acc;
    __syncthreads();
    for (int s = 128; s > 0; s >>= 1) {
        if ((int)threadIdx.x < s) sh[threadIdx.x] += sh[threadIdx.x + s];
        __syncthreads();
    }
    if (threadIdx.x == 0) partials[blockIdx.x] = sh[0];
}

__global__ void heat_sum_final_kernel(const double* __restrict__ partials, int n, double* __restrict__ out) {
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        double s = 0.0;
        for (int i = 0; i < n; ++i) s += partials[i];
        *out = s;
    }
}

}  // namespace heat
