// kernels_stream1.cuh — the row-marching kernel with one cell per lane (no alignment requirements).
// Shared by the handle API (STREAM on buffers that are not 16-byte friendly) and by the legacy launcher.
#pragma once
#include "stencil_ops.cuh"

namespace heat {

// ------------------------------------------------------------------------------------ STREAM, scalar
// The same marching scheme with one cell per lane and 8-/4-byte loads, for buffers that are not 16-byte
// friendly (odd pitch: caller-owned tensors, the reference's ghosted (nx+2) x (ny+2) arrays).  A warp-row is
// one coalesced 256-byte (fp64) request; nothing outside columns [x_begin, x_end) x rows [y_begin, y_end)
// is written.
template <typename T, int ORDER, bool FMA, int U>
__global__ void __launch_bounds__(256)
heat_stream1_kernel(const T* __restrict__ src, T* __restrict__ dst, int x_begin, int x_end, long pitch,
                    int y_begin, int y_end, int rows_per_block, T rx, T ry) {
    const int lane = threadIdx.x & 31;
    const long x = (long)x_begin + (long)blockIdx.x * blockDim.x + threadIdx.x;
    const bool active = x < x_end;
    const int yb = y_begin + blockIdx.y * rows_per_block;
    const int ye = min(yb + rows_per_block, y_end);
    if (yb >= ye) return;
    // lane 0 also fetches the cell left of the warp's span, lane 31 (or the last active lane's right
    // neighbour) the cell right of it; x_begin >= 1 and x_end <= nx - 1 keep both inside the row
    const long xe = (lane == 0) ? x - 1 : x + 1;
    const bool edge_lane = active && (lane == 0 || lane == 31);

    T win[U + 2], edg[U + 2];
    auto load_row = [&](int slot, int y) {
        const T* row = src + (long)y * pitch;
        win[slot] = active ? row[x] : T(0);
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
                const T c = win[u + 1];
                T w = __shfl_up_sync(0xffffffffu, c, 1);
                T e = __shfl_down_sync(0xffffffffu, c, 1);
                if (lane == 0) w = edg[u + 1];
                if (lane == 31) e = edg[u + 1];
                // the last active lane of a partial warp has no active right neighbour: fetch it directly
                if (active && lane != 31 && x + 1 >= x_end) e = src[(long)(y + u) * pitch + x + 1];
                if (active) dst[(long)(y + u) * pitch + x] = stencil<T, ORDER, FMA>(c, w, e, win[u], win[u + 2], rx, ry);
            }
        }
        win[0] = win[U];
        win[1] = win[U + 1];
        edg[0] = edg[U];
        edg[1] = edg[U + 1];
    }
}

}  // namespace heat
