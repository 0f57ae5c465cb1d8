// kernels_persist.cuh — launch-free stepping for fields that fit in the GPU's shared memory as a whole (sm_100a).
//
// Between the one-CTA RESIDENT kernel (fields <= 200 KB: the reference's 31^2 and 100^2 cases) and the TMA
// kernels (fields that live in HBM) sit grids like the reference's shipped 1024^2 deck: 8 MB per copy — too big
// for one SM, far too small to amortise one kernel launch per time step.  This kernel is launched ONCE for all
// nt steps (cooperative launch: every CTA is resident) with one CTA per SM.  The field is cut into row slabs,
// one per CTA, and each CTA keeps its slab — plus G ghost rows on either side — in both halves of a ping-pong
// pair in its own shared memory for the whole march; HBM is read once on the way in and written once on the way
// out.  A CTA advances G steps on a region that shrinks by one row per step (the multi-GPU ghost-depth scheme of
// heat_api.cu in miniature), then trades G rows with its two neighbours through a mailbox in global memory (L2):
//   write my edge rows -> bar.sync -> st.release.gpu flag[p] = epoch;   spin ld.acquire.gpu on flag[p-1], flag[p+1];
//   read their rows with ld.global.cg.
// Only neighbours wait for each other — there is no grid-wide barrier — and the mailbox is double-buffered by the
// epoch's parity, so a fast CTA can never overwrite rows a slow neighbour has not read yet (it cannot get two
// epochs ahead: it needs that neighbour's rows of the epoch in between).
// Inside a step a lane owns one 16-byte column group and marches down its rows with a 3-row register window,
// x-neighbours by warp shuffle — the STREAM kernel's inner loop reading shared memory instead of HBM.
// Per cell and step: the same stencil<T, ORDER, FMA> call as everywhere else, so the bits are the same.
#pragma once
#include <cuda_runtime.h>

#include <algorithm>

#include "stencil_ops.cuh"

namespace heat {

constexpr int kPersistThreads = 512;
constexpr int kPersistMaxG = 4;

struct PersistPlan {
    bool ok = false;
    int P = 0;          // CTAs = row slabs
    int G = 0;          // ghost depth = steps per neighbour exchange
    int spitch = 0;     // shared-memory row pitch, elements
    int lmax = 0;       // rows held per CTA (largest slab + 2G)
    size_t smem = 0;
};

__device__ __forceinline__ void st_release_gpu(unsigned* p, unsigned v) {
    asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned ld_acquire_gpu(const unsigned* p) {
    unsigned v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
template <typename T>
__device__ __forceinline__ Pack<T> ld_cg(const T* p) {          // L2 only: another SM wrote these bytes
    Pack<T> o;
    const float4 f = __ldcg(reinterpret_cast<const float4*>(p));
    *reinterpret_cast<float4*>(&o) = f;
    return o;
}

template <typename T, int ORDER, bool FMA>
__global__ void __launch_bounds__(kPersistThreads, 1)
heat_persist_kernel(const T* __restrict__ src, T* __restrict__ dst, int nx, int ny, long pitch, int nt, int G,
                    int spitch, int lmax, T rx, T ry, T* __restrict__ mailbox, unsigned* __restrict__ flags,
                    unsigned epoch0) {
    constexpr int V = Pack<T>::V;
    using Pk = Pack<T>;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    T* cur = reinterpret_cast<T*>(smem_raw);
    T* nxt = cur + (size_t)lmax * spitch;

    const int P = gridDim.x, p = blockIdx.x, tid = threadIdx.x, lane = tid & 31;
    const int base = ny / P, rem = ny % P;
    const int nr = base + (p < rem ? 1 : 0);
    const int r0 = p * base + min(p, rem);
    const int L = nr + 2 * G;                       // local row l <-> global row r0 - G + l
    const int nvec = (nx + V - 1) / V;              // 16-byte groups per row (the global pitch covers the last one)
    // local rows that exist in the field and local rows that may be updated (global rows 1 .. ny-2)
    const int l_first = max(0, G - r0), l_last = min(L - 1, ny - 1 - (r0 - G));
    const int lo_phys = 1 - (r0 - G), hi_phys = (ny - 2) - (r0 - G);

    // ---- load: own rows + ghosts into BOTH halves (the Dirichlet ring must be in both); a row's shared-memory
    // columns are shifted by V so that column -1 of the first group is addressable (zero pad)
    for (int i = tid; i < lmax * spitch; i += kPersistThreads) { cur[i] = T(0); nxt[i] = T(0); }
    __syncthreads();
    for (int i = tid; i < (l_last - l_first + 1) * nvec; i += kPersistThreads) {
        const int l = l_first + i / nvec, v = i % nvec;
        const Pk val = *reinterpret_cast<const Pk*>(src + (long)(r0 - G + l) * pitch + (long)v * V);
        *reinterpret_cast<Pk*>(cur + (size_t)l * spitch + V + v * V) = val;
        *reinterpret_cast<Pk*>(nxt + (size_t)l * spitch + V + v * V) = val;
    }
    __syncthreads();

    // thread -> (row group, column group): as many row groups as the block has room for
    const int lpr = min((nvec + 31) / 32 * 32, kPersistThreads);    // lanes working on one row (whole warps)
    const int groups = kPersistThreads / lpr;
    const int grp = tid / lpr, lir = tid % lpr;

    auto substep = [&](int s) {
        const int lo = max(s + 1, lo_phys), hi = min(L - 2 - s, hi_phys);      // inclusive
        const int nrows = hi - lo + 1;
        if (nrows > 0 && grp < groups) {
            const int per = (nrows + groups - 1) / groups;
            const int yb = lo + grp * per, ye = min(yb + per, hi + 1);
            for (int vb = 0; vb < nvec; vb += lpr) {
                const int v = vb + lir;
                const bool active = v < nvec;
                const int sx = V + (active ? v : 0) * V;             // column inside a shared-memory row
                const int sxe = (lane == 0) ? sx - 1 : sx + V;       // warp-edge lanes read their neighbour cell directly
                if (yb < ye) {
                    auto ldrow = [&](int y) { return *reinterpret_cast<const Pk*>(cur + (size_t)y * spitch + sx); };
                    // 4-row register window: the row after next is requested before the current row's arithmetic
                    // starts, so the shared-memory latency hides behind it (rows up to L-1 exist, y + 2 <= hi + 2 <= L)
                    Pk s_ = ldrow(yb - 1), c_ = ldrow(yb), n_ = ldrow(yb + 1);
                    T ec = (lane == 0 || lane == 31) ? cur[(size_t)yb * spitch + sxe] : T(0);
                    for (int y = yb; y < ye; ++y) {
                        const int y2 = min(y + 2, L - 1);
                        const Pk n2 = ldrow(y2);
                        const T en = (lane == 0 || lane == 31) ? cur[(size_t)(y + 1) * spitch + sxe] : T(0);
                        T left = __shfl_up_sync(0xffffffffu, c_.v[V - 1], 1);
                        T right = __shfl_down_sync(0xffffffffu, c_.v[0], 1);
                        if (lane == 0) left = ec;
                        if (lane == 31) right = ec;
                        Pk o;
#pragma unroll
                        for (int i = 0; i < V; ++i) {
                            const T w = (i == 0) ? left : c_.v[i - 1];
                            const T e = (i == V - 1) ? right : c_.v[i + 1];
                            const T r = stencil<T, ORDER, FMA>(c_.v[i], w, e, s_.v[i], n_.v[i], rx, ry);
                            const int x = v * V + i;
                            o.v[i] = (x >= 1 && x <= nx - 2) ? r : c_.v[i];
                        }
                        if (active) *reinterpret_cast<Pk*>(nxt + (size_t)y * spitch + sx) = o;
                        s_ = c_;
                        c_ = n_;
                        n_ = n2;
                        ec = en;
                    }
                }
            }
        }
        __syncthreads();
        T* t = cur; cur = nxt; nxt = t;
    };

    const size_t mrow = (size_t)nvec * V;                              // mailbox row length, elements
    auto mbox = [&](unsigned e, int slab, int side) {                  // G rows: side 0 = slab's top rows, 1 = bottom rows
        return mailbox + ((((size_t)(e & 1u) * P + slab) * 2 + side) * kPersistMaxG) * mrow;
    };

    unsigned epoch = epoch0;
    for (int t = 0; t < nt;) {
        const int adv = min(G, nt - t);
        for (int s = 0; s < adv; ++s) substep(s);
        t += adv;
        if (t >= nt) break;
        // ---- trade G rows with the neighbours through L2
        ++epoch;
        for (int i = tid; i < 2 * G * nvec; i += kPersistThreads) {
            const int side = i / (G * nvec), j = (i / nvec) % G, v = i % nvec;
            const int l = side == 0 ? G + j : nr + j;                  // my first / last G own rows
            *reinterpret_cast<Pk*>(mbox(epoch, p, side) + (size_t)j * mrow + (size_t)v * V) =
                *reinterpret_cast<const Pk*>(cur + (size_t)l * spitch + V + v * V);
        }
        __syncthreads();
        if (tid == 0) st_release_gpu(flags + p, epoch);      // release: everything the CTA wrote before the barrier is visible first
        if (tid == 32 && p > 0) while ((int)(ld_acquire_gpu(flags + p - 1) - epoch) < 0) { }
        if (tid == 64 && p + 1 < P) while ((int)(ld_acquire_gpu(flags + p + 1) - epoch) < 0) { }
        __syncthreads();
        for (int i = tid; i < 2 * G * nvec; i += kPersistThreads) {
            const int side = i / (G * nvec), j = (i / nvec) % G, v = i % nvec;
            if (side == 0 ? p == 0 : p + 1 == P) continue;             // no neighbour on that side
            // my top ghosts <- the bottom rows of the slab above; my bottom ghosts <- the top rows of the slab below
            const T* from = side == 0 ? mbox(epoch, p - 1, 1) : mbox(epoch, p + 1, 0);
            const int l = side == 0 ? j : G + nr + j;
            *reinterpret_cast<Pk*>(cur + (size_t)l * spitch + V + v * V) = ld_cg<T>(from + (size_t)j * mrow + (size_t)v * V);
        }
        __syncthreads();
    }

    // ---- store own rows
    for (int i = tid; i < nr * nvec; i += kPersistThreads) {
        const int l = G + i / nvec, v = i % nvec;
        *reinterpret_cast<Pk*>(dst + (long)(r0 + l - G) * pitch + (long)v * V) =
            *reinterpret_cast<const Pk*>(cur + (size_t)l * spitch + V + v * V);
    }
}

// ------------------------------------------------------------------------------------ host side
// Slabs and ghost depth for an nx x ny field of `esize`-byte cells on `sm_count` SMs with `smem_cap` bytes of
// dynamic shared memory per CTA; ok = false if no choice fits.
inline PersistPlan persist_plan(int nx, int ny, size_t esize, int sm_count, size_t smem_cap) {
    PersistPlan best;
    if (nx < 3 || ny < 3) return best;
    const int V = (int)(16 / esize);
    const int spitch = (nx + V - 1) / V * V + 2 * V;
    double best_cost = 0;
    for (int G = 1; G <= kPersistMaxG; ++G) {
        const int P = std::min(sm_count, ny / G);            // every slab must own at least G rows to give away
        if (P < 2) continue;
        const int lmax = ny / P + (ny % P ? 1 : 0) + 2 * G;
        const size_t smem = 2 * (size_t)lmax * spitch * esize;
        if (smem > smem_cap) continue;
        // rows a CTA computes per step (own + average redundant ghost rows) at ~8 cells/clk, plus the neighbour
        // exchange (~3000 clk) once per G steps
        const double cost = (ny / (double)P + (G - 1)) * nx / 8.0 + 3000.0 / G;
        if (!best.ok || cost < best_cost) {
            best.ok = true; best.P = P; best.G = G; best.spitch = spitch; best.lmax = lmax; best.smem = smem;
            best_cost = cost;
        }
    }
    return best;
}

inline size_t persist_mailbox_bytes(int nx, size_t esize, int sm_count) {
    const int V = (int)(16 / esize);
    return (size_t)2 * sm_count * 2 * kPersistMaxG * ((size_t)(nx + V - 1) / V * V) * esize;
}

}  // namespace heat
