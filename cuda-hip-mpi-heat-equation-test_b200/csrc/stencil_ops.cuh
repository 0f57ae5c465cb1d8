// stencil_ops.cuh — the per-cell arithmetic of the FTCS 5-point update, in each of the reference's
// three evaluation orders, with and without FMA contraction.
//
// Bit-exactness: in the exact mode every operation is an explicitly rounded intrinsic
// (__dadd_rn / __dmul_rn / __fadd_rn / ...), which nvcc never fuses or re-associates, whatever
// -fmad says.  That reproduces gfortran -O0 (fortran/serial/makefile:2) and NumPy ufunc
// evaluation (python/serial/heat.py:50-54) operation for operation.
#pragma once
#include <cuda_runtime.h>

namespace heat {

enum : int { ORDER_FORTRAN = 0, ORDER_PYCUDA = 1, ORDER_PYSERIAL = 2 };

template <typename T, bool FMA> struct Ops;

template <> struct Ops<double, false> {
    static __device__ __forceinline__ double add(double a, double b) { return __dadd_rn(a, b); }
    static __device__ __forceinline__ double sub(double a, double b) { return __dsub_rn(a, b); }
    static __device__ __forceinline__ double mul(double a, double b) { return __dmul_rn(a, b); }
};
template <> struct Ops<float, false> {
    static __device__ __forceinline__ float add(float a, float b) { return __fadd_rn(a, b); }
    static __device__ __forceinline__ float sub(float a, float b) { return __fsub_rn(a, b); }
    static __device__ __forceinline__ float mul(float a, float b) { return __fmul_rn(a, b); }
};

// c: centre, w/e: x-1/x+1 (contiguous index), s/n: y-1/y+1 (slow index).
template <typename T, int ORDER, bool FMA>
__device__ __forceinline__ T stencil(T c, T w, T e, T s, T n, T rx, T ry) {
    if constexpr (!FMA) {
        using O = Ops<T, false>;
        if constexpr (ORDER == ORDER_FORTRAN) {
            // ref: fortran/serial/heat.f90:66   T_old(j,k)+r*(T_old(j+1,k)+T_old(j,k+1)+T_old(j-1,k)+T_old(j,k-1)-4*T_old(j,k))
            return O::add(c, O::mul(rx, O::sub(O::add(O::add(O::add(e, n), w), s), O::mul(T(4), c))));
        } else if constexpr (ORDER == ORDER_PYCUDA) {
            // ref: python/cuda/cuda.py:78       u1 + (r * (ul + ur + utop + ubottom - (4 * u1)))
            return O::add(c, O::mul(rx, O::sub(O::add(O::add(O::add(w, e), n), s), O::mul(T(4), c))));
        } else {
            // ref: python/serial/heat.py:50-54  un + cx*(E - 2*un + W) + cy*(N - 2*un + S)
            T two_c = O::mul(T(2), c);
            T ax = O::mul(rx, O::add(O::sub(e, two_c), w));
            T ay = O::mul(ry, O::add(O::sub(n, two_c), s));
            return O::add(O::add(c, ax), ay);
        }
    } else {
        // FMA-contracted forms: same mathematical expression, 5 (or 6) issue slots instead of 7 (10).
        if constexpr (ORDER == ORDER_FORTRAN) {
            T t = ((e + n) + w) + s;
            return fma(rx, fma(T(-4), c, t), c);
        } else if constexpr (ORDER == ORDER_PYCUDA) {
            T t = ((w + e) + n) + s;
            return fma(rx, fma(T(-4), c, t), c);
        } else {
            T ax = fma(T(-2), c, e) + w;
            T ay = fma(T(-2), c, n) + s;
            return fma(ry, ay, fma(rx, ax, c));
        }
    }
}

// 16-byte vector of T: 2 doubles or 4 floats.  One lane moves one Pack per row.
template <typename T> struct PackInfo;
template <> struct PackInfo<double> { static constexpr int V = 2; };
template <> struct PackInfo<float>  { static constexpr int V = 4; };

template <typename T> struct alignas(16) Pack {
    static constexpr int V = PackInfo<T>::V;
    T v[V];
};

}  // namespace heat
