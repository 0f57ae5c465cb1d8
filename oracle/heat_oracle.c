/* CPU oracle (plain C) for the FTCS 5-point heat stencil — TEST INFRASTRUCTURE ONLY.
 *
 * A restatement of the reference's CPU time-march, used to CHECK the CUDA library and as the
 * timed "port" CPU baseline of bench.py.  Nothing in the product path links or calls it.
 * Build: see oracle/Makefile (gcc -O2 -ffp-contract=off -fno-fast-math; -fopenmp variant).
 * Parity pinning: checked bit-for-bit against oracle/heat_oracle.py, which in turn is pinned
 * against the reference's own outputs frozen in tests/golden/ (the reference has no tests).
 *
 * Layout: row-major [ny][pitch], x contiguous — byte-identical to Fortran T(i,j), i contiguous.
 *
 * Follows (paths relative to the reference root):
 *   fortran/serial/heat.f90:61-69   copy T_old = T, then the double loop          (order 0)
 *   python/cuda/cuda.py:70-79       ul + ur + utop + ubottom                      (order 1)
 *   python/serial/heat.py:49-54     un + cx*(E-2C+W) + cy*(N-2C+S)                (order 2)
 *   fortran/serial/heat.f90:15-17,28-48,59   parameters, coordinates, hat IC
 */
#include <stddef.h>
#include <string.h>
#include <stdlib.h>

/* Contracted evaluation (contract != 0), orders 0 and 1 only: what the reference's own GPU toolchains make
 * of the same source text — nvcc (-fmad=true by default, PyCUDA SourceModule, python/cuda/cuda.py:86) and
 * hipcc (-ffp-contract=fast, fortran/hip/Makefile) fuse  c + r*(t - 4*c)  into  fma(r, fma(-4, c, t), c).
 * Pinned against oracle/_ref/lib*ref_fma.so = the reference kernels compiled with gcc -mfma
 * -ffp-contract=fast (the _fma arrays under tests/golden/). */
#define FMA_double(a, b, c) __builtin_fma((a), (b), (c))
#define FMA_float(a, b, c) __builtin_fmaf((a), (b), (c))

#define DEFINE_STEP(T, SUF)                                                                       \
    static void step_##SUF(T *restrict dst, const T *restrict src, int nx, int ny, size_t pitch,  \
                           T rx, T ry, int order, int contract) {                                 \
        _Pragma("omp parallel for schedule(static)")                                              \
        for (int y = 1; y < ny - 1; ++y) {                                                        \
            const T *c = src + (size_t)y * pitch, *n = c + pitch, *s = c - pitch;                 \
            T *o = dst + (size_t)y * pitch;                                                       \
            if (contract) {                                                                       \
                for (int x = 1; x < nx - 1; ++x) {                                                \
                    T t = order == 0 ? ((c[x + 1] + n[x]) + c[x - 1]) + s[x]                      \
                                     : ((c[x - 1] + c[x + 1]) + n[x]) + s[x];                     \
                    o[x] = FMA_##T(rx, FMA_##T((T)-4, c[x], t), c[x]);                            \
                }                                                                                 \
            } else if (order == 0) {                                                                     \
                for (int x = 1; x < nx - 1; ++x)                                                  \
                    o[x] = c[x] + rx * ((((c[x + 1] + n[x]) + c[x - 1]) + s[x]) - (T)4 * c[x]);   \
            } else if (order == 1) {                                                              \
                for (int x = 1; x < nx - 1; ++x)                                                  \
                    o[x] = c[x] + rx * ((((c[x - 1] + c[x + 1]) + n[x]) + s[x]) - (T)4 * c[x]);   \
            } else {                                                                              \
                for (int x = 1; x < nx - 1; ++x)                                                  \
                    o[x] = (c[x] + rx * ((c[x + 1] - (T)2 * c[x]) + c[x - 1])) +                  \
                           ry * ((n[x] - (T)2 * c[x]) + s[x]);                                    \
            }                                                                                     \
        }                                                                                         \
    }                                                                                             \
    /* steps Jacobi steps in place on u; returns 0, or -1 on bad arguments / no memory. */        \
    int oracle_march_ex_##SUF(T *u, int nx, int ny, size_t pitch, T rx, T ry, int order,          \
                              int steps, int contract) {                                          \
        if (!u || nx < 1 || ny < 1 || pitch < (size_t)nx || order < 0 || order > 2) return -1;    \
        if (contract && order == 2) return -1;   /* NumPy never contracts (heat.py:50-54) */      \
        if (steps <= 0 || nx < 3 || ny < 3) return 0;                                             \
        size_t bytes = (size_t)ny * pitch * sizeof(T);                                            \
        T *old = (T *)malloc(bytes);                                                              \
        if (!old) return -1;                                                                      \
        for (int t = 0; t < steps; ++t) {                                                         \
            memcpy(old, u, bytes);             /* T_old = T   (heat.f90:63, heat.py:49) */        \
            step_##SUF(u, old, nx, ny, pitch, rx, ry, order, contract);                           \
        }                                                                                         \
        free(old);                                                                                \
        return 0;                                                                                 \
    }                                                                                             \
    int oracle_march_##SUF(T *u, int nx, int ny, size_t pitch, T rx, T ry, int order,             \
                           int steps) {                                                           \
        return oracle_march_ex_##SUF(u, nx, ny, pitch, rx, ry, order, steps, 0);                  \
    }

DEFINE_STEP(double, f64)
DEFINE_STEP(float, f32)

/* Threads the OpenMP build will use for a march (1 in the serial build); n > 0 sets the count first.
 * bench.py calls it so that the CPU arm's thread count is explicit and reported, whatever
 * OMP_NUM_THREADS the launcher exported. */
#ifdef _OPENMP
#include <omp.h>
int oracle_threads(int n) {
    if (n > 0) omp_set_num_threads(n);
    return omp_get_max_threads();
}
#else
int oracle_threads(int n) { (void)n; return 1; }
#endif

/* delta, dt, r of fortran/serial/heat.f90:15,17,59. */
void oracle_derive_fortran(int n, double sigma, double nu, double dom_len, double *delta,
                           double *dt, double *r) {
    double d = dom_len / (double)(float)(n - 1);
    double t = (sigma * (d * d)) / nu;
    *delta = d;
    *dt = t;
    *r = (nu * t) / (d * d);
}

/* Hat IC of fortran/serial/heat.f90:28-48 (y_hi = 1.5) or cuda_cuf/heat.F90:86 (y_hi = 1.0). */
int oracle_ic_fortran(double *u, int n, size_t pitch, double dom_len, double y_hi) {
    if (!u || n < 2 || pitch < (size_t)n) return -1;
    double delta = dom_len / (double)(float)(n - 1);
    double *c = (double *)malloc(sizeof(double) * (size_t)n);
    if (!c) return -1;
    c[0] = 0.0;
    for (int i = 1; i < n - 1; ++i) c[i] = c[i - 1] + delta;
    c[n - 1] = dom_len;
    for (int j = 0; j < n; ++j)
        for (int i = 0; i < n; ++i)
            u[(size_t)j * pitch + i] =
                (c[i] <= 1.5 && c[i] >= 0.5 && c[j] <= y_hi && c[j] >= 0.5) ? 2.0 : 1.0;
    free(c);
    return 0;
}

/* Sum of all cells in memory order with a plain double accumulator — the commented-out debug
 * check of fortran/mpi+cuda/heat.F90:267-272 (there i-inner is NOT the order used: the
 * reference sums j-inner; for a property test any fixed order will do and we document ours). */
double oracle_sum_f64(const double *u, int nx, int ny, size_t pitch) {
    double s = 0.0;
    for (int y = 0; y < ny; ++y)
        for (int x = 0; x < nx; ++x) s += u[(size_t)y * pitch + x];
    return s;
}
