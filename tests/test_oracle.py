"""CPU: pins the oracle (NumPy + C restatements) to outputs of the reference itself (tests/golden,
produced by tests/golden/make_golden.py) and the two restatements to each other, bit for bit."""
import os

import numpy as np
import pytest

from oracle import c_oracle, heat_oracle as ho, run_reference


def bits(a):
    return np.ascontiguousarray(a).view(np.uint64 if a.dtype == np.float64 else np.uint32)


def assert_bit_equal(a, b):
    assert a.shape == b.shape and a.dtype == b.dtype
    assert np.array_equal(bits(a), bits(b)), "max abs diff %g" % np.abs(a - b).max()


# ----------------------------------------------------------------- python/serial (unmodified run)
def test_pyserial_golden_known_answers(golden):
    u = golden("pyserial_31.npz")["u"]
    # SURVEY.md §4(v) / BASELINE.md §4
    assert u.shape == (31, 31) and u.dtype == np.float64
    assert float(u.sum()) == 1041.943124294281
    assert float(u.max()) == 1.8959236145019531
    assert float(u[15, 15]) == 1.3340044021606445
    assert float(u.min()) == 1.0


def test_numpy_oracle_matches_pyserial_golden(golden):
    assert_bit_equal(ho.pyserial_diffuse(31, 31, 10), golden("pyserial_31.npz")["u"])


def test_pyserial_params(golden):
    g = golden("pyserial_31.npz")
    dx, dy, dt, cx, cy = ho.derive_python(31, 31)
    assert dx == float(g["dx"]) and dy == float(g["dy"]) and dt == float(g["dt"])
    assert cx == 0.25 and cy == 0.25


@pytest.mark.skipif(not run_reference.available(), reason="reference tree only in the build container")
def test_reference_run_matches_committed_golden(golden):
    g = run_reference.run_python_serial()
    assert_bit_equal(g["u"], golden("pyserial_31.npz")["u"])
    g["diffuse"](60)                                   # the reference's own function, heat.py:45-58
    assert_bit_equal(g["u"], golden("pyserial_31_diffuse60.npz")["u"])


def test_order_p_pinned_by_rounding_sensitive_reference_run(golden):
    """pyserial_31_diffuse60.npz = the unmodified script, then its own diffuse(60): 72 steps in all, far past
    the ~26-step exact-arithmetic window, so the three op orders give DIFFERENT bits and only order P
    (heat.py:50-54) reproduces the reference."""
    g = golden("pyserial_31_diffuse60.npz")
    assert_bit_equal(g["u10"], golden("pyserial_31.npz")["u"])
    want = g["u"]
    assert_bit_equal(ho.pyserial_diffuse_calls([10, 60]), want)
    assert_bit_equal(ho.pyserial_diffuse_calls([10, 60], march_fn=c_oracle.march), want)
    # the same input through orders F and C: close, but not the reference's bits
    dx, dy, _, cx, cy = ho.derive_python(31, 31)
    v = g["u10"].copy()
    v[int(.5 / dy):int(1 / dy + 1), int(.5 / dx):int(1 / dx + 1)] = 2
    assert_bit_equal(ho.march(v, 61, ho.ORDER_PYSERIAL, cx, cy), want)
    for order in (ho.ORDER_FORTRAN, ho.ORDER_PYCUDA):
        other = ho.march(v, 61, order, cx, cy)
        assert not np.array_equal(bits(other), bits(want))
        assert np.abs(other - want).max() < 1e-14


# ----------------------------------------------------------------- HIP kernels (reference source on host)
@pytest.mark.parametrize("P", [1, 2])
def test_numpy_oracle_matches_hip_slabs(golden, P):
    g = golden("hip_slabs_n100.npz")
    assert float(g["r"]) == ho.derive_fortran(100, 0.25, 0.05, 2.0)[2] == 0.24999999999999997
    slabs = ho.mpi_slabs(100, P, 0.25, 0.05, 2.0, 10)
    for p, s in enumerate(slabs):
        assert_bit_equal(s, g["P%d_rank%d" % (P, p)])


def test_oracles_match_hip_random_field(golden):
    g = golden("hip_slabs_n100.npz")
    r = float(g["r"])
    want = g["rand_u5"]
    assert_bit_equal(ho.march(g["rand_u0"], 5, ho.ORDER_FORTRAN, r), want)
    assert_bit_equal(c_oracle.march(g["rand_u0"], 5, ho.ORDER_FORTRAN, r), want)


def test_mpi_slabs_equal_single_domain():
    """Decomposed run == undecomposed run (what SURVEY §8c asks of the multi-GPU path)."""
    one = ho.mpi_slabs(100, 1, 0.25, 0.05, 2.0, 10)[0]
    for P in (2, 4, 5):
        parts = ho.mpi_slabs(100, P, 0.25, 0.05, 2.0, 10)
        glued = np.concatenate([s[:, 1:-1] for s in parts], axis=1)
        assert_bit_equal(glued, one[:, 1:-1])


# ----------------------------------------------------------------- PyCUDA kernel string on host
@pytest.mark.parametrize("tag", ["r025", "rn100"])
def test_oracles_match_pycuda_kernel(golden, tag):
    g = golden("pycuda_kernel.npz")
    r = float(g["r_" + tag])
    assert_bit_equal(ho.march(g["u0"], 5, ho.ORDER_PYCUDA, r), g["u5_" + tag])
    assert_bit_equal(c_oracle.march(g["u0"], 5, ho.ORDER_PYCUDA, r), g["u5_" + tag])


def test_contracted_reading_of_gpu_kernels(golden):
    """nvcc (-fmad=true, PyCUDA's default) and hipcc (-ffp-contract=fast) fuse c + r*(t - 4c) into
    fma(r, fma(-4, c, t), c).  The *_fma goldens are the reference kernel sources compiled for the host
    with contraction on; the C oracle's `contract` mode must reproduce them, and they must differ from
    the uncontracted goldens exactly when r is not a power of two."""
    h = golden("hip_slabs_n100.npz")
    r = float(h["r"])
    assert_bit_equal(c_oracle.march(h["rand_u0"], 5, ho.ORDER_FORTRAN, r, contract=True), h["rand_u5_fma"])
    assert not np.array_equal(bits(h["rand_u5_fma"]), bits(h["rand_u5"]))
    assert np.max(np.abs(h["rand_u5_fma"] - h["rand_u5"]) / h["rand_u5"]) <= 1e-12
    p = golden("pycuda_kernel.npz")
    for tag, differs in (("r025", False), ("rn100", True)):
        rr = float(p["r_" + tag])
        assert_bit_equal(c_oracle.march(p["u0"], 5, ho.ORDER_PYCUDA, rr, contract=True), p["u5_" + tag + "_fma"])
        assert (not np.array_equal(bits(p["u5_" + tag + "_fma"]), bits(p["u5_" + tag]))) == differs
    u32 = p["u0"].astype(np.float32)
    a = c_oracle.march(u32, 5, ho.ORDER_PYCUDA, np.float32(0.2437), contract=True)
    b = c_oracle.march(u32, 5, ho.ORDER_PYCUDA, np.float32(0.2437))
    assert a.dtype == np.float32 and np.max(np.abs(a - b) / b) <= 1e-5


def test_pycuda_ic_is_uniform():
    """python/cuda/cuda.py:53 — empty slice; the shipped run is a fixed point."""
    u = ho.ic_pycuda(128, 128)
    assert (u == 1.0).all()
    assert_bit_equal(ho.march(u, 3, ho.ORDER_PYCUDA, 0.25), u)


@pytest.mark.skipif(not c_oracle.ref_available(), reason="oracle/_ref only where the reference tree exists")
def test_ref_builds_match_goldens(golden):
    g = golden("pycuda_kernel.npz")
    u = g["u0"]
    for _ in range(5):
        u = c_oracle.pycudaref_step(u, float(g["r_rn100"]))
    assert_bit_equal(u, g["u5_rn100"])
    h = golden("hip_slabs_n100.npz")
    v = h["rand_u0"]
    for _ in range(5):
        v = c_oracle.hipref_step(v, float(h["r"]), 56, 40, 1)
    assert_bit_equal(v, h["rand_u5"])
    u, v = g["u0"], h["rand_u0"]
    for _ in range(5):
        u = c_oracle.pycudaref_step(u, float(g["r_rn100"]), contract=True)
        v = c_oracle.hipref_step(v, float(h["r"]), 56, 40, 1, contract=True)
    assert_bit_equal(u, g["u5_rn100_fma"])
    assert_bit_equal(v, h["rand_u5_fma"])


# ----------------------------------------------------------------- C vs NumPy restatements
@pytest.mark.parametrize("order", [0, 1, 2])
@pytest.mark.parametrize("dtype", [np.float64, np.float32])
@pytest.mark.parametrize("shape", [(3, 3), (5, 17), (33, 64), (97, 131)])
def test_c_oracle_equals_numpy_oracle(order, dtype, shape):
    rng = np.random.default_rng(order * 100 + shape[0])
    u0 = (1.0 + rng.random(shape)).astype(dtype)
    rx, ry = 0.24999999999999997, 0.2437
    a = ho.march(u0, 7, order, rx, ry)
    b = c_oracle.march(u0, 7, order, dtype(rx), dtype(ry))
    assert_bit_equal(a, b)
    c = c_oracle.march(u0, 7, order, dtype(rx), dtype(ry), omp=True)
    assert_bit_equal(a, c)


def test_degenerate_shapes_are_noops():
    for shape in [(1, 1), (2, 9), (9, 2), (1, 50)]:
        u0 = np.random.default_rng(1).random(shape)
        assert_bit_equal(c_oracle.march(u0, 3, 0, 0.25), u0)
        assert_bit_equal(ho.march(u0, 3, 0, 0.25), u0)


def test_fortran_ic_and_params():
    for n in (31, 100, 1024):
        assert_bit_equal(c_oracle.ic_fortran(n, 2.0), ho.ic_fortran(n, 2.0))
        assert_bit_equal(c_oracle.ic_fortran(n, 2.0, 1.0), ho.ic_fortran(n, 2.0, 1.0))
        assert c_oracle.derive_fortran(n, 0.25, 0.05, 2.0) == ho.derive_fortran(n, 0.25, 0.05, 2.0)
    u = ho.ic_fortran(1024, 2.0)
    assert np.array_equal(u, u.T) and set(np.unique(u)) == {1.0, 2.0}
    # r is exactly 0.25 for the named benchmark sizes, not for n = 100
    for n in (31, 1024, 4096, 16384, 32768):
        assert ho.derive_fortran(n, 0.25, 0.05, 2.0)[2] == 0.25
    assert ho.derive_fortran(100, 0.25, 0.05, 2.0)[2] != 0.25


def test_exact_arithmetic_window_all_orders_agree():
    """r = 0.25 and few steps from a {1,2} field keeps every op exact, so the three
    op orders must agree bit for bit; the oracle has to reproduce that."""
    u0 = ho.ic_fortran(128, 2.0)
    ref = ho.march(u0, 12, ho.ORDER_FORTRAN, 0.25)
    assert_bit_equal(ho.march(u0, 12, ho.ORDER_PYCUDA, 0.25), ref)
    assert_bit_equal(ho.march(u0, 12, ho.ORDER_PYSERIAL, 0.25), ref)
    assert 1.0 <= ref.min() and ref.max() <= 2.0      # maximum principle, §4(iv)
    assert np.array_equal(ref, ref.T)                 # transpose symmetry, §4(iii)


def test_orders_differ_past_the_exact_window():
    u0 = ho.ic_fortran(100, 2.0)
    r = ho.derive_fortran(100, 0.25, 0.05, 2.0)[2]
    a = ho.march(u0, 40, ho.ORDER_FORTRAN, r)
    b = ho.march(u0, 40, ho.ORDER_PYSERIAL, r, r)
    assert not np.array_equal(bits(a), bits(b))
    assert np.abs(a - b).max() / np.abs(a).max() < 1e-12


# ----------------------------------------------------------------- input.dat + dumps
@pytest.mark.parametrize("text,want", [
    ("1024 0.25 0.05 2.0 30\n", (1024, 0.25, 0.05, 2.0, 30, None)),          # fortran/serial/input.dat
    ("100 0.25 0.05 2.0 10 1\n", (100, 0.25, 0.05, 2.0, 10, 1)),             # fortran/mpi+cuda/input.dat
    ("32768 0.25 0.05 1.0 25000 0", (32768, 0.25, 0.05, 1.0, 25000, 0)),     # fortran/hip/input.dat
    ("100, 2.5d-1 ,5.0D-2\n 2.0\n1000 /", (100, 0.25, 0.05, 2.0, 1000, None)),
])
def test_parse_input_dat(text, want):
    assert ho.parse_input_dat(text) == want


def test_parse_input_dat_short_record():
    with pytest.raises(ValueError):
        ho.parse_input_dat("100 0.25 0.05")


def test_soln_roundtrip(tmp_path):
    n = 12
    d, _, _ = ho.derive_fortran(n, 0.25, 0.05, 2.0)
    c = ho.fortran_coords(n, 2.0, d)
    T = ho.fortran_serial(n, 0.25, 0.05, 2.0, 3)
    p = os.path.join(tmp_path, "soln.dat")
    ho.write_soln(p, c, c, T)
    X, Y, Z = ho.read_soln(p)
    assert np.array_equal(Z, T.T) and np.array_equal(X[:, 0], c) and np.array_equal(Y[0], c)
