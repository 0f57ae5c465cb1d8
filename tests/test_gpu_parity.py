"""GPU parity tests: every call goes through the C ABI of libheat.so (via heatb200's ctypes layer) and is
compared with the oracle / the golden fixtures.  Bar: bit-exact for the default (no-FMA) mode in fp64 and
fp32; <= 1e-12 (fp64) / <= 1e-5 (fp32 vs the fp64 oracle) relative otherwise — BASELINE.json north_star."""
import numpy as np
import pytest

import heatb200
from heatb200 import _lib, config, ic
from oracle import c_oracle, heat_oracle as ho

pytestmark = pytest.mark.gpu

KERNELS = ["simple", "stream", "tma", "wtma"]
ORDERS = ["fortran", "pycuda", "pyserial"]
R100 = 0.24999999999999997          # r for n = 100 (fortran/mpi+cuda/input.dat): not a power of two


def bits(a):
    return np.ascontiguousarray(a).view(np.uint64 if a.dtype == np.float64 else np.uint32)


def assert_bit_equal(a, b):
    assert a.shape == b.shape and a.dtype == b.dtype
    if not np.array_equal(bits(a), bits(b)):
        bad = np.argwhere(bits(a) != bits(b))
        raise AssertionError("%d cells differ, first at %s, max abs diff %g" % (len(bad), bad[0], np.abs(a - b).max()))


def rand_field(shape, dtype=np.float64, seed=0):
    return (1.0 + np.random.default_rng(seed).random(shape)).astype(dtype)


# ------------------------------------------------------------------ golden fixtures (reference outputs)
def test_pyserial_dropin_matches_reference_run(golden):
    from heatb200 import serial
    serial.configure(n=31)
    serial.diffuse(10)
    assert_bit_equal(serial.u, golden("pyserial_31.npz")["u"])
    assert float(serial.u.sum()) == 1041.943124294281


@pytest.mark.parametrize("kernel", ["auto", "simple", "stream", "tma", "wtma"])
def test_pyserial_dropin_rounding_sensitive_reference_run(golden, kernel):
    """The reference script followed by ITS OWN diffuse(60) (tests/golden/make_golden.py): 72 steps, past the
    exact-arithmetic window, so only op order P (heat.py:50-54) reproduces these bits."""
    from heatb200 import serial
    g = golden("pyserial_31_diffuse60.npz")
    serial.configure(n=31)
    serial.kernel = kernel
    try:
        serial.diffuse(10)
        assert_bit_equal(serial.u, g["u10"])
        serial.diffuse(60)
        assert_bit_equal(serial.u, g["u"])
    finally:
        serial.kernel = "auto"
    # the same input through order F is close but not bit-equal: the golden does discriminate
    dx, dy = serial.dx, serial.dy
    v = g["u10"].copy()
    v[int(.5 / dy):int(1 / dy + 1), int(.5 / dx):int(1 / dx + 1)] = 2
    f = heatb200.march(v, 61, 0.25, order="fortran", kernel=kernel)
    assert not np.array_equal(bits(f), bits(g["u"])) and np.abs(f - g["u"]).max() < 1e-14


@pytest.mark.parametrize("kernel", KERNELS + ["resident", "auto"])
def test_pyserial_case_every_kernel(golden, kernel):
    u0 = ic.hat_pyserial(31, 31)
    _, _, _, cx, cy = config.derive_python(31, 31)
    got = heatb200.march(u0, 11, cx, cy, order="pyserial", kernel=kernel)
    assert_bit_equal(got, golden("pyserial_31.npz")["u"])


@pytest.mark.parametrize("kernel", KERNELS + ["resident"])
def test_hip_kernel_goldens(golden, kernel):
    g = golden("hip_slabs_n100.npz")
    r = float(g["r"])
    # the ghosted 102 x 102 array is a 102 x 102 field whose ring is the ghost layer
    got = heatb200.march(ho.ic_mpi_ghosted(100, 100), 10, r, order="fortran", kernel=kernel)
    assert_bit_equal(got, g["P1_rank0"])
    got = heatb200.march(g["rand_u0"], 5, r, order="fortran", kernel=kernel)
    assert_bit_equal(got, g["rand_u5"])


@pytest.mark.parametrize("kernel", KERNELS + ["resident"])
@pytest.mark.parametrize("tag", ["r025", "rn100"])
def test_pycuda_kernel_goldens(golden, kernel, tag):
    g = golden("pycuda_kernel.npz")
    got = heatb200.march(g["u0"], 5, float(g["r_" + tag]), order="pycuda", kernel=kernel)
    assert_bit_equal(got, g["u5_" + tag])


@pytest.mark.parametrize("kernel", KERNELS + ["resident"])
def test_contracted_goldens_bit_exact_with_fmad(golden, kernel):
    """HEAT_FLAG_FMAD + orders C / F = the reference's GPU kernel text as nvcc (-fmad=true) / hipcc
    (-ffp-contract=fast) compile it by default: bit-equal to the reference sources built for the host with
    contraction on (the *_fma goldens), and to the C oracle's contracted mode on other shapes."""
    p = golden("pycuda_kernel.npz")
    for tag in ("r025", "rn100"):
        got = heatb200.march(p["u0"], 5, float(p["r_" + tag]), order="pycuda", kernel=kernel, fmad=True)
        assert_bit_equal(got, p["u5_" + tag + "_fma"])
    h = golden("hip_slabs_n100.npz")
    got = heatb200.march(h["rand_u0"], 5, float(h["r"]), order="fortran", kernel=kernel, fmad=True)
    assert_bit_equal(got, h["rand_u5_fma"])
    if kernel != "resident":
        for dtype in (np.float64, np.float32):
            u0 = rand_field((130, 257), dtype, seed=12)
            for order in ("fortran", "pycuda"):
                want = c_oracle.march(u0, 9, ho.ORDERS[order], dtype(R100), contract=True)
                assert_bit_equal(heatb200.march(u0, 9, float(dtype(R100)), order=order, kernel=kernel, fmad=True), want)
                for k in (2, 4):
                    assert_bit_equal(heatb200.march(u0, 9, float(dtype(R100)), order=order, kernel="tma", tblock=k, fmad=True), want)


# ------------------------------------------------------------------ oracle, seeded inputs, ragged shapes
SHAPES = [(3, 3), (4, 5), (5, 17), (31, 31), (100, 100), (130, 257), (77, 1000), (515, 64)]


@pytest.mark.parametrize("kernel", KERNELS)
@pytest.mark.parametrize("order", ORDERS)
@pytest.mark.parametrize("dtype", [np.float64, np.float32])
def test_kernels_bit_exact_vs_oracle(kernel, order, dtype):
    for shape in SHAPES:
        u0 = rand_field(shape, dtype, seed=shape[0])
        rx, ry = dtype(R100), dtype(0.2437)
        want = ho.march(u0, 7, ho.ORDERS[order], rx, ry)
        got = heatb200.march(u0, 7, float(rx), float(ry), order=order, kernel=kernel)
        assert_bit_equal(got, want)


@pytest.mark.parametrize("order", ORDERS)
@pytest.mark.parametrize("dtype", [np.float64, np.float32])
def test_resident_kernel_bit_exact(order, dtype):
    for shape in [(3, 3), (31, 31), (100, 100), (64, 190)]:
        u0 = rand_field(shape, dtype, seed=3)
        want = ho.march(u0, 40, ho.ORDERS[order], dtype(R100), dtype(0.2437))
        got = heatb200.march(u0, 40, float(dtype(R100)), float(dtype(0.2437)), order=order, kernel="resident")
        assert_bit_equal(got, want)


@pytest.mark.parametrize("shape", [(1, 1), (1, 40), (2, 9), (9, 2), (40, 1)])
def test_degenerate_fields_are_untouched(shape):
    u0 = rand_field(shape)
    for kernel in ["auto", "simple"]:
        assert_bit_equal(heatb200.march(u0, 3, 0.25, kernel=kernel), u0)


def test_zero_steps_and_repeated_steps():
    u0 = rand_field((65, 130))
    with heatb200.HeatSolver(u0, R100, kernel="stream") as s:
        s.step(0)
        assert_bit_equal(s.gather(), u0)
        for _ in range(5):
            s.step(1)
        a = s.gather()
        t = s.timing()
        assert t.steps_total == 5 and t.kernel_launches == 5
    assert_bit_equal(a, ho.march(u0, 5, 0, R100))


def test_boundary_ring_is_preserved_exactly():
    u0 = rand_field((90, 140), seed=5)
    u0[0, :] = 3.25
    u0[:, -1] = -7.5
    for kernel in KERNELS:
        got = heatb200.march(u0, 9, R100, kernel=kernel)
        for sl in [(0, slice(None)), (-1, slice(None)), (slice(None), 0), (slice(None), -1)]:
            assert_bit_equal(got[sl], u0[sl])


# ------------------------------------------------------------------ on-chip persistent stepping (one launch for all steps)
@pytest.mark.parametrize("order", ORDERS)
@pytest.mark.parametrize("dtype", [np.float64, np.float32])
def test_persist_kernel_bit_exact(order, dtype):
    """PERSIST: the field lives in the shared memory of all SMs (a row slab per CTA), neighbour CTAs trade ghost
    rows through L2 every G steps.  Same bits as the oracle for ragged slabs, step counts that are and are not
    multiples of the ghost depth, very few rows per CTA, and odd widths."""
    for shape, nt in [((130, 257), 8), ((515, 300), 37), ((300, 1000), 9), ((1024, 1024), 30), ((149, 64), 21), ((2000, 130), 16)]:
        u0 = rand_field(shape, dtype, seed=shape[0] + nt)
        rx, ry = dtype(R100), dtype(0.2437)
        want = c_oracle.march(u0, nt, ho.ORDERS[order], rx, ry)
        with heatb200.HeatSolver(u0, float(rx), float(ry), order=order, kernel="persist") as s:
            assert s.kernel_name == "persist"
            s.step(nt)
            assert s.timing().kernel_launches == 1                     # all nt steps in ONE launch
            assert_bit_equal(s.gather(), want)


def test_persist_kernel_repeated_calls_short_calls_and_auto():
    u0 = rand_field((1024, 1024), seed=5)
    u0[0, :] = 3.5
    u0[:, -1] = -2.0                                                   # a ring that is not 1: must survive
    with heatb200.HeatSolver(u0, R100) as s:                           # AUTO: too big for one CTA, fits the SMs
        assert s.kernel_name == "persist"
        for _ in range(6):
            s.step(25)                                                 # flag epochs continue across launches
        s.step(3)                                                      # short call: per-step launches (TMA kernel)
        s.step(0)
        assert s.timing().kernel_launches == 6 + 3
        got = s.gather()
    assert_bit_equal(got, c_oracle.march(u0, 153, 0, R100, omp=True))
    for sl in [(0, slice(None)), (-1, slice(None)), (slice(None), 0), (slice(None), -1)]:
        assert_bit_equal(got[sl], u0[sl])
    got = heatb200.march(u0, 40, R100, order="pycuda", kernel="persist", fmad=True)
    assert_bit_equal(got, c_oracle.march(u0, 40, ho.ORDER_PYCUDA, R100, contract=True))
    with pytest.raises(heatb200.HeatError) as e:                       # 4096^2 does not fit on chip
        heatb200.HeatSolver(None, 0.25, nx=4096, ny=4096, dtype=np.float64, kernel="persist")
    assert e.value.code == _lib.ERR_UNSUPPORTED


# ------------------------------------------------------------------ temporal blocking (k fused steps)
@pytest.mark.parametrize("k", [2, 4])
@pytest.mark.parametrize("order", ORDERS)
@pytest.mark.parametrize("dtype", [np.float64, np.float32])
def test_temporal_blocking_bit_exact(k, order, dtype):
    """k steps per HBM round trip must give the same bits as k single steps (same per-cell op sequence);
    step counts that are not multiples of k finish with single steps."""
    for shape, nt in [((3, 3), 5), ((5, 17), 4), ((40, 130), 9), ((130, 257), 8), ((77, 1000), 7), ((515, 300), 10)]:
        u0 = rand_field(shape, dtype, seed=shape[1])
        rx, ry = dtype(R100), dtype(0.2437)
        want = ho.march(u0, nt, ho.ORDERS[order], rx, ry)
        got = heatb200.march(u0, nt, float(rx), float(ry), order=order, kernel="tma", tblock=k)
        assert_bit_equal(got, want)


@pytest.mark.parametrize("k", [2, 4])
def test_temporal_blocking_launch_count_and_ring(k):
    u0 = rand_field((300, 520), seed=8)
    u0[0, :] = 5.0
    u0[:, 0] = -3.0
    with heatb200.HeatSolver(u0, R100, kernel="tma", tblock=k) as s:
        s.step(4 * k + 1)
        assert s.timing().kernel_launches == 5           # 4 fused + 1 single
        got = s.gather()
    assert_bit_equal(got, ho.march(u0, 4 * k + 1, 0, R100))


def test_temporal_4096_matches_single_step_kernels():
    n = 4096
    delta, _, r = config.InputDeck(n, 0.25, 0.05, 2.0, 24).derive()
    u0 = ic.hat_fortran(n, 2.0, delta) + rand_field((n, n), seed=1) * 0.0 + 0.0
    u0 = u0 + 0.5 * (rand_field((n, n), seed=1) - 1.0) * (u0 > 1.5)   # roughen the hat so rounding is exercised
    base = heatb200.march(u0, 24, 0.2437, kernel="stream")
    for k in (2, 4):
        assert_bit_equal(heatb200.march(u0, 24, 0.2437, kernel="tma", tblock=k), base)


# ------------------------------------------------------------------ tolerance modes
@pytest.mark.parametrize("kernel", KERNELS)
@pytest.mark.parametrize("order", ORDERS)
def test_fmad_mode_within_1e12(kernel, order):
    u0 = rand_field((200, 333), seed=9)
    want = ho.march(u0, 50, ho.ORDERS[order], R100, 0.2437)
    got = heatb200.march(u0, 50, R100, 0.2437, order=order, kernel=kernel, fmad=True)
    assert np.max(np.abs(got - want) / np.abs(want)) <= 1e-12


@pytest.mark.parametrize("fmad", [False, True])
def test_fp32_within_1e5_of_fp64_oracle(fmad):
    u0 = rand_field((200, 333), seed=11)
    want = ho.march(u0, 50, 0, R100)
    got = heatb200.march(u0.astype(np.float32), 50, R100, kernel="stream", fmad=fmad)
    assert got.dtype == np.float32
    assert np.max(np.abs(got - want) / np.abs(want)) <= 1e-5


# ------------------------------------------------------------------ input.dat driven, shipped decks
@pytest.mark.parametrize("deck_text,y_hi", [
    ("100 0.25 0.05 2.0 1000", 1.0),        # fortran/cuda_cuf/input.dat, fortran/cuda_kernel/input.dat
    ("1024 0.25 0.05 2.0 30", 1.5),         # fortran/serial/input.dat
])
def test_shipped_input_decks(deck_text, y_hi):
    deck = config.parse_input_dat(deck_text)
    delta, _, r = deck.derive()
    assert (delta, r) == (ho.derive_fortran(deck.n, deck.sigma, deck.nu, deck.dom_len)[0],
                          ho.derive_fortran(deck.n, deck.sigma, deck.nu, deck.dom_len)[2])
    u0 = ic.hat_fortran(deck.n, deck.dom_len, delta, y_hi)
    assert_bit_equal(u0, ho.ic_fortran(deck.n, deck.dom_len, y_hi))
    want = c_oracle.march(u0, deck.ntime, 0, r)
    for kernel in ["auto", "stream", "tma"]:
        assert_bit_equal(heatb200.march(u0, deck.ntime, r, kernel=kernel), want)


def test_mpi_deck_dropin_single_rank(golden, tmp_path):
    """heatb200.mpi_deck on one GPU with the MPI programs' shipped deck: the interior equals what the reference's
    HIP kernels produce (golden), and soln00000.dat holds it with the reference's coordinates (heat.F90:136,277-288)."""
    from heatb200 import io as hio, mpi_deck
    deck, interior, _ = mpi_deck.run("100 0.25 0.05 2.0 10 1", str(tmp_path))
    gold = golden("hip_slabs_n100.npz")["P1_rank0"]
    assert deck.soln == 1 and interior.shape == (100, 100)
    assert_bit_equal(interior, np.ascontiguousarray(gold[1:-1, 1:-1]))
    x, y, T = hio.read_dump(str(tmp_path / "soln00000.dat"))
    delta = ho.derive_fortran(100, 0.25, 0.05, 2.0)[0]
    assert np.array_equal(T, interior) and np.array_equal(x, np.arange(100) * delta) and np.array_equal(y, x)
    # soln = 0 (fortran/hip/input.dat) writes nothing
    mpi_deck.run("64 0.25 0.05 1.0 5 0", str(tmp_path / "none"))
    assert not (tmp_path / "none").exists()


@pytest.mark.parametrize("deck_text,variant,y_hi", [("100 0.25 0.05 2.0 1000", "cuda", 1.0), ("257, 2.5d-1 5.0D-2 2.0 30 /", "serial", 1.5)])
def test_fortran_deck_dropin(deck_text, variant, y_hi, tmp_path):
    """heatb200.fortran_deck = `program heat` of fortran/serial (and cuda_cuf / cuda_kernel): int.dat holds the hat IC,
    soln.dat the field after ntime steps, both readable the way the reference's init.py / out.py read them."""
    from heatb200 import fortran_deck, io as hio
    deck, c, T = fortran_deck.run(deck_text, str(tmp_path), variant=variant)
    n = deck.n
    want = ho.fortran_serial(n, deck.sigma, deck.nu, deck.dom_len, deck.ntime, y_hi) if n <= 128 else \
        c_oracle.march(ho.ic_fortran(n, deck.dom_len, y_hi), deck.ntime, 0, ho.derive_fortran(n, deck.sigma, deck.nu, deck.dom_len)[2])
    assert_bit_equal(T, want)
    x, y, T2 = hio.read_dump(str(tmp_path / "soln.dat"))
    assert np.array_equal(T2, T) and np.array_equal(x, c) and np.array_equal(y, c)
    _, _, T0 = hio.read_dump(str(tmp_path / "int.dat"))
    assert_bit_equal(T0, ho.ic_fortran(n, deck.dom_len, y_hi))
    X, Y, Z = ho.read_soln(str(tmp_path / "soln.dat"))               # out.py's parsing
    assert np.array_equal(Z, T.T)


def test_cuda_dropin_small():
    from heatb200 import cuda
    cuda.configure(n=256, nt_=19, hat=True)
    u0 = cuda.u_new.copy()
    out = cuda.run()
    assert out.shape == (256, 256)
    assert_bit_equal(out, ho.march(u0, 20, ho.ORDER_PYCUDA, cuda.r))
    # the call-compatible single launch
    cuda.configure(n=64, nt_=0, hat=True)
    new, old = cuda.u_new.copy(), cuda.u_new.copy()
    cuda.heatCalculate(cuda.drv.InOut(new.ravel()), cuda.drv.In(old.ravel()), block=(32, 8, 1), grid=(2, 8))
    assert_bit_equal(new, ho.march(old, 1, ho.ORDER_PYCUDA, cuda.r))
    # as shipped (empty hat slice) the field is the fixed point 1
    cuda.configure(n=128, nt_=3)
    assert (cuda.run() == 1.0).all()


def test_cuda_dropin_reference_size_and_step_count():
    """python/cuda/cuda.py as shipped: 4096^2, nt = 10000 -> 10001 launches; its (empty-slice) field is the
    fixed point 1.  With the intended hat the 10001-step result must not depend on how steps are fused."""
    from heatb200 import cuda
    cuda.configure(n=4096, nt_=10000)
    out = cuda.run()
    assert out.shape == (4096, 4096) and (out == 1.0).all()
    cuda.configure(n=4096, nt_=10000, hat=True)
    u0 = cuda.u_new.copy()
    a = cuda.run()
    b = heatb200.march(u0, 10001, cuda.r, order="pycuda", kernel="tma", tblock=4)
    assert_bit_equal(a, b)
    assert 1.0 <= a.min() and a.max() < 2.0                      # maximum principle after 10001 steps
    assert np.abs(a - a.T).max() < 1e-11                         # symmetric up to rounding (W+E is summed before N+S)


def test_many_back_to_back_launches_small_grid():
    """2000 dependent launches of a grid far smaller than the GPU (programmatic dependent launch lets the
    next step's CTAs start early): any ordering bug shows up against the one-launch RESIDENT kernel."""
    u0 = rand_field((80, 120), seed=21)
    want = heatb200.march(u0, 2000, R100, kernel="resident")
    assert_bit_equal(want, c_oracle.march(u0, 2000, 0, R100))
    for kernel, k in [("tma", 1), ("wtma", 1), ("tma", 2), ("tma", 4), ("stream", 1)]:
        assert_bit_equal(heatb200.march(u0, 2000, R100, kernel=kernel, tblock=k), want)


# ------------------------------------------------------------------ full-size properties (BASELINE sizes)
def test_4096_kernels_agree_and_match_c_oracle():
    n = 4096
    deck = config.InputDeck(n, 0.25, 0.05, 2.0, 8)
    delta, _, r = deck.derive()
    assert r == 0.25
    u0 = ic.hat_fortran(n, 2.0, delta)
    want = c_oracle.march(u0, 8, 0, r, omp=True)
    outs = {k: heatb200.march(u0, 8, r, kernel=k) for k in KERNELS}
    for k in KERNELS:
        assert_bit_equal(outs[k], want)
    a = outs["tma"]
    assert np.array_equal(a, a.T)                      # transpose symmetry (SURVEY §4 iii)
    assert a.min() >= 1.0 and a.max() <= 2.0           # maximum principle (SURVEY §4 iv)
    # exact-arithmetic window: every op order gives the same bits (SURVEY §4 ii)
    for order in ["pycuda", "pyserial"]:
        assert_bit_equal(heatb200.march(u0, 8, r, order=order, kernel="stream"), a)


def test_16384_random_field_kernels_agree_and_linear():
    """At the north-star size the oracle is too slow; use kernel-vs-kernel bit equality (independent code
    paths) plus homogeneity of the update: march(a*u) == a*march(u) for a power of two (scaling by 2^k is
    exact in binary floating point, so every intermediate scales exactly)."""
    import torch
    n = 16384
    g = torch.Generator(device="cuda").manual_seed(0)
    u = 1.0 + torch.rand((n, n), dtype=torch.float64, device="cuda", generator=g)
    res = {}
    for k in ["stream", "tma"]:
        with heatb200.HeatSolver(None, 0.25, nx=n, ny=n, dtype=np.float64, kernel=k) as s:
            s.set_field_from_tensor(u)
            s.step(6)
            out = torch.empty_like(u)
            s.gather_into_tensor(out)
            res[k] = out
            ck = s.checksum()
            assert abs(ck - float(out.sum())) <= 1e-9 * abs(ck)
    assert torch.equal(res["stream"], res["tma"])
    with heatb200.HeatSolver(None, 0.25, nx=n, ny=n, dtype=np.float64, kernel="tma") as s:
        s.set_field_from_tensor(4.0 * u)
        s.step(6)
        out = torch.empty_like(u)
        s.gather_into_tensor(out)
    assert torch.equal(out, 4.0 * res["tma"])
    # spot-check a 64-row band against the C oracle (bands are independent for 6 steps if cut 6 rows deep)
    band = u[1000:1000 + 64 + 12].cpu().numpy()
    want = c_oracle.march(band, 6, 0, 0.25, omp=True)[6:-6]
    assert_bit_equal(res["tma"][1006:1006 + 64].cpu().numpy(), want)


@pytest.mark.parametrize("dtype,k", [(np.float64, 2), (np.float64, 4), (np.float32, 1), (np.float32, 2), (np.float32, 4)])
def test_16384_bands_vs_c_oracle_temporal_and_fp32(dtype, k):
    """What bench.py reports at 16384^2 beyond fp64 k=1: k = 2, 4 and fp32.  Three 256-row bands over the full
    width — the top one including the Dirichlet row, one in the middle, the bottom one including the last row —
    against the C oracle (a band cut `steps` rows deeper than what is compared is independent of the rest), and
    the whole field against the single-step STREAM kernel."""
    import torch
    n, steps, B = 16384, 8, 256
    tdt = torch.float64 if dtype == np.float64 else torch.float32
    g = torch.Generator(device="cuda").manual_seed(7)
    u = 1.0 + torch.rand((n, n), dtype=tdt, device="cuda", generator=g)
    u[0, :] = 1.0
    u[-1, :] = 1.0
    u[:, 0] = 1.0
    u[:, -1] = 1.0
    outs = {}
    for kernel, kk in (("tma", k), ("stream", 1)):
        with heatb200.HeatSolver(None, R100, nx=n, ny=n, dtype=dtype, kernel=kernel, tblock=kk) as s:
            s.set_field_from_tensor(u)
            s.step(steps)
            out = torch.empty_like(u)
            s.gather_into_tensor(out)
            outs[kernel] = out
    assert torch.equal(outs["tma"], outs["stream"])
    got = outs["tma"]
    r = dtype(R100)
    top = c_oracle.march(u[:B + steps].cpu().numpy(), steps, 0, r, omp=True)[:B]
    assert_bit_equal(got[:B].cpu().numpy(), top)
    m0 = 7000
    mid = c_oracle.march(u[m0 - steps:m0 + B + steps].cpu().numpy(), steps, 0, r, omp=True)[steps:-steps]
    assert_bit_equal(got[m0:m0 + B].cpu().numpy(), mid)
    bot = c_oracle.march(u[n - B - steps:].cpu().numpy(), steps, 0, r, omp=True)[steps:]
    assert_bit_equal(got[n - B:].cpu().numpy(), bot)


@pytest.mark.parametrize("dtype", [np.float64, np.float32])
def test_32768_wide_rows_every_kernel(dtype):
    """Rows as wide as the strong-scaling workload's (32768 columns = 256 strips of the TMA kernels: a grid shape
    of its own for their segment choice), 600 rows, against the C oracle."""
    u0 = rand_field((600, 32768), dtype, seed=33)
    want = c_oracle.march(u0, 9, 0, dtype(R100), omp=True)
    for kernel, k in (("tma", 1), ("tma", 2), ("tma", 4), ("wtma", 1), ("stream", 1)):
        assert_bit_equal(heatb200.march(u0, 9, float(dtype(R100)), kernel=kernel, tblock=k), want)


# ------------------------------------------------------------------ torch buffers, checksum, errors
def test_external_torch_buffers_and_unaligned_pitch():
    import torch
    u0 = rand_field((200, 301), seed=2)                 # too big for the one-CTA resident kernel
    b0 = torch.zeros((200, 301), dtype=torch.float64, device="cuda")
    b1 = torch.zeros_like(b0)
    with heatb200.HeatSolver(u0, R100, buffers=(b0, b1)) as s:
        assert s.kernel_name == "stream"            # odd pitch: the one-cell-per-lane variant of STREAM
        s.step(4)
        got = s.gather()
        cur = (b0, b1)[s.info().current]
    assert_bit_equal(got, ho.march(u0, 4, 0, R100))
    assert_bit_equal(cur.cpu().numpy(), got)       # the caller's tensors ARE the storage
    with pytest.raises(heatb200.HeatError) as e:
        heatb200.HeatSolver(u0, R100, buffers=(b0, b1), kernel="tma")   # TMA needs 16-byte friendly rows
    assert e.value.code == _lib.ERR_UNSUPPORTED
    for shape in [(3, 3), (40, 35), (130, 257), (77, 1001)]:            # odd pitches through the scalar STREAM kernel
        v0 = rand_field(shape, seed=shape[1])
        c0 = torch.zeros(shape, dtype=torch.float64, device="cuda")
        c1 = torch.zeros_like(c0)
        for kern in ("stream", "simple"):
            with heatb200.HeatSolver(v0, R100, 0.2437, order="pyserial", buffers=(c0, c1), kernel=kern, tblock=2) as s2:
                s2.step(7)
                assert_bit_equal(s2.gather(), ho.march(v0, 7, ho.ORDER_PYSERIAL, R100, 0.2437))


@pytest.mark.parametrize("shape,dtype,k", [((300, 520), np.float64, 1), ((131, 1001), np.float64, 2), ((260, 512), np.float32, 4)])
def test_job_pipeline_matches_blocking_calls(shape, dtype, k):
    """heat_submit / heat_drain: back-to-back jobs (upload, march, download overlapped across jobs) give the bits
    of the blocking set_field + step + gather sequence and of the oracle; pinned and pageable host memory."""
    import torch
    tdt = torch.float64 if dtype == np.float64 else torch.float32
    njobs = 4
    ins = [torch.empty(shape, dtype=tdt).pin_memory().numpy() for _ in range(njobs)]
    outs = [torch.empty(shape, dtype=tdt).pin_memory().numpy() for _ in range(njobs)]
    for i, a in enumerate(ins):
        a[...] = rand_field(shape, dtype, seed=40 + i)
    ins[3] = np.array(ins[3])                                   # one pageable input
    with heatb200.HeatSolver(ins[0], float(dtype(R100)), kernel="tma", tblock=k) as s:
        for i in range(njobs):
            s.submit(ins[i], 5 + i, outs[i])
        s.drain()
        for i in range(njobs):
            assert_bit_equal(outs[i], ho.march(ins[i], 5 + i, 0, dtype(R100)))
        # the handle is still an ordinary solver afterwards, and one output buffer may serve several jobs
        s.step(2)
        assert_bit_equal(s.gather(), ho.march(ins[3], 5 + 3 + 2, 0, dtype(R100)))
        s.submit(ins[0], 3, outs[0])
        s.submit(ins[1], 3, outs[0])
        s.drain()
        assert_bit_equal(outs[0], ho.march(ins[1], 3, 0, dtype(R100)))
        with pytest.raises(ValueError):
            s.submit(ins[0][:-1], 1, outs[0])


@pytest.mark.parametrize("dtype", [np.float64, np.float32])
def test_guard_bands_are_never_written_and_never_feed_a_result(dtype):
    """Our own memory check (compute-sanitizer is not available on this pool): the two field buffers are views into
    larger caller-owned allocations whose surrounding rows AND whose pitch padding are filled with NaN.  After a
    march with every kernel the guard rows must still hold that bit pattern (no out-of-bounds write), the result
    must equal the oracle bit for bit (so no NaN from a guard row or from the padding was ever used for an updated
    cell), and the padding columns inside the rows may only have been rewritten with what they held."""
    import torch
    tdt = torch.float64 if dtype == np.float64 else torch.float32
    G = 8                                               # guard rows above and below each buffer
    for (ny, nx, pitch) in [(70, 130, 144), (131, 1000, 1008), (300, 257, 272), (40, 64, 64)]:
        u0 = rand_field((ny, nx), dtype, seed=nx)
        want = ho.march(u0, 9, 0, dtype(R100))
        for kernel, k in [("simple", 1), ("stream", 1), ("tma", 1), ("wtma", 1), ("tma", 2), ("tma", 4), ("persist", 1)]:
            bigs = [torch.full((ny + 2 * G, pitch), float("nan"), dtype=tdt, device="cuda") for _ in range(2)]
            bufs = tuple(b[G:G + ny] for b in bigs)
            with heatb200.HeatSolver(u0, float(dtype(R100)), buffers=bufs, kernel=kernel, tblock=k) as s:
                s.step(9)
                got = s.gather()
            assert_bit_equal(got, want)
            for b in bigs:
                assert bool(torch.isnan(b[:G]).all()) and bool(torch.isnan(b[G + ny:]).all()), (kernel, k, ny, nx, "guard rows written")
                if pitch > nx:
                    assert bool(torch.isnan(b[G:G + ny, nx:]).all()), (kernel, k, ny, nx, "pitch padding changed")


def test_checksum_deterministic_and_close():
    u0 = rand_field((300, 500), seed=4)
    with heatb200.HeatSolver(u0, R100) as s:
        s.step(3)
        a, b = s.checksum(), s.checksum()
        assert a == b
        assert abs(a - s.gather().sum()) <= 1e-10 * abs(a)


def test_error_paths():
    u0 = rand_field((16, 16))
    with pytest.raises(heatb200.HeatError) as e:
        heatb200.HeatSolver(u0, 0.25, device=99)
    assert e.value.code == _lib.ERR_INVALID and "device" in str(e.value)
    with pytest.raises(heatb200.HeatError):
        heatb200.HeatSolver(u0, 0.25, nranks=2, rank=0, ny=32)             # no nccl id
    with pytest.raises(heatb200.HeatError):
        heatb200.HeatSolver(rand_field((400, 400)), 0.25, kernel="resident")  # does not fit one CTA
    with heatb200.HeatSolver(u0, 0.25) as s:
        with pytest.raises(heatb200.HeatError):
            s.step(-1)
    # decompositions every rank must refuse alike, before any collective (messages built from global values only)
    kw = dict(nx=64, ny=17, dtype=np.float64, nranks=2, rank=0, skip_exchange=True)
    with pytest.raises(heatb200.HeatError) as e:
        heatb200.HeatSolver(None, 0.25, ghost=9, **kw)                     # thinnest slab has 8 rows
    assert e.value.code == _lib.ERR_INVALID and "thinnest" in str(e.value) and "ghost depth 9" in str(e.value)
    with pytest.raises(heatb200.HeatError) as e:
        heatb200.HeatSolver(None, 0.25, ghost=3, tblock=2, **kw)
    assert e.value.code == _lib.ERR_INVALID and "ghost depth 3 must be a multiple of tblock 2" in str(e.value)
    with pytest.raises(heatb200.HeatError) as e:
        heatb200.HeatSolver(None, 0.25, ghost=2, ranks_x=2, peer_halo=True, **kw)
    assert e.value.code == _lib.ERR_UNSUPPORTED and "row slabs only" in str(e.value)


def test_legacy_launcher_large_ghosted_array():
    """launch_heat_eqn on a ghosted array big enough for its row-marching path (odd pitch nx + 2)."""
    import ctypes as C
    import torch
    nx, ny, ng = 1000, 300, 1
    g0 = rand_field((ny + 2, nx + 2), seed=77)
    old = torch.from_numpy(g0).cuda()
    new = old.clone()
    heatb200.lib().launch_heat_eqn(C.c_void_p(new.data_ptr()), C.c_void_p(old.data_ptr()), R100, nx, ny, ng)
    torch.cuda.synchronize()
    assert_bit_equal(new.cpu().numpy(), ho.march(g0, 1, ho.ORDER_FORTRAN, R100))


def test_legacy_launcher_contracted_and_single_precision(golden):
    """heat_legacy_set_fmad(1): launch_heat_eqn fuses as hipcc's default does -> the contracted hip golden;
    launch_*_sp: the -DSINGLE_PRECISION flavour of the reference's `real` (heat_kernel.cpp:5-9)."""
    import ctypes as C
    import torch
    L = heatb200.lib()
    h = golden("hip_slabs_n100.npz")
    r, nx, ny, ng = float(h["r"]), 56, 40, 1
    p = lambda t: C.c_void_p(t.data_ptr())
    for fm, key in ((1, "rand_u5_fma"), (0, "rand_u5")):
        prev = L.heat_legacy_set_fmad(fm)
        try:
            cur = torch.from_numpy(h["rand_u0"]).cuda()
            for _ in range(5):
                old = cur.clone()
                L.launch_heat_eqn(p(cur), p(old), r, nx, ny, ng)
            torch.cuda.synchronize()
        finally:
            L.heat_legacy_set_fmad(prev)
        assert_bit_equal(cur.cpu().numpy(), h[key])
    assert L.heat_legacy_set_fmad(0) == 0
    # fp32 flavour: update on small (thread-per-cell) and large (row-marching) ghosted arrays, then pack/unpack
    for nx, ny in ((56, 40), (1000, 300)):
        g0 = rand_field((ny + 2, nx + 2), np.float32, seed=nx)
        old = torch.from_numpy(g0).cuda()
        new = old.clone()
        L.launch_heat_eqn_sp(p(new), p(old), np.float32(R100), nx, ny, ng)
        torch.cuda.synchronize()
        assert_bit_equal(new.cpu().numpy(), ho.march(g0, 1, ho.ORDER_FORTRAN, np.float32(R100)))
        s1 = torch.zeros(ng * ny, dtype=torch.float32, device="cuda")
        s2 = torch.zeros_like(s1)
        L.launch_send_sp(p(new), p(s1), p(s2), nx, ny, ng)
        torch.cuda.synchronize()
        got = new.cpu().numpy()
        assert np.array_equal(s1.cpu().numpy(), got[1:-1, 1]) and np.array_equal(s2.cpu().numpy(), got[1:-1, nx])
        L.launch_recv1_sp(p(new), p(s2), nx, ny, ng)
        L.launch_recv2_sp(p(new), p(s1), nx, ny, ng)
        torch.cuda.synchronize()
        after = new.cpu().numpy()
        assert np.array_equal(after[1:-1, 0], got[1:-1, nx]) and np.array_equal(after[1:-1, nx + 1], got[1:-1, 1])


def test_legacy_hip_abi(golden):
    """launch_heat_eqn / launch_send / launch_recv1 / launch_recv2 with the reference's signatures on
    caller-owned ghosted device arrays, driven like fortran/hip/heat.F90:234-252,198-228 for P = 2."""
    import ctypes as C
    import torch
    g = golden("hip_slabs_n100.npz")
    L, r, nxl, ny, ng = heatb200.lib(), float(g["r"]), 50, 100, 1
    dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    p = lambda t: C.c_void_p(t.data_ptr())
    Td = [dev(ho.ic_mpi_ghosted(nxl, ny, q == 0, q == 1)) for q in range(2)]
    s1 = [torch.zeros(ng * ny, dtype=torch.float64, device="cuda") for _ in range(2)]
    s2 = [torch.zeros(ng * ny, dtype=torch.float64, device="cuda") for _ in range(2)]
    for _ in range(10):
        for q in range(2):
            old = Td[q].clone()                                    # hipMemcpy D2D, heat.F90:243
            L.launch_heat_eqn(p(Td[q]), p(old), r, nxl, ny, ng)
            torch.cuda.synchronize()
            L.launch_send(p(Td[q]), p(s1[q]), p(s2[q]), nxl, ny, ng)
        torch.cuda.synchronize()
        L.launch_recv1(p(Td[1]), p(s2[0]), nxl, ny, ng)            # rank 1's left ghost <- rank 0's last line
        L.launch_recv2(p(Td[0]), p(s1[1]), nxl, ny, ng)            # rank 0's right ghost <- rank 1's first line
        torch.cuda.synchronize()
    for q in range(2):
        assert_bit_equal(Td[q].cpu().numpy(), g["P2_rank%d" % q])
