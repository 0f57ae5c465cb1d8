"""CPU: the C-ABI library loads and exports every symbol include/heat_b200.h declares, the pure-host entry
points work without a GPU, compute entry points fail loudly (no CPU fallback), and the host layer
(input.dat parsing, parameter derivation, initial conditions, drop-in module globals) matches the oracle."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import heatb200
from heatb200 import _lib, config, ic
from oracle import heat_oracle as ho

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "heat_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b((?:heat|launch)_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    L = heatb200.lib()
    names = declared_symbols()
    assert len(names) >= 19
    for n in names:
        assert hasattr(L, n), "libheat.so does not export %s" % n
    assert sorted(_lib.PROTOTYPES) == names, "ctypes prototypes and header disagree"


def test_config_struct_matches_header_layout():
    # struct_size is checked by heat_init; a mismatch is reported as HEAT_ERR_INVALID, never UB
    cfg = _lib.HeatConfig()
    cfg.struct_size = 4
    h = C.c_void_p()
    rc = heatb200.lib().heat_init(C.byref(h), C.byref(cfg), None)
    assert rc == _lib.ERR_INVALID and b"struct_size" in heatb200.lib().heat_last_error()
    assert C.sizeof(_lib.HeatConfig) == 112


def test_tile_partition_and_argument_checks():
    from heatb200.solver import tile_partition
    nx, ny = 1000, 203
    for px, py in [(1, 4), (2, 2), (4, 2), (8, 1), (3, 5)]:
        tiles = [tile_partition(nx, ny, px, py, r) for r in range(px * py)]
        cover = np.zeros((ny, nx), dtype=np.int32)
        for r, (r0, nr, c0, nc) in enumerate(tiles):
            cover[r0:r0 + nr, c0:c0 + nc] += 1
            assert (r0, nr) == heatb200.slab_partition(ny, py, r // px) and (c0, nc) == heatb200.slab_partition(nx, px, r % px)
        assert (cover == 1).all()
    with pytest.raises(heatb200.HeatError):
        tile_partition(nx, ny, 2, 2, 4)
    # argument checks that need no GPU: tblock outside {0,1,2,4} is refused in one place, before any device work
    cfg = _lib.HeatConfig()
    cfg.struct_size = C.sizeof(_lib.HeatConfig)
    cfg.nx = cfg.ny = 64
    cfg.nranks = 1
    h = C.c_void_p()
    for bad in (3, 5, 8, -1):
        cfg.tblock = bad
        assert heatb200.lib().heat_init(C.byref(h), C.byref(cfg), None) == _lib.ERR_INVALID
        assert b"tblock" in heatb200.lib().heat_last_error()


def test_slab_partition():
    for ny, P in [(100, 1), (100, 3), (32768, 8), (7, 7), (10, 4)]:
        rows = [heatb200.slab_partition(ny, P, r) for r in range(P)]
        assert rows[0][0] == 0 and sum(n for _, n in rows) == ny
        for (a, n), (b, _) in zip(rows, rows[1:]):
            assert a + n == b
        assert max(n for _, n in rows) - min(n for _, n in rows) <= 1
    with pytest.raises(heatb200.HeatError):
        heatb200.slab_partition(10, 4, 4)


def test_derive_matches_oracle_bit_for_bit():
    for n in (31, 100, 1024, 4096, 16384, 32768, 131072, 77):
        for L in (1.0, 2.0, 3.7):
            d = config.InputDeck(n, 0.25, 0.05, L, 1).derive()
            assert d == ho.derive_fortran(n, 0.25, 0.05, L)
    assert config.derive_python(31, 31) == ho.derive_python(31, 31)
    assert config.derive_python(4096, 4096) == ho.derive_python(4096, 4096)


@pytest.mark.parametrize("text", [
    "1024 0.25 0.05 2.0 30\n", "100 0.25 0.05 2.0 10 1\n", "32768 0.25 0.05 1.0 25000 0",
    "100, 2.5d-1 ,5.0D-2\n 2.0\n1000 /", " 31 .25 .05 2. 11 / trailing junk 5 6"])
def test_parse_input_dat_matches_oracle(text):
    d = config.parse_input_dat(text)
    assert (d.n, d.sigma, d.nu, d.dom_len, d.ntime, d.soln) == ho.parse_input_dat(text)


def test_parse_input_dat_from_file_and_errors(tmp_path):
    p = tmp_path / "input.dat"
    p.write_text("100 0.25 0.05 2.0 1000\n")
    assert config.parse_input_dat(str(p)).ntime == 1000
    for bad in ["100 0.25 0.05", "", "1 0.25 0.05 2.0 10", "10.5 0.25 0.05 2.0 10", "10 0.25 0.05 2.0 -1"]:
        with pytest.raises(ValueError):
            config.parse_input_dat(bad + "\n")


def test_initial_conditions_match_oracle():
    for n in (31, 100, 257):
        delta = ho.derive_fortran(n, 0.25, 0.05, 2.0)[0]
        assert np.array_equal(ic.hat_fortran(n, 2.0, delta), ho.ic_fortran(n, 2.0))
        assert np.array_equal(ic.hat_fortran(n, 2.0, delta, 1.0), ho.ic_fortran(n, 2.0, 1.0))
        assert np.array_equal(ic.hat_pyserial(n, n), ho.ic_pyserial(n, n))
        assert np.array_equal(ic.hat_pycuda(n, n), ho.ic_pycuda(n, n))
        full = ic.hat_fortran(n, 2.0, delta)
        for P in (2, 3):
            parts = [ic.hat_fortran(n, 2.0, delta, rows=heatb200.slab_partition(n, P, r)) for r in range(P)]
            assert np.array_equal(np.concatenate(parts), full)
    g = ho.ic_mpi_ghosted(10, 12)
    assert np.array_equal(ic.mpi_uniform(12, 14), g)
    assert np.array_equal(np.concatenate([ic.mpi_uniform(12, 14, rows=(0, 5)), ic.mpi_uniform(12, 14, rows=(5, 9))]), g)


def test_dropin_module_globals_mirror_reference(golden):
    from heatb200 import serial, cuda
    serial.configure(n=31)
    g = golden("pyserial_31.npz")
    assert (serial.nx, serial.ny, serial.nt) == (31, 31, 17)
    assert serial.dx == float(g["dx"]) and serial.dt == float(g["dt"]) and serial.nu == .05 and serial.sigma == .25
    assert np.array_equal(serial.u, ho.ic_pyserial(31, 31)) and (serial.un == 1).all()
    assert np.array_equal(serial.x, np.linspace(0, 2, 31))
    assert (cuda.nx, cuda.ny, cuda.nt) == (4096, 4096, 10000) and cuda.r == 0.25
    assert callable(cuda.heatCalculate) and callable(serial.diffuse)


def test_no_cpu_fallback_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(heatb200.HeatError) as e:
        heatb200.march(np.ones((8, 8)), 1, 0.25)
    assert e.value.code == _lib.ERR_CUDA
    from heatb200 import serial
    serial.configure(n=31)
    with pytest.raises(heatb200.HeatError):
        serial.diffuse(10)


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "cuda-hip-mpi-heat-equation-test_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b|liboracle|oracle/", src, re.M), f


def test_serial_save_field_is_the_reference_dump_format(tmp_path):
    """serial.save_field writes `x y u` lines, x (contiguous index) outer — what fortran/serial/out.py parses."""
    from heatb200 import serial
    serial.configure(n=9)
    serial.u[:] = np.arange(81, dtype=float).reshape(9, 9) / 7.0
    p = tmp_path / "soln.dat"
    serial.save_field(str(p))
    X, Y, Z = ho.read_soln(str(p))
    assert np.array_equal(Z, serial.u.T) and np.array_equal(X[:, 0], serial.x) and np.array_equal(Y[0], serial.y)
    serial.configure(n=31)


def test_dropins_configure_from_input_deck(tmp_path):
    from heatb200 import serial, cuda
    p = tmp_path / "input.dat"
    p.write_text("100 0.25 0.05 2.0 1000\n")
    serial.configure(str(p))
    assert (serial.nx, serial.ny, serial.nt) == (100, 100, 1000) and serial.u.shape == (100, 100)
    assert serial.dx == 2.0 / 99
    cuda.configure(str(p))
    assert (cuda.nx, cuda.nt) == (100, 999) and cuda.u_new.shape == (100, 100)   # nt + 1 launches = ntime
    assert cuda.r == (cuda.nu * cuda.dt) / cuda.dx**2
    serial.configure(n=31)
    cuda.configure(n=4096, nt_=10000)


def test_io_dump_roundtrip_and_rank_files(tmp_path):
    from heatb200 import io as hio
    n = 14
    delta = ho.derive_fortran(n, 0.25, 0.05, 2.0)[0]
    c = ic.fortran_coords(n, 2.0, delta)
    T = ho.fortran_serial(n, 0.25, 0.05, 2.0, 4)
    p = str(tmp_path / "soln.dat")
    hio.write_dump(p, c, c, T)
    x, y, T2 = hio.read_dump(p)
    assert np.array_equal(x, c) and np.array_equal(y, c) and np.array_equal(T2, T)
    X, Y, Z = ho.read_soln(p)                       # the oracle's reader (out.py's logic) agrees
    assert np.array_equal(Z, T.T)
    # the oracle's writer and ours produce the same numbers
    q = str(tmp_path / "soln_oracle.dat")
    ho.write_soln(q, c, c, T)
    assert np.array_equal(np.loadtxt(q), np.loadtxt(p))
    # per-rank files tile the field
    parts = []
    for r in range(3):
        r0, nr = heatb200.slab_partition(n, 3, r)
        f = hio.write_rank_dump(str(tmp_path), r, c, c[r0:r0 + nr], T[r0:r0 + nr])
        assert f.endswith("soln%05d.dat" % r)
        parts.append(hio.read_dump(f)[2])
    assert np.array_equal(np.concatenate(parts), T)
    assert abs(hio.checksum_host(T) - T.sum()) < 1e-9


def test_deck_dropins_fail_loudly_without_gpu(tmp_path):
    """The program-level drop-ins have no CPU path either: on a box without a GPU they raise, and write no result file."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("needs a box without a GPU")
    from heatb200 import fortran_deck, mpi_deck
    with pytest.raises(heatb200.HeatError) as e:
        fortran_deck.run("64 0.25 0.05 2.0 5", str(tmp_path / "f"))
    assert e.value.code == _lib.ERR_CUDA
    assert (tmp_path / "f" / "int.dat").exists() and not (tmp_path / "f" / "soln.dat").exists()
    with pytest.raises(heatb200.HeatError) as e:
        mpi_deck.run("64 0.25 0.05 2.0 5 1", str(tmp_path / "m"))
    assert e.value.code == _lib.ERR_CUDA and not (tmp_path / "m").exists()
    with pytest.raises(ValueError):
        fortran_deck.run("64 0.25 0.05 2.0 5", str(tmp_path / "f"), variant="hip")


def test_input_dat_parser_agrees_with_the_oracle_on_random_decks():
    """Property test: the product's list-directed parser and the oracle's accept the same decks and return the same
    values, whatever the separators, exponent letters, optional 6th field and record terminator."""
    from hypothesis import given, settings, strategies as st

    seps = st.sampled_from([" ", "  ", ",", " , ", "\n", "\t"])
    expo = st.sampled_from(["e", "E", "d", "D"])

    @settings(max_examples=200, deadline=None)
    @given(n=st.integers(2, 200000), sigma=st.floats(1e-3, 1.0), nu=st.floats(1e-3, 10.0), L=st.floats(0.1, 100.0),
           ntime=st.integers(0, 10**6), soln=st.one_of(st.none(), st.integers(0, 1)), sep=st.lists(seps, min_size=5, max_size=5),
           ex=expo, slash=st.booleans())
    def check(n, sigma, nu, L, ntime, soln, sep, ex, slash):
        def f(v):
            return ("%.17e" % v).replace("e", ex)
        text = str(n) + sep[0] + f(sigma) + sep[1] + f(nu) + sep[2] + f(L) + sep[3] + str(ntime)
        if soln is not None:
            text += sep[4] + str(soln)
        if slash:
            text += " / trailing words 1 2 3"
        d = config.parse_input_dat(text + "\n")
        assert (d.n, d.sigma, d.nu, d.dom_len, d.ntime, d.soln) == ho.parse_input_dat(text) == (n, sigma, nu, L, ntime, soln)
        assert d.derive() == ho.derive_fortran(n, sigma, nu, L)

    check()
