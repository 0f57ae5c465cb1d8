"""The allocation-free checksum, heat_checksum_noalloc, called through the C ABI.

Who needs it: callers that cannot allocate device memory at the point of the call — host programs
that own all their device buffers and pass raw pointers (the convention of the reference's Fortran
drivers), processes whose device heap is already taken by another allocator, and multi-stream time
loops where cudaMalloc/cudaFree would synchronise the whole device.  Its partial sums sit in a fixed
array inside libheat.so rather than in a cudaMalloc'ed buffer.

Checked here: its result equals heat_checksum's (the allocating entry point) on the same device
data — the very same bits, since nblocks = 0 asks for heat_checksum's own reduction tree, which is
stricter than the 1e-10 relative tolerance the other checksum test uses — on two calls in a row with
different data; stale contents of the shared array do not leak into a later call; requests beyond the
array's capacity are refused with a clear error.  The file sorts last so that a `-x` run has already
executed every other test; it depends on no state left by them and also passes alone.
"""
import ctypes as C

import numpy as np
import pytest

import heatb200
from heatb200 import _lib

pytestmark = pytest.mark.gpu

TOL = 1e-10


def noalloc_sum(ptr, dtype_code, nx, ny, pitch, out, nblocks=0):
    rc = heatb200.lib().heat_checksum_noalloc(C.c_void_p(ptr), dtype_code, nx, ny, pitch, nblocks,
                                              C.c_void_p(out.data_ptr()), 0)
    _lib.check(rc)


@pytest.mark.parametrize("dtype", [np.float64, np.float32])
def test_noalloc_checksum_matches_allocating_one_twice_in_a_row(dtype):
    import torch
    rng = np.random.default_rng(42)
    u0 = (1.0 + rng.random((300, 500))).astype(dtype)
    code = _lib.F64 if dtype == np.float64 else _lib.F32
    out = torch.full((1,), -1.0, dtype=torch.float64, device="cuda")
    with heatb200.HeatSolver(u0, 0.2437, kernel="stream") as s:
        for call in range(2):                     # two calls in a row, on different field contents
            s.step(3)
            want = s.checksum()                   # existing entry point (allocates its scratch)
            info = s.info()
            s.sync()
            noalloc_sum(info.buf[info.current], code, info.nx, info.nrows, info.pitch, out)
            torch.cuda.synchronize()
            got = float(out.item())
            assert got == want, (call, got, want)          # same reduction tree, hence the same bits (header)
            assert abs(got - float(s.gather().astype(np.float64).sum())) <= TOL * abs(want)


def test_noalloc_checksum_ignores_stale_scratch_and_other_tree_sizes():
    import torch
    big = torch.full((700, 900), 1000.0, dtype=torch.float64, device="cuda")
    small = torch.full((5, 7), 0.5, dtype=torch.float64, device="cuda")
    out = torch.zeros(1, dtype=torch.float64, device="cuda")
    noalloc_sum(big.data_ptr(), _lib.F64, 900, 700, 900, out, nblocks=1024)   # fills every slot
    torch.cuda.synchronize()
    assert float(out.item()) == 700 * 900 * 1000.0
    for nb in (0, 1, 3):                          # fewer partials than the call before: stale slots unread
        noalloc_sum(small.data_ptr(), _lib.F64, 7, 5, 7, out, nblocks=nb)
        torch.cuda.synchronize()
        assert float(out.item()) == 5 * 7 * 0.5
    # a row pitch wider than nx: the padding must not be summed
    noalloc_sum(big.data_ptr(), _lib.F64, 100, 700, 900, out)
    torch.cuda.synchronize()
    assert float(out.item()) == 700 * 100 * 1000.0


def test_noalloc_checksum_rejects_what_does_not_fit():
    import torch
    t = torch.ones((8, 8), dtype=torch.float64, device="cuda")
    out = torch.zeros(1, dtype=torch.float64, device="cuda")
    L = heatb200.lib()
    rc = L.heat_checksum_noalloc(C.c_void_p(t.data_ptr()), _lib.F64, 8, 8, 8, 1025, C.c_void_p(out.data_ptr()), 0)
    assert rc == _lib.ERR_UNSUPPORTED
    msg = L.heat_last_error().decode()
    assert "1025" in msg and "1024" in msg
    assert L.heat_checksum_noalloc(None, _lib.F64, 8, 8, 8, 0, C.c_void_p(out.data_ptr()), 0) == _lib.ERR_INVALID
    assert L.heat_checksum_noalloc(C.c_void_p(t.data_ptr()), _lib.F64, 8, 8, 4, 0, C.c_void_p(out.data_ptr()), 0) == _lib.ERR_INVALID
    assert L.heat_checksum_noalloc(C.c_void_p(t.data_ptr()), 7, 8, 8, 8, 0, C.c_void_p(out.data_ptr()), 0) == _lib.ERR_INVALID
