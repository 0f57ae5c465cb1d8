"""CPU oracle (NumPy) for the FTCS 5-point heat stencil — TEST INFRASTRUCTURE ONLY.

This module restates, on the CPU, the arithmetic of the reference
cssrikanth/CUDA-HIP-MPI-Heat-equation-test for its one hot path.  It exists to CHECK
the CUDA library; only ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` /
``--impl reference`` legs of ``bench.py`` may import it.  The product package
(``heatb200``) never does, and fails loudly when ``libheat.so`` is missing.

Parity pinning: the reference ships no tests or golden vectors.  The oracle
is pinned instead against outputs of the reference itself, produced in the build container:
``python/serial/heat.py`` executed unmodified (``oracle/run_reference.py``), the reference's
HIP kernel source and PyCUDA kernel string compiled for the host (``oracle/_ref``), all
frozen as fixtures under ``tests/golden/`` by ``tests/golden/make_golden.py``.  Order P is pinned by
the script's own ``diffuse(60)`` (61 steps past its shipped run: rounding-sensitive, so it tells
order P from F and C); orders F and C by random fields through the host-compiled kernels — in their
uncontracted reading (gcc -ffp-contract=off) here, and in the contracted reading that nvcc / hipcc
give the same text by default in oracle/heat_oracle.c (``contract``), which NumPy cannot express.

Every function cites the reference file:line it follows (paths relative to the reference root).
Array convention everywhere: ``u[y, x]`` C-order, x contiguous — the same bytes as the Fortran
``T(i, j)`` with ``i`` (= x) contiguous.
"""
from __future__ import annotations

import re

import numpy as np

ORDER_FORTRAN = 0   # C + r*((((E+N)+W)+S) - 4*C)            fortran/serial/heat.f90:66
ORDER_PYCUDA = 1    # C + r*((((W+E)+N)+S) - 4*C)            python/cuda/cuda.py:78
ORDER_PYSERIAL = 2  # (C + cx*((E-2C)+W)) + cy*((N-2C)+S)    python/serial/heat.py:50-54
ORDERS = {"fortran": ORDER_FORTRAN, "pycuda": ORDER_PYCUDA, "pyserial": ORDER_PYSERIAL}


# --------------------------------------------------------------------------- input.dat
_NUM = re.compile(r"[+-]?(?:\d+\.?\d*|\.\d+)(?:[eEdD][+-]?\d+)?")


def parse_input_dat(text: str):
    """Fortran list-directed ``read(11,*) n, sigma, nu, dom_len, ntime [, soln]``.

    fortran/serial/heat.f90:13 (5 values), fortran/mpi+cuda/heat.F90:83 (6 values).
    Separators: blanks, commas, newlines; ``d``/``D`` exponents; a ``/`` ends the record.
    Returns (n, sigma, nu, dom_len, ntime, soln_or_None).
    """
    text = text.split("/")[0]
    toks = _NUM.findall(text.replace(",", " "))
    if len(toks) < 5:
        raise ValueError("input.dat: need at least 5 values, got %d" % len(toks))
    f = [float(t.replace("d", "e").replace("D", "e")) for t in toks]
    n, ntime = int(f[0]), int(f[4])
    soln = int(f[5]) if len(f) > 5 else None
    return n, f[1], f[2], f[3], ntime, soln


# --------------------------------------------------------------------------- parameters
def derive_fortran(n: int, sigma: float, nu: float, dom_len: float):
    """delta, dt, r exactly as fortran/serial/heat.f90:15,17,59 (all real*8 except real(n-1))."""
    delta = dom_len / float(np.float32(n - 1))      # real(n-1) is default (fp32) real
    dt = (sigma * (delta * delta)) / nu             # delta**2 with integer exponent = delta*delta
    r = (nu * dt) / (delta * delta)
    return delta, dt, r


def derive_python(nx: int, ny: int, nu: float = .05, sigma: float = .25, length: float = 2):
    """dx, dy, dt, cx, cy exactly as python/serial/heat.py:11-15,51,53 (and python/cuda/cuda.py:34-39)."""
    dx = length / (nx - 1)
    dy = length / (ny - 1)
    dt = sigma * dx * dy / nu
    cx = nu * dt / dx**2
    cy = nu * dt / dy**2
    return dx, dy, dt, cx, cy


# --------------------------------------------------------------------------- initial conditions
def ic_pyserial(nx: int, ny: int, length: float = 2):
    """python/serial/heat.py:20,25 — ones with u[a:b, a:b] = 2."""
    dx = length / (nx - 1)
    dy = length / (ny - 1)
    u = np.ones((ny, nx))
    u[int(.5 / dy):int(1 / dy + 1), int(.5 / dx):int(1 / dx + 1)] = 2
    return u


def ic_pycuda(nx: int, ny: int, length: float = 2):
    """python/cuda/cuda.py:46-53 — the slice start int(1.5/dy) exceeds its stop, so the field stays 1."""
    dx = length / (nx - 1)
    dy = length / (ny - 1)
    u = np.ones((ny, nx))
    u[int(1.5 / dy):int(1 / dy + 1), int(1.5 / dx):int(1 / dx + 1)] = 2
    return u


def fortran_coords(n: int, dom_len: float, delta: float):
    """Cumulative-addition coordinates, fortran/serial/heat.f90:28-36."""
    c = np.empty(n)
    c[0] = 0.0
    for i in range(1, n - 1):
        c[i] = c[i - 1] + delta
    c[n - 1] = dom_len
    return c


def ic_fortran(n: int, dom_len: float, y_hi: float = 1.5):
    """Hat IC of fortran/serial/heat.f90:40-48 (y_hi=1.5) or fortran/cuda_cuf/heat.F90:86 and
    fortran/cuda_kernel/heat.F90:99 (y_hi=1.0).  x varies along the contiguous index."""
    delta, _, _ = derive_fortran(n, 0.25, 0.05, dom_len)
    c = fortran_coords(n, dom_len, delta)
    inx = (c <= 1.5) & (c >= 0.5)
    iny = (c <= y_hi) & (c >= 0.5)
    u = np.ones((n, n))
    u[np.ix_(iny, inx)] = 2.0
    return u


def ic_mpi_ghosted(nx: int, ny: int, left_end: bool = True, right_end: bool = True):
    """Ghosted slab IC of fortran/mpi+cuda/heat.F90:243-251: T=2; ghost line 1 at physical ends
    of the decomposed (contiguous) direction; ghost lines 1 in the other direction always.
    Returned as [ny+2, nx+2] with the decomposed index contiguous (= x)."""
    u = np.full((ny + 2, nx + 2), 2.0)
    if left_end:
        u[:, 0] = 1.0
    if right_end:
        u[:, -1] = 1.0
    u[0, :] = 1.0
    u[-1, :] = 1.0
    return u


# --------------------------------------------------------------------------- the stencil
def step(u: np.ndarray, order: int, rx, ry=None) -> np.ndarray:
    """One Jacobi step on the interior of ``u`` (boundary ring copied through).

    ``rx``/``ry`` are scalars of ``u``'s dtype.  Op orders: see module constants.  NumPy evaluates
    every binary op as a separate correctly-rounded ufunc call — no FMA, no re-association.
    """
    ry = rx if ry is None else ry
    t = u.dtype.type
    rx, ry = t(rx), t(ry)
    C = u[1:-1, 1:-1]
    E = u[1:-1, 2:]      # +1 in the contiguous index
    W = u[1:-1, 0:-2]
    N = u[2:, 1:-1]      # +1 in the slow index
    S = u[0:-2, 1:-1]
    out = u.copy()
    if order == ORDER_FORTRAN:
        out[1:-1, 1:-1] = C + rx * ((((E + N) + W) + S) - t(4) * C)
    elif order == ORDER_PYCUDA:
        out[1:-1, 1:-1] = C + rx * ((((W + E) + N) + S) - t(4) * C)
    elif order == ORDER_PYSERIAL:
        out[1:-1, 1:-1] = (C + rx * ((E - t(2) * C) + W)) + ry * ((N - t(2) * C) + S)
    else:
        raise ValueError("bad op order %r" % (order,))
    return out


def march(u0: np.ndarray, steps: int, order: int, rx, ry=None) -> np.ndarray:
    """``steps`` Jacobi steps.  Callers apply the reference's own count convention:
    Fortran runs ``ntime`` (fortran/serial/heat.f90:61), both Python scripts run ``nt + 1``
    (python/serial/heat.py:48, python/cuda/cuda.py:96)."""
    u = np.array(u0, copy=True)
    for _ in range(steps):
        u = step(u, order, rx, ry)
    return u


def pyserial_diffuse(nx=31, ny=31, nt=10, nu=.05, sigma=.25, length=2):
    """Parameterised restatement of python/serial/heat.py:8-25,45-58 (module body + diffuse(nt))."""
    _, _, _, cx, cy = derive_python(nx, ny, nu, sigma, length)
    u = ic_pyserial(nx, ny, length)
    return march(u, nt + 1, ORDER_PYSERIAL, cx, cy)


def pyserial_diffuse_calls(calls, nx=31, ny=31, nu=.05, sigma=.25, length=2, march_fn=None):
    """The module body of python/serial/heat.py followed by ``diffuse(nt)`` for every nt in ``calls``, the
    way the script's own function behaves when called again (heat.py:45-48): each call writes the hat
    value 2 back into the CURRENT field's hat slice and then runs nt + 1 steps."""
    dx, dy, _, cx, cy = derive_python(nx, ny, nu, sigma, length)
    u = ic_pyserial(nx, ny, length)
    march_fn = march if march_fn is None else march_fn
    for nt in calls:
        u = np.array(u, copy=True)
        u[int(.5 / dy):int(1 / dy + 1), int(.5 / dx):int(1 / dx + 1)] = 2      # heat.py:46
        u = march_fn(u, nt + 1, ORDER_PYSERIAL, cx, cy)
    return u


def fortran_serial(n, sigma, nu, dom_len, ntime, y_hi=1.5):
    """fortran/serial/heat.f90:13-69 end to end (y_hi=1.0 gives the cuda_cuf / cuda_kernel IC)."""
    _, _, r = derive_fortran(n, sigma, nu, dom_len)
    return march(ic_fortran(n, dom_len, y_hi), ntime, ORDER_FORTRAN, r)


def mpi_slabs(n, nproc, sigma, nu, dom_len, ntime):
    """Emulates fortran/mpi+cuda/heat.F90:92-93,206-219,243-251 for ``nproc`` ranks on the CPU:
    per-rank ghosted slabs [ny+2, nxl+2] (decomposed index contiguous), update of all interior
    cells, then the two sendrecvs of swap() (:165-166).  Returns the list of per-rank slabs."""
    _, _, r = derive_fortran(n, sigma, nu, dom_len)
    nxl = n // nproc
    slabs = [ic_mpi_ghosted(nxl, n, p == 0, p == nproc - 1) for p in range(nproc)]
    for _ in range(ntime):
        slabs = [step(s, ORDER_FORTRAN, r) for s in slabs]
        first = [s[1:-1, 1].copy() for s in slabs]     # Td1s(1:ng, 1:ny)   :156
        last = [s[1:-1, nxl].copy() for s in slabs]    # Td2s(1:ng, 1:ny)   :157
        for p in range(nproc):
            if p + 1 < nproc:
                slabs[p][1:-1, nxl + 1] = first[p + 1]   # tag 1: right neighbour's first line
            if p > 0:
                slabs[p][1:-1, 0] = last[p - 1]          # tag 2: left neighbour's last line
    return slabs


# --------------------------------------------------------------------------- text dumps
def write_soln(path, x, y, T):
    """``x y T`` per line, contiguous index OUTER (fortran/serial/heat.f90:77-83).  T is [j, i]."""
    n_i, n_j = T.shape[1], T.shape[0]
    with open(path, "w") as f:
        for i in range(n_i):
            for j in range(n_j):
                f.write("  %-25.17g  %-25.17g  %-25.17g\n" % (x[i], y[j], T[j, i]))


def read_soln(path):
    """Parse an ``x y T`` dump the way fortran/serial/out.py:8-33 does; returns (X, Y, Z) squares."""
    rows = [ln.split() for ln in open(path) if len(ln.split()) == 3]
    a = np.array(rows, dtype=float)
    l = int(np.sqrt(len(a)))
    return a[:, 0].reshape(l, l), a[:, 1].reshape(l, l), a[:, 2].reshape(l, l)
