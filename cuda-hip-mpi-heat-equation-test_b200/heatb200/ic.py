"""Initial and boundary conditions of the reference's variants (layer L2), generated on the host,
whole-field or for one rank's row slab.  Arrays are [y, x], x contiguous.

ref: python/serial/heat.py:20,25; python/cuda/cuda.py:46-53; fortran/serial/heat.f90:28-48;
fortran/cuda_cuf/heat.F90:72-92; fortran/mpi+cuda/heat.F90:243-251.
"""
from __future__ import annotations

import numpy as np


def hat_pyserial(nx: int, ny: int, length: float = 2, dtype=np.float64, rows=None):
    """u = 1 with u[a:b, a:b] = 2, a = int(.5/d), b = int(1/d + 1)   (python/serial/heat.py:20,25)."""
    dx = length / (nx - 1)
    dy = length / (ny - 1)
    u = np.ones((ny, nx), dtype=dtype)
    u[int(.5 / dy):int(1 / dy + 1), int(.5 / dx):int(1 / dx + 1)] = 2
    return u if rows is None else np.ascontiguousarray(u[rows[0]:rows[0] + rows[1]])


def hat_pycuda(nx: int, ny: int, length: float = 2, dtype=np.float64):
    """python/cuda/cuda.py:46-53 as written: the slice [int(1.5/d):int(1/d+1)] is empty, the field is 1."""
    dx = length / (nx - 1)
    dy = length / (ny - 1)
    u = np.ones((ny, nx), dtype=dtype)
    u[int(1.5 / dy):int(1 / dy + 1), int(1.5 / dx):int(1 / dx + 1)] = 2
    return u


def fortran_coords(n: int, dom_len: float, delta: float) -> np.ndarray:
    """x(1)=0, x(i)=x(i-1)+delta, x(n)=dom_len   (fortran/serial/heat.f90:28-36)."""
    c = np.empty(n)
    acc = 0.0
    c[0] = 0.0
    for i in range(1, n - 1):
        acc = acc + delta
        c[i] = acc
    c[n - 1] = dom_len
    return c


def hat_fortran(n: int, dom_len: float, delta: float, y_hi: float = 1.5, dtype=np.float64, rows=None):
    """T = 2 where 0.5 <= x <= 1.5 and 0.5 <= y <= y_hi, else 1.  y_hi = 1.5: fortran/serial/heat.f90:42;
    y_hi = 1.0: fortran/cuda_cuf/heat.F90:86, fortran/cuda_kernel/heat.F90:99.
    `rows=(row0, nrows)` returns only that slab (what one rank of a row-slab run needs)."""
    c = fortran_coords(n, dom_len, delta)
    inx = (c <= 1.5) & (c >= 0.5)
    iny = (c <= y_hi) & (c >= 0.5)
    r0, nr = (0, n) if rows is None else rows
    u = np.ones((nr, n), dtype=dtype)
    u[np.ix_(iny[r0:r0 + nr], inx)] = 2.0
    return u


def mpi_uniform(nx: int, ny: int, dtype=np.float64, rows=None):
    """The MPI programs' IC seen as a global field with its boundary ring: interior 2, ring 1
    (fortran/mpi+cuda/heat.F90:243-251; there the ring is the ghost layer of an n x n interior,
    so a reference run with grid size n corresponds to nx = ny = n + 2 here)."""
    r0, nr = (0, ny) if rows is None else rows
    u = np.full((nr, nx), 2.0, dtype=dtype)
    u[:, 0] = 1.0
    u[:, -1] = 1.0
    if r0 == 0:
        u[0, :] = 1.0
    if r0 + nr == ny:
        u[-1, :] = 1.0
    return u
