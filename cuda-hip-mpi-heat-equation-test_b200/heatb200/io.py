"""Text dumps of the reference's Fortran programs: `x y T` per line, the contiguous (x) index outer.

ref: int.dat / soln.dat, fortran/serial/heat.f90:50-55,77-83; per-rank solnNNNNN.dat,
fortran/mpi+cuda/heat.F90:277-288 (there rank files hold the rank's own lines x(i), i = 1..nx_local, for all y);
parsed by fortran/serial/init.py / out.py (split on whitespace, three floats per line).
Arrays here are [y, x]; a row slab [row0, row0+nrows) therefore owns a range of the reference's INNER index.
"""
from __future__ import annotations

import numpy as np

_FMT = "  %-25.17g  %-25.17g  %-25.17g\n"      # gfortran list-directed real*8 prints 17 significant digits


def write_dump(path: str, x: np.ndarray, y: np.ndarray, T: np.ndarray) -> None:
    """soln.dat / int.dat of a whole [ny, nx] field."""
    ny, nx = T.shape
    with open(path, "w") as f:
        for i in range(nx):
            col = T[:, i]
            f.writelines(_FMT % (x[i], y[j], col[j]) for j in range(ny))


def rank_dump_name(rank: int) -> str:
    """'soln' // I5.5 // '.dat'   (fortran/mpi+cuda/heat.F90:278-281)."""
    return "soln%05d.dat" % rank


def write_rank_dump(directory: str, rank: int, x: np.ndarray, y_slab: np.ndarray, T_slab: np.ndarray) -> str:
    """One rank's file of a row-slab run: the slab's rows [nrows, nx] with their y coordinates."""
    import os
    path = os.path.join(directory, rank_dump_name(rank))
    write_dump(path, x, y_slab, T_slab)
    return path


def read_dump(path: str):
    """(x, y, T[y, x]) from a dump written by the reference or by write_dump (square or rectangular)."""
    a = np.loadtxt(path, ndmin=2)
    if a.shape[1] != 3:
        raise ValueError("%s: expected three columns `x y T`" % path)
    xs, ys = a[:, 0], a[:, 1]
    ny = 1
    while ny < len(xs) and xs[ny] == xs[0]:
        ny += 1
    nx = len(xs) // ny
    if nx * ny != len(xs):
        raise ValueError("%s: not a full nx*ny dump" % path)
    return xs[::ny].copy(), ys[:ny].copy(), a[:, 2].reshape(nx, ny).T.copy()


def checksum_host(T: np.ndarray) -> float:
    """Plain fp64 sum in the reference's order (i outer, j inner): fortran/mpi+cuda/heat.F90:267-272."""
    s = 0.0
    for v in np.asarray(T, dtype=np.float64).T.ravel():
        s += float(v)
    return s
