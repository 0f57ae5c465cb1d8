"""input.dat parsing and parameter derivation — the host side of the reference's layer L1.

ref: fortran/serial/heat.f90:11-17,59 (read + delta, dt, r); fortran/mpi+cuda/heat.F90:81-85 (6-value form);
python/serial/heat.py:8-15 and python/cuda/cuda.py:31-39 (the hard-coded Python parameters).
"""
from __future__ import annotations

import ctypes as C
import os
import re
from dataclasses import dataclass

from . import _lib

_TOKEN = re.compile(r"[+-]?(?:\d+\.?\d*|\.\d+)(?:[eEdDqQ][+-]?\d+)?")


@dataclass
class InputDeck:
    """One record of input.dat: grid size, sigma, nu, domain length, time iterations [, soln flag]
    (README.md:7 of the reference)."""
    n: int
    sigma: float
    nu: float
    dom_len: float
    ntime: int
    soln: int | None = None

    def derive(self):
        """(delta, dt, r) with the Fortran programs' exact operation sequence (computed by libheat's
        host helper so that Python and any other caller of the C ABI agree bit for bit)."""
        d, t, r = C.c_double(), C.c_double(), C.c_double()
        _lib.check(_lib.lib().heat_derive_fortran(self.n, self.sigma, self.nu, self.dom_len,
                                                  C.byref(d), C.byref(t), C.byref(r)))
        return d.value, t.value, r.value


def parse_input_dat(source: str) -> InputDeck:
    """Parse input.dat text or a path to it with Fortran list-directed rules, as
    `read (11,*) n, sigma, nu, dom_len, ntime[, soln]` does: values separated by blanks, commas or
    newlines; d/D exponents; `/` terminates the record; a 6th value is optional (5-value readers
    ignore it, fortran/serial/heat.f90:13; the MPI programs require it, fortran/mpi+cuda/heat.F90:83)."""
    text = source
    if "\n" not in source and os.path.isfile(source):
        with open(source) as f:
            text = f.read()
    text = text.split("/")[0].replace(",", " ")
    toks = _TOKEN.findall(text)
    if len(toks) < 5:
        raise ValueError("input.dat: expected `n sigma nu dom_len ntime [soln]`, found %d values" % len(toks))
    vals = [float(re.sub(r"[dDqQ]", "e", t)) for t in toks]
    if vals[0] != int(vals[0]) or vals[4] != int(vals[4]):
        raise ValueError("input.dat: grid size and time iterations must be integers")
    n, ntime = int(vals[0]), int(vals[4])
    if n < 2:
        raise ValueError("input.dat: grid size must be >= 2")
    if ntime < 0:
        raise ValueError("input.dat: time iterations must be >= 0")
    soln = int(vals[5]) if len(vals) > 5 else None
    return InputDeck(n, vals[1], vals[2], vals[3], ntime, soln)


def derive_python(nx: int, ny: int, nu: float = .05, sigma: float = .25, length: float = 2):
    """dx, dy, dt and the two stencil coefficients with the Python scripts' exact expressions
    (python/serial/heat.py:12-15,51,53; python/cuda/cuda.py:34-39)."""
    dx = length / (nx - 1)
    dy = length / (ny - 1)
    dt = sigma * dx * dy / nu
    cx = nu * dt / dx**2
    cy = nu * dt / dy**2
    return dx, dy, dt, cx, cy
