"""Drop-in for the reference's single-process Fortran programs (`program heat` of fortran/serial/heat.f90,
fortran/cuda_cuf/heat.F90, fortran/cuda_kernel/heat.F90): input.dat in, int.dat and soln.dat out.

    python -m heatb200.fortran_deck [input.dat [outdir [serial|cuda]]]

What it mirrors (fortran/serial/heat.f90): `read (11,*) n, sigma, nu, dom_len, ntime` (:13); delta, dt (:15-17);
coordinates by cumulative addition with the last point forced to dom_len (:28-36); the hat initial condition
0.5 <= x <= 1.5, 0.5 <= y <= 1.5 (:40-48) — the CUDA Fortran variants use y <= 1.0 instead
(fortran/cuda_cuf/heat.F90:86, fortran/cuda_kernel/heat.F90:99), select them with variant="cuda"; `int.dat` (:50-55);
r (:59); ntime steps of the update (:61-69) in operation order F; `soln.dat` (:77-83); "simulation completed!!!!" and
the total time (:71-74).  Deliberately absent: the per-step `print*,"time_it:",i` (:62), which serialises the loop on stdout.
The march runs in libheat.so on a B200; there is no CPU path.
"""
from __future__ import annotations

import os
import sys
import time

from . import config as _config, ic as _ic, io as _io
from .solver import HeatSolver as _HeatSolver


def run(deck_source="input.dat", outdir: str = ".", *, variant: str = "serial", kernel: str = "auto", device: int = 0,
        write_files: bool = True):
    """Returns (deck, coordinates, final field T[y, x]).  variant: "serial" (y <= 1.5) or "cuda" (y <= 1.0)."""
    if variant not in ("serial", "cuda"):
        raise ValueError("fortran_deck.run: variant must be 'serial' or 'cuda'")
    deck = _config.parse_input_dat(deck_source)
    delta, _, r = deck.derive()
    coords = _ic.fortran_coords(deck.n, deck.dom_len, delta)
    T0 = _ic.hat_fortran(deck.n, deck.dom_len, delta, y_hi=1.5 if variant == "serial" else 1.0)
    if write_files:
        os.makedirs(outdir, exist_ok=True)
        _io.write_dump(os.path.join(outdir, "int.dat"), coords, coords, T0)
    start = time.perf_counter()
    with _HeatSolver(T0, r, order="fortran", kernel=kernel, device=device) as s:
        s.step(deck.ntime)
        T = s.gather()
    finish = time.perf_counter()
    print(" simulation completed!!!!")
    print(" total time:", finish - start)
    if write_files:
        _io.write_dump(os.path.join(outdir, "soln.dat"), coords, coords, T)
    return deck, coords, T


def main(argv=None):
    argv = sys.argv[1:] if argv is None else argv
    run(argv[0] if argv else "input.dat", argv[1] if len(argv) > 1 else ".", variant=argv[2] if len(argv) > 2 else "serial")


if __name__ == "__main__":
    main()
