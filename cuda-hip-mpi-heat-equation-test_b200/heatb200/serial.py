"""Drop-in for the reference's python/serial/heat.py — same module globals, same `diffuse(nt)`,
same final field in the module global `u`; the march itself runs in libheat.so on a B200.

    from heatb200 import serial          # or: python -m heatb200.serial [input.dat]
    serial.diffuse(10)                   # nt + 1 = 11 steps, exactly like heat.py:48,70
    serial.u                             # [ny, nx] float64, bit-identical to the reference's u

Differences, all deliberate: importing does not plot or run (heat.py:29-42,61-70 need a GUI;
plotting is unavailable on this box) — call `main()` / run the module for the reference's script
behaviour, which ends by writing `soln.dat` (`x y u` lines) where the reference would open a window.
Optionally `configure(input.dat)` replaces the hard-coded parameters (heat.py:8-15) with a parsed
reference input deck; then `diffuse(ntime - 1)` performs the deck's `ntime` iterations.
"""
from __future__ import annotations

import sys

import numpy

from . import config as _config
from .solver import HeatSolver as _HeatSolver

###variable declarations — python/serial/heat.py:8-15
nx = 31
ny = 31
nt = 17
nu = .05
dx = 2 / (nx - 1)
dy = 2 / (ny - 1)
sigma = .25
dt = sigma * dx * dy / nu
_length = 2

x = numpy.linspace(0, 2, nx)
y = numpy.linspace(0, 2, ny)

u = numpy.ones((ny, nx))
un = numpy.ones((ny, nx))

# hat function I.C. — python/serial/heat.py:25
u[int(.5 / dy):int(1 / dy + 1), int(.5 / dx):int(1 / dx + 1)] = 2

kernel = "auto"     # which libheat kernel advances the field
device = 0


def configure(source=None, *, n=None, sigma_=None, nu_=None, length=None):
    """Re-derive every module global from an input.dat (path or text) or explicit values."""
    global nx, ny, nu, dx, dy, sigma, dt, x, y, u, un, nt, _length
    if source is not None:
        deck = _config.parse_input_dat(source)
        n, sigma_, nu_, length, nt = deck.n, deck.sigma, deck.nu, deck.dom_len, deck.ntime
    nx = ny = int(n if n is not None else nx)
    sigma = sigma if sigma_ is None else sigma_
    nu = nu if nu_ is None else nu_
    _length = _length if length is None else length
    dx = _length / (nx - 1)
    dy = _length / (ny - 1)
    dt = sigma * dx * dy / nu
    x = numpy.linspace(0, _length, nx)
    y = numpy.linspace(0, _length, ny)
    u = numpy.ones((ny, nx))
    un = numpy.ones((ny, nx))
    u[int(.5 / dy):int(1 / dy + 1), int(.5 / dx):int(1 / dx + 1)] = 2


def diffuse(nt):
    """python/serial/heat.py:45-58: re-apply the hat, run nt + 1 Jacobi steps, leave the result in `u`.
    (The four boundary writes of heat.py:55-58 restore the value 1 the ring already has; on the GPU the
    ring is simply never written.)"""
    u[int(.5 / dy):int(1 / dy + 1), int(.5 / dx):int(1 / dx + 1)] = 2
    cx = nu * dt / dx**2
    cy = nu * dt / dy**2
    # (`un` inside the reference's diffuse is a local; the module global `un` stays all ones)
    with _HeatSolver(u, cx, cy, order="pyserial", kernel=kernel, device=device) as s:
        s.step(nt + 1)
        s.gather(out=u)              # in place, like the reference's module global


def save_field(path="soln.dat"):
    """`x y u` per line, contiguous (x) index outer — the Fortran programs' dump format
    (fortran/serial/heat.f90:77-83), readable by the reference's out.py."""
    from . import io as _io
    _io.write_dump(path, x, y, u)


def main(argv=None):
    argv = sys.argv[1:] if argv is None else argv
    if argv:
        configure(argv[0])
        diffuse(nt - 1)
    else:
        diffuse(10)                  # python/serial/heat.py:70
    save_field("soln.dat")
    print("final field: shape %s sum %r max %r -> soln.dat" % (u.shape, float(u.sum()), float(u.max())))


if __name__ == "__main__":
    main()
