"""Runs the reference's python/serial/heat.py UNMODIFIED — TEST INFRASTRUCTURE ONLY.

Works only where /root/reference exists (the build container; not the GPU box).  matplotlib is
not installed and plotting is a graphics dependency that is unavailable on this box, so the
imports at python/serial/heat.py:2-3 are satisfied with inert stand-ins; everything numerical in
the script (heat.py:8-25 and diffuse(10), heat.py:45-58,70) runs as written.  The result is the
script's module global ``u``.  tests/golden/make_golden.py freezes it as a fixture.
"""
from __future__ import annotations

import os
import runpy
import sys
import types

REF = os.environ.get("HEAT_REFERENCE_ROOT", "/root/reference")


class _Inert:
    """Absorbs any attribute access / call chain (pyplot.figure().gca(...).plot_surface(...))."""

    def __getattr__(self, _):
        return self

    def __call__(self, *a, **k):
        return self


def _stub_modules():
    mods = {}
    for name in ("matplotlib", "matplotlib.pyplot", "matplotlib.cm", "mpl_toolkits",
                 "mpl_toolkits.mplot3d"):
        m = types.ModuleType(name)
        m.__getattr__ = lambda _n, _i=_Inert(): _i   # module-level __getattr__ (PEP 562)
        mods[name] = m
    mods["matplotlib"].pyplot = mods["matplotlib.pyplot"]
    mods["matplotlib"].cm = mods["matplotlib.cm"]
    mods["mpl_toolkits"].mplot3d = mods["mpl_toolkits.mplot3d"]
    return mods


def available() -> bool:
    return os.path.isfile(os.path.join(REF, "python", "serial", "heat.py"))


def run_python_serial():
    """Execute python/serial/heat.py; return its globals (``u`` is the final field, [ny, nx] fp64)."""
    saved = {k: sys.modules.get(k) for k in _stub_modules()}
    sys.modules.update(_stub_modules())
    try:
        return runpy.run_path(os.path.join(REF, "python", "serial", "heat.py"), run_name="__main__")
    finally:
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v


if __name__ == "__main__":
    g = run_python_serial()
    u = g["u"]
    print(u.shape, u.dtype, repr(float(u.sum())), repr(float(u.max())), repr(float(u[15, 15])))
