"""ctypes access to the C oracle (oracle/liboracle*.so) and to the host builds of the reference's
own kernel sources (oracle/_ref/*.so) — TEST INFRASTRUCTURE ONLY; see oracle/heat_oracle.c."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
_dp = C.POINTER(C.c_double)
_fp = C.POINTER(C.c_float)


def build(ref: bool = True) -> None:
    """make the oracle libraries; the _ref targets only where the reference tree exists."""
    subprocess.run(["make", "-s", "-C", HERE, "all"], check=True)
    if ref and os.path.isdir(os.environ.get("HEAT_REFERENCE_ROOT", "/root/reference")):
        subprocess.run(["make", "-s", "-C", HERE, "ref"], check=True)


def _load(name):
    path = os.path.join(HERE, name)
    if not os.path.isfile(path):
        build(ref=name.startswith("_ref"))
    return C.CDLL(path)


def lib(omp: bool = False):
    L = _load("liboracle_omp.so" if omp else "liboracle.so")
    L.oracle_march_f64.argtypes = [_dp, C.c_int, C.c_int, C.c_size_t, C.c_double, C.c_double, C.c_int, C.c_int]
    L.oracle_march_f32.argtypes = [_fp, C.c_int, C.c_int, C.c_size_t, C.c_float, C.c_float, C.c_int, C.c_int]
    L.oracle_march_ex_f64.argtypes = L.oracle_march_f64.argtypes + [C.c_int]
    L.oracle_march_ex_f32.argtypes = L.oracle_march_f32.argtypes + [C.c_int]
    L.oracle_threads.argtypes = [C.c_int]
    L.oracle_ic_fortran.argtypes = [_dp, C.c_int, C.c_size_t, C.c_double, C.c_double]
    L.oracle_derive_fortran.argtypes = [C.c_int, C.c_double, C.c_double, C.c_double, _dp, _dp, _dp]
    L.oracle_sum_f64.argtypes = [_dp, C.c_int, C.c_int, C.c_size_t]
    L.oracle_sum_f64.restype = C.c_double
    return L


def threads(n: int = 0, omp: bool = True) -> int:
    """Threads a march of the OpenMP build will use; n > 0 sets that count first (overriding whatever
    OMP_NUM_THREADS the launcher exported — torchrun sets it to 1)."""
    return int(lib(omp).oracle_threads(int(n)))


def march(u0: np.ndarray, steps: int, order: int, rx, ry=None, omp: bool = False, contract: bool = False) -> np.ndarray:
    """C oracle time-march on a copy of ``u0`` ([ny, nx] float64 or float32).  ``contract``: the FMA-fused
    reading of orders F and C that nvcc / hipcc give the reference's kernel text by default."""
    ry = rx if ry is None else ry
    u = np.ascontiguousarray(u0).copy()
    ny, nx = u.shape
    L = lib(omp)
    if u.dtype == np.float64:
        rc = L.oracle_march_ex_f64(u.ctypes.data_as(_dp), nx, ny, nx, rx, ry, order, steps, int(contract))
    elif u.dtype == np.float32:
        rc = L.oracle_march_ex_f32(u.ctypes.data_as(_fp), nx, ny, nx, rx, ry, order, steps, int(contract))
    else:
        raise TypeError(u.dtype)
    if rc != 0:
        raise RuntimeError("oracle_march failed")
    return u


def ic_fortran(n: int, dom_len: float, y_hi: float = 1.5) -> np.ndarray:
    u = np.empty((n, n))
    if lib().oracle_ic_fortran(u.ctypes.data_as(_dp), n, n, dom_len, y_hi) != 0:
        raise RuntimeError("oracle_ic_fortran failed")
    return u


def derive_fortran(n, sigma, nu, dom_len):
    d, t, r = C.c_double(), C.c_double(), C.c_double()
    lib().oracle_derive_fortran(n, sigma, nu, dom_len, C.byref(d), C.byref(t), C.byref(r))
    return d.value, t.value, r.value


# ------------------------------------------------------------------ reference kernels on the host
def ref_available() -> bool:
    return all(os.path.isfile(os.path.join(HERE, "_ref", f)) for f in
               ("libhipref.so", "libpycudaref.so", "libhipref_fma.so", "libpycudaref_fma.so"))


def hipref_step(Td_old: np.ndarray, r: float, nx: int, ny: int, ng: int = 1, contract: bool = False) -> np.ndarray:
    """One launch_heat_eqn (fortran/hip/heat_kernel.cpp:48-61) on a ghosted [(ny+2ng), (nx+2ng)]
    array, preceded by the Td_old = Td copy of fortran/hip/heat.F90:243.  ``contract``: the host build
    with FMA contraction on (what hipcc's default -ffp-contract=fast does to the same text)."""
    L = _load("_ref/libhipref_fma.so" if contract else "_ref/libhipref.so")
    L.launch_heat_eqn.argtypes = [_dp, _dp, C.c_double, C.c_int, C.c_int, C.c_int]
    L.launch_heat_eqn.restype = None
    old = np.ascontiguousarray(Td_old, dtype=np.float64)
    assert old.shape == (ny + 2 * ng, nx + 2 * ng)
    new = old.copy()
    L.launch_heat_eqn(new.ctypes.data_as(_dp), old.ctypes.data_as(_dp), r, nx, ny, ng)
    return new


def hipref_swap(Td: np.ndarray, nx: int, ny: int, ng: int = 1):
    """launch_send (heat_kernel.cpp:137-150): returns (Tds1, Tds2) packed lines of length ng*ny."""
    L = _load("_ref/libhipref.so")
    L.launch_send.argtypes = [_dp, _dp, _dp, C.c_int, C.c_int, C.c_int]
    L.launch_send.restype = None
    a = np.ascontiguousarray(Td, dtype=np.float64).copy()
    s1, s2 = np.zeros(ng * ny), np.zeros(ng * ny)
    L.launch_send(a.ctypes.data_as(_dp), s1.ctypes.data_as(_dp), s2.ctypes.data_as(_dp), nx, ny, ng)
    return s1, s2


def hipref_recv(Td: np.ndarray, Tdr: np.ndarray, which: int, nx: int, ny: int, ng: int = 1) -> np.ndarray:
    """launch_recv1 / launch_recv2 (heat_kernel.cpp:78-90,106-118) on a copy of Td."""
    L = _load("_ref/libhipref.so")
    f = L.launch_recv1 if which == 1 else L.launch_recv2
    f.argtypes = [_dp, _dp, C.c_int, C.c_int, C.c_int]
    f.restype = None
    a = np.ascontiguousarray(Td, dtype=np.float64).copy()
    r = np.ascontiguousarray(Tdr, dtype=np.float64)
    f(a.ctypes.data_as(_dp), r.ctypes.data_as(_dp), nx, ny, ng)
    return a


def pycudaref_step(u_old: np.ndarray, r: float, contract: bool = False) -> np.ndarray:
    """One heatCalculate launch as python/cuda/cuda.py:97-99 issues it (u_new starts as u_old).
    ``contract``: the host build with FMA contraction on (nvcc's default -fmad=true)."""
    L = _load("_ref/libpycudaref_fma.so" if contract else "_ref/libpycudaref.so")
    L.pycuda_ref_launch.argtypes = [_dp, _dp, C.c_int, C.c_int, C.c_double]
    old = np.ascontiguousarray(u_old, dtype=np.float64)
    ny, nx = old.shape
    new = old.copy()
    L.pycuda_ref_launch(new.ctypes.data_as(_dp), old.ctypes.data_as(_dp), nx, ny, r)
    return new
