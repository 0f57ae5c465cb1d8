"""ctypes binding of libheat.so (C ABI declared in include/heat_b200.h).

There is no CPU implementation behind this module: if the shared library is missing it raises,
and every compute entry point needs a CUDA device (heat_init returns HEAT_ERR_CUDA otherwise).
"""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libheat.so")

HEAT_OK, ERR_INVALID, ERR_CUDA, ERR_NCCL, ERR_NOMEM, ERR_UNSUPPORTED = 0, -1, -2, -3, -4, -5
F64, F32 = 0, 1
ORDER_FORTRAN, ORDER_PYCUDA, ORDER_PYSERIAL = 0, 1, 2
KERNEL_AUTO, KERNEL_SIMPLE, KERNEL_STREAM, KERNEL_TMA, KERNEL_RESIDENT, KERNEL_WTMA, KERNEL_PERSIST = 0, 1, 2, 3, 4, 5, 6
FLAG_FMAD, FLAG_OVERLAP, FLAG_SKIP_EXCHANGE, FLAG_PEER_HALO = 1, 2, 4, 8

ORDERS = {"fortran": ORDER_FORTRAN, "pycuda": ORDER_PYCUDA, "pyserial": ORDER_PYSERIAL}
KERNELS = {"auto": KERNEL_AUTO, "simple": KERNEL_SIMPLE, "stream": KERNEL_STREAM, "tma": KERNEL_TMA,
           "resident": KERNEL_RESIDENT, "wtma": KERNEL_WTMA, "persist": KERNEL_PERSIST}
KERNEL_NAMES = {v: k for k, v in KERNELS.items()}


class HeatError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__("libheat error %d: %s" % (code, msg))
        self.code = code


class HeatConfig(C.Structure):
    _fields_ = [
        ("struct_size", C.c_uint32), ("nx", C.c_int32), ("ny", C.c_int32), ("dtype", C.c_int32),
        ("op_order", C.c_int32), ("kernel", C.c_int32), ("tblock", C.c_int32), ("flags", C.c_uint32),
        ("rx", C.c_double), ("ry", C.c_double), ("device", C.c_int32), ("rank", C.c_int32),
        ("nranks", C.c_int32), ("ghost", C.c_int32), ("nccl_id", C.c_void_p),
        ("ext_buf", C.c_void_p * 2), ("ext_pitch", C.c_int64), ("stream", C.c_uint64),
        ("ranks_x", C.c_int32), ("reserved0", C.c_int32),
    ]


class HeatInfo(C.Structure):
    _fields_ = [
        ("nx", C.c_int32), ("ny", C.c_int32), ("row0", C.c_int32), ("nrows", C.c_int32),
        ("ghost", C.c_int32), ("kernel", C.c_int32), ("tblock", C.c_int32), ("pitch", C.c_int64),
        ("buf", C.c_void_p * 2), ("current", C.c_int32), ("device", C.c_int32),
        ("col0", C.c_int32), ("ncols", C.c_int32), ("ghost_left", C.c_int32), ("local_nx", C.c_int32),
    ]


class HeatTiming(C.Structure):
    _fields_ = [
        ("last_step_ms", C.c_double), ("last_halo_ms", C.c_double), ("steps_total", C.c_uint64),
        ("kernel_launches", C.c_uint64), ("halo_exchanges", C.c_uint64),
    ]


# name -> (restype, argtypes); every symbol include/heat_b200.h declares
_H = C.c_void_p
PROTOTYPES = {
    "heat_init": (C.c_int, [C.POINTER(_H), C.POINTER(HeatConfig), C.c_void_p]),
    "heat_set_field": (C.c_int, [_H, C.c_void_p, C.c_int64, C.c_int]),
    "heat_step": (C.c_int, [_H, C.c_int]),
    "heat_halo_exchange": (C.c_int, [_H]),
    "heat_peer_export": (C.c_int, [_H, C.c_void_p]),
    "heat_peer_attach": (C.c_int, [_H, C.c_void_p, C.c_void_p]),
    "heat_gather": (C.c_int, [_H, C.c_void_p, C.c_int, C.c_int]),
    "heat_submit": (C.c_int, [_H, C.c_void_p, C.c_int64, C.c_int, C.c_void_p]),
    "heat_drain": (C.c_int, [_H]),
    "heat_checksum": (C.c_int, [_H, C.POINTER(C.c_double)]),
    "heat_checksum_noalloc": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int64, C.c_int,
                                        C.c_void_p, C.c_uint64]),
    "heat_sync": (C.c_int, [_H]),
    "heat_get_info": (C.c_int, [_H, C.POINTER(HeatInfo)]),
    "heat_get_timing": (C.c_int, [_H, C.POINTER(HeatTiming)]),
    "heat_destroy": (C.c_int, [_H]),
    "heat_last_error": (C.c_char_p, []),
    "heat_version": (C.c_char_p, []),
    "heat_slab_partition": (C.c_int, [C.c_int, C.c_int, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "heat_tile_partition": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int] + [C.POINTER(C.c_int)] * 4),
    "heat_derive_fortran": (C.c_int, [C.c_int, C.c_double, C.c_double, C.c_double,
                                      C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_double)]),
    "heat_nccl_unique_id": (C.c_int, [C.c_void_p]),
    "launch_heat_eqn": (None, [C.c_void_p, C.c_void_p, C.c_double, C.c_int, C.c_int, C.c_int]),
    "launch_send": (None, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int]),
    "launch_recv1": (None, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int]),
    "launch_recv2": (None, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int]),
    "launch_heat_eqn_sp": (None, [C.c_void_p, C.c_void_p, C.c_float, C.c_int, C.c_int, C.c_int]),
    "launch_send_sp": (None, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int]),
    "launch_recv1_sp": (None, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int]),
    "launch_recv2_sp": (None, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int]),
    "heat_legacy_set_fmad": (C.c_int, [C.c_int]),
}

_lib = None


def lib() -> C.CDLL:
    """Load libheat.so once.  Raises (never falls back) if it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.isfile(LIB_PATH):
            raise ImportError(
                "%s is missing: build it with `python cuda-hip-mpi-heat-equation-test_b200/build.py` "
                "(or __graft_entry__.build()). There is no CPU fallback." % LIB_PATH)
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in PROTOTYPES.items():
            f = getattr(L, name)
            f.restype, f.argtypes = res, args
        _lib = L
    return _lib


def check(rc: int) -> None:
    if rc != HEAT_OK:
        raise HeatError(rc, lib().heat_last_error().decode())


def preload_nccl() -> str | None:
    """Map the NCCL that torch ships (if any) before libheat dlopen()s "libnccl.so.2", so both use one copy."""
    try:
        import nvidia.nccl  # type: ignore
        root = list(nvidia.nccl.__path__)[0]
        path = os.path.join(root, "lib", "libnccl.so.2")
        if os.path.isfile(path):
            C.CDLL(path, mode=C.RTLD_GLOBAL)
            return path
    except Exception:
        pass
    return None
