"""HeatSolver — thin object wrapper over the C ABI (heat_init / heat_step / heat_halo_exchange /
heat_gather).  All arithmetic happens in libheat.so on the GPU; this file only marshals arguments.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from ._lib import HeatConfig, HeatInfo, HeatTiming, check, lib

_DTYPES = {np.dtype(np.float64): _lib.F64, np.dtype(np.float32): _lib.F32}


def slab_partition(ny: int, nranks: int, rank: int):
    """(row0, nrows) of `rank` — the library's own partition (pure host code, works without a GPU)."""
    r0, nr = C.c_int(), C.c_int()
    check(lib().heat_slab_partition(ny, nranks, rank, C.byref(r0), C.byref(nr)))
    return r0.value, nr.value


def tile_partition(nx: int, ny: int, ranks_x: int, ranks_y: int, rank: int):
    """(row0, nrows, col0, ncols) of `rank` = ry * ranks_x + rx in a ranks_x x ranks_y decomposition."""
    v = [C.c_int() for _ in range(4)]
    check(lib().heat_tile_partition(nx, ny, ranks_x, ranks_y, rank, *[C.byref(x) for x in v]))
    return tuple(x.value for x in v)


def nccl_unique_id() -> bytes:
    _lib.preload_nccl()
    buf = C.create_string_buffer(128)
    check(lib().heat_nccl_unique_id(buf))
    return buf.raw


class HeatSolver:
    """One rank's share of an nx x ny FTCS march.

    field     : this rank's rows [row0, row0+nrows) x nx as a NumPy array (float64 or float32); with
                nranks == 1 that is the whole [ny, nx] field including its Dirichlet ring.
    rx, ry    : nu*dt/dx^2, nu*dt/dy^2 from the host (config.InputDeck.derive / config.derive_python).
    order     : 'fortran' | 'pycuda' | 'pyserial' — which reference expression to reproduce.
    kernel    : 'auto' | 'simple' | 'stream' | 'tma' | 'resident'.
    fmad      : allow FMA contraction (not bit-exact; <= 1e-12 relative).
    buffers   : optional pair of torch CUDA tensors [(nrows+2*ghost), pitch] used as the ping-pong storage.
    """

    def __init__(self, field, rx, ry=None, *, nx=None, ny=None, dtype=None, order="fortran", kernel="auto",
                 tblock=1, fmad=False, device=0, rank=0, nranks=1, ghost=0, overlap=False,
                 skip_exchange=False, peer_halo=False, nccl_id: bytes | None = None, buffers=None, stream=0,
                 ranks_x=1):
        self._h = C.c_void_p()
        ranks_x = max(1, int(ranks_x))
        if field is not None:
            field = np.ascontiguousarray(field)
            dtype = field.dtype
            if ranks_x == 1:
                nx = field.shape[1]
            if nranks == 1:
                ny = field.shape[0]
        if nx is None or ny is None or dtype is None:
            raise ValueError("HeatSolver: give a field, or nx, ny and dtype")
        dtype = np.dtype(dtype)
        if dtype not in _DTYPES:
            raise TypeError("HeatSolver: dtype must be float64 or float32, got %s" % dtype)
        self.dtype, self.nx, self.ny = dtype, int(nx), int(ny)
        self.rank, self.nranks, self.ranks_x = rank, nranks, ranks_x
        cfg = HeatConfig()
        cfg.struct_size = C.sizeof(HeatConfig)
        cfg.nx, cfg.ny, cfg.dtype = self.nx, self.ny, _DTYPES[dtype]
        cfg.op_order = _lib.ORDERS[order] if isinstance(order, str) else int(order)
        cfg.kernel = _lib.KERNELS[kernel] if isinstance(kernel, str) else int(kernel)
        cfg.tblock = int(tblock)
        cfg.flags = ((_lib.FLAG_FMAD if fmad else 0) | (_lib.FLAG_OVERLAP if overlap else 0) |
                     (_lib.FLAG_SKIP_EXCHANGE if skip_exchange else 0) | (_lib.FLAG_PEER_HALO if peer_halo else 0))
        cfg.rx = float(rx)
        cfg.ry = float(rx if ry is None else ry)
        cfg.device, cfg.rank, cfg.nranks, cfg.ghost = int(device), int(rank), int(nranks), int(ghost)
        self._id_buf = None
        if nranks > 1 and nccl_id is not None:
            _lib.preload_nccl()
            self._id_buf = C.create_string_buffer(bytes(nccl_id), 128)
            cfg.nccl_id = C.cast(self._id_buf, C.c_void_p)
        self._buffers = buffers
        if buffers is not None:
            b0, b1 = buffers
            if b0.shape != b1.shape or b0.dim() != 2 or not b0.is_cuda or not b0.is_contiguous() or not b1.is_contiguous():
                raise ValueError("HeatSolver: buffers must be two contiguous 2-D CUDA tensors of equal shape")
            cfg.ext_buf[0], cfg.ext_buf[1] = b0.data_ptr(), b1.data_ptr()
            cfg.ext_pitch = b0.shape[1]
        cfg.stream = int(stream)
        cfg.ranks_x = ranks_x
        if nranks % ranks_x:
            raise ValueError("HeatSolver: ranks_x %d does not divide nranks %d" % (ranks_x, nranks))
        r0, nr, c0, nc = tile_partition(self.nx, self.ny, ranks_x, nranks // ranks_x, rank)
        self.tile = (r0, nr, c0, nc)
        if field is not None and field.shape != (nr, nc):
            raise ValueError("HeatSolver: rank %d owns rows [%d, %d) x columns [%d, %d): expected field shape %s, got %s"
                             % (rank, r0, r0 + nr, c0, c0 + nc, (nr, nc), field.shape))
        ptr = field.ctypes.data_as(C.c_void_p) if field is not None else None
        check(lib().heat_init(C.byref(self._h), C.byref(cfg), ptr))

    # ------------------------------------------------------------------ the four named calls
    def step(self, nt: int) -> None:
        check(lib().heat_step(self._h, int(nt)))

    def halo_exchange(self) -> None:
        check(lib().heat_halo_exchange(self._h))

    def gather(self, out: np.ndarray | None = None, root: int | None = None):
        """This rank's rows as [nrows, nx]; with `root` set, the whole [ny, nx] field on that rank
        (None elsewhere).  Synchronises."""
        info = self.info()
        if root is None:
            if out is None:
                out = np.empty((info.nrows, info.ncols), dtype=self.dtype)
            check(lib().heat_gather(self._h, out.ctypes.data_as(C.c_void_p), 0, -1))
            return out
        if self.rank == root:
            if out is None:
                out = np.empty((self.ny, self.nx), dtype=self.dtype)
            check(lib().heat_gather(self._h, out.ctypes.data_as(C.c_void_p), 0, root))
            return out
        check(lib().heat_gather(self._h, None, 0, root))
        return None

    def set_field(self, field: np.ndarray) -> None:
        field = np.ascontiguousarray(field, dtype=self.dtype)
        check(lib().heat_set_field(self._h, field.ctypes.data_as(C.c_void_p), field.shape[1], 0))

    def submit(self, field: np.ndarray, nt: int, out: np.ndarray) -> None:
        """Queue one job — upload `field` (this rank's rows), advance nt steps, download into `out` — without
        waiting.  Back-to-back jobs overlap their PCIe copies with each other's march (heat_submit); give pinned
        arrays.  `field` and `out` belong to the library until drain() returns."""
        if field.dtype != self.dtype or out.dtype != self.dtype or not field.flags.c_contiguous or not out.flags.c_contiguous:
            raise ValueError("HeatSolver.submit: field and out must be C-contiguous arrays of dtype %s" % self.dtype)
        if field.shape != out.shape or field.shape != (self.tile[1], self.tile[3]):
            raise ValueError("HeatSolver.submit: field and out must both be this rank's tile %s" % ((self.tile[1], self.tile[3]),))
        check(lib().heat_submit(self._h, field.ctypes.data_as(C.c_void_p), field.shape[1], int(nt),
                                out.ctypes.data_as(C.c_void_p)))

    def drain(self) -> None:
        """Wait for every submitted job (uploads, marches, downloads)."""
        check(lib().heat_drain(self._h))

    def peer_export(self) -> bytes:
        """This rank's CUDA-IPC blob (buffers + flag words) for its neighbours (peer_halo=True only)."""
        buf = C.create_string_buffer(256)
        check(lib().heat_peer_export(self._h, buf))
        return buf.raw

    def peer_attach(self, up_blob: bytes | None, down_blob: bytes | None) -> None:
        """Map the neighbours' buffers and fill the ghosts (collective)."""
        ub = C.create_string_buffer(up_blob, 256) if up_blob else None
        db = C.create_string_buffer(down_blob, 256) if down_blob else None
        check(lib().heat_peer_attach(self._h, ub, db))

    # ------------------------------------------------------------------ extras
    def gather_into_tensor(self, t) -> None:
        """Copy this rank's rows into a contiguous torch CUDA tensor [nrows, nx] (device to device)."""
        check(lib().heat_gather(self._h, C.c_void_p(t.data_ptr()), 1, -1))

    def set_field_from_tensor(self, t) -> None:
        check(lib().heat_set_field(self._h, C.c_void_p(t.data_ptr()), t.stride(0), 1))

    def checksum(self) -> float:
        s = C.c_double()
        check(lib().heat_checksum(self._h, C.byref(s)))
        return s.value

    def sync(self) -> None:
        check(lib().heat_sync(self._h))

    def info(self) -> HeatInfo:
        i = HeatInfo()
        check(lib().heat_get_info(self._h, C.byref(i)))
        return i

    def timing(self) -> HeatTiming:
        t = HeatTiming()
        check(lib().heat_get_timing(self._h, C.byref(t)))
        return t

    @property
    def kernel_name(self) -> str:
        return _lib.KERNEL_NAMES[self.info().kernel]

    def close(self) -> None:
        if self._h:
            lib().heat_destroy(self._h)
            self._h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def march(field, steps, rx, ry=None, **kw):
    """Convenience: upload `field`, advance `steps`, return the final field (single GPU)."""
    with HeatSolver(field, rx, ry, **kw) as s:
        s.step(steps)
        return s.gather()
