"""Host-side plumbing for row-slab runs with one process per GPU (replaces the MPI set-up of
fortran/mpi+cuda/heat.F90:53-103).  torch.distributed only carries the 128-byte ncclUniqueId and test
traffic; halo rows move inside libheat.so (ncclSend/ncclRecv), never through torch.
Works with the `nccl` backend on GPUs and with `gloo` on CPUs (where heat_init then fails loudly).
"""
from __future__ import annotations

import os

import numpy as np

from . import solver as _solver


def env_rank():
    """(rank, world_size, local_rank) from the torchrun environment (defaults: single process)."""
    return (int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")),
            int(os.environ.get("LOCAL_RANK", "0")))


def broadcast_bytes(payload: bytes | None, nbytes: int, src: int = 0, device=None) -> bytes:
    """Broadcast a fixed-size byte string from `src` over the default process group."""
    import torch
    import torch.distributed as dist
    if device is None:
        device = "cuda" if dist.get_backend() == "nccl" else "cpu"
    t = torch.zeros(nbytes, dtype=torch.uint8, device=device)
    if dist.get_rank() == src:
        if payload is None or len(payload) != nbytes:
            raise ValueError("broadcast_bytes: source must supply exactly %d bytes" % nbytes)
        t.copy_(torch.frombuffer(bytearray(payload), dtype=torch.uint8))
    dist.broadcast(t, src)
    return t.cpu().numpy().tobytes()


def shared_nccl_id(make_id=_solver.nccl_unique_id) -> bytes:
    """Rank 0 creates the ncclUniqueId (heat_nccl_unique_id), everyone receives it.
    ref: MPI_Init + MPI_Cart_create, fortran/mpi+cuda/heat.F90:60-62,97."""
    import torch.distributed as dist
    payload = make_id() if dist.get_rank() == 0 else None
    return broadcast_bytes(payload, 128, 0)


def my_rows(ny: int):
    """This rank's (row0, nrows) under the library's partition."""
    import torch.distributed as dist
    return _solver.slab_partition(ny, dist.get_world_size(), dist.get_rank())


def my_tile(nx: int, ny: int, ranks_x: int):
    """This rank's (row0, nrows, col0, ncols) in a ranks_x x (world / ranks_x) decomposition."""
    import torch.distributed as dist
    world = dist.get_world_size()
    return _solver.tile_partition(nx, ny, ranks_x, world // ranks_x, dist.get_rank())


def tile_solver(field_tile: np.ndarray, nx_global: int, ny_global: int, ranks_x: int, rx, ry=None, *,
                nccl_id: bytes | None = None, device: int | None = None, **kw) -> "_solver.HeatSolver":
    """HeatSolver for this rank's tile of a [ny_global, nx_global] field cut into ranks_x x (world / ranks_x) tiles
    (rank = ry * ranks_x + rx); collective.  Halos travel through NCCL: rows directly, columns packed.
    ref: the second decomposition direction the reference reserves, fortran/mpi+cuda/heat.F90:28-29,87-97."""
    import torch.distributed as dist
    rank, world = dist.get_rank(), dist.get_world_size()
    if nccl_id is None and world > 1:
        nccl_id = shared_nccl_id()
    if device is None:
        device = int(os.environ.get("LOCAL_RANK", "0"))
    kw.pop("peer_halo", None)
    s = _solver.HeatSolver(field_tile, rx, ry, nx=nx_global, ny=ny_global, rank=rank, nranks=world, nccl_id=nccl_id,
                           device=device, ranks_x=ranks_x, **kw)
    s.halo_mode = "NCCL send/recv, %d x %d tiles" % (ranks_x, world // ranks_x)
    return s


def slab_solver(field_slab: np.ndarray, ny_global: int, rx, ry=None, *, nccl_id: bytes | None = None,
                device: int | None = None, **kw) -> "_solver.HeatSolver":
    """HeatSolver for this rank's slab of a [ny_global, nx] field; collective."""
    import torch.distributed as dist
    rank, world = dist.get_rank(), dist.get_world_size()
    if nccl_id is None and world > 1:
        nccl_id = shared_nccl_id()
    if device is None:
        device = int(os.environ.get("LOCAL_RANK", "0"))
    peer = kw.pop("peer_halo", False)            # False | True | "auto" (peer stores if CUDA IPC works, else NCCL)
    s = None
    if peer and world > 1:
        ok, err, blob = 1, None, b""
        try:
            s = _solver.HeatSolver(field_slab, rx, ry, ny=ny_global, rank=rank, nranks=world, nccl_id=nccl_id,
                                   device=device, peer_halo=True, **kw)
            blob = s.peer_export()
        except _solver._lib.HeatError as e:
            ok, err = 0, e
        # every rank publishes its CUDA-IPC blob; each takes its two neighbours' (replaces MPI_Cart_shift, :103)
        blobs = [None] * world
        dist.all_gather_object(blobs, blob)          # always matched, whatever happened above
        if ok:
            up = blobs[rank - 1] if rank > 0 else None
            down = blobs[rank + 1] if rank + 1 < world else None
            try:
                if (rank > 0 and not up) or (rank + 1 < world and not down):
                    raise _solver._lib.HeatError(-5, "a neighbour could not export its buffers")
                s.peer_attach(up, down)              # opens the handles; waits for nobody
            except _solver._lib.HeatError as e:
                ok, err = 0, e
        import torch
        flag = torch.tensor([ok], device="cuda" if dist.get_backend() == "nccl" else "cpu")
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        if int(flag.item()) == 1:
            s.halo_mode = "peer stores (CUDA IPC)"
            return s
        if s is not None:
            s.close()
        if peer != "auto":
            raise RuntimeError("peer halo unavailable on at least one rank" + ("" if ok else ": %s" % err))
        nccl_id = shared_nccl_id()                 # a fresh id: the first one may already be consumed
    s = _solver.HeatSolver(field_slab, rx, ry, ny=ny_global, rank=rank, nranks=world, nccl_id=nccl_id,
                           device=device, **kw)
    s.halo_mode = "NCCL send/recv" if world > 1 else "none"
    return s
