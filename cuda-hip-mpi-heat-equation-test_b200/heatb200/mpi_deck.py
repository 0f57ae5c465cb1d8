"""Drop-in for the reference's MPI programs (fortran/mpi+cuda/heat.F90, fortran/hip/heat.F90 — `program heat`):
one process per GPU, the 6-value input deck, the same initial field, the same time loop, the same per-rank dumps.

    python -m torch.distributed.run --nproc-per-node N -m heatb200.mpi_deck input.dat [outdir]

What it mirrors (fortran/mpi+cuda/heat.F90): `read(11,*) n, sigma, nu, dom_len, ntime, soln` (:83); delta, dt, r
(:125-129); T = 2 in the n x n interior with a ring of 1 around it (:243-251 — the ring is the reference's ghost
layer at the physical ends, so the field here is (n+2) x (n+2)); ntime steps of the update (:206-219) with the halo
exchange inside heat_step; with soln = 1 every rank writes `solnNNNNN.dat`, `x y T` per line over its own interior
cells with x(i) = (i-1)*delta (:136,277-288); rank 0 prints the time per iteration (:292).
Differences, deliberate: slabs are cut along the slow index (rows), not the contiguous one, so halos are contiguous;
the remainder n mod nproc is spread over the ranks instead of dropped (:92); NCCL or peer stores instead of MPI.
"""
from __future__ import annotations

import os
import sys
import time

import numpy as np

from . import config as _config, dist as _dist, ic as _ic, io as _io


def run(deck_source, outdir: str = ".", *, ghost: int = 1, kernel: str = "auto", peer_halo="auto", device=None):
    """Run the deck on the current torch.distributed world (or on one GPU when none is initialised).
    Returns (deck, interior_rows_of_this_rank [rows, n], seconds_per_iteration)."""
    import torch
    import torch.distributed as dist
    from .solver import HeatSolver, slab_partition
    deck = _config.parse_input_dat(deck_source)
    n = deck.n
    delta, _, r = deck.derive()
    multi = dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1
    rank, world = (dist.get_rank(), dist.get_world_size()) if multi else (0, 1)
    N = n + 2                                        # interior + the ring the reference keeps in its ghost layer
    r0, nr = slab_partition(N, world, rank)
    slab = _ic.mpi_uniform(N, N, rows=(r0, nr))
    if device is None:
        device = int(os.environ.get("LOCAL_RANK", "0"))
    if multi:
        s = _dist.slab_solver(slab, N, r, ghost=ghost, overlap=True, peer_halo=peer_halo, order="fortran", kernel=kernel, device=device)
        dist.barrier()                               # MPI_Barrier before the timer, :253
    else:
        s = HeatSolver(slab, r, order="fortran", kernel=kernel, device=device)
    t1 = time.perf_counter()
    s.step(deck.ntime)
    out = s.gather()
    if multi:
        dist.barrier()
    t2 = time.perf_counter()
    s.close()
    # this rank's interior cells: global rows 1..n, columns 1..n
    lo, hi = max(r0, 1), min(r0 + nr, N - 1)
    interior = out[lo - r0:hi - r0, 1:-1]
    if deck.soln == 1:
        coords = (np.arange(1, N - 1) - 1) * delta   # x(i) = (i-1)*delta for i = 1..n, :136
        os.makedirs(outdir, exist_ok=True)
        _io.write_rank_dump(outdir, rank, coords, coords[lo - 1:hi - 1], interior)
    per_it = (t2 - t1) / max(deck.ntime, 1)
    if rank == 0:
        print(" simulation completed!!!!")
        print(" total time:", per_it)
    return deck, interior, per_it


def main(argv=None):
    argv = sys.argv[1:] if argv is None else argv
    import torch
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world > 1:
        local = int(os.environ.get("LOCAL_RANK", "0"))
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    run(argv[0] if argv else "input.dat", argv[1] if len(argv) > 1 else ".")
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
