"""Worker for tests/test_multi_gpu.py: run under torch.distributed.run with one rank per GPU.
Each case marches a field decomposed into row slabs (NCCL halo exchange inside libheat.so) and compares the
globally gathered result on rank 0, bit for bit, with the same march on one GPU and with the oracle."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "cuda-hip-mpi-heat-equation-test_b200")):
    sys.path.insert(0, p)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import heatb200  # noqa: E402
from heatb200 import dist as hd  # noqa: E402
from oracle import c_oracle, heat_oracle as ho  # noqa: E402

R100 = 0.24999999999999997


def main():
    rank, world, local = hd.env_rank()
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    failures = []
    cases = [
        # (ny, nx, steps, ghost, overlap, tblock, kernel, dtype)
        (64, 96, 9, 1, False, 1, "simple", np.float64),      # the reference's protocol: exchange every step
        (64, 96, 9, 1, True, 1, "stream", np.float64),
        (203, 130, 23, 4, True, 1, "tma", np.float64),       # ragged rows (203 % world != 0), deep halo
        (203, 130, 23, 4, False, 2, "tma", np.float64),      # fused steps under a deep halo
        (512, 1000, 40, 8, True, 4, "tma", np.float64),
        (512, 1000, 37, 8, True, 2, "tma", np.float32),
        (1024, 2048, 30, 16, True, 1, "tma", np.float64),
        # peer-memory halo (CUDA IPC + copy engines + flag words) instead of NCCL
        (64, 96, 9, 1, False, 1, "simple", np.float64, True),
        (203, 130, 23, 4, True, 2, "tma", np.float64, True),
        (1024, 2048, 33, 16, True, 1, "tma", np.float64, True),
        (512, 1000, 37, 8, True, 4, "wtma", np.float32, True),
        # the bench's settings: ghost 32, and rows as wide as the strong-scaling workload's (32768 columns =
        # 256 strips, another grid shape for the TMA kernels' segment choice), k = 1, 2, 4, both halo modes
        (1024, 32768, 70, 32, True, 1, "tma", np.float64, True),
        (1024, 32768, 70, 32, True, 4, "tma", np.float64, True),
        (1024, 32768, 69, 32, True, 1, "tma", np.float64),
        (2048, 16384, 75, 32, True, 2, "tma", np.float32, True),
        (8191, 4096, 69, 32, True, 1, "tma", np.float64, True),      # the case bench.py re-checks after its timed region
        (8191, 4096, 70, 32, False, 4, "tma", np.float64, True),
    ]
    c_oracle.threads(max(1, (os.cpu_count() or 1) // max(world, 1)), omp=True)   # torchrun exports OMP_NUM_THREADS=1
    for ci, case in enumerate(cases):
        ny, nx, steps, g, overlap, k, kernel, dtype = case[:8]
        peer = len(case) > 8 and case[8]
        full = (1.0 + np.random.default_rng(100 + ci).random((ny, nx))).astype(dtype)
        r0, nr = hd.my_rows(ny)
        nccl_id = hd.shared_nccl_id()              # a ncclUniqueId serves exactly one communicator
        s = hd.slab_solver(full[r0:r0 + nr], ny, R100, 0.2437, nccl_id=nccl_id, device=local, order="pyserial",
                           kernel=kernel, tblock=k, ghost=g, overlap=overlap, peer_halo=peer)
        s.step(steps // 2)
        if ci % 2:
            s.halo_exchange()                      # an explicit, extra exchange must change nothing (heat_halo_exchange)
        s.step(steps - steps // 2)                 # split: exercises ghost state carried between calls
        got = s.gather(root=0)
        local_rows = s.gather()
        ck = s.checksum()
        t = s.timing()
        s.close()
        cks = torch.tensor([ck], dtype=torch.float64, device="cuda")
        dist.all_reduce(cks)
        if rank == 0:
            if ny * nx > (1 << 22):      # big cases: the C restatement (bit-equal to the NumPy one, tests/test_oracle.py)
                want = c_oracle.march(full, steps, ho.ORDER_PYSERIAL, dtype(R100), dtype(0.2437), omp=True)
            else:
                want = ho.march(full, steps, ho.ORDER_PYSERIAL, dtype(R100), dtype(0.2437))
            one = heatb200.march(full, steps, R100, 0.2437, order="pyserial", kernel=kernel, tblock=k, device=local)
            checks = {
                "global_vs_oracle": np.array_equal(got.view(np.uint8), want.view(np.uint8)),
                "global_vs_one_gpu": np.array_equal(got.view(np.uint8), one.view(np.uint8)),
                "local_rows": np.array_equal(local_rows, want[r0:r0 + nr]),
                "checksum": abs(float(cks.item()) - float(want.astype(np.float64).sum())) <= 1e-9 * abs(float(want.sum())),
            }
            ok = all(checks.values())
            if not ok:
                failures.append((ci, ny, nx, steps, g, overlap, k, kernel, str(dtype), checks, float(cks.item()),
                                 float(want.astype(np.float64).sum()),
                                 "one!=want at %s" % np.argwhere(one != want)[:6].tolist(),
                                 "local!=want at %s" % np.argwhere(local_rows != want[r0:r0 + nr])[:6].tolist()))
            print("case %d%s %s exchanges=%d launches=%d" % (ci, " (peer)" if peer else "", "ok" if ok else "FAIL", t.halo_exchanges, t.kernel_launches), flush=True)
    # job pipeline over the ranks (heat_submit / heat_drain): two different fields queued back to back; every rank
    # checks its own rows of both results against the oracle
    for peer in (False, True):
        ny, nx, steps = 512, 1000, 21
        fields = [1.0 + np.random.default_rng(300 + j).random((ny, nx)) for j in range(2)]
        r0, nr = hd.my_rows(ny)
        s = hd.slab_solver(fields[0][r0:r0 + nr], ny, R100, 0.2437, nccl_id=hd.shared_nccl_id(), device=local,
                           order="pyserial", kernel="tma", tblock=2, ghost=8, overlap=True, peer_halo=peer)
        outs = [np.empty((nr, nx)) for _ in range(2)]
        for j in range(2):
            s.submit(np.ascontiguousarray(fields[j][r0:r0 + nr]), steps, outs[j])
        s.drain()
        s.close()
        bad = sum(not np.array_equal(outs[j], ho.march(fields[j], steps, ho.ORDER_PYSERIAL, R100, 0.2437)[r0:r0 + nr])
                  for j in range(2))
        t = torch.tensor([bad], device="cuda")
        dist.all_reduce(t)
        if rank == 0:
            print("pipeline%s %s" % (" (peer)" if peer else "", "ok" if int(t.item()) == 0 else "FAIL"), " ", flush=True)
            if int(t.item()):
                failures.append(("pipeline", peer, int(t.item())))
    # 2-D (px x py) decomposition: tiles with ghost rows AND ghost columns (columns travel packed), every px > 1
    # that divides the world size — 2 x 1 on two GPUs, 2 x 2 and 4 x 1 on four, 2 x 4, 4 x 2 and 8 x 1 on eight
    tile_cases = [
        # (ny, nx, steps, ghost, tblock, kernel, dtype)
        (64, 96, 9, 1, 1, "simple", np.float64),
        (203, 260, 23, 4, 1, "tma", np.float64),
        (512, 1000, 37, 8, 2, "tma", np.float32),
        (1024, 2048, 35, 16, 4, "tma", np.float64),
        (777, 4096, 70, 32, 1, "stream", np.float64),
    ]
    for px in [d for d in range(2, world + 1) if world % d == 0]:
        for ci, (ny, nx, steps, g, k, kernel, dtype) in enumerate(tile_cases):
            if nx // px < g or ny // (world // px) < g:
                continue
            full = (1.0 + np.random.default_rng(500 + ci).random((ny, nx))).astype(dtype)
            r0, nr, c0, nc = hd.my_tile(nx, ny, px)
            s = hd.tile_solver(full[r0:r0 + nr, c0:c0 + nc], nx, ny, px, R100, 0.2437, nccl_id=hd.shared_nccl_id(), device=local,
                               order="pyserial", kernel=kernel, tblock=k, ghost=g)
            s.step(steps // 2)
            if ci % 2:
                s.halo_exchange()
            s.step(steps - steps // 2)
            got = s.gather(root=0)
            mine = s.gather()
            ck = s.checksum()
            s.close()
            want = c_oracle.march(full, steps, ho.ORDER_PYSERIAL, dtype(R100), dtype(0.2437), omp=True)
            bad = int(not np.array_equal(mine, want[r0:r0 + nr, c0:c0 + nc]))
            bad += int(abs(ck - float(want[r0:r0 + nr, c0:c0 + nc].astype(np.float64).sum())) > 1e-9 * abs(ck))
            if rank == 0:
                bad += int(not np.array_equal(got.view(np.uint8), want.view(np.uint8)))
            t = torch.tensor([bad], device="cuda")
            dist.all_reduce(t)
            if rank == 0:
                print("tiles %dx%d case %d %s" % (px, world // px, ci, "ok" if int(t.item()) == 0 else "FAIL"), " ", flush=True)
                if int(t.item()):
                    failures.append(("tiles", px, ci, int(t.item())))
    # the MPI programs' own deck (fortran/mpi+cuda/input.dat: 100 0.25 0.05 2.0 10 1) through heatb200.mpi_deck: every rank
    # writes its solnNNNNN.dat; stitched together they are, value for value, the field the reference's HIP kernels
    # produce for that deck (tests/golden/hip_slabs_n100.npz)
    import tempfile
    from heatb200 import io as hio, mpi_deck
    tmp = [tempfile.mkdtemp(prefix="heat_deck_") if rank == 0 else None]
    dist.broadcast_object_list(tmp, 0)
    deck, mine, _ = mpi_deck.run("100 0.25 0.05 2.0 10 1", tmp[0], ghost=2)
    dist.barrier()
    if rank == 0:
        gold = np.load(os.path.join(ROOT, "tests", "golden", "hip_slabs_n100.npz"))["P1_rank0"][1:-1, 1:-1]
        parts = [hio.read_dump(os.path.join(tmp[0], hio.rank_dump_name(r))) for r in range(world)]
        glued = np.concatenate([p[2] for p in parts], axis=0)
        ys = np.concatenate([p[1] for p in parts])
        ok = glued.shape == gold.shape and np.array_equal(glued, gold) and np.array_equal(ys, parts[0][0])
        print("mpi deck dumps %s" % ("ok" if ok else "FAIL"), " ", flush=True)
        if not ok:
            failures.append(("mpi_deck", glued.shape, gold.shape))
    flag = torch.tensor([len(failures)], device="cuda")
    dist.broadcast(flag, 0)
    dist.destroy_process_group()
    if rank == 0 and failures:
        print("FAILURES:", failures, flush=True)
    sys.exit(1 if int(flag.item()) else 0)


if __name__ == "__main__":
    main()
