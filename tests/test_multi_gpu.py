"""N > 1: (gpu) row slabs over 2 real GPUs with NCCL halo exchange, bit-identical to one GPU and to the
oracle; (cpu) world_size-2 gloo tests of the host-side plumbing the N > 1 path depends on."""
import json
import os
import socket
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def torchrun(script, nproc, args=(), timeout=600, env=None):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(nproc),
           "--master-addr", "127.0.0.1", "--master-port", str(free_port()), script, *args]
    return subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, env={**os.environ, **(env or {})})


@pytest.mark.gpu
def test_two_gpu_slabs_bit_identical_to_one_gpu():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (run with gpurun --gpus 2)")
    n = min(torch.cuda.device_count(), 4)
    r = torchrun(os.path.join(ROOT, "tests", "_mgpu_worker.py"), n, timeout=1500)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert r.stdout.count(" ok ") >= 25


# ------------------------------------------------------------------ CPU, gloo, world_size 2
_GLOO_WORKER = r'''
import os, sys
sys.path.insert(0, os.environ["HEAT_ROOT"]); sys.path.insert(0, os.path.join(os.environ["HEAT_ROOT"], "cuda-hip-mpi-heat-equation-test_b200"))
import numpy as np, torch, torch.distributed as dist
import heatb200
from heatb200 import dist as hd, ic, _lib
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
assert hd.env_rank()[:2] == (rank, world)
# 1. the id travels: rank 0's 128 bytes arrive everywhere (a stand-in id; creating a real one needs NCCL + a GPU)
fake = bytes(range(128))
got = hd.shared_nccl_id(make_id=lambda: fake)
assert got == fake and len(got) == 128
# 2. slabs tile the field: every rank builds only its rows; together they are the global IC
ny = nx = 101
delta = 2.0 / 100
r0, nr = hd.my_rows(ny)
mine = ic.hat_fortran(nx, 2.0, delta, rows=(r0, nr))
parts = [None] * world
dist.all_gather_object(parts, (r0, nr, mine))
parts.sort(key=lambda p: p[0])
assert parts[0][0] == 0 and sum(p[1] for p in parts) == ny
assert np.array_equal(np.concatenate([p[2] for p in parts]), ic.hat_fortran(nx, 2.0, delta))
# 3. a slab of the wrong shape is refused before any GPU work
try:
    hd.slab_solver(mine[:-1], ny, 0.25, nccl_id=got)
    raise SystemExit("short slab accepted")
except ValueError:
    pass
# 3b. 2-D tiles: my_tile agrees with the library's partition, the two ranks' tiles of a 2 x 1 cut tile the field,
#     and a tile of the wrong shape is refused before any GPU work
t0 = hd.my_tile(nx, ny, 2)
assert t0 == heatb200.solver.tile_partition(nx, ny, 2, 1, rank) and t0[0] == 0 and t0[1] == ny
tiles = [None] * world
dist.all_gather_object(tiles, t0)
assert tiles[0][2] == 0 and tiles[0][2] + tiles[0][3] == tiles[1][2] and tiles[1][2] + tiles[1][3] == nx
full = ic.hat_fortran(nx, 2.0, delta)
try:
    hd.tile_solver(full[:, t0[2]:t0[2] + t0[3] - 1], nx, ny, 2, 0.25, nccl_id=got)
    raise SystemExit("short tile accepted")
except ValueError:
    pass
# 4. with no GPU the collective init fails loudly on every rank — no CPU fallback
if not torch.cuda.is_available():
    try:
        hd.slab_solver(mine, ny, 0.25, nccl_id=got, ghost=2, overlap=True)
        raise SystemExit("heat_init succeeded without a GPU")
    except heatb200.HeatError as e:
        assert e.code == _lib.ERR_CUDA, e
    try:
        hd.tile_solver(np.ascontiguousarray(full[:, t0[2]:t0[2] + t0[3]]), nx, ny, 2, 0.25, nccl_id=got, ghost=2)
        raise SystemExit("tile_solver succeeded without a GPU")
    except heatb200.HeatError as e:
        assert e.code == _lib.ERR_CUDA, e
dist.barrier()
dist.destroy_process_group()
print("rank %d ok" % rank)
'''


def test_gloo_world2_host_plumbing(tmp_path):
    script = tmp_path / "gloo_worker.py"
    script.write_text(_GLOO_WORKER)
    r = torchrun(str(script), 2, timeout=300, env={"HEAT_ROOT": ROOT, "CUDA_VISIBLE_DEVICES": ""})
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
    assert "rank 0 ok" in r.stdout and "rank 1 ok" in r.stdout


def test_bench_reference_arm_only_rank0_prints():
    """bench.py --impl reference under torchrun: rank 0 prints one JSON line, the others exit 0 silently."""
    r = torchrun(os.path.join(ROOT, "bench.py"), 2, args=["--impl", "reference", "--gpus", "2", "--steps", "1",
                                                           "--warmup", "1", "--timesteps", "2"],
                 timeout=600, env={"CUDA_VISIBLE_DEVICES": ""})
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["n_gpus"] == 2 and d["e2e"]["h2d_bytes_per_step"] == 0
