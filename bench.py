#!/usr/bin/env python
"""bench.py — GLUPS of the FTCS heat-equation march on B200(s), with its HBM roofline, beside the CPU path.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload cfg2|cfg3|cfg4|cfg5] [--kernel ...]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N --steps K --warmup W
    python bench.py --impl reference ...      # the CPU arm (no GPU code on its path)

One "step" = one heat_step(nt) call of `timesteps` time steps over the workload's grid (resident in HBM).
Workloads (BASELINE.json configs): cfg2 4096^2 fp64 x 1000 steps (default; per GPU — weak scaling — when N > 1);
cfg3 16384^2 (--dtype f32 / --tblock k); cfg4 32768^2 strong scaling; cfg5 16384^2 per GPU weak scaling.
Prints ONE JSON line (rank 0).
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "cuda-hip-mpi-heat-equation-test_b200")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import numpy as np  # noqa: E402

METRIC = "GLUPS (lattice updates/s)"
UNIT = "GLUPS"
L2_BYTES = 126e6

WORKLOADS = {
    # name: (nx, ny_per_gpu or ny_global, timesteps per step, scaling)
    "cfg2": dict(nx=4096, ny=4096, timesteps=1000, scaling="weak",
                 desc="4096x4096 fp64, 1000 timesteps per step (BASELINE.json configs[1]); per GPU when N>1"),
    "cfg3": dict(nx=16384, ny=16384, timesteps=500, scaling="weak",
                 desc="16384x16384, 500 timesteps per step (configs[2])"),
    "cfg4": dict(nx=32768, ny=32768, timesteps=100, scaling="strong",
                 desc="32768x32768 fp64 strong scaling over row slabs, 100 timesteps per step (configs[3])"),
    "cfg5": dict(nx=16384, ny=16384, timesteps=100, scaling="weak",
                 desc="16384x16384 per GPU fp64 weak scaling, 100 timesteps per step (configs[4])"),
}


def peaks():
    """Measured HBM peak (driver-written MEASURED_PEAKS.json) or the profiling guide's fallback."""
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """Samples SM clock and throttle reasons of one GPU while the timed region runs (NVML)."""

    def __init__(self, index: int):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thr = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _loop(self):
        nv = self.nv
        names = {"hw_slowdown": 0x8, "sw_power_cap": 0x4, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20,
                 "hw_power_brake": 0x80, "sync_boost": 0x10, "applications_clocks_setting": 0x2}
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    mask = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    mask = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, bit in names.items():
                    if mask & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            self._stop.wait(0.02)

    def __enter__(self):
        if self.nv is not None:
            self._thr = threading.Thread(target=self._loop, daemon=True)
            self._thr.start()
        return self

    def __exit__(self, *exc):
        self._stop.set()
        if self._thr:
            self._thr.join()

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": ["unavailable"]}
        return {"sm_mhz": float(np.median(self.samples)), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(self.samples)}


# --------------------------------------------------------------------------------------- CPU arm
def cpu_march_glups(nx, ny, budget_s, threads_all=True):
    """The oracle's C port of the fortran/serial loop (copy + double loop, op order F) on the host cores.
    bench.py is allowed to execute oracle/ only here: as the timed CPU baseline, never on the GPU path."""
    from oracle import c_oracle
    from heatb200 import ic, config
    deck = config.InputDeck(nx, 0.25, 0.05, 2.0, 1)
    delta, _, r = deck.derive()
    u0 = ic.hat_fortran(nx, 2.0, delta) if nx == ny else np.ones((ny, nx))
    cores = os.cpu_count() or 1
    try:
        cores = len(os.sched_getaffinity(0))
    except Exception:
        pass
    t0 = time.perf_counter()
    c_oracle.march(u0, 2, 0, r, omp=threads_all)                 # calibration + page touch
    per = max((time.perf_counter() - t0) / 2, 1e-4)
    steps = max(3, min(2000, int(budget_s / per)))
    t0 = time.perf_counter()
    c_oracle.march(u0, steps, 0, r, omp=threads_all)
    dt = time.perf_counter() - t0
    glups = (nx - 2) * (ny - 2) * steps / dt / 1e9
    return glups, cores if threads_all else 1, steps, dt


def numpy_reference_glups(nx, ny, budget_s):
    """The reference's own python/serial expression (heat.py:49-58) restated in NumPy, 1 thread."""
    from oracle import heat_oracle as ho
    u = ho.ic_pyserial(nx, ny)
    t0 = time.perf_counter()
    u = ho.step(u, ho.ORDER_PYSERIAL, 0.25, 0.25)
    per = max(time.perf_counter() - t0, 1e-4)
    steps = max(2, min(200, int(budget_s / per)))
    t0 = time.perf_counter()
    for _ in range(steps):
        u = ho.step(u, ho.ORDER_PYSERIAL, 0.25, 0.25)
    dt = time.perf_counter() - t0
    return (nx - 2) * (ny - 2) * steps / dt / 1e9, steps


def run_reference_arm(args, wl):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    from oracle import c_oracle
    from heatb200 import ic, config
    nx, ny = wl["nx"], wl["ny"]
    if nx * ny > 4096 * 4096:                 # bounded sample: a 4096-row band of the workload's width
        ny = 4096
    delta, _, r = config.InputDeck(nx, 0.25, 0.05, 2.0, 1).derive()
    u0 = ic.hat_fortran(nx, 2.0, delta) if nx == ny else ic.mpi_uniform(nx, ny)
    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    t0 = time.perf_counter()
    c_oracle.march(u0, 2, 0, r, omp=True)
    per = max((time.perf_counter() - t0) / 2, 1e-4)
    total_budget = 60.0
    sample_ts = max(1, min(wl["timesteps"], int(total_budget / (args.steps + args.warmup) / per)))
    for _ in range(args.warmup):
        c_oracle.march(u0, sample_ts, 0, r, omp=True)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        c_oracle.march(u0, sample_ts, 0, r, omp=True)
    dt = time.perf_counter() - t0
    glups = (nx - 2) * (ny - 2) * sample_ts * args.steps / dt / 1e9
    sample = ("%dx%d fp64, %d timesteps per step (bounded sample of the workload's %d), C port of the "
              "fortran/serial loop (copy + double loop, op order F), OpenMP over rows" % (nx, ny, sample_ts, wl["timesteps"]))
    line = {
        "impl": "reference", "metric": METRIC, "value": glups, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
        "higher_is_better": True, "scaling": wl["scaling"], "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": args.workload + ": " + wl["desc"], "cpu_sample": sample},
        "cpu_baseline": {"value": glups, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": glups, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)
    return 0


# --------------------------------------------------------------------------------------- GPU arm
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="cfg2", choices=sorted(WORKLOADS))
    ap.add_argument("--kernel", default=os.environ.get("HEAT_BENCH_KERNEL", "auto"))
    ap.add_argument("--dtype", default="f64", choices=["f64", "f32"])
    ap.add_argument("--tblock", type=int, default=1)
    ap.add_argument("--fmad", action="store_true")
    ap.add_argument("--ghost", type=int, default=int(os.environ.get("HEAT_BENCH_GHOST", "32")))
    ap.add_argument("--no-overlap", action="store_true")
    ap.add_argument("--skip-exchange", action="store_true", help="diagnostic: no halo traffic (results wrong)")
    ap.add_argument("--halo", default="auto", choices=["auto", "peer", "nccl"],
                    help="halo exchange: peer stores from the edge-strip kernels over CUDA IPC (auto falls back to NCCL "
                         "if IPC is unavailable) or ncclSend/ncclRecv")
    ap.add_argument("--timesteps", type=int, default=0, help="override time steps per bench step")
    ap.add_argument("--field", default="hat", choices=["hat", "random"],
                    help="synthetic input: the reference's hat IC, or seeded uniform [1,2) (guards against value-dependent shortcuts)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3                       # timing rules: W >= 3
    wl = dict(WORKLOADS[args.workload])
    if args.timesteps > 0:
        wl["timesteps"] = args.timesteps
    if args.impl == "reference":
        return run_reference_arm(args, wl)

    # stdout carries exactly one JSON line: anything libraries print there (e.g. NCCL's version banner)
    # is sent to stderr for the duration of the run
    json_fd = os.dup(1)
    os.dup2(2, 1)

    import torch
    import heatb200
    from heatb200 import config, ic

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            sys.exit("bench.py --gpus %d must be launched with torch.distributed.run --nproc-per-node %d" % (args.gpus, args.gpus))
        args.gpus = world
    if not torch.cuda.is_available():
        sys.exit("bench.py: no CUDA device; the product has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dist = None
    nccl_id = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        from heatb200 import dist as hdist
        nccl_id = hdist.shared_nccl_id()     # rank 0: heat_nccl_unique_id(); broadcast over torch.distributed

    dtype = np.float64 if args.dtype == "f64" else np.float32
    esize = np.dtype(dtype).itemsize
    nx = wl["nx"]
    ny_global = wl["ny"] * world if wl["scaling"] == "weak" else wl["ny"]
    timesteps = wl["timesteps"]
    row0, nrows = heatb200.slab_partition(ny_global, world, rank)
    # synthetic input: the reference's hat IC (fortran/serial/heat.f90:40-48) on the unit-spacing-derived grid;
    # r from the input.dat derivation for a square grid of side nx (exactly 0.25 for all named sizes)
    delta, _, r = config.InputDeck(nx, 0.25, 0.05, 2.0, timesteps).derive()
    if ny_global == nx:
        field = ic.hat_fortran(nx, 2.0, delta, dtype=dtype, rows=(row0, nrows))
    else:                                     # non-square global grids (weak scaling): hat in x, per-rank in y
        field = ic.mpi_uniform(nx, ny_global, dtype=dtype, rows=(row0, nrows))
        field[:, nx // 4: 3 * nx // 4] = 2.0
        field[:, 0] = field[:, -1] = 1.0
        if row0 == 0:
            field[0] = 1.0
        if row0 + nrows == ny_global:
            field[-1] = 1.0
    if args.field == "random":
        interior = (1.0 + np.random.default_rng(1234 + rank).random(field.shape)).astype(dtype)
        interior[:, 0] = interior[:, -1] = 1.0
        if row0 == 0:
            interior[0] = 1.0
        if row0 + nrows == ny_global:
            interior[-1] = 1.0
        field = interior
    pinned = torch.empty(field.shape, dtype=torch.float64 if dtype == np.float64 else torch.float32).pin_memory()
    host_in = pinned.numpy()
    host_in[...] = field
    host_out = torch.empty_like(pinned).pin_memory().numpy()

    # a real (non-default) torch stream: the library launches on it, so torch's events bracket its kernels
    tstream = torch.cuda.Stream()
    torch.cuda.set_stream(tstream)
    stream = tstream.cuda_stream
    common = dict(order="fortran", kernel=args.kernel, tblock=args.tblock, fmad=args.fmad, device=local, stream=stream)
    if world > 1:
        solver = hdist.slab_solver(host_in, ny_global, r, nccl_id=nccl_id, ghost=args.ghost, overlap=not args.no_overlap,
                                   skip_exchange=args.skip_exchange,
                                   peer_halo={"auto": "auto", "peer": True, "nccl": False}[args.halo], **common)
    else:
        solver = heatb200.HeatSolver(host_in, r, **common)
    class solver_halo_mode:                                    # remembered for the JSON line (solver is closed by then)
        v = getattr(solver, "halo_mode", "none")
    lups_per_step = (nx - 2) * (ny_global - 2) * timesteps      # whole job, all ranks

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, k):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(k):
            fn()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        if dist is not None:
            t = torch.tensor([ms], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        barrier()
        return ms

    # ---- device-resident throughput ("value")
    for _ in range(args.warmup):
        solver.step(timesteps)
    l0 = solver.timing().kernel_launches
    with ClockSampler(local) as clk:
        ms = timed(lambda: solver.step(timesteps), args.steps)
    tm = solver.timing()
    launches = tm.kernel_launches - l0
    value = lups_per_step * args.steps / (ms * 1e-3) / 1e9

    # ---- roofline of the dominant kernel: algorithmic bytes per launch / average launch duration over the
    # timed region (CUDA events on the stream the kernels are launched on; launch gaps are charged to the kernel)
    launches_per_step = launches / args.steps
    avg_kernel_ms = ms / max(launches, 1)
    local_lups_per_launch = (nx - 2) * (min(row0 + nrows, ny_global - 1) - max(row0, 1)) * timesteps / max(launches_per_step, 1)
    # algorithmic DRAM bytes per lattice update: read + write once per HBM round trip of `tblock` fused steps
    bytes_per_lup = 2 * esize / max(args.tblock, 1)
    achieved = local_lups_per_launch * bytes_per_lup / (avg_kernel_ms * 1e-3) / 1e9
    peak, peak_src = peaks()
    traffic = None
    try:      # measured DRAM bytes per launch of this kernel/config, from the committed ncu capture (if one exists)
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            traffic = json.load(f).get("%s:%s:%s:k%d" % (args.workload, solver.kernel_name, args.dtype, max(args.tblock, 1)))
    except Exception:
        pass
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": traffic, "algorithmic_bytes_per_launch": local_lups_per_launch * bytes_per_lup, "peak_source": peak_src, "kernel": solver.kernel_name,
                "algorithmic_bytes_per_lup": bytes_per_lup, "avg_kernel_ms": avg_kernel_ms,
                "glups_roof": peak / bytes_per_lup,
                "effective_gbs_at_single_step_traffic": achieved * max(args.tblock, 1)}

    # ---- end to end through the public API with HOST buffers: H2D of the field, the march, D2H of the result
    def e2e_step():
        solver.set_field(host_in)
        solver.step(timesteps)
        solver.gather(out=host_out)
    e2e_step()
    e2e_ms = timed(e2e_step, max(2, min(args.steps, 5)))
    e2e_k = max(2, min(args.steps, 5))
    e2e_value = lups_per_step * e2e_k / (e2e_ms * 1e-3) / 1e9
    slab_bytes = int(nrows * nx * esize)
    e2e = {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": slab_bytes, "d2h_bytes_per_step": slab_bytes,
           "steps": e2e_k, "ms_per_step": e2e_ms / e2e_k}

    # ---- extra, N=1 only: the same workload with 4 time steps fused per HBM round trip (bit-identical results)
    extras = None
    if world == 1 and args.tblock == 1 and not args.no_extras:
        s4 = heatb200.HeatSolver(host_in, r, **{**common, "tblock": 4})
        for _ in range(2):
            s4.step(timesteps)
        k4 = max(2, min(args.steps, 5))
        ms4 = timed(lambda: s4.step(timesteps), k4)

        def e2e4():
            s4.set_field(host_in)
            s4.step(timesteps)
            s4.gather(out=host_out)
        e2e4()
        ems4 = timed(e2e4, k4)
        s4.close()
        extras = {"tblock4": {"value": lups_per_step * k4 / (ms4 * 1e-3) / 1e9, "e2e": lups_per_step * k4 / (ems4 * 1e-3) / 1e9,
                              "unit": UNIT, "note": "temporal blocking k=4, 4 B/LUP algorithmic; same field bits as k=1"}}

    total_launches = launches
    if dist is not None:
        t = torch.tensor([float(launches)], dtype=torch.float64, device="cuda")
        dist.all_reduce(t)
        total_launches = int(t.item())
    halo_ms = tm.last_halo_ms
    solver.close()

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": wl["scaling"], "vs_baseline": None,
            "dtype": args.dtype, "data": "synthetic",
            "config": {"workload": args.workload + ": " + wl["desc"], "nx": nx, "ny_global": ny_global,
                       "timesteps_per_step": timesteps, "kernel": roofline["kernel"], "tblock": args.tblock,
                       "fmad": bool(args.fmad), "op_order": "fortran", "field": args.field,
                       "parallelism": "row-slabs x%d" % world + (", ghost %d, overlap %s, halo via %s" % (args.ghost, not args.no_overlap, solver_halo_mode.v) if world > 1 else ""),
                       "l2": "working set %.0f MB per GPU vs 126 MB L2: inputs larger than L2, no flush" % (2 * nrows * nx * esize / 1e6)},
            "roofline": roofline, "e2e": e2e, "gpu_launches": int(total_launches),
            "clocks": clk.summary(),
        }
        if extras:
            line["extras"] = extras
        if world > 1:
            line["halo"] = {"last_exchange_ms": halo_ms, "exchanges": int(tm.halo_exchanges)}
        if world == 1 and not args.no_cpu_baseline:
            cnx = min(nx, 4096)
            g, cores, st, dt = cpu_march_glups(cnx, cnx, budget_s=12.0)
            gn, stn = numpy_reference_glups(cnx, cnx, budget_s=5.0)
            g1, _, st1, dt1 = cpu_march_glups(cnx, cnx, budget_s=4.0, threads_all=False)
            line["cpu_baseline"] = {
                "value": g, "unit": UNIT, "cores": cores, "kind": "port",
                "sample": "%dx%d fp64, %d timesteps in %.1f s: C port of the fortran/serial loop, OpenMP on %d cores" % (cnx, cnx, st, dt, cores),
                "numpy_1thread": {"value": gn, "sample": "python/serial expression (heat.py:49-58) in NumPy, %dx%d, %d steps" % (cnx, cnx, stn)},
                "c_port_1thread": {"value": g1, "sample": "same C port, 1 thread (the reference's CPU programs are serial), %dx%d, %d steps in %.1f s" % (cnx, cnx, st1, dt1)}}
        sys.stdout.flush()
        os.write(json_fd, (json.dumps(line) + "\n").encode())
    if dist is not None:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
