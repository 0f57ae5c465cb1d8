#!/usr/bin/env python
"""bench.py — GLUPS of the FTCS heat-equation march on B200(s), with its HBM roofline, beside the CPU path.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload auto|cfg2|cfg3|cfg4|cfg5] [--kernel ...]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N --steps K --warmup W
    python bench.py --impl reference ...      # the CPU arm (no GPU code, no libheat.so on its path)

One "step" = one heat_step(nt) call of `timesteps` time steps over the workload's grid (resident in HBM).
Workloads are BASELINE.json's configs:
  cfg3  16384^2 fp64, 500 time steps per step, k = 1 — the configuration the >= 85 %-of-HBM-roofline target is
        quoted on; what `--workload auto` runs at N = 1 on a single-GPU box
  cfg4  32768^2 fp64 STRONG scaling over row slabs, 100 time steps per step — the >= 90 %-efficiency target;
        what `--workload auto` runs at every N (N = 1 included) when more than one GPU is visible, so that a
        1 -> 2 -> 4 -> 8 sweep on one box is a single strong-scaling series
  cfg2  4096^2 fp64, 1000 time steps (its 268 MB working set is L2-assisted; reported under `extras`)
  cfg5  16384^2 per GPU, weak scaling
Prints ONE JSON line (rank 0).  For N > 1 the line carries "parity": after the timed region a ragged mid-size
field is marched over the N ranks (same halo mode, k = 1 and k = 4) and compared bit for bit with the 1-GPU
march, and a band of it with the CPU oracle; a mismatch makes every rank exit non-zero.
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
    "cfg2": dict(nx=4096, ny=4096, timesteps=1000, scaling="weak",
                 desc="4096x4096 fp64, 1000 timesteps per step (BASELINE.json configs[1]); per GPU when N>1"),
    "cfg3": dict(nx=16384, ny=16384, timesteps=500, scaling="weak",
                 desc="16384x16384, 500 timesteps per step, single B200 (BASELINE.json configs[2])"),
    "cfg4": dict(nx=32768, ny=32768, timesteps=100, scaling="strong",
                 desc="32768x32768 fp64 strong scaling over row slabs, 100 timesteps per step (BASELINE.json configs[3])"),
    "cfg5": dict(nx=16384, ny=16384, timesteps=100, scaling="weak",
                 desc="16384x16384 per GPU fp64 weak scaling, 100 timesteps per step (BASELINE.json configs[4])"),
}


def peaks():
    """Measured HBM peak (driver-written MEASURED_PEAKS.json) or the profiling guide's fallback."""
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def visible_gpus() -> int:
    """GPUs this process could use, without creating a CUDA context (NVML; honours CUDA_VISIBLE_DEVICES)."""
    cvd = os.environ.get("CUDA_VISIBLE_DEVICES")
    if cvd is not None:
        return len([t for t in cvd.split(",") if t.strip()])
    try:
        import pynvml
        pynvml.nvmlInit()
        return int(pynvml.nvmlDeviceGetCount())
    except Exception:
        return 1


def choose_workload(name: str, world: int) -> tuple[str, str]:
    if name != "auto":
        return name, "--workload " + name
    if world > 1:
        return "cfg4", "auto: N > 1 runs the strong-scaling workload"
    if visible_gpus() > 1:
        return "cfg4", "auto: more than one GPU is visible, so N = 1 is the first point of the strong-scaling series"
    return "cfg3", "auto: single-GPU box runs the roofline workload"


def workload_config(args, name, why, wl, world):
    """The `config` object — what is run, derived from the command line and the box only, so the GPU arm and
    the CPU arm (--impl reference) print the same one.  What the run resolved to (kernel, halo mode) is under `run`."""
    ny_global = wl["ny"] * world if wl["scaling"] == "weak" else wl["ny"]
    esize = 8 if args.dtype == "f64" else 4
    ws = 2 * -(-ny_global // world) * wl["nx"] * esize
    return {"workload": name + ": " + wl["desc"], "workload_choice": why, "nx": wl["nx"], "ny_global": ny_global,
            "timesteps_per_step": wl["timesteps"], "dtype": args.dtype, "tblock": args.tblock, "fmad": bool(args.fmad),
            "op_order": "fortran", "field": args.field, "kernel_requested": args.kernel,
            "parallelism": "row-slabs x%d" % world + (", ghost %d, overlap %s, halo %s" % (args.ghost, not args.no_overlap, args.halo) if world > 1 else ""),
            "l2": "working set %.0f MB per GPU = %.1f x the 126 MB L2, no flush between steps; consecutive steps sweep in "
                  "opposite directions, so at most L2/working-set = %.0f %% of a step's bytes can be L2 hits "
                  "(roofline.traffic vs algorithmic_bytes_per_launch shows what reached DRAM)"
                  % (ws / 1e6, ws / L2_BYTES, min(1.0, L2_BYTES / ws) * 100)}


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
# bench.py executes oracle/ only below: as the timed CPU baseline / reference arm and as the CHECKER of the
# N > 1 parity cases — never as the thing measured on the GPU path.  Nothing here imports heatb200.
def host_threads() -> int:
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def cpu_field(nx, ny):
    """The workload's synthetic field from the oracle's own generators (hat IC of fortran/serial/heat.f90:40-48 on
    square grids; a band of an n-wide field otherwise) and r from its input.dat derivation."""
    from oracle import heat_oracle as ho
    _, _, r = ho.derive_fortran(nx, 0.25, 0.05, 2.0)
    if nx == ny:
        return ho.ic_fortran(nx, 2.0), r
    u = np.ones((ny, nx))
    u[1:-1, nx // 4: 3 * nx // 4] = 2.0
    return u, r


def cpu_march_glups(nx, ny, budget_s, threads):
    """The oracle's C port of the fortran/serial loop (copy + double loop, op order F) on `threads` host threads."""
    from oracle import c_oracle
    u0, r = cpu_field(nx, ny)
    omp = threads > 1
    used = c_oracle.threads(threads, omp=True) if omp else 1
    t0 = time.perf_counter()
    c_oracle.march(u0, 2, 0, r, omp=omp)                 # calibration + page touch
    per = max((time.perf_counter() - t0) / 2, 1e-4)
    steps = max(3, min(2000, int(budget_s / per)))
    t0 = time.perf_counter()
    c_oracle.march(u0, steps, 0, r, omp=omp)
    dt = time.perf_counter() - t0
    return (nx - 2) * (ny - 2) * steps / dt / 1e9, used, steps, dt


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


def run_reference_arm(args, name, why, wl, world):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    from oracle import c_oracle
    nx, ny = wl["nx"], wl["ny"]
    if nx * ny > 4096 * 4096:                 # bounded sample: a band of the workload's full width
        ny = max(1024, 4096 * 4096 // nx)
    u0, r = cpu_field(nx, ny)
    # explicit thread count: launchers export OMP_NUM_THREADS=1 (torchrun does), which must not shrink the CPU arm
    threads = c_oracle.threads(host_threads(), omp=True)
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
    sample = ("%dx%d fp64 band, %d timesteps per step (bounded sample of the workload's %dx%d x %d), C port of the "
              "fortran/serial loop (copy + double loop, op order F), OpenMP on %d threads"
              % (nx, ny, sample_ts, wl["nx"], wl["ny"], wl["timesteps"], threads))
    line = {
        "impl": "reference", "metric": METRIC, "value": glups, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
        "higher_is_better": True, "scaling": wl["scaling"], "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(args, name, why, wl, world),
        "cpu_baseline": {"value": glups, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": glups, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)
    return 0


# --------------------------------------------------------------------------------------- N > 1 parity
def multi_gpu_parity(args, hdist, heatb200, dist, torch, rank, world, local, stream):
    """Ragged mid-size field over the N ranks with the bench's own halo settings, k = 1 and k = 4: the gathered
    field must equal the 1-GPU march bit for bit, and a band across a slab boundary must equal the CPU oracle.
    ref: the decomposed run equals the undecomposed one, fortran/mpi+cuda/heat.F90:145-193,206-219."""
    ny, nx = 8191, 4096                       # 8191 rows: ragged over 2, 4 and 8 ranks
    r = 0.24999999999999997                   # n = 100's coefficient: not a power of two, every rounding matters
    full = 1.0 + np.random.default_rng(2024).random((ny, nx))
    full[0] = full[-1] = 1.0
    full[:, 0] = full[:, -1] = 1.0
    r0, nr = heatb200.slab_partition(ny, world, rank)
    cases, ok, detail = 0, True, []
    for k, steps in ((1, 2 * args.ghost + 5), (4, 2 * args.ghost + 6)):
        nccl_id = hdist.shared_nccl_id()
        s = hdist.slab_solver(full[r0:r0 + nr], ny, r, nccl_id=nccl_id, ghost=args.ghost, overlap=not args.no_overlap,
                              peer_halo={"auto": "auto", "peer": True, "nccl": False}[args.halo], order="fortran",
                              kernel=args.kernel, tblock=k, device=local, stream=stream)
        s.step(steps)
        got = s.gather(root=0)
        mode = s.halo_mode
        s.close()
        cases += 1
        if rank == 0:
            one = heatb200.march(full, steps, r, order="fortran", kernel=args.kernel, tblock=k, device=local)
            same = bool(np.array_equal(got.view(np.uint64), one.view(np.uint64)))
            # band across the first slab boundary, cut `steps` rows deeper than what is compared
            from oracle import c_oracle
            b0 = heatb200.slab_partition(ny, world, 1)[0]
            lo, hi = b0 - 32 - steps, b0 + 32 + steps
            want = c_oracle.march(full[lo:hi], steps, 0, r, omp=False)[steps:-steps]
            band = bool(np.array_equal(got[lo + steps:hi - steps].view(np.uint64), want.view(np.uint64)))
            ok = ok and same and band
            detail.append({"k": k, "steps": steps, "ny": ny, "nx": nx, "ghost": args.ghost, "halo": mode,
                           "equals_1gpu": same, "band_equals_cpu_oracle": band})
    flag = torch.tensor([1 if ok else 0], device="cuda")
    dist.broadcast(flag, 0)
    return {"cases": cases, "ok": bool(int(flag.item())), "detail": detail}


# --------------------------------------------------------------------------------------- GPU arm
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="auto", choices=["auto"] + sorted(WORKLOADS))
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
    ap.add_argument("--no-parity", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3                       # timing rules: W >= 3
    world = int(os.environ.get("WORLD_SIZE", "1"))
    name, why = choose_workload(args.workload, max(world, args.gpus))
    wl = dict(WORKLOADS[name])
    if args.timesteps > 0:
        wl["timesteps"] = args.timesteps
    if args.impl == "reference":
        return run_reference_arm(args, name, why, wl, max(world, args.gpus))

    # stdout carries exactly one JSON line: anything libraries print there (e.g. NCCL's version banner)
    # is sent to stderr for the duration of the run
    json_fd = os.dup(1)
    os.dup2(2, 1)

    import torch
    import heatb200
    from heatb200 import config, ic

    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            sys.exit("bench.py --gpus %d must be launched with torch.distributed.run --nproc-per-node %d" % (args.gpus, args.gpus))
        args.gpus = world
    if not torch.cuda.is_available():
        sys.exit("bench.py: no CUDA device; the product has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dist = None
    hdist = None
    nccl_id = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        from heatb200 import dist as hdist
        nccl_id = hdist.shared_nccl_id()     # rank 0: heat_nccl_unique_id(); broadcast over torch.distributed

    dtype = np.float64 if args.dtype == "f64" else np.float32
    esize = np.dtype(dtype).itemsize
    nx = wl["nx"]
    cfg = workload_config(args, name, why, wl, world)
    ny_global = cfg["ny_global"]
    timesteps = wl["timesteps"]
    row0, nrows = heatb200.slab_partition(ny_global, world, rank)

    def make_field(nx_, ny_, r0_, nr_, dt_):
        # synthetic input: the reference's hat IC (fortran/serial/heat.f90:40-48); r from the input.dat derivation
        # for a square grid of side nx (exactly 0.25 for all named sizes)
        delta_, _, r_ = config.InputDeck(nx_, 0.25, 0.05, 2.0, 1).derive()
        if ny_ == nx_:
            f = ic.hat_fortran(nx_, 2.0, delta_, dtype=dt_, rows=(r0_, nr_))
        else:                                 # non-square global grids (weak scaling): hat in x, per-rank in y
            f = ic.mpi_uniform(nx_, ny_, dtype=dt_, rows=(r0_, nr_))
            f[:, nx_ // 4: 3 * nx_ // 4] = 2.0
            f[:, 0] = f[:, -1] = 1.0
            if r0_ == 0:
                f[0] = 1.0
            if r0_ + nr_ == ny_:
                f[-1] = 1.0
        if args.field == "random":
            f = (1.0 + np.random.default_rng(1234 + rank).random(f.shape)).astype(dt_)
            f[:, 0] = f[:, -1] = 1.0
            if r0_ == 0:
                f[0] = 1.0
            if r0_ + nr_ == ny_:
                f[-1] = 1.0
        return f, r_

    def pinned_pair(f):
        tdt = torch.float64 if f.dtype == np.float64 else torch.float32
        a = torch.empty(f.shape, dtype=tdt).pin_memory()
        a.numpy()[...] = f
        b = torch.empty(f.shape, dtype=tdt).pin_memory()
        return a.numpy(), b.numpy(), (a, b)

    field, r = make_field(nx, ny_global, row0, nrows, dtype)
    host_in, host_out, _keep = pinned_pair(field)
    del field

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
    halo_mode = getattr(solver, "halo_mode", "none")
    kernel_name = solver.kernel_name
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
    traffic, ncu = None, None
    if world == 1:
        try:      # measured DRAM bytes per launch of this kernel/config, from the committed ncu capture (if one exists)
            with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
                ncu = json.load(f).get("%s:%s:%s:k%d" % (name, kernel_name, args.dtype, max(args.tblock, 1)))
            traffic = ncu["bytes"] if ncu else None
        except Exception:
            pass
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": traffic, "ncu": ncu,
                "algorithmic_bytes_per_launch": local_lups_per_launch * bytes_per_lup, "peak_source": peak_src, "kernel": kernel_name,
                "algorithmic_bytes_per_lup": bytes_per_lup, "avg_kernel_ms": avg_kernel_ms,
                "glups_roof": peak / bytes_per_lup,
                "effective_gbs_at_single_step_traffic": achieved * max(args.tblock, 1),
                # `peak` is the pool's measured copy bandwidth (torch b.copy_(a)); it is not the ceiling of the HBM
                # interface — this kernel's read and write streams are more compact than a flat copy's and pass it.
                # Against the interface rate ncu reports its DRAM throughput against (8.18 TB/s):
                "frac_of_hbm_interface_8180": achieved / 8180.0,
                "note": "frac = achieved / peak with peak = the pool's measured torch copy bandwidth; a value above 1 means this "
                        "kernel moves its algorithmic bytes faster than that copy does, not that bytes were skipped: see `traffic` "
                        "(ncu DRAM bytes per launch) next to algorithmic_bytes_per_launch and ncu.dram_pct (share of the HBM "
                        "interface rate)"}
    if traffic:
        # what actually crossed the HBM pins per launch (ncu) over the same launch time: below `achieved` when part
        # of a step is served by the 126 MB L2 (4096^2), above it when halo columns / rows are re-read
        roofline["dram_gbs_from_ncu_traffic"] = traffic / (avg_kernel_ms * 1e-3) / 1e9
        roofline["dram_frac_from_ncu_traffic"] = roofline["dram_gbs_from_ncu_traffic"] / peak

    # ---- end to end through the public API with HOST (pinned) buffers: every step uploads the field, marches it and
    # downloads the result.  `value`: jobs queued with HeatSolver.submit (heat_submit) — step i+1's upload and step
    # i-1's download overlap step i's march; `serial`: the blocking set_field / step / gather sequence.
    e2e = None
    if not args.no_e2e:
        slab_bytes = int(nrows * nx * esize)

        def timed_jobs(k):
            barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(k):
                solver.submit(host_in, timesteps, host_out)
            solver.drain()                    # the last download has landed in host memory
            e1.record()
            torch.cuda.synchronize()
            m = e0.elapsed_time(e1)
            if dist is not None:
                t = torch.tensor([m], dtype=torch.float64, device="cuda")
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                m = float(t.item())
            barrier()
            return m
        timed_jobs(1)
        e2e_ms = timed_jobs(args.steps)
        e2e = {"value": lups_per_step * args.steps / (e2e_ms * 1e-3) / 1e9, "unit": UNIT,
               "h2d_bytes_per_step": slab_bytes, "d2h_bytes_per_step": slab_bytes,
               "steps": args.steps, "ms_per_step": e2e_ms / args.steps,
               "mode": "pipelined jobs (heat_submit/heat_drain): upload, march and download of consecutive steps overlap"}

        def e2e_step():
            solver.set_field(host_in)
            solver.step(timesteps)
            solver.gather(out=host_out)
        ks = max(2, min(args.steps, 5))
        ser_ms = timed(e2e_step, ks)
        e2e["serial"] = {"value": lups_per_step * ks / (ser_ms * 1e-3) / 1e9, "steps": ks, "ms_per_step": ser_ms / ks,
                         "mode": "blocking set_field + step + gather per step"}
        # where the serial time goes (separately timed, 2 rounds, max over ranks): upload, march, download
        split = {}
        for nm, fn in (("set_field_ms", lambda: solver.set_field(host_in)), ("step_ms", lambda: solver.step(timesteps)),
                       ("gather_ms", lambda: solver.gather(out=host_out))):
            split[nm] = timed(fn, 2) / 2
        e2e["serial"]["split"] = split

    total_launches = launches
    if dist is not None:
        t = torch.tensor([float(launches)], dtype=torch.float64, device="cuda")
        dist.all_reduce(t)
        total_launches = int(t.item())
    halo_ms = tm.last_halo_ms
    halo_exchanges = int(tm.halo_exchanges)
    solver.close()

    # ---- extras, N = 1 only: short runs of the other kernels / precisions / sizes BASELINE.json names
    extras = None
    if world == 1 and args.tblock == 1 and args.dtype == "f64" and not args.no_extras and name == "cfg3":
        extras = {}

        def short_run(tag, nx_, dt_, k, ts, fmad=False, note=""):
            f, r_ = make_field(nx_, nx_, 0, nx_, dt_)
            s = heatb200.HeatSolver(f, r_, **{**common, "tblock": k, "fmad": fmad})
            for _ in range(3):
                s.step(ts)
            n = 3
            m = timed(lambda: s.step(ts), n)
            kn = s.kernel_name
            s.close()
            g = (nx_ - 2) ** 2 * ts * n / (m * 1e-3) / 1e9
            bpl = 2 * np.dtype(dt_).itemsize / k
            extras[tag] = {"value": g, "unit": UNIT, "kernel": kn, "timesteps_per_step": ts, "steps": n,
                           "algorithmic_bytes_per_lup": bpl, "frac_of_measured_hbm_peak": g * bpl / peak, "note": note}
        short_run("cfg3_f64_k2", nx, np.float64, 2, 200, note="temporal blocking, 2 steps per HBM round trip; same bits as k=1")
        short_run("cfg3_f64_k4", nx, np.float64, 4, 200, note="temporal blocking, 4 steps per HBM round trip; same bits as k=1")
        short_run("cfg3_f64_k4_fmad", nx, np.float64, 4, 200, fmad=True, note="k=4 with HEAT_FLAG_FMAD (5 instead of 7 DP issues per update)")
        short_run("cfg3_f32_k1", nx, np.float32, 1, 200)
        short_run("cfg3_f32_k4", nx, np.float32, 4, 200)
        short_run("cfg2_f64_k1", 4096, np.float64, 1, 1000,
                  note="268 MB working set vs 126 MB L2: serpentine sweeping serves part of every step from L2, so this "
                       "fraction is EFFECTIVE bandwidth, not HBM utilisation (ncu: 0.81 of the algorithmic bytes reach DRAM)")
        short_run("cfg2_f64_k4", 4096, np.float64, 4, 1000)

    parity = None
    if world > 1 and not args.no_parity and not args.skip_exchange:
        parity = multi_gpu_parity(args, hdist, heatb200, dist, torch, rank, world, local, stream)

    rc = 0
    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": wl["scaling"], "vs_baseline": None,
            "dtype": args.dtype, "data": "synthetic",
            "config": cfg,
            "run": {"kernel": kernel_name, "halo": halo_mode},
            "roofline": roofline, "gpu_launches": int(total_launches),
            "clocks": clk.summary(),
        }
        if e2e:
            line["e2e"] = e2e
        if extras:
            line["extras"] = extras
        if world > 1:
            line["halo"] = {"last_exchange_ms": halo_ms, "exchanges": halo_exchanges, "mode": halo_mode}
        if parity is not None:
            line["parity"] = parity
        if world == 1 and not args.no_cpu_baseline:
            nthreads = host_threads()
            cnx = nx
            cny = max(1024, 4096 * 4096 // nx)                    # a band of the workload's width, 16 Mi cells
            g, cores, st, dt = cpu_march_glups(cnx, cny, budget_s=12.0, threads=nthreads)
            gn, stn = numpy_reference_glups(4096, 4096, budget_s=5.0)
            g1, _, st1, dt1 = cpu_march_glups(cnx, cny, budget_s=4.0, threads=1)
            line["cpu_baseline"] = {
                "value": g, "unit": UNIT, "cores": cores, "kind": "port",
                "sample": "%dx%d fp64 band of the workload, %d timesteps in %.1f s: C port of the fortran/serial loop, OpenMP on %d threads" % (cnx, cny, st, dt, cores),
                "numpy_1thread": {"value": gn, "sample": "python/serial expression (heat.py:49-58) in NumPy, 4096x4096, %d steps" % stn},
                "c_port_1thread": {"value": g1, "sample": "same C port, 1 thread (the reference's CPU programs are serial), %dx%d, %d steps in %.1f s" % (cnx, cny, st1, dt1)}}
        sys.stdout.flush()
        os.write(json_fd, (json.dumps(line) + "\n").encode())
    if parity is not None and not parity["ok"]:
        rc = 3
        if rank == 0:
            sys.stderr.write("bench.py: MULTI-GPU PARITY FAILED: %s\n" % json.dumps(parity))
    if dist is not None:
        dist.destroy_process_group()
    return rc


if __name__ == "__main__":
    sys.exit(main())
