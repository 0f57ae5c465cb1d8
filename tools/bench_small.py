#!/usr/bin/env python
"""Mid-size grids (the reference's shipped 1024^2 deck and neighbours): GLUPS of one heat_step(nt) call per kernel.
    python tools/bench_small.py [nt]
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "cuda-hip-mpi-heat-equation-test_b200"))
import numpy as np
import heatb200

nt = int(sys.argv[1]) if len(sys.argv) > 1 else 2000
for n, dtype in ((256, np.float64), (512, np.float64), (768, np.float64), (1024, np.float64), (1100, np.float64), (1024, np.float32), (2048, np.float32)):
    u0 = (1.0 + np.random.default_rng(0).random((n, n))).astype(dtype)
    for kernel, k in (("persist", 1), ("tma", 1), ("tma", 4), ("stream", 1)):
        try:
            s = heatb200.HeatSolver(u0, 0.25, kernel=kernel, tblock=k)
        except heatb200.HeatError as e:
            print("%4d %s %-8s k=%d  unavailable (%s)" % (n, np.dtype(dtype).name, kernel, k, str(e)[:60]))
            continue
        s.step(nt); s.sync()
        best = 1e9
        for _ in range(3):
            s.step(nt); s.sync()
            best = min(best, s.timing().last_step_ms)
        print("%4d %s %-8s k=%d  %8.3f us/step  %8.1f GLUPS" % (n, np.dtype(dtype).name, kernel, k, best * 1e3 / nt, (n - 2) ** 2 * nt / best / 1e6), flush=True)
        s.close()
