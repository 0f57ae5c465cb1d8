"""Small, fast march through every kernel — the program to put under `compute-sanitizer --tool memcheck`
(one tool per gpurun call; run it plainly first).  Exits non-zero on any mismatch with the oracle."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "cuda-hip-mpi-heat-equation-test_b200")):
    sys.path.insert(0, p)
import heatb200  # noqa: E402
from oracle import heat_oracle as ho  # noqa: E402

bad = 0
for shape in [(3, 3), (31, 31), (70, 130), (130, 257), (40, 1000)]:
    for dtype in (np.float64, np.float32):
        u0 = (1.0 + np.random.default_rng(shape[1]).random(shape)).astype(dtype)
        want = ho.march(u0, 6, ho.ORDER_FORTRAN, dtype(0.2437))
        for kernel, k in [("simple", 1), ("stream", 1), ("tma", 1), ("wtma", 1), ("tma", 2), ("tma", 4), ("auto", 1), ("persist", 1)]:
            if kernel == "persist" and shape[0] < 8:
                continue
            got = heatb200.march(u0, 6 if kernel != "persist" else 9, float(dtype(0.2437)), kernel=kernel, tblock=k)
            if kernel == "persist":
                got, want9 = got, ho.march(u0, 9, ho.ORDER_FORTRAN, dtype(0.2437))
                if not np.array_equal(got.view(np.uint8), want9.view(np.uint8)):
                    bad += 1
                    print("MISMATCH", shape, dtype, kernel, k)
                continue
            if not np.array_equal(got.view(np.uint8), want.view(np.uint8)):
                bad += 1
                print("MISMATCH", shape, dtype, kernel, k)
        with heatb200.HeatSolver(u0, 0.2437) as s:
            s.step(2)
            s.checksum()
print("sanitize_small: %d mismatches" % bad)
sys.exit(1 if bad else 0)
