"""Throughput of the kernels that accept caller-owned, odd-pitch buffers (torch tensors of width nx = 8193)."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "cuda-hip-mpi-heat-equation-test_b200")):
    sys.path.insert(0, p)
import torch
import heatb200
import ctypes as C

nx, ny, steps = 8193, 8192, 200
b0 = torch.ones((ny, nx), dtype=torch.float64, device="cuda")
b0[2000:6000, 2000:6000] = 2.0
b1 = b0.clone()
for kern in ("simple", "stream"):
    s = heatb200.HeatSolver(None, 0.25, nx=nx, ny=ny, dtype=np.float64, kernel=kern, buffers=(b0, b1))
    s.step(20); s.sync()
    s.step(steps)
    ms = s.timing().last_step_ms
    print("%-7s odd pitch %d: %.1f GLUPS (%.0f GB/s algorithmic)" % (kern, nx, (nx - 2) * (ny - 2) * steps / ms / 1e6, (nx - 2) * (ny - 2) * steps * 16 / ms / 1e6))
    s.close()
# legacy launcher on a ghosted (n+2)^2 array, as fortran/hip/heat.F90 would call it
n = 8192
old = torch.ones((n + 2, n + 2), dtype=torch.float64, device="cuda"); new = old.clone()
L = heatb200.lib()
for _ in range(5):
    L.launch_heat_eqn(C.c_void_p(new.data_ptr()), C.c_void_p(old.data_ptr()), 0.25, n, n, 1)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(100):
    L.launch_heat_eqn(C.c_void_p(new.data_ptr()), C.c_void_p(old.data_ptr()), 0.25, n, n, 1)
    old, new = new, old
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1)
print("legacy launch_heat_eqn %d^2 ghosted: %.1f GLUPS" % (n, n * n * 100 / ms / 1e6))
