"""Regenerates tests/golden/*.npz from the REFERENCE ITSELF (run in the build container, where
/root/reference exists; the GPU box only reads the committed .npz files).

Sources of truth, none of them ours:
  pyserial_31.npz       python/serial/heat.py executed unmodified (oracle/run_reference.py)
  hip_slabs_n100.npz    fortran/hip/heat_kernel.cpp (launch_heat_eqn/send/recv1/recv2) compiled
                        for the host (oracle/_ref/libhipref.so), driven in the order of
                        fortran/hip/heat.F90:234-252,198-228 with fortran/mpi+cuda/input.dat
                        (100 0.25 0.05 2.0 10 1) for P = 1 and P = 2 ranks; plus 5 launches of
                        launch_heat_eqn on a seeded random ghosted 56x40 field
  pycuda_kernel.npz     the kernel string of python/cuda/cuda.py:60-83 compiled for the host
                        (oracle/_ref/libpycudaref.so), 5 launches on a seeded 64x64 field, for
                        r = 0.25 and for r = 0.24999999999999997 (n = 100)

  pyserial_31_diffuse60.npz   the same unmodified script, then ITS OWN diffuse(60) called from the
                        script's globals (heat.py:45-58: hat re-applied to the diffused field,
                        61 more steps) — far outside the exact-arithmetic window, so the bits
                        depend on operation order P
  *_fma arrays          the same reference kernel sources compiled with gcc -mfma
                        -ffp-contract=fast (oracle/_ref/lib*ref_fma.so): the contracted reading
                        nvcc (-fmad=true) and hipcc (-ffp-contract=fast) give that text by default

Run:  python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import c_oracle, heat_oracle as ho, run_reference  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


def golden_pyserial():
    g = run_reference.run_python_serial()
    np.savez_compressed(os.path.join(OUT, "pyserial_31.npz"), u=g["u"], nx=g["nx"], ny=g["ny"],
                        nt_called=10, dx=g["dx"], dy=g["dy"], dt=g["dt"], nu=g["nu"], sigma=g["sigma"])
    print("pyserial_31: sum", repr(float(g["u"].sum())))


def golden_pyserial_diffuse60():
    g = run_reference.run_python_serial()       # module body + diffuse(10), as shipped
    u10 = g["u"].copy()
    g["diffuse"](60)                            # the reference's own function, heat.py:45-58
    np.savez_compressed(os.path.join(OUT, "pyserial_31_diffuse60.npz"), u10=u10, u=g["u"], calls=np.array([10, 60]))
    print("pyserial_31_diffuse60: sum", repr(float(g["u"].sum())))


def hip_program(n, nproc, ntime, dom_len=2.0, sigma=0.25, nu=0.05):
    """The time loop of fortran/hip/heat.F90:234-252 + swap() :198-228 with the reference kernels."""
    _, _, r = ho.derive_fortran(n, sigma, nu, dom_len)
    nxl, ny, ng = n // nproc, n, 1
    slabs = [ho.ic_mpi_ghosted(nxl, ny, p == 0, p == nproc - 1) for p in range(nproc)]
    for _ in range(ntime):
        slabs = [c_oracle.hipref_step(s, r, nxl, ny, ng) for s in slabs]
        sends = [c_oracle.hipref_swap(s, nxl, ny, ng) for s in slabs]
        for p in range(nproc):
            if p > 0:                      # ileftx /= MPI_PROC_NULL: recv1 <- left's Tds2 (tag 2)
                slabs[p] = c_oracle.hipref_recv(slabs[p], sends[p - 1][1], 1, nxl, ny, ng)
            if p + 1 < nproc:              # irightx: recv2 <- right's Tds1 (tag 1)
                slabs[p] = c_oracle.hipref_recv(slabs[p], sends[p + 1][0], 2, nxl, ny, ng)
    return r, slabs


def golden_hip():
    out = {}
    for P in (1, 2):
        r, slabs = hip_program(100, P, 10)
        out["r"] = r
        for p, s in enumerate(slabs):
            out["P%d_rank%d" % (P, p)] = s
    # a seeded random ghosted field, one rank, 5 steps: every cell exercises the rounding of order F
    rng = np.random.default_rng(7)
    g0 = 1.0 + rng.random((40 + 2, 56 + 2))
    g = g0
    for _ in range(5):
        g = c_oracle.hipref_step(g, out["r"], 56, 40, 1)
    out["rand_u0"], out["rand_u5"] = g0, g
    g = g0
    for _ in range(5):
        g = c_oracle.hipref_step(g, out["r"], 56, 40, 1, contract=True)
    out["rand_u5_fma"] = g
    np.savez_compressed(os.path.join(OUT, "hip_slabs_n100.npz"), **out)
    print("hip_slabs_n100: r", repr(out["r"]), "sum P1", repr(float(out["P1_rank0"].sum())))


def golden_pycuda():
    rng = np.random.default_rng(20261106)
    u0 = 1.0 + rng.random((64, 64))   # square, as the reference always is: its bounds test swaps
    # POINT_SIZE_X/Y (cuda.py:77,85), which only matters for non-square fields
    out = {"u0": u0}
    for tag, r in (("r025", 0.25), ("rn100", ho.derive_fortran(100, 0.25, 0.05, 2.0)[2])):
        u = u0
        for _ in range(5):
            u = c_oracle.pycudaref_step(u, r)
        out["r_" + tag] = r
        out["u5_" + tag] = u
        u = u0
        for _ in range(5):
            u = c_oracle.pycudaref_step(u, r, contract=True)
        out["u5_" + tag + "_fma"] = u
    np.savez_compressed(os.path.join(OUT, "pycuda_kernel.npz"), **out)
    print("pycuda_kernel: sum", repr(float(out["u5_rn100"].sum())))


if __name__ == "__main__":
    if not run_reference.available():
        sys.exit("reference tree not found; fixtures can only be regenerated in the build container")
    c_oracle.build(ref=True)
    golden_pyserial()
    golden_pyserial_diffuse60()
    golden_hip()
    golden_pycuda()
