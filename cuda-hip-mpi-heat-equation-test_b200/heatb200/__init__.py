"""heatb200 — B200-native FTCS 5-point heat-equation march behind the C ABI of libheat.so.

Drop-in for the two Python entry points of cssrikanth/CUDA-HIP-MPI-Heat-equation-test:
  heatb200.serial  <->  python/serial/heat.py      heatb200.cuda  <->  python/cuda/cuda.py
"""
from ._lib import HeatError, lib  # noqa: F401
from .config import InputDeck, derive_python, parse_input_dat  # noqa: F401
from .solver import HeatSolver, march, nccl_unique_id, slab_partition  # noqa: F401
