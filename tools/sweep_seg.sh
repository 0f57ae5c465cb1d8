#!/bin/bash
# TMA-kernel segment length sweep at cfg3 and cfg4 (N=1)
mkdir -p gpurun_out
for wl in cfg3 cfg4; do
for seg in 0 64 128 256 512 1024 2048; do
  ts=100; [ $wl = cfg4 ] && ts=40
  HEAT_TMA_SEG=$seg python bench.py --workload $wl --steps 5 --warmup 3 --timesteps $ts --no-extras --no-cpu-baseline --no-e2e 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$wl seg=$seg', round(d['value'],1), round(d['roofline']['frac'],4))"
done; done | tee gpurun_out/sweep_seg.txt
