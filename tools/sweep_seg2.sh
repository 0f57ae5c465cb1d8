#!/bin/bash
mkdir -p gpurun_out
run() { # env... -- args
  python bench.py "$@" --steps 5 --warmup 3 --no-extras --no-cpu-baseline --no-e2e 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(round(d['value'],1), round(d['roofline']['frac'],4))"
}
{
for seg in 16 32 48 64 80 96; do echo -n "cfg3 f64 k1 seg=$seg "; HEAT_TMA_SEG=$seg run --workload cfg3 --timesteps 100; done
for seg in 32 48 64 96; do echo -n "cfg4 f64 k1 seg=$seg "; HEAT_TMA_SEG=$seg run --workload cfg4 --timesteps 40; done
for seg in 0 32 48 64 96 128; do echo -n "cfg2 f64 k1 seg=$seg "; HEAT_TMA_SEG=$seg run --workload cfg2 --timesteps 1000; done
for seg in 0 32 64 128 256; do echo -n "cfg3 f32 k1 seg=$seg "; HEAT_TMA_SEG=$seg run --workload cfg3 --dtype f32 --timesteps 100; done
for k in 2 4; do for seg in 0 32 64 128 256 512 1024; do echo -n "cfg3 f64 k$k tseg=$seg "; HEAT_TEMPORAL_SEG=$seg run --workload cfg3 --tblock $k --timesteps 100; done; done
for seg in 0 64 128 256 512; do echo -n "cfg3 f32 k4 tseg=$seg "; HEAT_TEMPORAL_SEG=$seg run --workload cfg3 --dtype f32 --tblock 4 --timesteps 100; done
for seg in 0 32 64 128 256; do echo -n "cfg2 f64 k4 tseg=$seg "; HEAT_TEMPORAL_SEG=$seg run --workload cfg2 --tblock 4 --timesteps 1000; done
} | tee gpurun_out/sweep_seg2.txt
