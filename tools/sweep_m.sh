#!/bin/bash
mkdir -p gpurun_out
run() {
  python bench.py "$@" --steps 5 --warmup 3 --no-extras --no-cpu-baseline --no-e2e 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(round(d['value'],1), round(d['roofline']['frac'],4))"
}
{
for m in 2 1; do
for seg in 0 88 160 256; do echo -n "cfg3 f64 k4 M=$m tseg=$seg "; HEAT_TEMPORAL_M=$m HEAT_TEMPORAL_SEG=$seg run --workload cfg3 --tblock 4 --timesteps 100; done
for seg in 0 64 88 160; do echo -n "cfg3 f32 k4 M=$m tseg=$seg "; HEAT_TEMPORAL_M=$m HEAT_TEMPORAL_SEG=$seg run --workload cfg3 --dtype f32 --tblock 4 --timesteps 100; done
for seg in 0 32 48 82 128; do echo -n "cfg2 f64 k4 M=$m tseg=$seg "; HEAT_TEMPORAL_M=$m HEAT_TEMPORAL_SEG=$seg run --workload cfg2 --tblock 4 --timesteps 1000; done
echo -n "cfg3 f64 k2 M=$m "; HEAT_TEMPORAL_M=$m run --workload cfg3 --tblock 2 --timesteps 100
echo -n "cfg3 f32 k2 M=$m "; HEAT_TEMPORAL_M=$m run --workload cfg3 --dtype f32 --tblock 2 --timesteps 100
echo -n "cfg3 f64 k4 fmad M=$m "; HEAT_TEMPORAL_M=$m run --workload cfg3 --tblock 4 --timesteps 100 --fmad
echo -n "cfg3 f32 k4 fmad M=$m "; HEAT_TEMPORAL_M=$m run --workload cfg3 --dtype f32 --tblock 4 --timesteps 100 --fmad
done
} | tee gpurun_out/sweep_m.txt
HEAT_TEMPORAL_M=1 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "temporal" 2>&1 | tail -3
