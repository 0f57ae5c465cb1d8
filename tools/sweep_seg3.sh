#!/bin/bash
mkdir -p gpurun_out
run() {
  python bench.py "$@" --steps 5 --warmup 3 --no-extras --no-cpu-baseline --no-e2e 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(round(d['value'],1), round(d['roofline']['frac'],4))"
}
{
for seg in 30 46 62 78 94; do echo -n "cfg3 f64 k1 seg=$seg "; HEAT_TMA_SEG=$seg run --workload cfg3 --timesteps 100; done
for seg in 46 62; do echo -n "cfg4 f64 k1 seg=$seg "; HEAT_TMA_SEG=$seg run --workload cfg4 --timesteps 40; done
for seg in 46 62 78; do echo -n "cfg2 f64 k1 seg=$seg "; HEAT_TMA_SEG=$seg run --workload cfg2 --timesteps 1000; done
for seg in 46 62 78 94; do echo -n "cfg3 f32 k1 seg=$seg "; HEAT_TMA_SEG=$seg run --workload cfg3 --dtype f32 --timesteps 100; done
for seg in 14 20 26 32 44; do echo -n "cfg3 f64 k2 tseg=$seg "; HEAT_TEMPORAL_SEG=$seg run --workload cfg3 --tblock 2 --timesteps 100; done
for seg in 88 112 160 184 208; do echo -n "cfg3 f64 k4 tseg=$seg "; HEAT_TEMPORAL_SEG=$seg run --workload cfg3 --tblock 4 --timesteps 100; done
for seg in 40 52 64 88; do echo -n "cfg3 f32 k4 tseg=$seg "; HEAT_TEMPORAL_SEG=$seg run --workload cfg3 --dtype f32 --tblock 4 --timesteps 100; done
for seg in 20 32 64; do echo -n "cfg3 f32 k2 tseg=$seg "; HEAT_TEMPORAL_SEG=$seg run --workload cfg3 --dtype f32 --tblock 2 --timesteps 100; done
for seg in 0 16 32 64; do echo -n "cfg2 f64 k2 tseg=$seg "; HEAT_TEMPORAL_SEG=$seg run --workload cfg2 --tblock 2 --timesteps 1000; done
echo -n "cfg3 e2e pipeline seg=46: "; HEAT_TMA_SEG=46 python bench.py --workload cfg3 --steps 5 --warmup 3 --no-extras --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(round(d['value'],1), json.dumps(d['e2e']))"
} | tee gpurun_out/sweep_seg3.txt
python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "pipeline or legacy or contracted" 2>&1 | tail -3
