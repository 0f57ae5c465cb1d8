#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
python tools/bench_small.py 2000 2>&1 | tee gpurun_out/bench_small.txt
run() {
  python bench.py "$@" --steps 5 --warmup 3 --no-extras --no-cpu-baseline --no-e2e 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(round(d['value'],1), round(d['roofline']['frac'],4))"
}
{
for seg in 0 5 8 11; do echo -n "cfg3 f64 k2 tseg=$seg "; HEAT_TEMPORAL_SEG=$seg run --workload cfg3 --tblock 2 --timesteps 100; done
for seg in 8 14; do echo -n "cfg3 f32 k2 tseg=$seg "; HEAT_TEMPORAL_SEG=$seg run --workload cfg3 --dtype f32 --tblock 2 --timesteps 100; done
for seg in 0 14 30 46; do echo -n "cfg3 f64 wtma k1 tseg=$seg "; HEAT_TEMPORAL_SEG=$seg run --workload cfg3 --kernel wtma --timesteps 100; done
for seg in 0 30 46; do echo -n "cfg3 f32 wtma k1 tseg=$seg "; HEAT_TEMPORAL_SEG=$seg run --workload cfg3 --kernel wtma --dtype f32 --timesteps 100; done
echo -n "cfg3 f64 k4 "; run --workload cfg3 --tblock 4 --timesteps 100
echo -n "cfg3 f64 k4 fmad "; run --workload cfg3 --tblock 4 --timesteps 100 --fmad
echo -n "cfg3 f32 k4 "; run --workload cfg3 --tblock 4 --dtype f32 --timesteps 100
echo -n "cfg3 f32 k4 fmad "; run --workload cfg3 --tblock 4 --dtype f32 --timesteps 100 --fmad
echo -n "cfg3 f64 k1 random field "; run --workload cfg3 --timesteps 100 --field random
} | tee gpurun_out/sweep4.txt
python bench.py --steps 10 --warmup 3 > gpurun_out/bench_cfg3_b.json 2> gpurun_out/bench_cfg3_b.err; cat gpurun_out/bench_cfg3_b.json
# ncu captures (after the plain runs above)
cap() { # name skip args...
  name=$1; skip=$2; shift 2
  ncu --set full --clock-control none --import-source on -k regex:heat_ --launch-skip $skip -c 2 -f -o gpurun_out/$name python bench.py "$@" --steps 1 --warmup 3 --no-extras --no-cpu-baseline --no-e2e > gpurun_out/$name.log 2>&1
  ncu -i gpurun_out/$name.ncu-rep --page raw --csv > gpurun_out/${name}_raw.csv 2>/dev/null
}
cap r02_prof_k1_16384 30 --workload cfg3 --timesteps 10
cap r02_prof_k2_16384 12 --workload cfg3 --timesteps 8 --tblock 2
cap r02_prof_k4_16384 6 --workload cfg3 --timesteps 8 --tblock 4
cap r02_prof_f32_k1_16384 30 --workload cfg3 --timesteps 10 --dtype f32
cap r02_prof_f32_k4_16384 6 --workload cfg3 --timesteps 8 --tblock 4 --dtype f32
cap r02_prof_k4_4096 6 --workload cfg2 --timesteps 8 --tblock 4
cap r02_prof_k1_4096 30 --workload cfg2 --timesteps 10
ls -la gpurun_out/*.ncu-rep
