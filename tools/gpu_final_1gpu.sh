#!/bin/bash
# final single-GPU evidence: full GPU test log, the driver's bench command, the CPU arm, launch list, ncu captures
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/r02_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02_pytest_gpu.log
tail -4 gpurun_out/r02_pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r02_bench_default_cfg3.json 2> gpurun_out/r02_bench_default_cfg3.err; echo "bench rc=$?"
cat gpurun_out/r02_bench_default_cfg3.json
python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > gpurun_out/r02_bench_reference_cfg3.json 2>/dev/null
cat gpurun_out/r02_bench_reference_cfg3.json
python bench.py --field random --steps 5 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/r02_bench_cfg3_random.json 2>/dev/null
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches_cfg3.csv python bench.py --steps 2 --warmup 3 --timesteps 20 --no-extras --no-cpu-baseline --no-e2e > gpurun_out/ncu_launch.log 2>&1
tail -2 gpurun_out/r02_launches_cfg3.csv
cap() { # name skip args...
  name=$1; skip=$2; shift 2
  python bench.py "$@" --steps 1 --warmup 3 --no-extras --no-cpu-baseline --no-e2e > /dev/null 2>&1 || { echo "$name: plain run failed"; return; }
  ncu --set full --clock-control none --import-source on -k regex:heat_ --launch-skip $skip -c 2 -f -o /tmp/$name python bench.py "$@" --steps 1 --warmup 3 --no-extras --no-cpu-baseline --no-e2e > gpurun_out/$name.log 2>&1
  ncu -i /tmp/$name.ncu-rep --page raw --csv > gpurun_out/${name}_raw.csv 2>/dev/null
  ncu -i /tmp/$name.ncu-rep --page details --csv > gpurun_out/${name}_details.csv 2>/dev/null
  rm -f /tmp/$name.ncu-rep
  echo "$name done"
}
cap r02_prof_k2_16384 12 --workload cfg3 --timesteps 8 --tblock 2
cap r02_prof_k4_16384 6 --workload cfg3 --timesteps 8 --tblock 4
cap r02_prof_f32_k4_16384 6 --workload cfg3 --timesteps 8 --tblock 4 --dtype f32
cap r02_prof_k4_4096 6 --workload cfg2 --timesteps 8 --tblock 4
