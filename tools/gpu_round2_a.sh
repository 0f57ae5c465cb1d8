#!/bin/bash
# round-2 first GPU pass: tests, default bench (cfg3), reference arm, launch list
set -x
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/smi.txt
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
python bench.py --steps 10 --warmup 3 > gpurun_out/bench_cfg3.json 2> gpurun_out/bench_cfg3.err; echo "bench rc=$?"
cat gpurun_out/bench_cfg3.json
python bench.py --impl reference --steps 5 --warmup 3 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
cat gpurun_out/bench_ref.json
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_cfg3.csv python bench.py --steps 2 --warmup 3 --timesteps 20 --no-extras --no-cpu-baseline --no-e2e > gpurun_out/ncu_launch.log 2>&1
tail -3 gpurun_out/launches_cfg3.csv
