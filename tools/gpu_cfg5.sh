#!/bin/bash
N=8
tr_() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400)) "$@"; }
mkdir -p gpurun_out
tr_ bench.py --gpus $N --steps 5 --warmup 3 --workload cfg5 --no-e2e --no-parity > gpurun_out/r02_bench_n8_cfg5.json 2>/dev/null; cut -c1-200 gpurun_out/r02_bench_n8_cfg5.json
tr_ bench.py --gpus $N --steps 5 --warmup 3 --workload cfg5 --no-e2e --no-parity --no-overlap > gpurun_out/r02_bench_n8_cfg5_noov.json 2>/dev/null; cut -c1-200 gpurun_out/r02_bench_n8_cfg5_noov.json
tr_ bench.py --gpus $N --steps 5 --warmup 3 --workload cfg5 --no-e2e --no-parity --tblock 2 > gpurun_out/r02_bench_n8_cfg5_k2.json 2>/dev/null; cut -c1-200 gpurun_out/r02_bench_n8_cfg5_k2.json
