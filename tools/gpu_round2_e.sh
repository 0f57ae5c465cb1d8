#!/bin/bash
N=${1:-4}
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 tests/_mgpu_worker.py > gpurun_out/parity_worker_n$N.log 2>&1; echo "worker rc=$?" >> gpurun_out/parity_worker_n$N.log
grep -v "^W\|^\[\|^\*\|^Setting\|^$" gpurun_out/parity_worker_n$N.log | tail -50
python tools/bench_small.py 2000 2>&1 | grep "persist\|tma      k=1" | tee gpurun_out/bench_small_b.txt
python -m pytest tests -m gpu -x -q -k "persist or multi" 2>&1 | tail -3
