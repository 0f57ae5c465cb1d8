#!/bin/bash
# N-GPU pass: parity worker, then bench (auto -> cfg4 strong) at N and at N=1 on the same box
N=${1:-2}
set -x
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 tests/_mgpu_worker.py > gpurun_out/parity_worker_n$N.log 2>&1; echo "worker rc=$?" >> gpurun_out/parity_worker_n$N.log
grep -v "^W\|^\[" gpurun_out/parity_worker_n$N.log | tail -25
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_n${N}_cfg4.json 2> gpurun_out/bench_n${N}_cfg4.err; echo "bench rc=$?"
cat gpurun_out/bench_n${N}_cfg4.json; tail -5 gpurun_out/bench_n${N}_cfg4.err
python bench.py --gpus 1 --steps 10 --warmup 3 > gpurun_out/bench_n1_cfg4.json 2> gpurun_out/bench_n1_cfg4.err; echo "bench1 rc=$?"
cat gpurun_out/bench_n1_cfg4.json; tail -5 gpurun_out/bench_n1_cfg4.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --impl reference --gpus $N --steps 3 --warmup 3 > gpurun_out/bench_ref_n${N}.json 2> gpurun_out/bench_ref_n${N}.err
cat gpurun_out/bench_ref_n${N}.json
