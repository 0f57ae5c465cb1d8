#!/bin/bash
N=${1:-8}
mkdir -p gpurun_out
trun() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400)) "$@"; }
trun bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_n${N}_cfg4.json 2> gpurun_out/bench_n${N}_cfg4.err; echo "cfg4 rc=$?"
cat gpurun_out/bench_n${N}_cfg4.json
trun bench.py --gpus $N --steps 10 --warmup 3 --workload cfg2 --no-parity > gpurun_out/bench_n${N}_cfg2.json 2> gpurun_out/bench_n${N}_cfg2.err; echo "cfg2 rc=$?"
cat gpurun_out/bench_n${N}_cfg2.json
trun bench.py --gpus $N --steps 5 --warmup 3 --workload cfg2 --no-parity --no-e2e --ghost 16 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('cfg2 ghost16', d['value'], d['halo'])"
trun bench.py --gpus $N --steps 5 --warmup 3 --workload cfg2 --no-parity --no-e2e --ghost 8 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('cfg2 ghost8', d['value'], d['halo'])"
trun bench.py --gpus $N --steps 5 --warmup 3 --workload cfg2 --no-parity --no-e2e --skip-exchange 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('cfg2 skip-exchange', d['value'])"
python bench.py --gpus 1 --workload cfg2 --steps 5 --warmup 3 --no-e2e --no-extras --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('cfg2 n1', d['value'])"
