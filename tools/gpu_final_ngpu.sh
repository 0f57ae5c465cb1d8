#!/bin/bash
# final multi-GPU evidence on one box: parity worker, the driver's bench command at N and at 1, host<->device ceiling
N=${1:-8}
mkdir -p gpurun_out
trun() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400)) "$@"; }
if [ -z "$SKIP_WORKER" ]; then
trun tests/_mgpu_worker.py > gpurun_out/r02_parity_worker_${N}gpu.log 2>&1; echo "worker rc=$?" >> gpurun_out/r02_parity_worker_${N}gpu.log
grep -c " ok " gpurun_out/r02_parity_worker_${N}gpu.log; grep "FAIL\|rc=" gpurun_out/r02_parity_worker_${N}gpu.log
fi
trun bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/r02_bench_n${N}_cfg4.json 2> gpurun_out/r02_bench_n${N}_cfg4.err; echo "bench N=$N rc=$?"
cut -c1-400 gpurun_out/r02_bench_n${N}_cfg4.json
if [ -z "$SKIP_N1" ]; then
python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r02_bench_n1_cfg4.json 2> gpurun_out/r02_bench_n1_cfg4.err; echo "bench N=1 rc=$?"
cut -c1-400 gpurun_out/r02_bench_n1_cfg4.json
trun tools/pcie_ceiling.py > gpurun_out/r02_pcie_ceiling_n${N}.json 2>/dev/null
cat gpurun_out/r02_pcie_ceiling_n${N}.json | head -c 1500; echo
fi
