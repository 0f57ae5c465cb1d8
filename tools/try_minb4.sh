#!/bin/bash
L=cuda-hip-mpi-heat-equation-test_b200/heatb200/libheat.so
run() {
  python bench.py "$@" --steps 5 --warmup 3 --no-extras --no-cpu-baseline --no-e2e 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(round(d['value'],1), round(d['roofline']['frac'],4))"
}
for variant in base minb4; do
  [ $variant = minb4 ] && cp scratch/libheat_minb4.so $L
  echo -n "$variant cfg3 f64 k4 "; run --workload cfg3 --tblock 4 --timesteps 100
  echo -n "$variant cfg3 f32 k4 "; run --workload cfg3 --dtype f32 --tblock 4 --timesteps 100
  echo -n "$variant cfg2 f64 k4 "; run --workload cfg2 --tblock 4 --timesteps 1000
  echo -n "$variant cfg2 f64 k4 fmad "; run --workload cfg2 --tblock 4 --timesteps 1000 --fmad
  echo -n "$variant cfg3 f64 k2 "; run --workload cfg3 --tblock 2 --timesteps 100
done
