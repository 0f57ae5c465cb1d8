#!/bin/bash
# ncu --set full captures of the dominant kernels (each only after the same command ran clean without ncu).
# Only the CSV pages come back (the .ncu-rep files would exceed gpurun's 64 MiB return limit).
mkdir -p gpurun_out
rm -f gpurun_out/*.ncu-rep
cap() { # name skip args...
  name=$1; skip=$2; shift 2
  python bench.py "$@" --steps 1 --warmup 3 --no-extras --no-cpu-baseline --no-e2e > /dev/null 2>&1 || { echo "$name: plain run failed"; return; }
  ncu --set full --clock-control none --import-source on -k regex:heat_ --launch-skip $skip -c 2 -f -o /tmp/$name python bench.py "$@" --steps 1 --warmup 3 --no-extras --no-cpu-baseline --no-e2e > gpurun_out/$name.log 2>&1
  ncu -i /tmp/$name.ncu-rep --page raw --csv > gpurun_out/${name}_raw.csv 2>/dev/null
  ncu -i /tmp/$name.ncu-rep --page details --csv > gpurun_out/${name}_details.csv 2>/dev/null
  if [ -n "$KEEP_SOURCE" ]; then ncu -i /tmp/$name.ncu-rep --page source --csv > gpurun_out/${name}_source.csv 2>/dev/null; fi
  rm -f /tmp/$name.ncu-rep
  echo "$name done"
}
cap r02_prof_k1_16384 30 --workload cfg3 --timesteps 10
cap r02_prof_k2_16384 12 --workload cfg3 --timesteps 8 --tblock 2
KEEP_SOURCE=1 cap r02_prof_k4_16384 6 --workload cfg3 --timesteps 8 --tblock 4
cap r02_prof_f32_k1_16384 30 --workload cfg3 --timesteps 10 --dtype f32
KEEP_SOURCE=1 cap r02_prof_f32_k4_16384 6 --workload cfg3 --timesteps 8 --tblock 4 --dtype f32
cap r02_prof_k4_4096 6 --workload cfg2 --timesteps 8 --tblock 4
cap r02_prof_k1_4096 30 --workload cfg2 --timesteps 10
cap r02_prof_k1_32768 9 --workload cfg4 --timesteps 3
ls -la gpurun_out/ | head -40
