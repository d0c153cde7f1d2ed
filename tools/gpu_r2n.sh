#!/bin/bash
# ncu --set full at HEAD: cone kernel (rocket) and the float TMEM kernel
mkdir -p gpurun_out
cap() { name=$1; kern=$2; shift; shift
  timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu "$@" > gpurun_out/plain_$name.log 2>&1 && \
  timeout 1500 ncu --set full --clock-control none --import-source on -k regex:$kern -s 4 -c 1 -o gpurun_out/prof_$name -f python bench.py --steps 2 --warmup 3 --no-cpu "$@" > gpurun_out/ncu_$name.log 2>&1
  echo "ncu $name rc=$?"; }
cap rocket admm_cone --workload rocket
cap quadrotor_easy_f32 admm_tmem --precision 32
