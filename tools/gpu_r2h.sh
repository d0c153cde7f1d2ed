#!/bin/bash
# ncu --set full captures (one launch each) of the solve kernel of every bench workload, after a plain run of the same command
mkdir -p gpurun_out
cap() { name=$1; kern=$2; shift; shift
  timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu "$@" > gpurun_out/plain_$name.log 2>&1 && \
  timeout 1500 ncu --set full --clock-control none --import-source on -k regex:$kern -s 4 -c 1 -o gpurun_out/prof_$name -f python bench.py --steps 2 --warmup 3 --no-cpu "$@" > gpurun_out/ncu_$name.log 2>&1
  echo "ncu $name rc=$?"; }
cap quadrotor_easy admm_tmem
cap quadrotor_hard admm_tmem --workload quadrotor_hard
cap cartpole admm_tmem --workload cartpole
cap rocket admm_cone --workload rocket
cap rollout admm_tmem --workload rollout --rollout-steps 50
timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/ncu_launch.log 2>&1
echo "ncu launches rc=$?"
ls -la gpurun_out/*.ncu-rep
