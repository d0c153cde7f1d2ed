#!/bin/bash
# round 2: the new bench.py on every workload (N = 1), queue order on / off, ncu launch list of the default command
mkdir -p gpurun_out
run() { name=$1; shift; timeout 900 python bench.py "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err; echo "$name rc=$?"; head -c 1500 gpurun_out/$name.json; echo; tail -2 gpurun_out/$name.err; }
run bench_easy_o1 --no-cpu --steps 48
run bench_easy_o0 --no-cpu --steps 48 --order 0
run bench_hard_o1 --no-cpu --steps 16 --workload quadrotor_hard
run bench_hard_o0 --no-cpu --steps 16 --workload quadrotor_hard --order 0
run bench_full
run bench_cartpole --workload cartpole
run bench_rocket --workload rocket --steps 4
run bench_rollout --workload rollout --steps 4
timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"; cat gpurun_out/bench_ref.json
timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/ncu_launch.log 2>&1
echo "ncu launches rc=$?"
grep -c . gpurun_out/launches.csv
