#!/bin/bash
mkdir -p gpurun_out
timeout 600 python bench.py --workload rocket --steps 1 --warmup 3 --no-cpu --batch 16384 > gpurun_out/plain_rocket.log 2>&1 && \
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:admm_cone -s 2 -c 1 -o gpurun_out/prof_cone -f python bench.py --workload rocket --steps 1 --warmup 3 --no-cpu --batch 16384 > gpurun_out/ncu_cone.log 2>&1
echo "ncu rc=$?"; ls -la gpurun_out/prof_cone.ncu-rep
