#!/bin/bash
# end-of-round profiler evidence: launch list of the default bench command, ncu --set full of the headline kernel and of the
# latency kernel (each after the same command has exited 0 without ncu)
mkdir -p gpurun_out
timeout 600 python bench.py --steps 2 --warmup 3 > gpurun_out/prof_plain_default.log 2>&1; echo "plain default rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/r02_p_ncu_launches_default_bench.csv python bench.py --steps 2 --warmup 3 > gpurun_out/prof_ncu_launches.log 2>&1; echo "launch list rc=$?"
bash tools/gpu_prof.sh 3 r02_p_easy
timeout 120 python tools/gpu_warp_one.py > gpurun_out/prof_plain_warp.log 2>&1 && timeout 300 ncu --set full --clock-control none --import-source on -k regex:admm_warp -s 1 -c 1 -o gpurun_out/prof_r02_p_warp -f python tools/gpu_warp_one.py > gpurun_out/prof_ncu_warp.log 2>&1; echo "ncu warp rc=$?"
ls -la gpurun_out/*.ncu-rep
