#!/bin/bash
# One GPU-box session: parity tests, pipe microbenchmarks, bench (both arms), smoke, ncu launch list + full capture.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.csv 2>&1
nproc > gpurun_out/nproc.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -5 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/smoke.log
timeout 600 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; cat gpurun_out/bench.json; tail -5 gpurun_out/bench.err
timeout 300 python bench.py --workload quadrotor_hard --no-cpu > gpurun_out/bench_hard.json 2> gpurun_out/bench_hard.err; echo "bench hard rc=$?"; cat gpurun_out/bench_hard.json
timeout 300 python bench.py --kernel 2 --no-cpu --steps 10 > gpurun_out/bench_subwarp.json 2> gpurun_out/bench_subwarp.err; echo "bench subwarp rc=$?"; timeout 300 python bench.py --kernel 1 --no-cpu --steps 10 > gpurun_out/bench_thread.json 2> gpurun_out/bench_thread.err; echo "bench thread rc=$?"
timeout 300 python bench.py --kernel 4 --no-cpu --steps 20 > gpurun_out/bench_tmem2.json 2> gpurun_out/bench_tmem2.err; echo "bench tmem2 rc=$?"
timeout 300 python bench.py --kernel 4 --no-cpu --steps 20 --workload quadrotor_hard > gpurun_out/bench_tmem2_hard.json 2> gpurun_out/bench_tmem2_hard.err
timeout 300 python bench.py --precision 32 --steps 20 > gpurun_out/bench_f32.json 2> gpurun_out/bench_f32.err; echo "bench f32 rc=$?"; cat gpurun_out/bench_f32.json
timeout 300 python bench.py --precision 32 --steps 20 --no-cpu --workload quadrotor_hard > gpurun_out/bench_f32_hard.json 2> gpurun_out/bench_f32_hard.err
timeout 600 python tools/gpu_float.py > gpurun_out/float.json 2> gpurun_out/float.err; echo "float rc=$?"
timeout 600 python tools/gpu_rollout.py > gpurun_out/rollout.json 2> gpurun_out/rollout.err; echo "rollout rc=$?"; cat gpurun_out/rollout.json
timeout 600 python tools/gpu_perinstance.py > gpurun_out/perinst.json 2> gpurun_out/perinst.err; echo "perinst rc=$?"; cat gpurun_out/perinst.json
timeout 300 python tools/gpu_steady.py > gpurun_out/steady.json 2> gpurun_out/steady.err; echo "steady rc=$?"; cat gpurun_out/steady.json
timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "bench ref rc=$?"; cat gpurun_out/bench_ref.json
timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/ncu_launch.log 2>&1
echo "ncu launches rc=$?"
timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/plain2.log 2>&1 && \
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:admm_tmem -s 3 -c 2 -o gpurun_out/prof_tmem -f python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"
ls -la gpurun_out
