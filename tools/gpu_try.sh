mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
timeout 300 python bench.py --no-cpu > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; cat gpurun_out/bench.json; tail -3 gpurun_out/bench.err
timeout 300 python bench.py --workload quadrotor_hard --no-cpu > gpurun_out/bench_hard.json 2> gpurun_out/bench_hard.err; echo "bench hard rc=$?"; cat gpurun_out/bench_hard.json
timeout 300 python bench.py --kernel 4 --no-cpu --steps 20 > gpurun_out/bench_tmem2.json 2> gpurun_out/bench_tmem2.err; echo "bench tmem2 rc=$?"; cat gpurun_out/bench_tmem2.json
timeout 600 python tools/gpu_rollout.py > gpurun_out/rollout.json 2> gpurun_out/rollout.err; echo "rollout rc=$?"; cat gpurun_out/rollout.json
timeout 600 python tools/gpu_perinstance.py > gpurun_out/perinst.json 2> gpurun_out/perinst.err; echo rc=$?; cat gpurun_out/perinst.json
