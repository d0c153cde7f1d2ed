mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "rollout" > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -30 gpurun_out/pytest_gpu.log
timeout 300 python bench.py --steps 5 --no-cpu > gpurun_out/b.json 2>gpurun_out/b.err; echo "bench rc=$? lines=$(wc -l < gpurun_out/b.json)"
