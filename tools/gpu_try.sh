mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "per_instance_dynamics" > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -25 gpurun_out/pytest_gpu.log
timeout 600 python tools/gpu_perinstance.py > gpurun_out/perinst.json 2> gpurun_out/perinst.err; echo rc=$?; cat gpurun_out/perinst.json; tail -3 gpurun_out/perinst.err
