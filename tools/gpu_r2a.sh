#!/bin/bash
# round 2, first GPU session: all GPU tests (no -x), the queue-order experiment, default bench
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -30 gpurun_out/pytest_gpu.log
timeout 600 python tools/gpu_order.py > gpurun_out/order.log 2>&1; echo "order rc=$?"; tail -20 gpurun_out/order.log
timeout 600 python bench.py --no-cpu --steps 20 > gpurun_out/bench_a.json 2> gpurun_out/bench_a.err; echo "bench rc=$?"; cat gpurun_out/bench_a.json; tail -3 gpurun_out/bench_a.err
