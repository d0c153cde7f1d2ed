#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_parity.py tests/test_gpu_dropin.py -m gpu -q -x > gpurun_out/warp_pytest.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/warp_pytest.log
timeout 120 python tools/gpu_b1.py 2>&1 | tail -5
timeout 300 python tools/gpu_small_batch.py > gpurun_out/small_batch.jsonl 2> gpurun_out/small_batch.err; echo "small rc=$?"; cat gpurun_out/small_batch.jsonl; tail -3 gpurun_out/small_batch.err
timeout 120 python tools/gpu_warp_one.py && timeout 200 ncu --set full --clock-control none --import-source on -k regex:admm_warp -s 1 -c 1 -o gpurun_out/warp_one -f python tools/gpu_warp_one.py > gpurun_out/warp_one_ncu.log 2>&1; tail -2 gpurun_out/warp_one_ncu.log
