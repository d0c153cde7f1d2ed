#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_parity.py tests/test_gpu_dropin.py -m gpu -q -x > gpurun_out/warp_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/warp_pytest.log
timeout 120 python tools/gpu_b1.py 2>&1 | tail -5 | head -3
timeout 300 python tools/gpu_small_batch.py > gpurun_out/small_batch.jsonl 2> gpurun_out/small_batch.err; echo "small rc=$?"; cat gpurun_out/small_batch.jsonl; tail -3 gpurun_out/small_batch.err
cp tinympc_julia_b200/lib/libtinympc_jl.so build/default_lib.so
for v in build/variants/wp_*.so; do
  [ -f "$v" ] || continue
  cp $v tinympc_julia_b200/lib/libtinympc_jl.so
  echo "== $v"; timeout 120 python tools/gpu_warp_one.py 2>&1 | grep "warp kernel cycles" | tail -1
done
cp build/default_lib.so tinympc_julia_b200/lib/libtinympc_jl.so
