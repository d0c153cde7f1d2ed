#!/bin/bash
# per-section cycle counts of the latency kernel for the build variants under build/variants/wp_*.so (-DTMPC_WARP_PROFILE)
mkdir -p gpurun_out
cp tinympc_julia_b200/lib/libtinympc_jl.so build/default_lib.so
for v in build/variants/wp_*.so; do
  [ -f "$v" ] || continue
  cp $v tinympc_julia_b200/lib/libtinympc_jl.so
  echo "== $v"; timeout 120 python tools/gpu_warp_one.py 2>&1 | grep "warp kernel cycles" | tail -1
done
cp build/default_lib.so tinympc_julia_b200/lib/libtinympc_jl.so
