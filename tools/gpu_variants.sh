# builds were prepared as lib variants under build/variants/*.so ; time each with gpu_quick (kernel 3 only)
mkdir -p gpurun_out
for v in build/variants/*.so; do
  cp $v tinympc_julia_b200/lib/libtinympc_jl.so
  echo "== $v"
  QONLY=3 timeout 300 python tools/gpu_quick.py 2>&1 | grep -E "^quad|^cartpole"
done
