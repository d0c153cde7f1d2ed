# builds were prepared as lib variants under build/variants/*.so ; run $VCMD (default: gpu_quick, kernel 3 only) with each
mkdir -p build
mkdir -p gpurun_out
cp tinympc_julia_b200/lib/libtinympc_jl.so build/default_lib.so
for v in build/default_lib.so build/variants/*.so; do
  [ -f "$v" ] || continue
  [ "$v" = build/default_lib.so ] || cp $v tinympc_julia_b200/lib/libtinympc_jl.so
  echo "== $v"
  if [ -n "$VCMD" ]; then timeout 300 bash -c "$VCMD" 2>&1 | tail -4; else QONLY=3 timeout 300 python tools/gpu_quick.py 2>&1 | grep -E "^quad|^cartpole"; fi
done
cp build/default_lib.so tinympc_julia_b200/lib/libtinympc_jl.so
