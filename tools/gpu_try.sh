mkdir -p gpurun_out
timeout 600 python tools/gpu_quick.py > gpurun_out/quick.log 2>&1; echo "quick rc=$?"; grep -E "kernel[ =][34] " gpurun_out/quick.log | tail -40
