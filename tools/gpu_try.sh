mkdir -p gpurun_out
QMI=20 timeout 600 python tools/gpu_perinstance.py 2>/dev/null | tee gpurun_out/perinst_mi20.json
QMI=100 timeout 600 python tools/gpu_perinstance.py 2>/dev/null | tee gpurun_out/perinst.json
