#!/bin/bash
# per-instance dynamics (PI variant of the TMEM kernel) at HEAD: timing with the ordered / natural queue, max_iter 100 / 20, and one ncu --set full capture
mkdir -p gpurun_out
python tools/gpu_perinstance.py > gpurun_out/pi_100.json 2>gpurun_out/pi.err; cat gpurun_out/pi_100.json
QMI=20 python tools/gpu_perinstance.py > gpurun_out/pi_20.json 2>>gpurun_out/pi.err; cat gpurun_out/pi_20.json
QMI=20 timeout 1200 ncu --set full --clock-control none --import-source on -k regex:admm_tmem -s 1 -c 1 -o gpurun_out/prof_pi -f python tools/gpu_perinstance.py > gpurun_out/ncu_pi.log 2>&1; echo "ncu rc=$?"
