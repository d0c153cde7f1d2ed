#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/pytest_gpu.log
for wl in quadrotor_easy quadrotor_hard cartpole; do
st=48; [ $wl = quadrotor_hard ] && st=16
timeout 900 python bench.py --workload $wl --steps $st --no-cpu > gpurun_out/o_$wl.json 2> gpurun_out/o_$wl.err; python - <<PY
import json
d=json.load(open("gpurun_out/o_$wl.json")); u=d["e2e"]["u0_only"]; print("$wl ms %.4f value %.4e e2e %.4e u0 %.4e (%.3f ms) ratio %.3f"%(d["ms_per_step"], d["value"], d["e2e"]["value"], u["value"], u["ms_per_step"], u["vs_device_value"]))
PY
tail -2 gpurun_out/o_$wl.err
done
