#!/bin/bash
mkdir -p gpurun_out
for top in 4 16 64 256; do
for wl in quadrotor_easy quadrotor_hard cartpole; do
  steps=48; [ $wl = quadrotor_hard ] && steps=16
  TMPC_ORDER_TOP=$top timeout 600 python bench.py --no-cpu --steps $steps --workload $wl > gpurun_out/top_${top}_$wl.json 2> gpurun_out/top_${top}_$wl.err
  python - <<PY
import json
d=json.load(open("gpurun_out/top_${top}_$wl.json"))
print("top $top $wl  ms %.4f  min %.4f max %.4f"%(d["ms_per_step"], d["per_rank"][0]["kernel_ms_min"], d["per_rank"][0]["kernel_ms_max"]))
PY
done; done
