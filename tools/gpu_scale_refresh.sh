#!/bin/bash
# 8-GPU session (late round 2): weak scaling of the workloads whose kernels changed since r02_b (cartpole on 12 warps, the
# cone kernel with the compile-time constraint set), and the headline workload again at 1 and 8 GPUs
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
for wl in quadrotor_easy cartpole rocket; do
  for n in 1 8; do
  st=16; [ $wl = rocket ] && st=4
  timeout 600 $TR --nproc-per-node $n --master-port $((29800+n)) bench.py --gpus $n --steps $st --warmup 3 --no-cpu --workload $wl > gpurun_out/scale_${wl}_${n}.json 2> gpurun_out/scale_${wl}_${n}.err; echo "$wl $n rc=$?"
  done
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob("gpurun_out/scale_*.json")):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        pr=d["per_rank"]
        print(f, "gpus", d["n_gpus"], "value %.4e ms %.4f e2e %.4e"%(d["value"], d["ms_per_step"], d["e2e"]["value"]), "u0", d["e2e"].get("u0_only",{}).get("value"), "rank ms min/max %.4f %.4f"%(min(p["kernel_ms_mean"] for p in pr), max(p["kernel_ms_mean"] for p in pr)))
    except Exception as e: print(f, "ERR", e)
PY
