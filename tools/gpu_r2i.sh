#!/bin/bash
# 8-GPU session: host-link ceiling at 1/2/4/8 ranks, weak scaling of every workload
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
for n in 1 2 4 8; do
  timeout 300 $TR --nproc-per-node $n --master-port $((29600+n)) tools/d2h_ceiling.py > gpurun_out/d2h_${n}.json 2> gpurun_out/d2h_${n}.err; echo "d2h $n rc=$?"; cat gpurun_out/d2h_${n}.json | cut -c1-700
done
for n in 1 2 4 8; do
  timeout 600 $TR --nproc-per-node $n --master-port $((29700+n)) bench.py --gpus $n --steps 24 --warmup 5 --no-cpu > gpurun_out/scale_easy_${n}.json 2> gpurun_out/scale_easy_${n}.err; echo "easy $n rc=$?"
done
for wl in rollout rocket cartpole quadrotor_hard; do
  for n in 1 8; do
  st=16; [ $wl = rollout ] && st=3; [ $wl = rocket ] && st=4
  timeout 900 $TR --nproc-per-node $n --master-port $((29800+n)) bench.py --gpus $n --steps $st --warmup 3 --no-cpu --workload $wl > gpurun_out/scale_${wl}_${n}.json 2> gpurun_out/scale_${wl}_${n}.err; echo "$wl $n rc=$?"
  done
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob("gpurun_out/scale_*.json")):
    try:
        d=json.load(open(f))
        pr=d["per_rank"]
        print(f, "value %.4e ms %.4f e2e %.4e"%(d["value"], d["ms_per_step"], d["e2e"]["value"]), "u0", d["e2e"].get("u0_only",{}).get("value"), "rank ms min/max %.4f %.4f"%(min(p["kernel_ms_mean"] for p in pr), max(p["kernel_ms_mean"] for p in pr)))
    except Exception as e: print(f, "ERR", e)
PY
nvidia-smi topo -m > gpurun_out/topo.txt 2>&1
