#!/bin/bash
# Weak-scaling run on one box: bench.py at N = 1, 2, 4, 8 (as many as the box has), one JSON line each.
mkdir -p gpurun_out
NG=$(nvidia-smi -L | wc -l)
for N in 1 2 4 8; do
  [ $N -gt $NG ] && break
  if [ $N -eq 1 ]; then
    timeout 600 python bench.py --no-cpu > gpurun_out/scale_n1.json 2> gpurun_out/scale_n1.err
  else
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + N)) \
      bench.py --gpus $N --no-cpu > gpurun_out/scale_n$N.json 2> gpurun_out/scale_n$N.err
  fi
  echo "N=$N rc=$?"; python - <<P
import json
try:
    d = json.loads(open("gpurun_out/scale_n$N.json").read().strip().splitlines()[-1])
    print($N, round(d["value"] / 1e6, 1), "M solves/s", round(d["ms_per_step"], 3), "ms; e2e", round(d["e2e"]["value"] / 1e6, 1))
except Exception as e:
    print("no line:", e)
P
done
