#!/bin/bash
# driver-style check on 2 GPUs: both arms under torchrun, then the GPU tests
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29911 bench.py --impl reference --gpus 2 --steps 5 --warmup 1 > gpurun_out/q_ref2.json 2> gpurun_out/q_ref2.err; echo "ref rc=$?"; head -c 300 gpurun_out/q_ref2.json; echo
timeout 600 $TR --master-port 29912 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/q_gpu2.json 2> gpurun_out/q_gpu2.err; echo "gpu rc=$?"
python - <<'PY'
import json
d=json.load(open("gpurun_out/q_gpu2.json"))
print("value %.4e ms %.4f e2e %.4e u0 %.4e"%(d["value"], d["ms_per_step"], d["e2e"]["value"], d["e2e"]["u0_only"]["value"]), [ (p["rank"], round(p["kernel_ms_mean"],4)) for p in d["per_rank"]], d["collective"])
PY
timeout 600 $TR --master-port 29913 bench.py --gpus 2 --workload rollout --steps 2 --warmup 3 --no-cpu > gpurun_out/q_roll2.json 2> gpurun_out/q_roll2.err; echo "rollout rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/q_roll2.json')); print('rollout value %.4e ms %.3f'%(d['value'], d['ms_per_step']), d['collective'])"
