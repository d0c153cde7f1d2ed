#!/bin/bash
# quick check of a TMEM-kernel change: its parity tests + the four bench lines it moves (no CPU legs)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/quick_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/quick_pytest.log
for a in "quadrotor_easy 64" "quadrotor_hard 64" "cartpole 64" "quadrotor_easy 32" "rollout 64"; do
  set -- $a
  st=24; [ $1 = rollout ] && st=3
  timeout 600 python bench.py --workload $1 --precision $2 --steps $st --warmup 5 --no-cpu > gpurun_out/quick_$1_$2.json 2> gpurun_out/quick_$1_$2.err; echo "$a rc=$?"
  python -c "
import json; d=json.load(open('gpurun_out/quick_$1_$2.json')); print('  value %.4e ms %.4f frac %s'%(d['value'], d['ms_per_step'], d.get('roofline',{}).get('frac')), d.get('parity_full_batch'))"
done
