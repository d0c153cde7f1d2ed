#!/bin/bash
# the four bench lines a TMEM-kernel change moves, one line each (VCMD of tools/gpu_variants.sh)
for a in "quadrotor_easy 64" "quadrotor_hard 64" "cartpole 64" "quadrotor_easy 32"; do
  set -- $a
  timeout 300 python bench.py --workload $1 --precision $2 --steps 24 --warmup 5 --no-cpu 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('$1 $2 ms %.4f frac %.4f'%(d['ms_per_step'], d['roofline']['frac']))"
done
