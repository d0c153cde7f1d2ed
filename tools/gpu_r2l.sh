#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/pytest_gpu.log
run() { name=$1; shift; timeout 900 python bench.py "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err; echo "$name rc=$?"; python - <<PY
import json
d=json.load(open("gpurun_out/$name.json"))
print("  value %.4e ms %.4f frac %.4f iters/s %.4e kernel %s"%(d["value"], d["ms_per_step"], d["roofline"]["frac"], d["admm_iters_per_s"], d["impl_config"]["kernel"]), d.get("parity_full_batch"))
PY
tail -2 gpurun_out/$name.err; }
run bench_easy --steps 48 --no-cpu
run bench_f32 --precision 32 --steps 48
run bench_f32_hard --precision 32 --steps 16 --workload quadrotor_hard --no-cpu
