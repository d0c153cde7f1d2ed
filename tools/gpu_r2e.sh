#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_gpu.log
timeout 900 python bench.py --workload rocket --steps 6 > gpurun_out/bench_rocket.json 2> gpurun_out/bench_rocket.err; echo "rocket rc=$?"
python - <<PY
import json
d=json.load(open("gpurun_out/bench_rocket.json"))
print("rocket value %.4e ms %.4f frac %.4f kernel %s iters/s %.4e"%(d["value"], d["ms_per_step"], d["roofline"]["frac"], d["impl_config"]["kernel"], d["admm_iters_per_s"]))
print(d.get("parity_full_batch")); print(d.get("cpu_baseline")); print(d["e2e"])
PY
tail -3 gpurun_out/bench_rocket.err
