#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
run() { name=$1; shift; timeout 900 python bench.py "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err; echo "$name rc=$?"; python - <<PY
import json
d=json.load(open("gpurun_out/$name.json"))
print("  value %.4e ms %.4f frac %.4f per_rank %s"%(d["value"], d["ms_per_step"], d["roofline"]["frac"], {k:round(v,4) for k,v in d["per_rank"][0].items() if k.startswith("kernel")}))
print("  e2e %.4e (%.3f ms) unpipelined %.3f ms u0 %s"%(d["e2e"]["value"], d["e2e"]["ms_per_step"], d["e2e"].get("unpipelined_ms_per_step",0), d["e2e"].get("u0_only",{}).get("vs_device_value")))
PY
tail -2 gpurun_out/$name.err; }
run bench_easy_o1 --no-cpu --steps 48
run bench_easy_o0 --no-cpu --steps 48 --order 0
run bench_hard_o1 --no-cpu --steps 16 --workload quadrotor_hard
run bench_hard_o0 --no-cpu --steps 16 --workload quadrotor_hard --order 0
run bench_cartpole --no-cpu --workload cartpole
