#!/bin/bash
# the bench lines DESIGN.md section 7 quotes (one GPU, CPU legs included) + smoke + GPU tests
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"
run() { name=$1; shift; timeout 900 python bench.py "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err; echo "$name rc=$?"; python - <<PY
import json
d=json.load(open("gpurun_out/$name.json"))
r=d.get("roofline",{})
print("  value %.4e ms %.4f frac %s iters/s %.4e"%(d["value"], d["ms_per_step"], r.get("frac"), d.get("admm_iters_per_s",0)), d.get("parity_full_batch"))
e=d["e2e"]; print("  e2e %.4e u0 %s"%(e["value"], e.get("u0_only",{}).get("value")), d.get("single_solve",{}).get("gpu"))
PY
}
run r02_p_bench_quadrotor_easy
run r02_p_bench_quadrotor_easy_natural_order --order 0 --no-cpu
run r02_p_bench_quadrotor_hard --workload quadrotor_hard --steps 16
run r02_p_bench_cartpole --workload cartpole
run r02_p_bench_rocket --workload rocket
run r02_p_bench_rollout --workload rollout
run r02_p_bench_float_easy --precision 32
run r02_p_bench_float_hard --precision 32 --workload quadrotor_hard --steps 16 --no-cpu
run r02_p_bench_float_cartpole --precision 32 --workload cartpole --no-cpu
run r02_p_bench_reference_arm --impl reference --steps 5 --warmup 1
run r02_p_bench_reference_arm_rocket --impl reference --workload rocket
run r02_p_bench_reference_arm_rollout --impl reference --workload rollout
timeout 300 python tools/gpu_small_batch.py > gpurun_out/r02_p_small_batch.jsonl 2> gpurun_out/r02_p_small_batch.err; echo "small batch rc=$?"; cat gpurun_out/r02_p_small_batch.jsonl
timeout 120 python tools/gpu_b1.py > gpurun_out/r02_p_b1_split.txt 2>&1; tail -4 gpurun_out/r02_p_b1_split.txt
