"""Steady-state rate of the solve loop alone: quadrotor, B = 65,536, tolerances 0 so that every instance runs exactly
max_iter iterations (no refills, no write-out passes inside the loop, residual tests fail at the first knot point).
Development aid: separates the sweep rate from the per-solve overheads."""
import sys, os, json
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import numpy as np, torch
import cases
import tinympc_julia_b200 as t
from tinympc_julia_b200 import problems
B = int(os.environ.get("QB", 65536)); K = int(os.environ.get("QK", 3)); MI = int(os.environ.get("QMI", 100))
c = cases.quadrotor_batch(B, hard=True, max_iter=MI)
c["abs_pri_tol"] = 0.0; c["abs_dua_tol"] = 0.0
sol = cases.make_cuda_solver(c, kernel=K)
t.set_x0(sol, c["x0"])
ts = torch.cuda.Stream()
best = None
for rep in range(5):
    sol.reset(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(ts); sol.solve_async(ts.cuda_stream); e1.record(ts); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1); best = ms if best is None else min(best, ms)
st, it = sol.status()
F = problems.flops_per_iteration(12, 4, 20)["F"]
print(json.dumps({"workload": f"quadrotor steady state B={B} max_iter={MI} tol=0", "kernel": sol.kernel, "ms": best,
                  "iters_per_s": float(it.sum()) / (best * 1e-3), "tflops": float(it.sum()) * F / (best * 1e-3) / 1e12,
                  "mean_iters": float(it.mean())}))
