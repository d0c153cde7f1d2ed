"""Per-instance dynamics at scale: batched setup time and solve throughput, quadrotor B = 65,536."""
import json, os, sys, time
ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import cases
import tinympc_julia_b200 as t
from tinympc_julia_b200 import problems

B = int(os.environ.get("QB", 65536))
base = problems.quadrotor()
n, m, N = base.nx, base.nu, 20
rng = np.random.default_rng(3)
A = np.asfortranarray(base.A[:, :, None] * (1.0 + 0.05 * rng.uniform(-1, 1, (n, n, B))))
PB = float(os.environ.get("QPB", 0.2))      # relative perturbation of B (0.2: ~0.1 % of the systems need all max_iter iterations)
Bm = np.asfortranarray(base.B[:, :, None] * (1.0 + PB * rng.uniform(-1, 1, (n, m, B))))
x0 = np.asfortranarray(0.1 * rng.uniform(-1, 1, (n, B)))
MI = int(os.environ.get("QMI", 100))
c = cases._base(base, N, bounds=base.bounds(N), x0=x0, max_iter=MI)
sol = cases.make_cuda_solver(c)
KPI = int(os.environ.get("QK", 0))
t0 = time.perf_counter()
t.set_instance_dynamics(sol, A, Bm, None, base.Q, base.R)
t_setup = time.perf_counter() - t0
sol.set_kernel(KPI)
t.set_x0(sol, x0)
best = 1e9
for rep in range(3):
    sol.reset()
    t0 = time.perf_counter(); sol.solve_async(0); sol.synchronize(); best = min(best, time.perf_counter() - t0)
st, it = sol.status()
print(json.dumps({"workload": f"quadrotor per-instance dynamics B={B} (A +-5 %, B +-{100 * PB:.0f} %)", "max_iter": MI, "max_iters_seen": int(it.max()),
                  "admm_iters_per_s": float(it.sum()) / best, "instances_at_max_iter": int((st != 0).sum()), "setup_s_incl_upload": t_setup, "solve_ms": best * 1e3,
                  "solves_per_s": B / best, "mean_iters": float(it.mean()), "solved": float((st == 0).mean()), "kernel": sol.kernel}))
