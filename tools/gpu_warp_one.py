"""Three cold solves of ONE hard quadrotor instance on the latency kernel (profiling target)."""
import sys
sys.path.insert(0, "/root/repo/tests"); sys.path.insert(0, "/root/repo")
import numpy as np, torch
import cases
import tinympc_julia_b200 as t
c = cases.quadrotor_batch(1, hard=True, seed=3)
sol = cases.make_cuda_solver(c, kernel=t.KERNEL_WARP)
t.set_x0(sol, c["x0"])
for _ in range(3):
    sol.reset(); t.solve(sol)
print(sol.status())
