"""Long-horizon form of the TMEM kernel (quadrotor 12/4, N = 30 and 40, B = 65,536, hard workload): device time against the
thread kernel (development aid)."""
import json, os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import numpy as np, torch
import cases
import tinympc_julia_b200 as t
from tinympc_julia_b200 import problems
B = int(os.environ.get("QB", 65536))
stream = torch.cuda.Stream()
for N in (20, 30, 40):
    for hard in (False, True):
        c = cases.quadrotor_batch(B, hard=hard, N=N)
        row = {"N": N, "hard": hard}
        for name, k in (("tmem", t.KERNEL_TMEM), ("thread", t.KERNEL_THREAD)):
            sol = cases.make_cuda_solver(c, kernel=k)
            t.set_x0(sol, c["x0"])
            ts = []
            for rep in range(4):
                sol.reset()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                with torch.cuda.stream(stream):
                    e0.record(stream); sol.solve_async(stream.cuda_stream); e1.record(stream)
                torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
            st, it = sol.status()
            F = problems.flops_per_iteration(12, 4, N)["F"]
            ms = min(ts[1:])
            row[name] = {"ms": round(ms, 3), "admm_iters_per_s": float(it.sum()) / ms * 1e3, "tflops": float(it.sum()) * F / ms / 1e9, "mean_iters": float(it.mean())}
        print(json.dumps(row), flush=True)
