"""Cone-kernel timing (development aid): config 4 (rocket 6/3/50, box + cones + half-spaces), B = 65,536, cold, every instance
runs the 100 iterations; device time with CUDA events; a 512-instance sample against the oracle."""
import json, os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import numpy as np
import torch
import cases, oracle
import tinympc_julia_b200 as t

B = int(os.environ.get("QB", 65536))
c = cases.rocket(B, N=50)
sol = cases.make_cuda_solver(c, kernel=int(os.environ.get('QK', t.KERNEL_CONE)))
t.set_x_ref(sol, c["Xref"]); t.set_u_ref(sol, c["Uref"]); t.set_x0(sol, c["x0"])
stream = torch.cuda.Stream()
ts = []
for rep in range(int(os.environ.get("QREPS", 4))):
    sol.reset()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(stream):
        e0.record(stream); sol.solve_async(stream.cuda_stream); e1.record(stream)
    torch.cuda.synchronize()
    ts.append(e0.elapsed_time(e1))
st, it = sol.status()
X, U = t.get_solution(sol)
S = 512
cs = dict(c); cs["x0"] = np.asfortranarray(c["x0"][:, :S])
want = cases.run_checker(cs, oracle.load_oracle(), nthreads=os.cpu_count() or 1)
print(json.dumps({"ms": [round(x, 3) for x in ts], "best_ms": min(ts[1:]), "iters_sum": int(it.sum()),
                  "sample_iter_mismatches": int((it[:S] != want["iters"][0]).sum()),
                  "sample_max_abs_dU": float(np.abs(U[..., :S] - want["U"]).max())}))
