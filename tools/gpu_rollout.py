"""Config 5 (BASELINE.json): closed-loop batched MPC rollout on ONE GPU — quadrotor, warm-started ADMM, per-instance
sliding line references, everything device-resident (tmpc_batch_rollout_async).  Prints one JSON line: instance-steps/s,
ADMM iterations/s, achieved FP64 FMA fraction.  Development/measurement aid; the judged contract is bench.py."""
import argparse, json, os, sys, time
ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import torch
import cases
import tinympc_julia_b200 as t
from tinympc_julia_b200 import problems

ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, default=65536)
ap.add_argument("--steps", type=int, default=200)
ap.add_argument("--kernel", type=int, default=0)
ap.add_argument("--reps", type=int, default=3)
args = ap.parse_args()
B, S = args.batch, args.steps
c = cases.quadrotor_closed_loop(B, S)
s, N = c["sys"], c["N"]
sol = cases.make_cuda_solver(c, kernel=args.kernel)
xlong = torch.from_numpy(np.ascontiguousarray(np.asfortranarray(c["Xref"]).transpose(2, 1, 0))).cuda()
st_out = torch.full((S, B), -1, dtype=torch.int32, device="cuda")
it_out = torch.full((S, B), -1, dtype=torch.int32, device="cuda")
best = None
ts = torch.cuda.Stream()
for rep in range(args.reps):
    t.set_x0(sol, c["x0"])
    sol.reset()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(ts)
    sol.rollout_async(S, xlong.data_ptr(), st_out.data_ptr(), it_out.data_ptr(), ts.cuda_stream)
    e1.record(ts)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    best = ms if best is None else min(best, ms)
iters = int(it_out.sum().item())
F = problems.flops_per_iteration(s.nx, s.nu, N)["F"]
print(json.dumps({"workload": f"closed loop quadrotor B={B} steps={S} warm start sliding per-instance refs",
                  "kernel": sol.kernel, "ms": best, "instance_steps_per_s": B * S / (best * 1e-3),
                  "admm_iters_per_s": iters / (best * 1e-3), "mean_iters_per_step": iters / (B * S),
                  "solved_fraction": float((st_out == 0).float().mean().item()),
                  "tflops": iters * F / (best * 1e-3) / 1e12, "launches": sol.launch_count}))
