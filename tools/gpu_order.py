"""Queue-order experiment (development aid): config 3 at B = 65,536, seeds 3..10, easy and hard — device time of one cold
solve with the natural queue order against the longest-first order (queue_order.cu), CUDA events, L2 flushed."""
import json, os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import numpy as np
import torch
import cases
import tinympc_julia_b200 as t

B = int(os.environ.get("QB", 65536))
REPS = int(os.environ.get("QREPS", 8))
stream = torch.cuda.Stream()
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
out = {"B": B, "reps": REPS, "rows": []}
for hard in (False, True):
    for seed in range(3, 11):
        c = cases.quadrotor_batch(B, hard=hard, seed=seed)
        sol = cases.make_cuda_solver(c)
        t.set_x0(sol, c["x0"])
        row = {"hard": hard, "seed": seed}
        for mode, name in ((0, "natural"), (1, "longest_first")):
            sol.set_order(mode)
            ts = []
            for rep in range(REPS + 2):
                flush.zero_()
                sol.reset()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                with torch.cuda.stream(stream):
                    e0.record(stream)
                    sol.solve_async(stream.cuda_stream)
                    e1.record(stream)
                torch.cuda.synchronize()
                if rep >= 2:
                    ts.append(e0.elapsed_time(e1))
            st, it = sol.status()
            row[name + "_ms"] = float(np.median(ts))
            row[name + "_min_ms"] = float(np.min(ts))
            row["iters_sum"] = int(it.sum()); row["iters_max"] = int(it.max())
        row["gain"] = row["natural_ms"] / row["longest_first_ms"]
        out["rows"].append(row)
        print(json.dumps(row), flush=True)
for hard in (False, True):
    rows = [r for r in out["rows"] if r["hard"] == hard]
    print("hard" if hard else "easy", "mean natural %.4f ms, longest-first %.4f ms" %
          (np.mean([r["natural_ms"] for r in rows]), np.mean([r["longest_first_ms"] for r in rows])), flush=True)
