"""Small batches: device time of one cold solve of B quadrotor (easy / hard x0) or cartpole instances on the latency kernel (one warp per instance) and on
the throughput kernel (TMEM), CUDA events on the launching stream, median of 30 after 10; plus the B = 1 path through the
reference's own symbols (tools/gpu_b1.py prints its split).  One JSON line per B."""
import sys, json
sys.path.insert(0, "/root/repo/tests"); sys.path.insert(0, "/root/repo")
import numpy as np, torch
import cases
import tinympc_julia_b200 as t

st = torch.cuda.Stream()
for wl in ("easy", "hard", "cartpole"):
    for B in (1, 64, 296, 297, 1024, 2048, 4096):
        c = cases.cartpole_batch(B, seed=3) if wl == "cartpole" else cases.quadrotor_batch(B, hard=(wl == "hard"), seed=3)
        row = {"B": B, "x0": wl}
        for name, k in (("warp", t.KERNEL_WARP), ("tmem", t.KERNEL_TMEM)):
            sol = cases.make_cuda_solver(c, kernel=k)
            t.set_x0(sol, c["x0"])
            ev = []
            for _ in range(40):
                sol.reset()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                with torch.cuda.stream(st):
                    e0.record(st); sol.solve_async(st.cuda_stream); e1.record(st)
                torch.cuda.synchronize(); ev.append(e0.elapsed_time(e1) * 1e3)
            it = sol.status()[1]
            row[name + "_us"] = round(float(np.median(ev[10:])), 2)
            row[name + "_iters_mean"] = round(float(np.mean(it)), 2)
            assert sol.kernel == k
        print(json.dumps(row), flush=True)
