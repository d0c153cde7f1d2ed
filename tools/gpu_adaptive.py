"""Adaptive rho at the size of config 3 (quadrotor 12/4/20, B = 65,536, cold, easy and hard): device time of one solve on
the thread kernel with per-instance adapted cache (CUDA events on the launching stream, L2 flushed), the plain thread
kernel beside it, the compiled reference on the host threads on a sample, and parity of the sample."""
import json, os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import numpy as np
import torch
import cases
import oracle
import tinympc_julia_b200 as t

B = int(os.environ.get("QB", 65536))
REPS = int(os.environ.get("QREPS", 5))
SAMPLE = int(os.environ.get("QSAMPLE", 4096))
stream = torch.cuda.Stream()
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
ref = oracle.load_ref() or oracle.load_oracle()
threads = os.cpu_count() or 1
for hard in (False, True):
    base = cases.quadrotor_batch(B, hard=hard, seed=3)
    row = {"workload": "quadrotor_" + ("hard" if hard else "easy"), "B": B}
    for name, c, kernel in (("adaptive", cases.with_adaptive_rho(cases.quadrotor_batch(B, hard=hard, seed=3)), t.KERNEL_AUTO),
                            ("plain_thread", base, t.KERNEL_THREAD)):
        sol = cases.make_cuda_solver(c, kernel=kernel)
        t.set_x0(sol, c["x0"])
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
        ms = float(np.median(ts))
        row[name] = {"ms": ms, "solves_per_s": B / ms * 1e3, "admm_iters_per_s": float(it.sum()) / ms * 1e3,
                     "mean_iters": float(it.mean()), "unconverged": int(st.sum()), "kernel": sol.kernel}
        if name == "adaptive":
            X, U = t.get_solution(sol)
            rho, _, _ = t.get_adapted_cache(sol)
            cs = dict(c)
            cs["x0"] = np.asfortranarray(c["x0"][:, :SAMPLE])
            want = cases.run_checker(cs, ref, nthreads=threads)
            row["cpu_reference"] = {"kind": ref.kind, "threads": threads, "sample": SAMPLE, "seconds": want["seconds"],
                                    "solves_per_s": SAMPLE / want["seconds"]}
            row["parity_sample"] = {"instances": SAMPLE,
                                    "iteration_mismatches": int((it[:SAMPLE] != want["iters"][0]).sum()),
                                    "status_mismatches": int((st[:SAMPLE] != want["ret"][0]).sum()),
                                    "max_abs_dU": float(np.abs(U[..., :SAMPLE] - want["U"]).max()),
                                    "max_rel_drho": float(np.abs(rho[:SAMPLE] / want["adapted"]["rho"] - 1).max())}
    print(json.dumps(row), flush=True)
