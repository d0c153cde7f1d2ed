"""Quick GPU sanity + timing driver used during development (not a test, not the bench)."""
import sys, time, os
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import numpy as np
import cases, oracle
import tinympc_julia_b200 as t

orc = oracle.load_oracle()
KS = tuple(int(x) for x in os.environ.get('QONLY', '1,2,3').split(','))
FAST = os.environ.get('QONLY') is not None
for name in () if FAST else ("cartpole_one_solve", "cartpole_test_basic_bounds", "cartpole_batch_64", "quadrotor_easy_64", "quadrotor_hard_64", "quadrotor_hover", "quadrotor_per_instance_32", "quadrotor_closed_loop_16x40", "rocket_julia_example", "random_5x2_16"):
    c = cases.GOLDEN_CASES[name]()
    want = cases.load_golden(name)
    for k in KS:
        sol = cases.make_cuda_solver(c, kernel=0)
        sol.set_kernel(k)
        if sol.kernel != k:
            continue
        got = cases.run_cuda(c, kernel=k)
        print(name, "kernel", k, "iter mismatches", int((got["iters"] != want["iters"]).sum()), "of", want["iters"].size,
              "max|dX|", float(np.abs(got["X"] - want["X"]).max()), "max|dU|", float(np.abs(got["U"] - want["U"]).max()), flush=True)

import torch
B = int(os.environ.get("QB", 65536))
for hard in (False, True):
    c = cases.quadrotor_batch(B, hard=hard)
    for k in KS:
        sol = cases.make_cuda_solver(c, kernel=0)
        sol.set_kernel(k)
        if sol.kernel != k:
            continue
        t.set_x0(sol, c["x0"])
        ts = []
        for rep in range(4):
            sol.reset()
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            sol.solve_async(0)
            sol.synchronize()
            torch.cuda.synchronize()
            ts.append(time.perf_counter() - t0)
        st, it = sol.status()
        F = 24612
        best = min(ts)
        print(f"quad hard={hard} kernel={k} B={B} best {best*1e3:.3f} ms  solves/s {B/best:.3e} iters/s {it.sum()/best:.3e} "
              f"TFLOP/s {it.sum()*F/best/1e12:.2f} mean iters {it.mean():.2f} solved {(st==0).mean():.3f}", flush=True)

# config 2: batched cartpole, B = 65,536
from tinympc_julia_b200 import problems
Fc = problems.flops_per_iteration(4, 1, 20)["F"]
c = cases.cartpole_batch(B)
for k in KS:
    sol = cases.make_cuda_solver(c, kernel=0)
    sol.set_kernel(k)
    if sol.kernel != k:
        continue
    t.set_x0(sol, c["x0"])
    ts = []
    for rep in range(4):
        sol.reset()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        sol.solve_async(0)
        sol.synchronize()
        ts.append(time.perf_counter() - t0)
    st, it = sol.status()
    best = min(ts)
    print(f"cartpole kernel={k} B={B} best {best*1e3:.3f} ms  solves/s {B/best:.3e} iters/s {it.sum()/best:.3e} "
          f"TFLOP/s {it.sum()*Fc/best/1e12:.2f} mean iters {it.mean():.2f} solved {(st==0).mean():.3f}", flush=True)

# config 4: batched rocket landing, nx=6 nu=3 N=50, box + both cones + half-spaces (thread kernel), B = 65,536
if not FAST or os.environ.get("QROCKET"):
    Fr = problems.flops_per_iteration(6, 3, 50)["F"]
    c = cases.rocket(B, N=50)
    sol = cases.make_cuda_solver(c, kernel=0)
    t.set_x_ref(sol, c["Xref"]); t.set_u_ref(sol, c["Uref"])
    t.set_x0(sol, c["x0"])
    ts = []
    for rep in range(3):
        sol.reset()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        sol.solve_async(0)
        sol.synchronize()
        ts.append(time.perf_counter() - t0)
    st, it = sol.status()
    best = min(ts)
    print(f"rocket kernel={sol.kernel} B={B} best {best*1e3:.3f} ms  solves/s {B/best:.3e} iters/s {it.sum()/best:.3e} "
          f"TFLOP/s(box part) {it.sum()*Fr/best/1e12:.2f} mean iters {it.mean():.2f} solved {(st==0).mean():.3f}", flush=True)
