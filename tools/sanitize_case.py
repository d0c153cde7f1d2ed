"""Small solves of every on-chip kernel variant for compute-sanitizer (one tool per run):
    compute-sanitizer --tool memcheck|racecheck|synccheck python tools/sanitize_case.py [which ...]
which: tmem tmem_warm roll pi subwarp32 cone (default: all).  Results are compared with the oracle so a run that passes the
tool but computes garbage still fails."""
import os
import sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import numpy as np
import cases
import oracle
import tinympc_julia_b200 as t
from tinympc_julia_b200 import problems

which = sys.argv[1:] or ["tmem", "tmem_warm", "roll", "pi", "subwarp32", "cone"]
orc = oracle.load_oracle()

if "tmem" in which or "tmem_warm" in which:
    c = cases.quadrotor_batch(130, hard=True, max_iter=40)
    want = cases.run_checker(c, orc, nthreads=4)
    got = cases.run_cuda(c, kernel=t.KERNEL_TMEM)
    cases.compare(got, want, what="tmem")
    print("tmem ok", flush=True)
    if "tmem_warm" in which:       # warm start on the same handle: reads the carried blocks, REFS = 0
        sol = got["solver"]
        t.set_x0(sol, np.asfortranarray(c["x0"] * 0.9))
        assert t.solve(sol) in (0, 1)
        print("tmem warm ok", flush=True)
if "roll" in which:
    c = cases.quadrotor_closed_loop(40, 6)
    want = cases.run_checker(c, orc, nthreads=4)
    import torch
    sol = cases.make_cuda_solver(c, kernel=t.KERNEL_TMEM)
    t.set_x0(sol, c["x0"])
    xl = torch.from_numpy(np.ascontiguousarray(np.asfortranarray(c["Xref"]).transpose(2, 1, 0))).cuda()
    st = torch.full((6, 40), -1, dtype=torch.int32, device="cuda"); it = torch.full((6, 40), -1, dtype=torch.int32, device="cuda")
    sol.rollout_async(6, xl.data_ptr(), st.data_ptr(), it.data_ptr(), 0)
    sol.synchronize()
    assert np.array_equal(it.cpu().numpy(), want["iters"]) and np.array_equal(st.cpu().numpy(), want["ret"])
    print("roll ok", flush=True)
if "pi" in which:
    B = 70
    c = cases.quadrotor_batch(B, hard=True, max_iter=30)
    base = problems.quadrotor()
    rng = np.random.default_rng(3)
    A = np.asfortranarray(base.A[:, :, None] * (1.0 + 0.05 * rng.uniform(-1, 1, (12, 12, B))))
    Bm = np.asfortranarray(base.B[:, :, None] * (1.0 + 0.1 * rng.uniform(-1, 1, (12, 4, B))))
    outs = []
    for k in (t.KERNEL_TMEM, t.KERNEL_THREAD):
        sol = cases.make_cuda_solver(c, kernel=0)
        t.set_instance_dynamics(sol, A, Bm, None, base.Q, base.R)
        sol.set_kernel(k)
        t.set_x0(sol, c["x0"])
        assert t.solve(sol) in (0, 1)
        outs.append((sol.status(), t.get_solution(sol)))
    assert np.array_equal(outs[0][0][1], outs[1][0][1]) and np.array_equal(outs[0][1].controls, outs[1][1].controls)
    print("pi ok", flush=True)
if "subwarp32" in which:
    c32, want32 = cases.golden_f32_case("quadrotor_hard_64")
    got = cases.run_cuda(c32, kernel=t.KERNEL_SUBWARP, precision=32)
    cases.compare_f32(got, want32, what="subwarp f32")
    print("subwarp32 ok", flush=True)
if "cone" in which:
    c = cases.rocket(20, N=50)
    c["max_iter"] = 30
    want = cases.run_checker(c, orc, nthreads=4)
    got = cases.run_cuda(c, kernel=t.KERNEL_CONE)
    cases.compare(got, want, what="cone")
    t.set_x0(got["solver"], np.asfortranarray(c["x0"] * 1.01))      # warm start: reads the carried element-major state
    assert t.solve(got["solver"]) in (0, 1)
    print("cone ok", flush=True)
