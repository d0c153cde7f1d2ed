"""Adaptive rho on the GPU — reference: solve(), tinympc/TinyMPC/src/tinympc/admm.cpp:331-357, and
benchmark_rho_adaptation, rho_benchmark.cpp:213-249.

The sensitivity matrices are loaded the way the Julia package loads them (tests/test_sensitivity.jl:40-53:
compute_sensitivity_autograd -> codegen_with_sensitivity, which stores them in the live cache, codegen.cpp:90-97).
Checkers: the golden fixtures captured from the compiled reference (tests/golden/adaptive_*.npz, also run by
test_gpu_parity.py::test_cuda_matches_golden_fixture), the oracle restatement on seeded batches, and the reference's
own Julia-facing library driven call for call in a fresh process.
Bar: identical status and iteration count per instance, states / controls within 1e-5 rel / 1e-6 abs; the adapted rho
and Kinf of every instance within 1e-9 relative.
"""
import ctypes as C
import json
import os
import subprocess
import sys

import numpy as np
import pytest

import cases
import oracle
import tinympc_julia_b200 as t
from tinympc_julia_b200 import _lib

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))


def _check_adapted(sol, want, what):
    rho, K, Pm = t.get_adapted_cache(sol)
    np.testing.assert_allclose(rho, want["rho"], rtol=1e-9, err_msg=f"{what}: adapted rho")
    np.testing.assert_allclose(K, want["Kinf"], rtol=1e-9, atol=1e-12, err_msg=f"{what}: adapted Kinf")
    if "Pinf" in want:
        np.testing.assert_allclose(Pm, want["Pinf"], rtol=1e-9, atol=1e-10, err_msg=f"{what}: adapted Pinf")


@pytest.mark.parametrize("name", sorted(n for n in cases.GOLDEN_CASES if n.startswith("adaptive_")))
def test_adapted_cache_matches_golden(name):
    """rho and Kinf every instance ends with, against the compiled reference's (the fixtures keep them)."""
    c = cases.GOLDEN_CASES[name]()
    g = cases.load_golden(name)
    out = cases.run_cuda(c)
    cases.compare(out, g, what=name)
    _check_adapted(out["solver"], {"rho": g["adapted_rho"], "Kinf": g["adapted_Kinf"]}, name)
    assert (g["adapted_rho"] != c["sys"].rho).any(), "the fixture must exercise the branch"


@pytest.mark.parametrize("make,B", [(lambda B: cases.quadrotor_batch(B), 3000),
                                    (lambda B: cases.quadrotor_batch(B, hard=True, max_iter=40), 1500),
                                    (lambda B: cases.cartpole_batch(B), 3000),
                                    (lambda B: cases.quadrotor_per_instance(B), 600),
                                    (lambda B: cases.random_system(7, 3, 15, B, seed=21), 400)])
def test_adaptive_matches_oracle_seeded(orc, make, B):
    c = cases.with_adaptive_rho(make(B))
    want = cases.run_checker(c, orc, nthreads=8)
    out = cases.run_cuda(c)
    assert out["solver"].kernel == t.KERNEL_THREAD
    cases.compare(out, want, what="adaptive rho vs oracle")
    _check_adapted(out["solver"], want["adapted"], "seeded batch")
    assert (want["iters"] > 5).any() and np.ptp(want["adapted"]["rho"]) > 0


def test_unclipped_and_narrow_clip(orc):
    """adaptive_rho_enable_clipping = 0, and a clip interval so narrow that every update saturates."""
    for kw in (dict(clip=False), dict(lo=0.9, hi=1.1)):
        c = cases.with_adaptive_rho(cases.quadrotor_batch(500, seed=8), **kw)
        want = cases.run_checker(c, orc, nthreads=8)
        out = cases.run_cuda(c)
        cases.compare(out, want, what=f"adaptive rho {kw}")
        _check_adapted(out["solver"], want["adapted"], str(kw))


def test_adapted_cache_is_carried_state(orc):
    """The reference adapts its cache in place: a second solve starts from the adapted rho / Kinf / Pinf (closed loop of
    the oracle, nsteps = 2); tmpc_batch_reset puts every instance back on the setup's cache; clearing the flag keeps
    solving with the adapted cache (reference: nothing restores cache->Kinf)."""
    c = cases.with_adaptive_rho(cases.quadrotor_closed_loop(300, 3))
    want = cases.run_checker(c, orc, nthreads=8)
    out = cases.run_cuda(c)
    cases.compare(out, want, what="three warm adaptive steps")
    np.testing.assert_allclose(out["x_final"], want["x_final"], rtol=1e-5, atol=1e-6)
    sol = out["solver"]
    _check_adapted(sol, want["adapted"], "after three steps")
    # reset: cold again, same results as the first pass
    sol.reset()
    rho, K, _ = t.get_adapted_cache(sol)
    assert np.all(rho == c["sys"].rho) and np.ptp(K, axis=2).max() == 0
    again = cases.run_cuda(c, sol=sol)
    assert np.array_equal(again["iters"], out["iters"]) and np.array_equal(again["U"], out["U"])
    # flag cleared on an adapted batch: the thread kernel keeps reading every instance's own cache and stops adapting
    rho1, K1, P1 = t.get_adapted_cache(sol)
    t.update_settings(sol, max_iter=c["max_iter"], en_state_bound=True, en_input_bound=True, adaptive_rho=False)
    assert sol.kernel == t.KERNEL_THREAD
    t.set_x0(sol, np.asfortranarray(again["x_final"]))
    assert t.solve(sol) in (0, 1)
    rho2, K2, P2 = t.get_adapted_cache(sol)
    assert np.array_equal(rho1, rho2) and np.array_equal(K1, K2) and np.array_equal(P1, P2)
    st, it = sol.status()
    assert (it >= 1).all()


def test_device_rollout_with_adaptive_rho(orc):
    """tmpc_batch_rollout_async on an adaptive batch: the generic device loop (solve kernel + plant-step kernel per step)."""
    import torch
    nsteps, B = 6, 200
    c = cases.with_adaptive_rho(cases.quadrotor_closed_loop(B, nsteps))
    want = cases.run_checker(c, orc, nthreads=8)
    sol = cases.make_cuda_solver(c)
    t.set_x0(sol, c["x0"])
    xlong = torch.from_numpy(np.ascontiguousarray(np.transpose(c["Xref"], (2, 1, 0)))).cuda()     # [B][N + nsteps - 1][nx]
    st = torch.zeros((nsteps, B), dtype=torch.int32, device="cuda")
    it = torch.zeros((nsteps, B), dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    sol.rollout_async(nsteps, xlong.data_ptr(), st.data_ptr(), it.data_ptr())
    sol.synchronize()
    assert np.array_equal(it.cpu().numpy(), want["iters"]) and np.array_equal(st.cpu().numpy(), want["ret"])
    X, U = t.get_solution(sol)
    ok = np.abs(U - want["U"]) <= 1e-6 + 1e-5 * np.abs(want["U"])
    assert ok.all()
    _check_adapted(sol, want["adapted"], "device rollout")


def test_adaptive_refusals():
    """flag on without sensitivity matrices (the reference crashes there: 0 x 0 matrices, rho_benchmark.cpp:204-207),
    float build, an explicit on-chip kernel: refused with a message; clearing the flag restores the solver."""
    c = cases.GOLDEN_CASES["quadrotor_easy_64"]()
    sol = cases.make_cuda_solver(c)
    t.update_settings(sol, max_iter=100, en_state_bound=True, en_input_bound=True, adaptive_rho=True)
    t.set_x0(sol, c["x0"])
    assert t.solve(sol) not in (0, 1)
    assert "sensitivity" in _lib.last_error()
    t.update_settings(sol, max_iter=100, en_state_bound=True, en_input_bound=True, adaptive_rho=False)
    cases.compare(cases.run_cuda(c, sol=sol), cases.load_golden("quadrotor_easy_64"), what="after clearing adaptive_rho")
    # with the matrices: an explicit TMEM request cannot be honoured
    ca = cases.with_adaptive_rho(cases.GOLDEN_CASES["quadrotor_easy_64"]())
    sol = cases.make_cuda_solver(ca, kernel=t.KERNEL_TMEM)
    t.set_x0(sol, ca["x0"])
    assert t.solve(sol) not in (0, 1)
    sol.set_kernel(t.KERNEL_AUTO)
    cases.compare(cases.run_cuda(ca, sol=sol), cases.load_golden("adaptive_quadrotor_easy_64"), what="AUTO after TMEM refusal")
    # float build
    sol = cases.make_cuda_solver(ca, precision=32)
    t.set_x0(sol, ca["x0"])
    assert t.solve(sol) not in (0, 1)
    assert "64-bit" in _lib.last_error()
    # wrong shapes / non-finite entries
    lib = _lib.load()
    v = ca["adaptive"]["sens"]
    bad = np.asfortranarray(v["dK"].copy())
    bad[0, 0] = np.nan
    P = lambda a: np.asfortranarray(a).ctypes.data_as(C.POINTER(C.c_double))
    assert lib.tmpc_batch_set_sensitivity(sol._h, P(bad), P(v["dP"]), P(v["dC1"]), P(v["dC2"])) != 0
    assert lib.tmpc_batch_set_sensitivity(sol._h, None, P(v["dP"]), P(v["dC1"]), P(v["dC2"])) != 0


# ------------------------------------------------------------------ the Julia-facing symbols against the reference's own library

_WORKER = r"""
import ctypes as C, json, sys, os, tempfile
import numpy as np
sys.path.insert(0, sys.argv[3]); sys.path.insert(0, os.path.join(sys.argv[3], "tests"))
import cases
from test_gpu_dropin import JL, F, P, capture_stdout
from tinympc_julia_b200 import lqr_sensitivity
j = JL(sys.argv[1])
scrub = None
if sys.argv[2] != "-":
    scrub = C.CDLL(sys.argv[2]).ref_scrub_stack
out = []
with capture_stdout():
    for which in ("cartpole", "quadrotor"):
        c = cases.cartpole_batch(4, seed=12) if which == "cartpole" else cases.quadrotor_batch(4, seed=12)
        s, N = c["sys"], c["N"]
        dK, dP, dC1, dC2 = lqr_sensitivity(s.A, s.B, s.Q, s.R, s.rho)
        rc = [j.setup(s.A, s.B, s.f, s.Q, s.R, s.rho, s.nx, s.nu, N, adaptive_rho=True, adaptive_rho_min=0.1 * s.rho,
                      adaptive_rho_max=10 * s.rho)]
        rc.append(j.set_bounds(*c["bounds"]))
        d = tempfile.mkdtemp()
        mats = [F(dK), F(dP), F(dC1), F(dC2)]
        args = []
        for m in mats:
            args += [P(m), m.shape[0], m.shape[1]]
        rc_cg = j.lib.codegen_with_sensitivity(d.encode(), *args, 0)
        for b in range(4):
            x0 = c["x0"][:, b] * (1.0 + 0.5 * b)
            rc.append(j.set_x0(x0))
            if b == 3:      # last solve with the flag cleared: the adapted cache stays in use
                rc.append(j.update_settings(en=(1, 1, 0, 0, 0, 0), adaptive_rho=False))
            if scrub: scrub()
            r = j.solve()
            X, U = j.solution()
            out.append(dict(which=which, b=b, rc=rc, ret=r, iters=j.iters(), X=X.tolist(), U=U.tolist()))
            rc = []
print("RESULT" + json.dumps(out))
"""


def _run_worker(lib_path, scrub_lib):
    root = os.path.dirname(HERE)
    p = subprocess.run([sys.executable, "-c", _WORKER, lib_path, scrub_lib, root], capture_output=True, text=True, timeout=600)
    if p.returncode != 0:
        return None, p.stderr[-2000:]
    line = [ln for ln in p.stdout.splitlines() if ln.startswith("RESULT")]
    return (json.loads(line[-1][6:]) if line else None), p.stderr[-2000:]


def test_julia_symbols_adaptive_against_reference_library():
    """setup(adaptive_rho = true) -> set_bound_constraints -> codegen_with_sensitivity -> (set_x0 -> solve_mpc) x 4, the
    last one after update_settings(adaptive_rho = false): our library and the reference's own libtinympc_jl.so, each in
    a fresh process (one solver per process on both sides; the reference also needs a clean stack)."""
    if not os.path.exists(oracle.REFJL_SO) or not os.path.exists(oracle.REFBATCH_SO):
        pytest.skip("oracle/_ref not built (needs /root/reference at build time)")
    ours, err = _run_worker(_lib.LIB_PATH, "-")
    assert ours is not None, err
    ref, err = _run_worker(oracle.REFJL_SO, oracle.REFBATCH_SO)
    if ref is None:
        pytest.skip("the reference library crashed in its adaptive branch (uninitialised RhoAdapter, Q12): " + err[-300:])
    assert len(ours) == len(ref) == 8
    adapted = False
    for a, b in zip(ours, ref):
        what = f"{a['which']} solve {a['b']}"
        assert a["rc"] == b["rc"], what
        assert a["ret"] == b["ret"] and a["iters"] == b["iters"], f"{what}: {a['ret']}, {a['iters']} vs reference {b['ret']}, {b['iters']}"
        np.testing.assert_allclose(np.array(a["U"]), np.array(b["U"]), rtol=1e-5, atol=1e-6, err_msg=what)
        np.testing.assert_allclose(np.array(a["X"]), np.array(b["X"]), rtol=1e-5, atol=1e-6, err_msg=what)
        adapted = adapted or a["iters"][0] > 5
    assert adapted, "no solve reached the first rho update"
