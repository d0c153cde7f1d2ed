"""Parity of the CUDA solve path against the oracle and the reference's golden fixtures — GPU only.

Every call goes through the C ABI (ctypes -> lib/libtinympc_jl.so).  Bar (BASELINE.json north_star):
identical solve status and iteration count per instance; states and controls within 1e-5 relative /
1e-6 absolute (tinytype = double).
"""
import numpy as np
import pytest

import cases
import tinympc_julia_b200 as t
from tinympc_julia_b200 import problems as problems_mod

pytestmark = pytest.mark.gpu

THREAD, SUBWARP, TMEM, CONE, WARP = t.KERNEL_THREAD, t.KERNEL_SUBWARP, t.KERNEL_TMEM, t.KERNEL_CONE, t.KERNEL_WARP


def kernels_for(c):
    """Kernels able to run a case: the thread kernel always; the on-chip kernels where they apply."""
    ks = [THREAD]
    sol = cases.make_cuda_solver(c, kernel=t.KERNEL_AUTO)
    for k in (SUBWARP, TMEM, CONE, WARP):
        sol.set_kernel(k)
        if sol.kernel == k:
            ks.append(k)
    del sol
    return ks


@pytest.mark.parametrize("name", sorted(cases.GOLDEN_CASES))
def test_cuda_matches_golden_fixture(name):
    c = cases.GOLDEN_CASES[name]()
    want = cases.load_golden(name)
    for k in kernels_for(c):
        got = cases.run_cuda(c, kernel=k)
        cases.compare(got, want, what=f"cuda[kernel {k}] vs golden[{name}]")
        np.testing.assert_allclose(got["x_final"], want["x_final"], rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("make,B", [(cases.cartpole_batch, 2000), (cases.quadrotor_batch, 1500),
                                    (lambda B: cases.quadrotor_batch(B, hard=True), 1000),
                                    (cases.quadrotor_per_instance, 500), (lambda B: cases.rocket(B, N=50), 200),
                                    (lambda B: cases.random_system(7, 3, 15, B, seed=21), 300)])
def test_cuda_matches_oracle_seeded(orc, make, B):
    c = make(B)
    want = cases.run_checker(c, orc, nthreads=8)
    for k in kernels_for(c):
        got = cases.run_cuda(c, kernel=k)
        cases.compare(got, want, what=f"cuda[kernel {k}] vs oracle")


@pytest.mark.parametrize("nx,nu,N,lin", [(20, 8, 12, True), (32, 16, 6, False), (1, 1, 5, False), (3, 2, 40, True)])
def test_thread_kernel_shape_limits(orc, nx, nu, N, lin):
    """The general kernel at and near its limits (nx <= 32, nu <= 16; run-time dimensions, rolled loops), with and without
    half-spaces, cold and warm, against the oracle."""
    c = cases.random_system(nx, nu, N, 96, seed=100 + nx, with_linear=lin)
    c["nsteps"] = 2
    want = cases.run_checker(c, orc, nthreads=8)
    got = cases.run_cuda(c)
    assert got["solver"].kernel == THREAD
    cases.compare(got, want, what=f"thread kernel {nx}x{nu}, N = {N}")
    np.testing.assert_allclose(got["x_final"], want["x_final"], rtol=1e-5, atol=1e-6)


def test_tmem_kernel_long_horizon_form(orc):
    """Horizons 22 .. 42 of the quadrotor shape run the four-warp form of the TMEM kernel (all 512 Tensor-Memory columns of a
    lane quarter per warp): cold and warm solves against the oracle and, bit for bit, the thread kernel; the device closed
    loop on such a horizon (launch sequence, not the fused kernel); N = 43 is beyond it."""
    import torch
    for N in (22, 30, 42):
        c = cases.quadrotor_batch(700, hard=True, seed=40 + N, N=N, max_iter=30)
        c["nsteps"] = 2                                        # second step = warm start from the carried state
        want = cases.run_checker(c, orc, nthreads=8)
        a = cases.run_cuda(c, kernel=TMEM)
        assert a["solver"].kernel == TMEM, N
        b = cases.run_cuda(c, kernel=THREAD)
        cases.compare(a, want, what=f"long-horizon TMEM form, N = {N}")
        for k in ("ret", "iters", "X", "U"):
            assert np.array_equal(a[k], b[k]), f"N = {N}: {k} differs between the TMEM and the thread kernel"
    c = cases.quadrotor_closed_loop(200, 5)                    # per-instance sliding references
    c["N"] = 20
    # same closed loop on a 30-knot horizon: references of that length
    x0, Xlong = problems_mod.quadrotor_line_refs(200, 30, 5, 7)
    c = cases.quadrotor_batch(200, N=30)
    c.update(x0=x0, Xref=Xlong, ref_mode=cases.oracle.REF_SLIDING, nsteps=5)
    want = cases.run_checker(c, orc, nthreads=8)
    sol = cases.make_cuda_solver(c, kernel=TMEM)
    assert sol.kernel == TMEM
    t.set_x0(sol, c["x0"])
    st, it, xf = _device_rollout(sol, 5, xlong=c["Xref"])
    assert np.array_equal(it, want["iters"]) and np.array_equal(st, want["ret"])
    np.testing.assert_allclose(xf, want["x_final"], rtol=1e-5, atol=1e-6)
    c = cases.quadrotor_batch(64, N=43)
    sol = cases.make_cuda_solver(c)
    assert sol.kernel != TMEM
    sol.set_kernel(TMEM)
    assert sol.kernel == -1


def test_cone_kernel_every_compiled_form(orc):
    """admm_cone.cu compiles the rocket shape five times: box only, cones only (the reference's rocket example), one half-space
    per side, cones + one half-space per side (config 4), and the general form with run-time flags.  Each against the oracle and, bit for bit,
    against the thread kernel; the general form with constraint sets the fixed forms do not cover (cone on one side only,
    a cone that does not start at element 0, two half-spaces, per-instance references, no box)."""
    from tinympc_julia_b200 import problems
    e = problems.rocket().extra
    variants = []
    variants.append(("cones only", cases.rocket(150, N=30, linear=False)))
    c = cases.rocket(150, N=30, linear=False)
    c["cones"] = None
    variants.append(("box only", c))
    c = cases.rocket(150, N=30)
    c["cones"] = None
    variants.append(("one half-space per side, no cones", c))
    variants.append(("cones + one half-space per side", cases.rocket(150, N=50)))
    c = cases.rocket(150, N=30, linear=False)
    c["cones"] = dict(Acu=e["Acu"], qcu=e["qcu"], cu=e["cu"], Acx=[], qcx=[], cx=[])       # Julia order: lands on the state side only (Q1)
    variants.append(("cone on one side", c))
    c = cases.rocket(150, N=30, linear=False)
    c["cones"] = dict(Acu=[3], qcu=[3], cu=[0.4], Acx=e["Acx"], qcx=e["qcx"], cx=e["cx"])  # state-side cone on elements 3..5
    variants.append(("cone at element 3", c))
    c = cases.rocket(150, N=30)
    Ax, bx, Au, bu = c["linear"]
    c["linear"] = (np.asfortranarray(np.vstack([Ax, -0.5 * Ax + 0.1])), np.array([bx[0], 40.0]), Au, bu)
    variants.append(("two state half-spaces", c))
    c = cases.rocket(150, N=30)
    rng = np.random.default_rng(2)
    c["Xref"] = np.asfortranarray(c["Xref"][:, :, None] + 0.05 * rng.standard_normal((6, 30, 150)))
    c["Uref"] = np.asfortranarray(np.repeat(c["Uref"][:, :, None], 150, axis=2))
    c["ref_mode"] = cases.oracle.REF_PER_INSTANCE
    variants.append(("per-instance references", c))
    c = cases.rocket(150, N=30)
    c["bounds"] = None
    variants.append(("no box", c))
    for what, c in variants:
        c["max_iter"] = 40
        want = cases.run_checker(c, orc, nthreads=8)
        a = cases.run_cuda(c, kernel=CONE)
        assert a["solver"].kernel == CONE, what
        b = cases.run_cuda(c, kernel=THREAD)
        cases.compare(a, want, what=f"cone kernel, {what}")
        for k in ("ret", "iters", "X", "U"):
            assert np.array_equal(a[k], b[k]), f"{what}: {k} differs between the cone and the thread kernel"


def test_warp_kernel_is_what_small_batches_run_on(orc):
    """TMPC_KERNEL_AUTO: the latency kernel (one warp per instance) up to 2048 instances, the throughput kernels beyond; a
    long horizon (shared memory sized by N), per-instance bounds and references, and a warm-started sequence on it."""
    for B, want_k in ((1, WARP), (2048, WARP), (2049, TMEM)):
        sol = cases.make_cuda_solver(cases.quadrotor_batch(B), kernel=t.KERNEL_AUTO)
        assert sol.kernel == want_k, (B, sol.kernel)
    for B, want_k in ((4096, WARP), (4097, TMEM)):                                  # cartpole shape: light blocks, higher limit
        sol = cases.make_cuda_solver(cases.cartpole_batch(B), kernel=t.KERNEL_AUTO)
        assert sol.kernel == want_k, (B, sol.kernel)
    sol = cases.make_cuda_solver(cases.rocket(4, N=10), kernel=t.KERNEL_AUTO)      # cones: not this kernel
    assert sol.kernel != WARP
    for N in (2, 3, 45, 120):
        c = cases.quadrotor_batch(40, hard=True, seed=N, N=N, max_iter=30)
        cases.compare(cases.run_cuda(c, kernel=WARP), cases.run_checker(c, orc), what=f"warp kernel N={N}")
    c = cases.quadrotor_per_instance(37)
    cases.compare(cases.run_cuda(c, kernel=WARP), cases.run_checker(c, orc), what="warp kernel, per-instance bounds and references")


def test_cuda_ragged_batch_sizes(orc):
    """B not a multiple of the tile (8) / warp (32) / block: tail lanes must stay inert."""
    for B in (1, 2, 7, 9, 31, 33, 130):
        c = cases.quadrotor_batch(B, hard=True, seed=100 + B, max_iter=40)
        want = cases.run_checker(c, orc)
        for k in kernels_for(c):
            cases.compare(cases.run_cuda(c, kernel=k), want, what=f"B={B} kernel {k}")


def test_cuda_closed_loop_warm_start(orc):
    """config 5 pattern: carried duals/slacks/cost terms must reproduce the reference's iteration counts."""
    c = cases.quadrotor_closed_loop(64, 60)
    want = cases.run_checker(c, orc, nthreads=8)
    for k in kernels_for(c):
        got = cases.run_cuda(c, kernel=k)
        cases.compare(got, want, what=f"closed loop kernel {k}")
        np.testing.assert_allclose(got["x_final"], want["x_final"], rtol=1e-5, atol=1e-6)


def test_kernels_interchangeable_mid_rollout(orc):
    """The carried state has one HBM format: switching kernels between solves must not change results."""
    c = cases.quadrotor_closed_loop(24, 12)
    want = cases.run_checker(c, orc)
    sol = cases.make_cuda_solver(c)
    order = kernels_for(c)
    if len(order) < 2:
        pytest.skip("only one kernel available for this shape")
    import oracle
    x = np.asfortranarray(c["x0"]).copy(order="F")
    s, N = c["sys"], c["N"]
    iters = np.zeros((c["nsteps"], x.shape[1]), dtype=np.int32)
    for k in range(c["nsteps"]):
        sol.set_kernel(order[k % len(order)])
        t.set_x0(sol, x)
        t.set_x_ref(sol, np.asfortranarray(c["Xref"][:, k:k + N, :]))
        t.solve(sol)
        iters[k] = sol.status()[1]
        U = t.get_solution(sol).controls
        x = np.asfortranarray(cases._plant(s.A, s.B, s.f, x, U[:, 0, :]))
    assert np.array_equal(iters, want["iters"])


def test_cuda_check_termination_and_max_iter(orc):
    c = cases.quadrotor_batch(100, hard=True)
    for chk, mi in ((3, 50), (1, 1), (7, 20)):
        c["check_termination"], c["max_iter"] = chk, mi
        want = cases.run_checker(c, orc)
        for k in kernels_for(c):
            cases.compare(cases.run_cuda(c, kernel=k), want, what=f"check={chk} max_iter={mi} kernel {k}")


def test_cuda_unconstrained_and_flag_subsets(orc):
    """Q3 of SURVEY.md: with the box flags off the primal state residual is identically zero."""
    c = cases.quadrotor_batch(64)
    c["bounds"] = None
    want = cases.run_checker(c, orc)
    for k in kernels_for(c):
        cases.compare(cases.run_cuda(c, kernel=k), want, what=f"unconstrained kernel {k}")
    # input bounds only (state bounds infinite), as tests/test_basic.jl:47-70 does
    c = cases.quadrotor_batch(64, hard=True)
    xm, xM, um, uM = c["bounds"]
    c["bounds"] = (np.full_like(xm, -np.inf), np.full_like(xM, np.inf), um, uM)
    want = cases.run_checker(c, orc)
    for k in kernels_for(c):
        cases.compare(cases.run_cuda(c, kernel=k), want, what=f"input-bounds-only kernel {k}")


def test_cuda_set_cache_terms_override(orc):
    """set_cache_terms replaces Kinf/Pinf/Quu_inv/AmBKt only."""
    c = cases.cartpole_batch(32)
    p = cases.to_oracle_problem(c)
    k0 = orc.get_cache(p)
    c["cache"] = {k: k0[k] * (1.0 + 1e-3) for k in ("Kinf", "Pinf", "Quu_inv", "AmBKt")}
    want = cases.run_checker(c, orc)
    for k in kernels_for(c):
        cases.compare(cases.run_cuda(c, kernel=k), want, what=f"cache override kernel {k}")


def test_full_size_batch_properties():
    """BASELINE.json config 3 at full size (B = 65,536): properties that do not need the oracle —
    bounds respected exactly by the returned slacks (Q4), x(:,0) = clamp(x0), determinism across
    kernels, iteration counts within [1, max_iter], converged instances report status 0."""
    B = 65536
    c = cases.quadrotor_batch(B)
    res = {}
    for k in kernels_for(c):
        got = cases.run_cuda(c, kernel=k)
        X, U = got["X"], got["U"]
        assert np.all(U <= 0.5) and np.all(U >= -0.5) and np.all(np.abs(X) <= 5.0)
        assert np.array_equal(X[:, 0, :], np.clip(c["x0"], -5, 5))
        it, st = got["iters"][0], got["ret"][0]
        assert it.min() >= 1 and it.max() <= 100
        assert np.all((st == 0) | (it == 100))
        assert (st == 0).mean() > 0.99
        res[k] = got
    a = res[THREAD]
    for k, b in res.items():
        assert np.array_equal(a["iters"], b["iters"]) and np.array_equal(a["ret"], b["ret"]), f"kernel {k}"
        np.testing.assert_allclose(a["X"], b["X"], rtol=1e-5, atol=1e-6)
        np.testing.assert_allclose(a["U"], b["U"], rtol=1e-5, atol=1e-6)


def test_full_size_parity_against_compiled_reference(ref, orc):
    """BASELINE.json's full sizes inside pytest: every instance of a B = 65,536 batch of configs 2 and 3 (AUTO kernel)
    against the compiled, unmodified reference core (oracle/_ref, the checker of the fixtures; the C restatement where
    it did not travel) — identical status and iteration count per instance, states / controls within 1e-5 / 1e-6 —
    and the N = 50 rocket with both cones and half-spaces (config 4) at B = 4,096 on the on-chip cone kernel."""
    chk = ref if ref is not None else orc
    for c, kern in ((cases.quadrotor_batch(65536, seed=4), TMEM), (cases.quadrotor_batch(65536, hard=True, seed=6), TMEM),
                    (cases.cartpole_batch(65536, seed=3), TMEM), (cases.rocket(4096, N=50, seed=5), CONE)):
        want = cases.run_checker(c, chk, nthreads=16)
        got = cases.run_cuda(c, kernel=t.KERNEL_AUTO)
        assert got["solver"].kernel == kern
        cases.compare(got, want, what=f"full size vs {chk.kind} ({c['sys'].name}, B={c['x0'].shape[1]})")


def test_full_size_closed_loop_against_compiled_reference(ref, orc):
    """config 5 at a size the CPU finishes in seconds: 2,048 instances x 100 warm-started MPC steps with per-instance
    sliding references, ONE fused launch, per-step status / iteration counts and the final plant state against the
    reference's closed loop."""
    import torch
    chk = ref if ref is not None else orc
    B, nsteps = 2048, 100
    c = cases.quadrotor_closed_loop(B, nsteps, seed=8)
    want = cases.run_checker(c, chk, nthreads=16)
    sol = cases.make_cuda_solver(c)
    t.set_x0(sol, c["x0"])
    xlong = torch.from_numpy(np.ascontiguousarray(np.asfortranarray(c["Xref"]).transpose(2, 1, 0))).cuda()
    st_out = torch.full((nsteps, B), -1, dtype=torch.int32, device="cuda")
    it_out = torch.full((nsteps, B), -1, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    sol.rollout_async(nsteps, xlong.data_ptr(), st_out.data_ptr(), it_out.data_ptr(), 0)
    sol.synchronize()
    assert sol.launch_count <= 3
    assert np.array_equal(it_out.cpu().numpy(), want["iters"]) and np.array_equal(st_out.cpu().numpy(), want["ret"])
    import ctypes as C
    from tinympc_julia_b200 import _lib
    xf = torch.empty((B, 12), dtype=torch.float64, device="cuda")
    assert _lib.load().tmpc_batch_get_x0(sol._h, C.cast(xf.data_ptr(), C.c_void_p), 1) == 0
    np.testing.assert_allclose(xf.cpu().numpy().T, want["x_final"], rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("B,nsteps", [(48, 30), (9, 7)])
def test_device_rollout_matches_oracle(orc, B, nsteps):
    """tmpc_batch_rollout_async (config 5 on the device: solve -> apply u0 -> plant step -> slide the reference window,
    nothing leaves the GPU between steps) against the oracle's closed loop: per-step iteration counts and status,
    final plant state, last solution."""
    import torch
    c = cases.quadrotor_closed_loop(B, nsteps)
    want = cases.run_checker(c, orc, nthreads=4)
    s, N = c["sys"], c["N"]
    for k in kernels_for(c):
        sol = cases.make_cuda_solver(c, kernel=k)
        t.set_x0(sol, c["x0"])
        # (nx, N+nsteps-1, B) Fortran == (B, N+nsteps-1, nx) C-contiguous
        xlong = torch.from_numpy(np.ascontiguousarray(np.asfortranarray(c["Xref"]).transpose(2, 1, 0))).cuda()
        st_out = torch.full((nsteps, B), -1, dtype=torch.int32, device="cuda")
        it_out = torch.full((nsteps, B), -1, dtype=torch.int32, device="cuda")
        torch.cuda.synchronize()
        sol.rollout_async(nsteps, xlong.data_ptr(), st_out.data_ptr(), it_out.data_ptr(), 0)
        sol.synchronize()
        assert np.array_equal(it_out.cpu().numpy(), want["iters"]), f"kernel {k}: per-step iteration counts"
        assert np.array_equal(st_out.cpu().numpy(), want["ret"]), f"kernel {k}: per-step status"
        X, U = t.get_solution(sol)
        np.testing.assert_allclose(X, want["X"], rtol=1e-5, atol=1e-6)
        np.testing.assert_allclose(U, want["U"], rtol=1e-5, atol=1e-6)
        xf = torch.empty((B, s.nx), dtype=torch.float64, device="cuda")
        import ctypes as C
        from tinympc_julia_b200 import _lib
        rc = _lib.load().tmpc_batch_get_x0(sol._h, C.cast(xf.data_ptr(), C.c_void_p), 1)
        assert rc == 0
        np.testing.assert_allclose(xf.cpu().numpy().T, want["x_final"], rtol=1e-5, atol=1e-6)


def test_solve_host_pipelined_matches_stepwise(orc):
    """tmpc_batch_solve_host (x0 up, chunked solve, solution down in one call) == set_x0 / reset / solve / get_*:
    bit-identical results cold and warm, with the batch large enough to be cut into chunks."""
    B = 148 * 64 * 3 + 37
    c = cases.quadrotor_batch(B, hard=True, max_iter=25)
    ref = cases.make_cuda_solver(c)
    pip = cases.make_cuda_solver(c)
    rng = np.random.default_rng(5)
    for step, cold in enumerate((True, False, False)):
        x0 = np.asfortranarray(c["x0"] * (1.0 + 0.1 * step) + 0.01 * rng.standard_normal(c["x0"].shape))
        t.set_x0(ref, x0)
        if cold:
            ref.reset()
        r0 = t.solve(ref)
        sol0 = t.get_solution(ref)
        st0, it0 = ref.status()
        r1, sol1, st1, it1 = pip.solve_host(x0, cold=cold, chunks=0 if step < 2 else 3)
        assert r0 == r1
        assert np.array_equal(st0, st1) and np.array_equal(it0, it1)
        assert np.array_equal(sol0.states, sol1.states) and np.array_equal(sol0.controls, sol1.controls)
    # the same with per-instance dynamics (each chunk reads its own slice of the operand blocks)
    base = problems_mod.quadrotor()
    A = np.asfortranarray(base.A[:, :, None] * (1.0 + 0.05 * rng.uniform(-1, 1, (12, 12, B))))
    Bm = np.asfortranarray(base.B[:, :, None] * (1.0 + 0.1 * rng.uniform(-1, 1, (12, 4, B))))
    for sol in (ref, pip):
        t.set_instance_dynamics(sol, A, Bm, None, base.Q, base.R)
        assert sol.kernel == TMEM
    for step, cold in enumerate((True, False)):
        x0 = np.asfortranarray(c["x0"] * (1.0 - 0.1 * step))
        t.set_x0(ref, x0)
        if cold:
            ref.reset()
        r0 = t.solve(ref)
        sol0 = t.get_solution(ref)
        st0, it0 = ref.status()
        r1, sol1, st1, it1 = pip.solve_host(x0, cold=cold, chunks=3)
        assert r0 == r1 and np.array_equal(st0, st1) and np.array_equal(it0, it1)
        assert np.array_equal(sol0.states, sol1.states) and np.array_equal(sol0.controls, sol1.controls)
    # small batch / other kernels: single chunk path
    c = cases.rocket(40, N=50)
    want = cases.run_checker(c, orc)
    sol = cases.make_cuda_solver(c)
    if c["Xref"] is not None:
        t.set_x_ref(sol, c["Xref"])
    if c["Uref"] is not None:
        t.set_u_ref(sol, c["Uref"])
    r, s1, st1, it1 = sol.solve_host(c["x0"], cold=True)
    assert np.array_equal(it1, want["iters"][0]) and np.array_equal(st1, want["ret"][0])
    np.testing.assert_allclose(s1.controls, want["U"], rtol=1e-5, atol=1e-6)


def _device_rollout(sol, nsteps, xlong=None, fused=True):
    import ctypes as C
    import torch
    from tinympc_julia_b200 import _lib
    B = sol.B
    assert _lib.load().tmpc_batch_set_rollout_fused(sol._h, 1 if fused else 0) == 0
    st_out = torch.full((nsteps, B), -1, dtype=torch.int32, device="cuda")
    it_out = torch.full((nsteps, B), -1, dtype=torch.int32, device="cuda")
    xl = None
    if xlong is not None:     # (nx, cols, B) Fortran == (B, cols, nx) C-contiguous
        xl = torch.from_numpy(np.ascontiguousarray(np.asfortranarray(xlong).transpose(2, 1, 0))).cuda()
    torch.cuda.synchronize()
    sol.rollout_async(nsteps, xl.data_ptr() if xl is not None else 0, st_out.data_ptr(), it_out.data_ptr(), 0)
    sol.synchronize()
    xf = torch.empty((B, sol.nx), dtype=torch.float64, device="cuda")
    assert _lib.load().tmpc_batch_get_x0(sol._h, C.cast(xf.data_ptr(), C.c_void_p), 1) == 0
    return st_out.cpu().numpy(), it_out.cpu().numpy(), xf.cpu().numpy().T


def test_fused_rollout_constant_references(orc):
    """Fused closed loop without a sliding window: no references (cartpole, quadrotor hover) and shared references;
    fused and launch-sequence forms against the oracle's closed loop."""
    for c, nsteps in ((cases.cartpole_batch(40), 12), (cases.quadrotor_batch(33, hard=True, max_iter=30), 9),
                      (cases.GOLDEN_CASES["quadrotor_hover"](), 5)):
        c["nsteps"] = nsteps
        want = cases.run_checker(c, orc, nthreads=4)
        for fused in (True, False):
            sol = cases.make_cuda_solver(c, kernel=TMEM)
            if c["ref_mode"] == 1:
                if c["Xref"] is not None:
                    t.set_x_ref(sol, c["Xref"])
                if c["Uref"] is not None:
                    t.set_u_ref(sol, c["Uref"])
            t.set_x0(sol, c["x0"])
            st, it, xf = _device_rollout(sol, nsteps, fused=fused)
            assert np.array_equal(it, want["iters"]) and np.array_equal(st, want["ret"]), f"fused={fused}"
            np.testing.assert_allclose(xf, want["x_final"], rtol=1e-5, atol=1e-6)
            X, U = t.get_solution(sol)
            np.testing.assert_allclose(U, want["U"], rtol=1e-5, atol=1e-6)


def test_fused_rollout_split_in_two_launches(orc):
    """A rollout continued by a second call: the instances leave the chip after the first launch (write-out with the
    lagging slacks of converged solves) and are warm-installed by the second.  Includes solves that stop at max_iter."""
    n1, n2, B = 6, 8, 24
    c = cases.quadrotor_closed_loop(B, n1 + n2)
    c["x0"] = np.asfortranarray(c["x0"] * 4.0)
    c["max_iter"] = 12
    want = cases.run_checker(c, orc, nthreads=4)
    assert (want["ret"] != 0).any() and (want["ret"] == 0).any()
    N = c["N"]
    sol = cases.make_cuda_solver(c, kernel=TMEM)
    t.set_x0(sol, c["x0"])
    st1, it1, _ = _device_rollout(sol, n1, xlong=c["Xref"][:, :N + n1 - 1, :])
    st2, it2, xf = _device_rollout(sol, n2, xlong=c["Xref"][:, n1:n1 + N + n2 - 1, :])
    assert np.array_equal(np.vstack([it1, it2]), want["iters"])
    assert np.array_equal(np.vstack([st1, st2]), want["ret"])
    np.testing.assert_allclose(xf, want["x_final"], rtol=1e-5, atol=1e-6)


def test_per_instance_dynamics(orc):
    """B different systems in one batch.  The device setup kernel must reproduce the host cache bit
    for bit (same operation order), and every instance must solve exactly like a single-instance reference solver set up
    with that instance's A, B, f, Q, R."""
    import ctypes as C
    import dataclasses
    from tinympc_julia_b200 import _lib
    lib = _lib.load()
    for base, B, scale in ((problems_mod.quadrotor(), 9, 0.3), (problems_mod.cartpole(), 7, 0.5)):
        rng = np.random.default_rng(17)
        n, m, N = base.nx, base.nu, 20 if base.nx == 12 else 15
        A = np.stack([base.A * (1.0 + 0.05 * rng.uniform(-1, 1, base.A.shape)) for _ in range(B)], axis=2)
        Bm = np.stack([base.B * (1.0 + 0.2 * rng.uniform(-1, 1, base.B.shape)) for _ in range(B)], axis=2)
        f = 0.01 * rng.uniform(-1, 1, (n, B))
        Q = np.stack([np.diag(np.diag(base.Q) * (1.0 + 0.3 * rng.uniform(-1, 1, n))) for _ in range(B)], axis=2)
        R = np.stack([np.diag(np.diag(base.R) * (1.0 + 0.3 * rng.uniform(-1, 1, m))) for _ in range(B)], axis=2)
        x0 = np.asfortranarray(scale * rng.uniform(-1, 1, (n, B)))
        c = cases._base(base, N, bounds=base.bounds(N), x0=x0, max_iter=60)
        sol = cases.make_cuda_solver(c)
        t.set_instance_dynamics(sol, A, Bm, f, Q, R)
        assert sol.kernel == TMEM                 # AUTO: the on-chip kernel reading per-instance blocks
        res = {}
        for k in (TMEM, THREAD):
            sol.set_kernel(k)
            assert sol.kernel == k
            sol.reset()
            t.set_x0(sol, x0)
            t.solve(sol)
            res[k] = (sol.status(), t.get_solution(sol))
            # warm-started second solve from a nearby state: carried state per instance
            t.set_x0(sol, np.asfortranarray(0.9 * x0))
            t.solve(sol)
            res[k] += (sol.status(), t.get_solution(sol))
        for a, bb in zip(res[TMEM], res[THREAD]):
            assert np.array_equal(a[0], bb[0]) and np.array_equal(a[1], bb[1]), "TMEM and thread kernels disagree"
        (st, it), (X, U) = res[TMEM][0], res[TMEM][1]
        for b in range(B):
            sb = dataclasses.replace(base, A=np.asfortranarray(A[:, :, b]), B=np.asfortranarray(Bm[:, :, b]),
                                     f=f[:, b].copy(), Q=np.asfortranarray(Q[:, :, b]), R=np.asfortranarray(R[:, :, b]))
            cb = cases._base(sb, N, bounds=sb.bounds(N), x0=np.asfortranarray(x0[:, b:b + 1]), max_iter=60)
            want = cases.run_checker(cb, orc)
            assert it[b] == want["iters"][0, 0] and st[b] == want["ret"][0, 0], f"instance {b}"
            np.testing.assert_allclose(X[:, :, b], want["X"][:, :, 0], rtol=1e-5, atol=1e-6)
            np.testing.assert_allclose(U[:, :, b], want["U"][:, :, 0], rtol=1e-5, atol=1e-6)
            # device setup kernel == host setup (tmpc_compute_cache), bit for bit
            dev = [np.zeros((m, n), order="F"), np.zeros((n, n), order="F"), np.zeros((m, m), order="F"),
                   np.zeros((n, n), order="F"), np.zeros(n), np.zeros(m)]
            dp = lambda a: a.ctypes.data_as(C.POINTER(C.c_double))
            assert lib.tmpc_batch_get_instance_cache(sol._h, b, *[dp(a) for a in dev]) == 0
            host = [np.zeros_like(a) for a in dev] + [np.zeros(n), np.zeros(m)]
            assert lib.tmpc_compute_cache(dp(sb.A), dp(sb.B), dp(np.ascontiguousarray(sb.f)), dp(sb.Q), dp(sb.R), base.rho, n, m,
                                          *[dp(a) for a in host]) == 0
            for d, hh in zip(dev, host):
                assert np.array_equal(d, hh), f"instance {b}: device cache differs from the host cache"
    # warm start keeps working per instance
    t.set_x0(sol, np.asfortranarray(0.9 * x0))
    assert t.solve(sol) in (0, 1)


def test_per_instance_dynamics_closed_loop(orc):
    """Closed loop over B different systems (domain-randomised plant = the instance's own A, B, f): the fused rollout of
    the TMEM kernel's PI variant, the launch sequence (TMEM and thread kernels) and, per instance, the oracle's closed
    loop of a single-instance problem with that instance's matrices."""
    import dataclasses
    base = problems_mod.quadrotor()
    B, N, nsteps = 7, 20, 6
    rng = np.random.default_rng(23)
    n, m = base.nx, base.nu
    A = np.stack([base.A * (1.0 + 0.05 * rng.uniform(-1, 1, base.A.shape)) for _ in range(B)], axis=2)
    Bm = np.stack([base.B * (1.0 + 0.2 * rng.uniform(-1, 1, base.B.shape)) for _ in range(B)], axis=2)
    f = 0.005 * rng.uniform(-1, 1, (n, B))
    x0 = np.asfortranarray(0.3 * rng.uniform(-1, 1, (n, B)))
    want = []
    for b in range(B):
        sb = dataclasses.replace(base, A=np.asfortranarray(A[:, :, b]), B=np.asfortranarray(Bm[:, :, b]), f=f[:, b].copy())
        cb = cases._base(sb, N, bounds=sb.bounds(N), x0=np.asfortranarray(x0[:, b:b + 1]), max_iter=25, nsteps=nsteps)
        want.append(cases.run_checker(cb, orc))
    w_it = np.hstack([w["iters"] for w in want]); w_st = np.hstack([w["ret"] for w in want])
    w_xf = np.hstack([w["x_final"] for w in want])
    assert (w_st == 0).any()
    c = cases._base(base, N, bounds=base.bounds(N), x0=x0, max_iter=25)
    for kernel, fused in ((TMEM, True), (TMEM, False), (THREAD, False)):
        sol = cases.make_cuda_solver(c)
        t.set_instance_dynamics(sol, A, Bm, f, base.Q, base.R)
        sol.set_kernel(kernel)
        assert sol.kernel == kernel
        t.set_x0(sol, x0)
        n0 = sol.launch_count
        st, it, xf = _device_rollout(sol, nsteps, fused=fused)
        if fused:
            assert sol.launch_count - n0 <= 2, "the fused closed loop is one solve launch (+ at most a state conversion)"
        assert np.array_equal(it, w_it) and np.array_equal(st, w_st), f"kernel {kernel} fused={fused}"
        np.testing.assert_allclose(xf, w_xf, rtol=1e-5, atol=1e-6)


# ------------------------------------------------------------------------------------------ float build
# `typedef float tinytype` (reference types.hpp:15).  Checker: the reference itself compiled with the
# float typedef (oracle/_ref/libtinympc_refbatch32.so) — committed outputs in tests/golden/f32_*.npz, the live library
# where oracle/_ref travelled to this box.  Tolerance: cases.compare_f32.

def kernels32_for(c):
    """Float build: the thread kernel always, the shared-memory sub-warp kernel and the TMEM kernel where they apply."""
    sol = cases.make_cuda_solver(c, kernel=t.KERNEL_AUTO, precision=32)
    ks = [THREAD]
    for k in (SUBWARP, TMEM):
        sol.set_kernel(k)
        if sol.kernel == k:
            ks.append(k)
    del sol
    return ks


@pytest.mark.parametrize("name", cases.GOLDEN_F32_CASES)
def test_float_build_matches_float_reference_fixture(name):
    c, want = cases.golden_f32_case(name)
    outs = {}
    for kern in kernels32_for(c):
        got = outs[kern] = cases.run_cuda(c, kernel=kern, precision=32)
        for k in ("X", "U"):
            assert np.array_equal(got[k], got[k].astype(np.float32).astype(np.float64)), "results must be floats widened to double"
        cases.compare_f32(got, want, what=f"cuda float[kernel {kern}] vs float reference[{name}]")
    for kern in (SUBWARP, TMEM):      # same sequence of rounded operations: the float kernels agree bit for bit
        if kern in outs:
            for k in ("ret", "iters", "X", "U"):
                assert np.array_equal(outs[THREAD][k], outs[kern][k]), f"float thread vs kernel {kern}: {k} differs"


def test_float_build_seeded_batches_against_live_float_reference():
    import oracle
    ref32 = oracle.load_ref32()
    if ref32 is None:
        pytest.skip("oracle/_ref/libtinympc_refbatch32.so did not travel to this box")
    for make, B in ((cases.cartpole_batch, 3000), (cases.quadrotor_batch, 3000),
                    (lambda B: cases.quadrotor_batch(B, hard=True), 2000), (lambda B: cases.rocket(B, N=50), 200)):
        c = make(B)
        cache = ref32.get_cache(cases.to_oracle_problem(c))
        c["cache"] = {k: cache[k] for k in ("Kinf", "Pinf", "Quu_inv", "AmBKt")}
        want = cases.run_checker(c, ref32, nthreads=8)
        for kern in kernels32_for(c):
            got = cases.run_cuda(c, kernel=kern, precision=32)
            cases.compare_f32(got, want, what=f"cuda float[kernel {kern}] vs live float reference (B={B})")


def test_float_build_warm_start_and_guards():
    """A second solve warm-starts from the float carried state, whichever float kernel wrote it (the TMEM kernel keeps it
    as instance-major float blocks, the others element-major: converted on the way); the cone kernel and per-instance
    dynamics are 64-bit only and say so; switching precision needs a reset batch."""
    c, _ = cases.golden_f32_case("quadrotor_easy_64")
    ref_sol = cases.make_cuda_solver(c, kernel=THREAD, precision=32)      # three solves on the thread kernel alone
    ref_runs = [cases.run_cuda(c, sol=ref_sol) for _ in range(3)]
    sol = cases.make_cuda_solver(c, kernel=t.KERNEL_AUTO, precision=32)
    assert sol.precision == 32 and sol.kernel == TMEM      # AUTO: the on-chip float kernel where it applies
    first = cases.run_cuda(c, sol=sol)
    sol.set_kernel(THREAD)
    again = cases.run_cuda(c, sol=sol)                      # same x0, warm start from the other kernel's state
    assert again["iters"].max() <= first["iters"].min()
    sol.set_kernel(SUBWARP)
    third = cases.run_cuda(c, sol=sol)
    for got, want in ((first, ref_runs[0]), (again, ref_runs[1]), (third, ref_runs[2])):
        for k in ("ret", "iters", "X", "U"):
            assert np.array_equal(got[k], want[k]), f"warm start across float kernels: {k} differs"
    sol.set_kernel(TMEM)
    fourth = cases.run_cuda(c, sol=sol)                     # and back onto the TMEM kernel (element-major -> blocks)
    fourth_ref = cases.run_cuda(c, sol=ref_sol)
    for k in ("ret", "iters", "X", "U"):
        assert np.array_equal(fourth[k], fourth_ref[k]), f"warm start back on the float TMEM kernel: {k} differs"
    with pytest.raises(t.TinyMPCError):
        sol.set_precision(64)                               # carries state
    sol.set_kernel(CONE)
    assert sol.kernel == -1                                 # not this shape, and 64-bit only
    assert t.solve(sol) not in (0, 1)                       # refused, not silently run on another kernel
    sol.set_kernel(THREAD)
    sol.reset()
    sol.set_precision(64)
    sol.set_kernel(t.KERNEL_AUTO)
    back = cases.run_cuda(cases.GOLDEN_CASES["quadrotor_easy_64"](), sol=sol)
    cases.compare(back, cases.load_golden("quadrotor_easy_64"), what="double again after set_precision(64)")

