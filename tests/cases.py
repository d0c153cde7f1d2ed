"""Shared test workloads: one description, three executors (oracle port, compiled reference, CUDA).

A case is a plain dict so that the same seeded inputs reach
  * ``oracle.load_oracle()``  — the plain-C restatement,
  * ``oracle.load_ref()``     — the unmodified reference core (when oracle/_ref is built),
  * the CUDA library through the Julia-mirror API in ``tinympc_julia_b200.solver``.
Workloads follow SURVEY.md §8d (systems, seeds, distributions).
"""
from __future__ import annotations

import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import oracle  # noqa: E402
from tinympc_julia_b200 import problems  # noqa: E402

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

# tolerance north_star states for tinytype = double
RTOL, ATOL = 1e-5, 1e-6
# ... and for tinytype = float
RTOL_F32, ATOL_F32 = 1e-4, 1e-5


def _base(sys_, N, **kw):
    c = dict(sys=sys_, N=N, abs_pri_tol=1e-3, abs_dua_tol=1e-3, max_iter=100, check_termination=1,
             bounds=None, bounds_per_instance=False, cones=None, linear=None,
             x0=None, Xref=None, Uref=None, ref_mode=oracle.REF_ZERO, nsteps=1, cache=None, adaptive=None)
    c.update(kw)
    return c


def cartpole_one_solve():
    """config 1: examples/cartpole_example_one_solve.jl:11-27."""
    return _base(problems.cartpole(), 20, max_iter=10, x0=np.array([[0.5], [0.0], [0.0], [0.0]]))


def cartpole_test_basic_bounds():
    """tests/test_basic.jl:47-70: N=10, u in [-1,1], x unbounded, x0=[1,0,0,0]."""
    s = problems.cartpole()
    N = 10
    b = (np.full((4, N), -np.inf), np.full((4, N), np.inf), np.full((1, N - 1), -1.0), np.full((1, N - 1), 1.0))
    return _base(s, N, bounds=b, x0=np.array([[1.0], [0.0], [0.0], [0.0]]))


def cartpole_batch(B, seed=2):
    """config 2."""
    s = problems.cartpole()
    return _base(s, 20, bounds=s.bounds(20), x0=problems.cartpole_batch_x0(B, seed))


def quadrotor_batch(B, hard=False, seed=3, N=20, max_iter=100):
    """config 3."""
    s = problems.quadrotor()
    return _base(s, N, bounds=s.bounds(N), x0=problems.quadrotor_batch_x0(B, hard, seed), max_iter=max_iter)


def quadrotor_hover_golden():
    """Xref z = 2 for all knots, x0 of quadrotor_hovering.cpp:61, saturates."""
    s = problems.quadrotor()
    N = 20
    Xref = np.zeros((12, N), order="F")
    Xref[2, :] = 2.0
    x0 = np.array([0, 1, 0, 0.2, 0, 0, 0.1, 0, 0, 0, 0, 0], dtype=np.float64).reshape(12, 1)
    return _base(s, N, bounds=s.bounds(N), x0=x0, Xref=Xref, Uref=np.zeros((4, N - 1), order="F"), ref_mode=oracle.REF_SHARED)


def quadrotor_per_instance(B, seed=11):
    """per-instance references AND per-instance bounds (the TMA-per-instance style inputs)."""
    s = problems.quadrotor()
    N = 20
    rng = np.random.default_rng(seed)
    x0 = np.asfortranarray(rng.uniform(-0.2, 0.2, size=(12, B)))
    Xref = np.zeros((12, N, B), order="F")
    Xref[0:3, :, :] = rng.uniform(-0.3, 0.3, size=(3, 1, B))
    Uref = np.asfortranarray(rng.uniform(-0.05, 0.05, size=(4, N - 1, B)))
    umax = rng.uniform(0.2, 0.6, size=(4, 1, B)) * np.ones((4, N - 1, B))
    xmax = rng.uniform(0.5, 5.0, size=(12, 1, B)) * np.ones((12, N, B))
    b = tuple(np.asfortranarray(a) for a in (-xmax, xmax, -umax, umax))
    return _base(s, N, bounds=b, bounds_per_instance=True, x0=x0, Xref=Xref, Uref=Uref, ref_mode=oracle.REF_PER_INSTANCE)


def quadrotor_closed_loop(B, nsteps, seed=5):
    """config 5: warm-started rollout with per-instance sliding line references."""
    s = problems.quadrotor()
    N = 20
    x0, Xlong = problems.quadrotor_line_refs(B, N, nsteps, seed)
    return _base(s, N, bounds=s.bounds(N), x0=x0, Xref=Xlong, ref_mode=oracle.REF_SLIDING, nsteps=nsteps)


def with_adaptive_rho(c, lo=0.1, hi=10.0, clip=True):
    """the case solved with adaptive rho.  rho is kept inside [lo, hi] x the setup rho; the
    sensitivity matrices are the Julia package's own finite differences (src/TinyMPC.jl:301-351), loaded the way
    tests/test_sensitivity.jl:40-53 loads them (codegen_with_sensitivity stores them in the live cache)."""
    from tinympc_julia_b200 import lqr_sensitivity
    s = c["sys"]
    dK, dP, dC1, dC2 = lqr_sensitivity(s.A, s.B, s.Q, s.R, s.rho)
    c["adaptive"] = dict(rho_min=lo * s.rho, rho_max=hi * s.rho, clip=bool(clip), sens=dict(dK=dK, dP=dP, dC1=dC1, dC2=dC2))
    return c


def rocket(B, N=50, seed=4, linear=True, tol_pri=2e-3):
    """config 4: box + both cones (Julia argument order, so SURVEY.md Q1 applies) + synthetic half-spaces."""
    s = problems.rocket()
    x0, Xref, Uref = problems.rocket_batch(B, N, seed)
    e = s.extra
    cones = dict(Acu=e["Acu"], qcu=e["qcu"], cu=e["cu"], Acx=e["Acx"], qcx=e["qcx"], cx=e["cx"])
    lin = problems.rocket_linear_constraints() if linear else None
    return _base(s, N, bounds=s.bounds(N), cones=cones, linear=lin, x0=x0, Xref=Xref, Uref=Uref,
                 ref_mode=oracle.REF_SHARED, abs_pri_tol=tol_pri)


def rocket_julia_example():
    """examples/rocket_landing_constraints.jl:17-68 — N=10, both cones, tol 2e-3 (golden: iter 37)."""
    s = problems.rocket()
    N = 10
    xinit = s.extra["xinit"]
    x0 = (1.1 * xinit).reshape(6, 1)
    Xref = np.zeros((6, N), order="F")
    for i in range(N):
        Xref[:, i] = xinit + (0.0 - xinit) * i / 99.0   # NTOTAL - 1, as the example does
    Uref = np.zeros((3, N - 1), order="F")
    Uref[2, :] = 10.0
    e = s.extra
    cones = dict(Acu=e["Acu"], qcu=e["qcu"], cu=e["cu"], Acx=e["Acx"], qcx=e["qcx"], cx=e["cx"])
    return _base(s, N, bounds=s.bounds(N), cones=cones, x0=x0, Xref=Xref, Uref=Uref, ref_mode=oracle.REF_SHARED,
                 abs_pri_tol=2e-3, abs_dua_tol=1e-3)


def random_system(nx, nu, N, B, seed=7, with_linear=True):
    """A stable-ish random system of arbitrary shape: exercises the runtime-dimension kernel."""
    rng = np.random.default_rng(seed)
    A = np.eye(nx) + 0.05 * rng.standard_normal((nx, nx))
    Bm = 0.1 * rng.standard_normal((nx, nu))
    f = 0.01 * rng.standard_normal(nx)
    s = problems.System(f"rand{nx}x{nu}", A, Bm, f, np.diag(rng.uniform(1, 10, nx)), np.diag(rng.uniform(0.5, 2, nu)), 1.5,
                        x_min=np.full(nx, -3.0), x_max=np.full(nx, 3.0), u_min=np.full(nu, -0.8), u_max=np.full(nu, 0.8))
    x0 = np.asfortranarray(rng.uniform(-1, 1, size=(nx, B)))
    Xref = np.asfortranarray(0.1 * rng.standard_normal((nx, N)))
    Uref = np.asfortranarray(0.01 * rng.standard_normal((nu, N - 1)))
    lin = None
    if with_linear:
        Ax = np.asfortranarray(rng.standard_normal((2, nx)))
        Au = np.asfortranarray(rng.standard_normal((1, nu)))
        lin = (Ax, np.array([0.5, 0.8]), Au, np.array([0.3]))
    return _base(s, N, bounds=s.bounds(N), linear=lin, x0=x0, Xref=Xref, Uref=Uref, ref_mode=oracle.REF_SHARED, max_iter=60)


# ------------------------------------------------------------------------------------------ executors

def to_oracle_problem(c) -> oracle.Problem:
    s = c["sys"]
    p = oracle.Problem(A=s.A, B=s.B, Q=s.Q, R=s.R, rho=s.rho, N=c["N"], f=s.f,
                       abs_pri_tol=c["abs_pri_tol"], abs_dua_tol=c["abs_dua_tol"], max_iter=c["max_iter"],
                       check_termination=c["check_termination"], cache=c["cache"])
    if c.get("adaptive"):
        ad = c["adaptive"]
        p.adaptive_rho, p.adaptive_rho_min, p.adaptive_rho_max = 1, ad["rho_min"], ad["rho_max"]
        p.adaptive_rho_enable_clipping, p.sensitivity = int(ad["clip"]), ad["sens"]
    if c["bounds"] is not None:
        p.with_bounds(*c["bounds"], per_instance=c["bounds_per_instance"])
    if c["cones"] is not None:
        k = c["cones"]
        # Julia argument order -> core slots: first triple (the caller's "u") into the STATE slots
        p.Acx, p.qcx, p.cx = k["Acu"], k["qcu"], k["cu"]
        p.Acu, p.qcu, p.cu = k["Acx"], k["qcx"], k["cx"]
        p.en_state_soc = int(len(k["Acx"]) > 0)
        p.en_input_soc = int(len(k["Acu"]) > 0)
    if c["linear"] is not None:
        Ax, bx, Au, bu = c["linear"]
        p.Alin_x, p.blin_x, p.Alin_u, p.blin_u = Ax, bx, Au, bu
        p.en_state_linear = int(np.size(bx) > 0)
        p.en_input_linear = int(np.size(bu) > 0)
    return p


def run_checker(c, lib: oracle.BatchLib, nthreads=1):
    return lib.batch_run(to_oracle_problem(c), c["x0"], c["Xref"], c["Uref"], c["ref_mode"], c["nsteps"], nthreads)


def make_cuda_solver(c, kernel=0, device=-1, precision=64):
    """Drive the CUDA library through the Julia-mirror API, in the order a Julia user would."""
    import tinympc_julia_b200 as t
    s = c["sys"]
    B = c["x0"].shape[1]
    sol = t.TinyMPCBatchSolver()
    ad = c.get("adaptive")
    adkw = dict(adaptive_rho=True, adaptive_rho_min=ad["rho_min"], adaptive_rho_max=ad["rho_max"],
                adaptive_rho_clipping=ad["clip"]) if ad else {}
    t.setup(sol, s.A, s.B, s.f, s.Q, s.R, s.rho, s.nx, s.nu, c["N"], batch=B, device=device,
            abs_pri_tol=c["abs_pri_tol"], abs_dua_tol=c["abs_dua_tol"], max_iter=c["max_iter"],
            check_termination=bool(c["check_termination"]), **adkw)
    if ad:
        v = ad["sens"]
        t.set_sensitivity(sol, v["dK"], v["dP"], v["dC1"], v["dC2"])
    if precision != 64:
        sol.set_precision(precision)
    if c["cache"] is not None:
        k = c["cache"]
        t.set_cache_terms(sol, k["Kinf"], k["Pinf"], k["Quu_inv"], k["AmBKt"])
    if c["bounds"] is not None:
        t.set_bound_constraints(sol, *c["bounds"])
    if c["cones"] is not None:
        k = c["cones"]
        t.set_cone_constraints(sol, k["Acu"], k["qcu"], k["cu"], k["Acx"], k["qcx"], k["cx"])
    if c["linear"] is not None:
        t.set_linear_constraints(sol, *c["linear"])
    if c["check_termination"] not in (0, 1):
        # the Julia API only carries a Bool (src/TinyMPC.jl:185); integer periods go through the C ABI directly
        import ctypes as C
        from tinympc_julia_b200 import _lib
        lib = _lib.load()
        tols, ints = (C.c_double * 2)(), (C.c_int * 8)()
        assert lib.tmpc_batch_get_settings(sol._h, tols, ints) == 0
        assert lib.tmpc_batch_update_settings(sol._h, tols[0], tols[1], ints[0], int(c["check_termination"]), ints[2], ints[3],
                                              ints[4], ints[5], ints[6], ints[7], 0, 0.1, 10.0, 1) == 0
    sol.set_kernel(kernel)
    return sol


def run_cuda(c, kernel=0, sol=None, precision=64):
    """Returns the same dict layout as the checkers (ret, iters [nsteps,B], X, U, x_final)."""
    import tinympc_julia_b200 as t
    s = c["sys"]
    nx, nu, N = s.nx, s.nu, c["N"]
    x0 = np.asfortranarray(c["x0"])
    B = x0.shape[1]
    nsteps = c["nsteps"]
    if sol is None:
        sol = make_cuda_solver(c, kernel, precision=precision)
    mode = c["ref_mode"]
    if mode in (oracle.REF_SHARED, oracle.REF_PER_INSTANCE):
        if c["Xref"] is not None:
            t.set_x_ref(sol, c["Xref"])
        if c["Uref"] is not None:
            t.set_u_ref(sol, c["Uref"])
    elif mode == oracle.REF_SLIDING and c["Uref"] is not None:
        t.set_u_ref(sol, c["Uref"])
    ret = np.zeros((nsteps, B), dtype=np.int32)
    iters = np.zeros((nsteps, B), dtype=np.int32)
    x = x0.copy(order="F")
    A, Bm, f = s.A, s.B, s.f
    for k in range(nsteps):
        t.set_x0(sol, x)
        if mode == oracle.REF_SLIDING and c["Xref"] is not None:
            t.set_x_ref(sol, np.asfortranarray(c["Xref"][:, k:k + N, :]))
        rc = t.solve(sol)
        assert rc in (0, 1), f"solve returned {rc}"
        st, it = sol.status()
        ret[k], iters[k] = st, it
        X, U = t.get_solution(sol)
        # host plant step in the reference's operation order: (A x + B u) + f per row, ascending-k sums
        x = np.asfortranarray(_plant(A, Bm, f, x, U[:, 0, :]))
    return {"ret": ret, "iters": iters, "X": X, "U": U, "x_final": x, "solver": sol}


def _plant(A, Bm, f, x, u):
    s1 = np.zeros_like(x)
    for k in range(A.shape[1]):
        s1 = s1 + A[:, k:k + 1] * x[k:k + 1, :]
    s2 = np.zeros_like(x)
    for k in range(Bm.shape[1]):
        s2 = s2 + Bm[:, k:k + 1] * u[k:k + 1, :]
    return (s1 + s2) + f.reshape(-1, 1)


def compare(got, want, what="", rtol=RTOL, atol=ATOL, allow_iter_mismatch=0):
    """status + iteration count identical per instance; states/controls within tolerance."""
    nbad_ret = int((got["ret"] != want["ret"]).sum())
    nbad_it = int((got["iters"] != want["iters"]).sum())
    assert nbad_ret <= allow_iter_mismatch, f"{what}: {nbad_ret} status mismatches"
    assert nbad_it <= allow_iter_mismatch, f"{what}: {nbad_it} iteration-count mismatches " \
        f"(first at {np.argwhere(got['iters'] != want['iters'])[:5].tolist()})"
    same = (got["iters"][-1] == want["iters"][-1]) & (got["ret"][-1] == want["ret"][-1])
    for key in ("X", "U"):
        g, w = got[key][..., same], want[key][..., same]
        err = np.abs(g - w) - (atol + rtol * np.abs(w))
        assert err.size == 0 or err.max() <= 0, f"{what}: {key} out of tolerance by {err.max():.3e}"
    return nbad_ret, nbad_it


def golden_f32_case(name):
    """(case with the float reference's cache installed, fixture) for one of GOLDEN_F32_CASES."""
    g = load_golden("f32_" + name)
    c = GOLDEN_CASES[name]()
    n, m = c["sys"].nx, c["sys"].nu
    c["cache"] = {"Kinf": g["cache_Kinf"].reshape(m, n, order="F"), "Pinf": g["cache_Pinf"].reshape(n, n, order="F"),
                  "Quu_inv": g["cache_Quu_inv"].reshape(m, m, order="F"), "AmBKt": g["cache_AmBKt"].reshape(n, n, order="F")}
    return c, g


def compare_f32(got, want, what="", max_flip_frac=0.02):
    """Float build.  The CUDA kernel and the float reference round differently inside the mat-vecs (Eigen sums float
    packets of 8, the kernel one ascending FMA chain), so an instance whose residual lands within a few float ulps of
    the tolerance can stop one iteration apart: at most `max_flip_frac` of the instances (and at least one allowed),
    each by at most 2 iterations.  Instances that stop at the same iteration: states / controls within north_star's
    float tolerance (1e-4 relative / 1e-5 absolute) when the solve converged within 20 iterations; within 10x that
    for longer solves and for those that ran into max_iter (the two roundings differ by a float ulp per operation and
    the difference compounds over every iteration: measured on B200, quadrotor x0 ~ U(+-0.5), a solve that converges
    after 40-90 iterations ends up to 1.2e-5 away in absolute terms).  Returns (flips, worst excess over the tight
    tolerance among the short converged solves)."""
    gi, wi = got["iters"][-1], want["iters"][-1]
    same = (gi == wi) & (got["ret"][-1] == want["ret"][-1])
    nflip = int((~same).sum())
    assert nflip <= max(1, int(max_flip_frac * same.size)), f"{what}: {nflip} of {same.size} instances stop at another iteration"
    assert np.abs(gi - wi).max() <= 2, f"{what}: iteration counts differ by {np.abs(gi - wi).max()}"
    conv = same & (want["ret"][-1] == 0) & (wi <= 20)
    unconv = same & ~conv
    worst = 0.0
    for key in ("X", "U"):
        for sel, k in ((conv, 1.0), (unconv, 10.0)):
            g, w = got[key][..., sel], want[key][..., sel]
            if g.size == 0:
                continue
            err = np.abs(g - w) - k * (ATOL_F32 + RTOL_F32 * np.abs(w))
            assert err.max() <= 0, f"{what}: {key} out of the float tolerance (x{k:g}) by {err.max():.3e}"
            if k == 1.0:
                worst = max(worst, float(err.max()))
    return nflip, worst


# ------------------------------------------------------------------------------------------ golden fixtures
# name -> zero-argument constructor.  tests/golden/make_golden.py runs the COMPILED REFERENCE on each and
# stores its outputs in tests/golden/<name>.npz; the tests re-create the inputs from the same seeds.
GOLDEN_CASES = {
    "cartpole_one_solve": cartpole_one_solve,
    "cartpole_test_basic_bounds": cartpole_test_basic_bounds,
    "cartpole_batch_64": lambda: cartpole_batch(64),
    "quadrotor_hover": quadrotor_hover_golden,
    "quadrotor_easy_64": lambda: quadrotor_batch(64),
    "quadrotor_hard_64": lambda: quadrotor_batch(64, hard=True),
    "quadrotor_per_instance_32": lambda: quadrotor_per_instance(32),
    "quadrotor_closed_loop_16x40": lambda: quadrotor_closed_loop(16, 40),
    "rocket_julia_example": rocket_julia_example,
    "rocket_soc_linear_16": lambda: rocket(16),
    "rocket_soc_16_N20": lambda: rocket(16, N=20, linear=False),
    "random_5x2_16": lambda: random_system(5, 2, 8, 16),
    "random_9x3_16": lambda: random_system(9, 3, 12, 16, seed=9, with_linear=False),
    # adaptive rho: fixtures also hold the adapted rho of every instance
    "adaptive_quadrotor_easy_64": lambda: with_adaptive_rho(quadrotor_batch(64)),
    "adaptive_quadrotor_hard_64": lambda: with_adaptive_rho(quadrotor_batch(64, hard=True)),
    "adaptive_cartpole_64": lambda: with_adaptive_rho(cartpole_batch(64)),
    "adaptive_quadrotor_closed_loop_16x12": lambda: with_adaptive_rho(quadrotor_closed_loop(16, 12)),
    "adaptive_rocket_soc_linear_8": lambda: with_adaptive_rho(rocket(8, N=20)),
    "adaptive_random_5x2_16": lambda: with_adaptive_rho(random_system(5, 2, 8, 16), clip=False),
}

# values quoted in SURVEY.md §8c / BASELINE.md §2 (captured from the compiled reference during the survey)
SURVEY_GOLDEN = {
    "cartpole_one_solve": dict(ret=0, iter=7, u_first=[1.1780759329, 0.9569844550, 0.7578020204, 0.5784400825, 0.4170123365],
                               u_last=-0.6794357189),
    "cartpole_test_basic_bounds": dict(ret=0, iter=41, controls=[1, 1, 1, 1, 1, 1, 0.7331404, 0.43364709, 0.16293939]),
    "quadrotor_hover": dict(ret=1, iter=100, u0=[0.5, 0.5, 0.5, 0.5]),
    "rocket_julia_example": dict(ret=0, iter=37, u0=[12.34232183, -10.0, 65.2407144]),
}


# float build (`typedef float tinytype`): single cold solves of these cases through oracle/_ref/libtinympc_refbatch32.so;
# the fixtures (tests/golden/f32_<name>.npz) also hold the float reference's Riccati cache, which the tests install
# through set_cache_terms (the cache is only 1e-5-converged, so parity needs identical cache inputs)
GOLDEN_F32_CASES = ["cartpole_one_solve", "cartpole_test_basic_bounds", "cartpole_batch_64", "quadrotor_hover",
                    "quadrotor_easy_64", "quadrotor_hard_64", "rocket_julia_example", "rocket_soc_linear_16",
                    "rocket_soc_16_N20", "random_5x2_16", "random_9x3_16"]


def load_golden(name):
    path = os.path.join(GOLDEN_DIR, name + ".npz")
    with np.load(path) as z:
        return {k: z[k] for k in z.files}
