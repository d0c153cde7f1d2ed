"""The sixteen drop-in symbols (include/tinympc_b200.h PART 1 == reference src/bindings.cpp:21-490), VALUE-checked on the GPU.

Two libraries are loaded side by side with the same ctypes prototypes:
  * tinympc_julia_b200/lib/libtinympc_jl.so   — this repository (CUDA, B = 1 singleton),
  * oracle/_ref/libtinympc_jl.so              — the unmodified reference bindings + core compiled here (oracle/Makefile),
and every scenario is replayed, call for call, the way src/TinyMPC.jl drives them (`setup` = setup_solver followed by
update_settings with every en_* = 0, src/TinyMPC.jl:77-104).  Return codes, get_states / get_controls (1e-5 rel /
1e-6 abs) and the iteration count (`Solution: iter=..` of print_problem_data, the only place the reference shows it;
get_solve_info on our side) must agree.  Scenarios follow the reference's own tests and examples:
tests/test_basic.jl:24-70, tests/test_settings.jl:17-92, tests/test_cache.jl:18-50,
examples/cartpole_example_one_solve.jl, examples/cartpole_example_mpc.jl:35-51,
examples/rocket_landing_constraints.jl:17-68, and a second setup_solver replacing the first.
"""
import ctypes as C
import os
import re
import sys
import tempfile

import numpy as np
import pytest

import cases
import oracle
from tinympc_julia_b200 import _lib, problems

pytestmark = pytest.mark.gpu

RTOL, ATOL = cases.RTOL, cases.ATOL
P = lambda a: a.ctypes.data_as(C.POINTER(C.c_double))
IP = lambda a: a.ctypes.data_as(C.POINTER(C.c_int))
F = lambda a: np.asfortranarray(np.asarray(a, dtype=np.float64))

PART1 = ["setup_solver", "set_x0", "set_x_ref", "set_u_ref", "solve_mpc", "get_states", "get_controls", "cleanup_solver",
         "codegen", "codegen_with_sensitivity", "print_problem_data", "set_cache_terms", "update_settings",
         "set_bound_constraints", "set_linear_constraints", "set_cone_constraints"]


class capture_stdout:
    """fd-level capture: both libraries print from C (printf / std::cout)."""

    def __enter__(self):
        sys.stdout.flush()
        self.saved = os.dup(1)
        self.tmp = tempfile.TemporaryFile()
        os.dup2(self.tmp.fileno(), 1)
        return self

    def __exit__(self, *exc):
        libc = C.CDLL(None)
        libc.fflush(None)
        os.dup2(self.saved, 1)
        os.close(self.saved)
        self.tmp.seek(0)
        self.text = self.tmp.read().decode(errors="replace")
        self.tmp.close()


class JL:
    """What src/TinyMPC.jl does with one library (same call sequence, same argument marshalling)."""

    def __init__(self, path):
        self.lib = C.CDLL(path)
        for name in PART1:
            fn = getattr(self.lib, name)
            fn.restype, fn.argtypes = _lib.SIGNATURES[name]
        self.n = self.m = self.N = 0

    def setup(self, A, B, f, Q, R, rho, n, m, N, abs_pri_tol=1e-3, abs_dua_tol=1e-3, max_iter=100, check_termination=True,
              adaptive_rho=False, adaptive_rho_min=0.1, adaptive_rho_max=10.0):
        A, B, f, Q, R = F(A), F(np.asarray(B).reshape(n, m)), F(np.asarray(f).reshape(n, 1)), F(Q), F(R)
        self.n, self.m, self.N = n, m, N
        rc = self.lib.setup_solver(P(A), n, n, P(B), n, m, P(f), n, 1, P(Q), n, n, P(R), m, m, float(rho), n, m, N, 0)
        if rc != 0:
            return rc
        # src/TinyMPC.jl:89-104
        return self.update_settings(abs_pri_tol=abs_pri_tol, abs_dua_tol=abs_dua_tol, max_iter=max_iter,
                                    check_termination=check_termination, adaptive_rho=adaptive_rho,
                                    adaptive_rho_min=adaptive_rho_min, adaptive_rho_max=adaptive_rho_max)

    def update_settings(self, abs_pri_tol=1e-3, abs_dua_tol=1e-3, max_iter=100, check_termination=True, en=(0,) * 6,
                        adaptive_rho=False, adaptive_rho_min=0.1, adaptive_rho_max=10.0, clipping=True):
        return self.lib.update_settings(abs_pri_tol, abs_dua_tol, max_iter, int(check_termination), *[int(e) for e in en],
                                        int(adaptive_rho), adaptive_rho_min, adaptive_rho_max, int(clipping), 0)

    def set_x0(self, x0):
        a = F(np.asarray(x0).reshape(-1, 1))
        return self.lib.set_x0(P(a), a.shape[0], 1, 0)

    def set_x_ref(self, X):
        a = F(X)
        return self.lib.set_x_ref(P(a), a.shape[0], a.shape[1], 0)

    def set_u_ref(self, U):
        a = F(U)
        return self.lib.set_u_ref(P(a), a.shape[0], a.shape[1], 0)

    def set_bounds(self, xm, xM, um, uM):
        xm, xM, um, uM = F(xm), F(xM), F(um), F(uM)
        return self.lib.set_bound_constraints(P(xm), *xm.shape, P(xM), *xM.shape, P(um), *um.shape, P(uM), *uM.shape, 0)

    def set_cones(self, Acu, qcu, cu, Acx, qcx, cx):
        i32 = lambda a: np.ascontiguousarray(np.asarray(a, dtype=np.int32))
        Acu, qcu, Acx, qcx = i32(Acu), i32(qcu), i32(Acx), i32(qcx)
        cu, cx = np.ascontiguousarray(np.asarray(cu, dtype=np.float64)), np.ascontiguousarray(np.asarray(cx, dtype=np.float64))
        return self.lib.set_cone_constraints(IP(Acu), len(Acu), IP(qcu), len(qcu), P(cu), len(cu),
                                             IP(Acx), len(Acx), IP(qcx), len(qcx), P(cx), len(cx), 0)

    def set_linear(self, Ax, bx, Au, bu):
        Ax, Au = F(Ax), F(Au)
        bx, bu = np.ascontiguousarray(bx, dtype=np.float64), np.ascontiguousarray(bu, dtype=np.float64)
        return self.lib.set_linear_constraints(P(Ax), *Ax.shape, P(bx), len(bx), P(Au), *Au.shape, P(bu), len(bu), 0)

    def set_cache(self, K, Pm, Qi, Am):
        K, Pm, Qi, Am = F(K), F(Pm), F(Qi), F(Am)
        return self.lib.set_cache_terms(P(K), *K.shape, P(Pm), *Pm.shape, P(Qi), *Qi.shape, P(Am), *Am.shape, 0)

    def solve(self):
        return self.lib.solve_mpc(0)

    def solution(self):
        X, U = np.zeros((self.n, self.N), order="F"), np.zeros((self.m, self.N - 1), order="F")
        r, c = C.c_int(), C.c_int()
        assert self.lib.get_states(P(X), C.byref(r), C.byref(c)) == 0 and (r.value, c.value) == (self.n, self.N)
        assert self.lib.get_controls(P(U), C.byref(r), C.byref(c)) == 0 and (r.value, c.value) == (self.m, self.N - 1)
        return X, U

    def iters(self):
        """`Solution: iter=K, solved=S` as print_problem_data prints it (src/bindings.cpp:236-237)."""
        with capture_stdout() as cap:
            assert self.lib.print_problem_data(0) == 0
        m = re.search(r"iter=(\d+), solved=(\d+)", cap.text)
        assert m, cap.text
        return int(m.group(1)), int(m.group(2))


@pytest.fixture(scope="module")
def pair():
    if not os.path.exists(oracle.REFJL_SO):
        pytest.skip("oracle/_ref/libtinympc_jl.so not built (needs /root/reference at build time)")
    ours, ref = JL(_lib.LIB_PATH), JL(oracle.REFJL_SO)
    yield ours, ref
    with capture_stdout():
        ours.lib.cleanup_solver()
        ref.lib.cleanup_solver()


def both(pair, fn):
    """Run fn(lib) on both libraries (stdout silenced: the reference prints per converged solve); return both results."""
    with capture_stdout():
        a = fn(pair[0])
        b = fn(pair[1])
    return a, b


def assert_same_solution(pair, what):
    (Xa, Ua), (Xb, Ub) = pair[0].solution(), pair[1].solution()
    np.testing.assert_allclose(Xa, Xb, rtol=RTOL, atol=ATOL, err_msg=f"{what}: states")
    np.testing.assert_allclose(Ua, Ub, rtol=RTOL, atol=ATOL, err_msg=f"{what}: controls")
    ia, ib = pair[0].iters(), pair[1].iters()
    assert ia == ib, f"{what}: (iter, solved) {ia} vs reference {ib}"
    return Xa, Ua, ia


def cartpole():
    s = problems.cartpole()
    return s.A, s.B, np.zeros(4), s.Q, s.R


def test_one_solve_cartpole(pair):
    """examples/cartpole_example_one_solve.jl:11-27 — SURVEY.md §8c: iter 7, controls[1] = 1.1780759329."""
    A, B, f, Q, R = cartpole()
    N = 20
    assert both(pair, lambda j: j.setup(A, B, f, Q, R, 1.0, 4, 1, N, max_iter=10)) == (0, 0)
    assert both(pair, lambda j: j.set_x0([0.5, 0, 0, 0])) == (0, 0)
    assert both(pair, lambda j: j.solve()) == (0, 0)
    X, U, (it, solved) = assert_same_solution(pair, "one solve")
    assert (it, solved) == (7, 1)
    np.testing.assert_allclose(U[0, :5], [1.1780759329, 0.9569844550, 0.7578020204, 0.5784400825, 0.4170123365], atol=1e-9)
    np.testing.assert_allclose(U[0, 18], -0.6794357189, atol=1e-9)
    # our extra symbol agrees with what print_problem_data shows
    it2, sv2 = C.c_int(), C.c_int()
    ours = C.CDLL(_lib.LIB_PATH)
    assert ours.get_solve_info(C.byref(it2), C.byref(sv2)) == 0 and (it2.value, sv2.value) == (7, 1)


def test_basic_reference_setting_and_bounds(pair):
    """tests/test_basic.jl:24-70: references set explicitly; then input bounds u in [-1, 1] (iter 41)."""
    A, B, f, Q, R = cartpole()
    N = 10
    assert both(pair, lambda j: j.setup(A, B, f, Q, R, 1.0, 4, 1, N)) == (0, 0)
    assert both(pair, lambda j: (j.set_x0([0.5, 0, 0, 0]), j.set_x_ref(np.zeros((4, N))), j.set_u_ref(np.zeros((1, N - 1))))) == ((0, 0, 0),) * 2
    assert both(pair, lambda j: j.solve()) == (0, 0)
    assert_same_solution(pair, "test_basic reference setting")

    assert both(pair, lambda j: j.setup(A, B, f, Q, R, 1.0, 4, 1, N)) == (0, 0)          # fresh solver, as the testset does
    b = (np.full((4, N), -np.inf), np.full((4, N), np.inf), np.full((1, N - 1), -1.0), np.full((1, N - 1), 1.0))
    assert both(pair, lambda j: j.set_bounds(*b)) == (0, 0)
    assert both(pair, lambda j: (j.set_x0([1.0, 0, 0, 0]), j.set_x_ref(np.zeros((4, N))), j.set_u_ref(np.zeros((1, N - 1))))) == ((0, 0, 0),) * 2
    assert both(pair, lambda j: j.solve()) == (0, 0)
    X, U, (it, solved) = assert_same_solution(pair, "test_basic bounds")
    assert (it, solved) == (41, 1)
    assert U.min() >= -1.0 and U.max() <= 1.0
    np.testing.assert_allclose(U[0], [1, 1, 1, 1, 1, 1, 0.7331404, 0.43364709, 0.16293939], atol=1e-6)


def test_settings(pair):
    """tests/test_settings.jl:17-92: custom tolerance, the full settings call, adaptive-rho BINDING, max_iter = 1."""
    A, B, f, Q, R = cartpole()
    N = 2

    def run(what):
        assert both(pair, lambda j: (j.set_x0([0.1, 0, 0, 0]), j.set_x_ref(np.zeros((4, N))), j.set_u_ref(np.zeros((1, N - 1))))) == ((0, 0, 0),) * 2
        ra, rb = both(pair, lambda j: j.solve())
        assert ra == rb, f"{what}: solve returned {ra}, reference {rb}"
        assert_same_solution(pair, what)
        return ra

    assert both(pair, lambda j: j.setup(A, B, f, Q, R, 1.0, 4, 1, N, abs_pri_tol=5.0)) == (0, 0)
    assert run("abs_pri_tol = 5") == 0
    assert both(pair, lambda j: j.setup(A, B, f, Q, R, 1.0, 4, 1, N)) == (0, 0)
    assert both(pair, lambda j: j.update_settings(abs_pri_tol=1e-4, abs_dua_tol=1e-4, max_iter=50)) == (0, 0)
    assert run("all settings") == 0
    # "Adaptive Rho Settings": the reference test only covers the binding (setup must succeed); it never solves
    assert both(pair, lambda j: j.setup(A, B, f, Q, R, 1.0, 4, 1, N, adaptive_rho=True, adaptive_rho_min=0.5, adaptive_rho_max=5.0)) == (0, 0)
    assert both(pair, lambda j: j.setup(A, B, f, Q, R, 1.0, 4, 1, N, max_iter=1)) == (0, 0)
    assert run("max_iter = 1") >= 0


def test_cache_terms(pair):
    """tests/test_cache.jl:18-50: arbitrary matrices are accepted; installing the computed cache leaves the solve unchanged."""
    A, B, f, Q, R = cartpole()
    N = 10
    rng = np.random.default_rng(0)
    assert both(pair, lambda j: j.setup(A, B, f, Q, R, 1.0, 4, 1, N)) == (0, 0)
    K, Pm, Qi, Am = rng.random((1, 4)), rng.random((4, 4)), rng.random((1, 1)), rng.random((4, 4))
    assert both(pair, lambda j: j.set_cache(K, Pm, Qi, Am)) == (0, 0)
    # the overridden (nonsense) cache drives both solvers through the same iterations (Q11: APf, BPf, rho untouched)
    assert both(pair, lambda j: j.update_settings(max_iter=5)) == (0, 0)
    assert both(pair, lambda j: j.set_x0([0.2, 0, 0, 0])) == (0, 0)
    ra, rb = both(pair, lambda j: j.solve())
    assert ra == rb
    assert_same_solution(pair, "overridden cache")
    # computed cache (what TinyMPC.solve_lqr would hand back): taken from the reference core itself
    ref = oracle.load_ref()
    c = ref.get_cache(oracle.Problem(A=A, B=B, Q=Q, R=R, rho=1.0, N=N, f=f))
    assert both(pair, lambda j: j.setup(A, B, f, Q, R, 1.0, 4, 1, N)) == (0, 0)
    assert both(pair, lambda j: j.set_cache(c["Kinf"], c["Pinf"], c["Quu_inv"], c["AmBKt"])) == (0, 0)
    assert both(pair, lambda j: j.set_x0([0.5, 0, 0, 0])) == (0, 0)
    assert both(pair, lambda j: j.solve()) == (0, 0)
    assert_same_solution(pair, "computed cache")


def test_rocket_cones_through_julia_order(pair):
    """examples/rocket_landing_constraints.jl:17-68 through set_cone_constraints (Q1 crossing): iter 37."""
    c = cases.rocket_julia_example()
    s, N = c["sys"], c["N"]
    assert both(pair, lambda j: j.setup(s.A, s.B, s.f, s.Q, s.R, s.rho, 6, 3, N, abs_pri_tol=2e-3, abs_dua_tol=1e-3)) == (0, 0)
    assert both(pair, lambda j: j.set_bounds(*c["bounds"])) == (0, 0)
    k = c["cones"]
    assert both(pair, lambda j: j.set_cones(k["Acu"], k["qcu"], k["cu"], k["Acx"], k["qcx"], k["cx"])) == (0, 0)
    assert both(pair, lambda j: (j.set_x0(c["x0"]), j.set_x_ref(c["Xref"]), j.set_u_ref(c["Uref"]))) == ((0, 0, 0),) * 2
    assert both(pair, lambda j: j.solve()) == (0, 0)
    X, U, (it, solved) = assert_same_solution(pair, "rocket example")
    assert (it, solved) == (37, 1)
    np.testing.assert_allclose(U[:, 0], [12.34232183, -10.0, 65.2407144], rtol=1e-7)


def test_linear_constraints_binding(pair):
    """set_linear_constraints through the reference's symbol (src/bindings.cpp:414-450), random 5x2 system."""
    c = cases.random_system(5, 2, 8, 1)
    s, N = c["sys"], c["N"]
    assert both(pair, lambda j: j.setup(s.A, s.B, s.f, s.Q, s.R, s.rho, 5, 2, N, max_iter=c["max_iter"])) == (0, 0)
    assert both(pair, lambda j: j.set_bounds(*c["bounds"])) == (0, 0)
    assert both(pair, lambda j: j.set_linear(*c["linear"])) == (0, 0)
    assert both(pair, lambda j: (j.set_x0(c["x0"][:, 0]), j.set_x_ref(c["Xref"]), j.set_u_ref(c["Uref"]))) == ((0, 0, 0),) * 2
    ra, rb = both(pair, lambda j: j.solve())
    assert ra == rb
    assert_same_solution(pair, "linear constraints")


def test_warm_started_mpc_loop(pair):
    """examples/cartpole_example_mpc.jl:35-51, 50 steps: solve -> controls[1] -> x = A x + B u -> set_x0, nothing reset."""
    A, B, f, Q, R = cartpole()
    N = 20
    assert both(pair, lambda j: j.setup(A, B, f, Q, R, 1.0, 4, 1, N, max_iter=10)) == (0, 0)
    assert both(pair, lambda j: (j.set_x0([0.5, 0, 0, 0]), j.set_x_ref(np.zeros((4, N))), j.set_u_ref(np.zeros((1, N - 1))))) == ((0, 0, 0),) * 2
    xa = xb = np.array([0.5, 0, 0, 0])
    for k in range(50):
        ra, rb = both(pair, lambda j: j.solve())
        assert ra == rb, f"step {k}: solve returned {ra}, reference {rb}"
        (Xa, Ua), (Xb, Ub) = pair[0].solution(), pair[1].solution()
        np.testing.assert_allclose(Ua, Ub, rtol=RTOL, atol=ATOL, err_msg=f"step {k}")
        np.testing.assert_allclose(Xa, Xb, rtol=RTOL, atol=ATOL, err_msg=f"step {k}")
        assert pair[0].iters() == pair[1].iters(), f"step {k}"
        xa = A @ xa + (np.asarray(B).reshape(4, 1) * Ua[0, 0]).ravel()
        xb = A @ xb + (np.asarray(B).reshape(4, 1) * Ub[0, 0]).ravel()
        with capture_stdout():
            assert pair[0].set_x0(xa) == 0 and pair[1].set_x0(xb) == 0
    np.testing.assert_allclose(xa, xb, rtol=1e-5, atol=1e-6)


def test_second_setup_replaces_first(pair):
    """Q9: one solver per process; a second setup_solver (another shape) replaces the first and starts from zeros."""
    A, B, f, Q, R = cartpole()
    assert both(pair, lambda j: j.setup(A, B, f, Q, R, 1.0, 4, 1, 20, max_iter=10)) == (0, 0)
    assert both(pair, lambda j: j.set_x0([0.5, 0, 0, 0])) == (0, 0)
    assert both(pair, lambda j: j.solve()) == (0, 0)
    s = problems.quadrotor()
    N = 20
    assert both(pair, lambda j: j.setup(s.A, s.B, s.f, s.Q, s.R, s.rho, 12, 4, N)) == (0, 0)
    assert both(pair, lambda j: j.set_bounds(*s.bounds(N))) == (0, 0)
    x0 = problems.quadrotor_batch_x0(1, hard=False, seed=3)[:, 0]
    assert both(pair, lambda j: j.set_x0(x0)) == (0, 0)
    ra, rb = both(pair, lambda j: j.solve())
    assert ra == rb == 0
    X, U, (it, solved) = assert_same_solution(pair, "quadrotor after cartpole")
    assert X.shape == (12, N) and solved == 1
    # dimension errors leave the solver in place and return the core's status 1 on both sides (tiny_api.cpp:53-66)
    bad = both(pair, lambda j: j.lib.setup_solver(P(F(s.A)), 12, 12, P(F(s.B)), 12, 4, P(F(s.f.reshape(-1, 1))), 12, 1,
                                                   P(F(s.Q)), 12, 12, P(F(s.R)), 4, 4, 5.0, 11, 4, N, 0))
    assert bad[0] != 0 and bad[1] != 0
