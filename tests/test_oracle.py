"""Pins the plain-C oracle (oracle/tinympc_oracle.c) — CPU only.

(a) against the values quoted in SURVEY.md §8c, (b) against the committed golden fixtures
(tests/golden/*.npz = outputs of the compiled reference), (c) against the compiled reference itself
when oracle/_ref is present.  Bar: status and iteration count identical, states/controls within the
double-precision tolerance north_star states (1e-5 relative / 1e-6 absolute).
"""
import numpy as np
import pytest

import cases


@pytest.mark.parametrize("name", sorted(cases.SURVEY_GOLDEN))
def test_oracle_reproduces_survey_values(orc, name):
    g = cases.SURVEY_GOLDEN[name]
    out = cases.run_checker(cases.GOLDEN_CASES[name](), orc)
    assert int(out["ret"][0, 0]) == g["ret"]
    assert int(out["iters"][0, 0]) == g["iter"]
    U = out["U"][:, :, 0]
    if "u_first" in g:
        np.testing.assert_allclose(U[0, :5], g["u_first"], rtol=0, atol=1e-9)
        np.testing.assert_allclose(U[0, 18], g["u_last"], rtol=0, atol=1e-9)
    if "controls" in g:
        np.testing.assert_allclose(U[0, :], g["controls"], rtol=0, atol=1e-7)
    if "u0" in g:
        np.testing.assert_allclose(U[:, 0], g["u0"], rtol=0, atol=1e-7)


@pytest.mark.parametrize("name", sorted(cases.GOLDEN_CASES))
def test_oracle_matches_golden_fixture(orc, name):
    want = cases.load_golden(name)
    got = cases.run_checker(cases.GOLDEN_CASES[name](), orc)
    cases.compare(got, want, what=f"oracle vs golden[{name}]")
    np.testing.assert_allclose(got["x_final"], want["x_final"], rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("name", sorted(cases.GOLDEN_CASES))
def test_oracle_matches_compiled_reference(orc, ref, name):
    if ref is None:
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    c = cases.GOLDEN_CASES[name]()
    cases.compare(cases.run_checker(c, orc), cases.run_checker(c, ref), what=f"oracle vs reference[{name}]")


def test_oracle_cache_matches_reference(orc, ref):
    if ref is None:
        pytest.skip("oracle/_ref not built")
    for make in (cases.cartpole_one_solve, cases.quadrotor_hover_golden, cases.rocket_julia_example):
        p = cases.to_oracle_problem(make())
        a, b = orc.get_cache(p), ref.get_cache(p)
        for k in a:
            np.testing.assert_allclose(a[k], b[k], rtol=1e-9, atol=1e-11 * max(1.0, np.abs(b[k]).max()), err_msg=k)


def test_oracle_multithreaded_equals_single(orc):
    c = cases.quadrotor_batch(48, hard=True)
    a, b = cases.run_checker(c, orc, nthreads=1), cases.run_checker(c, orc, nthreads=4)
    assert np.array_equal(a["iters"], b["iters"]) and np.array_equal(a["X"], b["X"]) and np.array_equal(a["U"], b["U"])


def test_reference_rejects_zero_check_termination(orc):
    c = cases.cartpole_one_solve()
    c["check_termination"] = 0
    with pytest.raises(RuntimeError):
        cases.run_checker(c, orc)


# ------------------------------------------------------------------ adaptive rho

def test_sensitivity_matrices_like_the_julia_package():
    """tests/test_sensitivity.jl:18-85 on the host-side mirror of compute_sensitivity_autograd (src/TinyMPC.jl:301-351):
    shapes, finite, deterministic, different at another rho; and the forward difference is a derivative (it agrees with
    a central difference of the same fixed point to first order in the step)."""
    from tinympc_julia_b200 import lqr_sensitivity, problems
    s = problems.cartpole()
    A, B, Q, R = s.A, s.B.reshape(4, 1), s.Q, s.R
    dK, dP, dC1, dC2 = lqr_sensitivity(A, B, Q, R, 1.0)
    assert dK.shape == (1, 4) and dP.shape == (4, 4) and dC1.shape == (1, 1) and dC2.shape == (4, 4)
    assert all(np.isfinite(m).all() for m in (dK, dP, dC1, dC2))
    again = lqr_sensitivity(A, B, Q, R, 1.0)
    assert all(np.array_equal(a, b) for a, b in zip((dK, dP, dC1, dC2), again))
    other = lqr_sensitivity(A, B, Q, R, 2.0)
    assert all(not np.allclose(a, b) for a, b in zip((dK, dP, dC1, dC2), other))
    from tinympc_julia_b200.solver import _solve_lqr
    h = 1e-4
    central = [(y - x) / (2 * h) for x, y in zip(_solve_lqr(A, B, Q, R, 1.0 - h), _solve_lqr(A, B, Q, R, 1.0 + h))]
    for a, b in zip((dK, dP, dC1, dC2), central):
        np.testing.assert_allclose(a, b, rtol=2e-3, atol=1e-6 * max(1.0, np.abs(b).max()))


def test_adaptive_branch_changes_rho_and_matches_reference_cache(orc, ref):
    """The adapted cache (rho, Kinf, Pinf of every instance after the solve) of the oracle against the compiled reference's;
    the branch is live (rho leaves its setup value) and stays inside the clip interval."""
    c = cases.with_adaptive_rho(cases.quadrotor_batch(96, seed=17))
    a = cases.run_checker(c, orc)
    rho0 = c["sys"].rho
    assert (a["adapted"]["rho"] != rho0).any()
    assert a["adapted"]["rho"].min() >= 0.1 * rho0 - 1e-12 and a["adapted"]["rho"].max() <= 10 * rho0 + 1e-12
    short = a["iters"][0] <= 5                                 # solves that stop before iteration 6 never adapt
    assert np.all(a["adapted"]["rho"][short] == rho0)
    if ref is None:
        pytest.skip("oracle/_ref not built")
    b = cases.run_checker(c, ref)
    for k in ("rho", "Kinf", "Pinf"):
        np.testing.assert_allclose(a["adapted"][k], b["adapted"][k], rtol=1e-9, atol=1e-11, err_msg=k)


def test_adaptive_without_sensitivity_is_rejected(orc, ref):
    """0 x 0 sensitivity matrices: the reference crashes (rho_benchmark.cpp:204-207); both checkers refuse instead."""
    c = cases.with_adaptive_rho(cases.cartpole_batch(4))
    p = cases.to_oracle_problem(c)
    p.sensitivity = None
    for lib in (orc, ref):
        if lib is None:
            continue
        with pytest.raises(RuntimeError):
            lib.batch_run(p, c["x0"])
