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
