"""The float build (`typedef float tinytype`, reference types.hpp:15) — CPU part.

The checker here is the REFERENCE ITSELF compiled with the float typedef (oracle/Makefile target ref32 ->
oracle/_ref/libtinympc_refbatch32.so); its outputs are committed as tests/golden/f32_*.npz by
tests/golden/make_golden.py.  These tests pin the fixtures to the live float reference where it is built, bound the
distance between the two builds of the reference, and check the library's float setup (tmpc_compute_cache_f32)
against the float reference's cache.  The GPU part is tests/test_gpu_parity.py::test_float_*.
"""
import ctypes as C

import numpy as np
import pytest

import cases
import oracle
from tinympc_julia_b200 import _lib


@pytest.fixture(scope="module")
def ref32():
    return oracle.load_ref32()


@pytest.mark.parametrize("name", cases.GOLDEN_F32_CASES)
def test_float_fixture_matches_live_float_reference(ref32, name):
    if ref32 is None:
        pytest.skip("oracle/_ref/libtinympc_refbatch32.so not built (needs /root/reference)")
    want = cases.load_golden("f32_" + name)
    got = cases.run_checker(cases.GOLDEN_CASES[name](), ref32)
    assert np.array_equal(got["ret"], want["ret"]) and np.array_equal(got["iters"], want["iters"])
    assert np.array_equal(got["X"], want["X"]) and np.array_equal(got["U"], want["U"])


@pytest.mark.parametrize("name", cases.GOLDEN_F32_CASES)
def test_float_fixture_values_are_floats_near_the_double_build(name):
    """Every number the float build returns is a float widened to double, and it stays near the double build's
    answer where both stopped at the same iteration (the two builds solve the same problem)."""
    f32, f64 = cases.load_golden("f32_" + name), cases.load_golden(name)
    for k in ("X", "U"):
        assert np.array_equal(f32[k], f32[k].astype(np.float32).astype(np.float64)), k
    same = (f32["iters"][-1] == f64["iters"][-1]) & (f32["ret"][-1] == f64["ret"][-1])
    assert same.mean() >= 0.9, f"{name}: only {same.mean():.2f} of the instances stop at the same iteration in both builds"
    for k in ("X", "U"):
        d = np.abs(f32[k][..., same] - f64[k][..., same])
        scale = 1.0 + np.abs(f64[k][..., same])
        assert (d / scale).max() < 2e-3, f"{name}: {k} differs by {(d / scale).max():.2e} between the builds"


@pytest.mark.parametrize("name", ["cartpole_one_solve", "quadrotor_hover", "rocket_julia_example", "random_5x2_16"])
def test_library_float_cache_close_to_float_reference_cache(name):
    """tmpc_compute_cache_f32 (host setup in float) against the cache the float reference computed (in the fixture).
    Both are float Riccati recursions stopped at max|dKinf| < 1e-5 with different summation orders: they agree to
    the recursion's own accuracy, not bit for bit (which is why parity tests install the reference's cache)."""
    lib = _lib.load()
    c = cases.GOLDEN_CASES[name]()
    s = c["sys"]
    n, m = s.nx, s.nu
    f64 = lambda a: np.asfortranarray(np.asarray(a, dtype=np.float64))
    A, B, f, Q, R = [f64(a) for a in (s.A, s.B, s.f, s.Q, s.R)]
    out = {k: np.zeros(sh, order="F") for k, sh in dict(Kinf=(m, n), Pinf=(n, n), Quu_inv=(m, m), AmBKt=(n, n), APf=(n,),
                                                         BPf=(m,), Qd=(n,), Rd=(m,)).items()}
    P = lambda a: a.ctypes.data_as(_lib.c_double_p)
    rc = lib.tmpc_compute_cache_f32(P(A), P(B), P(f), P(Q), P(R), s.rho, n, m,
                                    *[P(out[k]) for k in ("Kinf", "Pinf", "Quu_inv", "AmBKt", "APf", "BPf", "Qd", "Rd")])
    assert rc == 0
    g = cases.load_golden("f32_" + name)
    for k in ("Kinf", "Pinf", "Quu_inv", "AmBKt", "APf", "BPf"):
        want = g["cache_" + k].reshape(out[k].shape, order="F")
        assert np.array_equal(out[k], out[k].astype(np.float32).astype(np.float64)), f"{k} is not float-valued"
        np.testing.assert_allclose(out[k], want, rtol=2e-3, atol=2e-3 * max(1e-3, np.abs(want).max()), err_msg=k)
    np.testing.assert_array_equal(out["Qd"], (np.diag(s.Q).astype(np.float32) + np.float32(s.rho)).astype(np.float64))
