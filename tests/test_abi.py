"""C-ABI surface and host-side logic — CPU only (no compute calls, there is no GPU here).

* the library loads and exports every symbol include/tinympc_b200.h declares;
* the sixteen reference symbols (src/bindings.cpp) are among them, with the ctypes prototypes the
  Julia ccalls imply (src/TinyMPC.jl);
* host logic that needs no device: the LQR cache (reference tiny_api.cpp:244-318) against the oracle,
  argument validation, and the loud failure when no CUDA device exists.
"""
import ctypes as C
import os
import re

import numpy as np
import pytest

import cases
from tinympc_julia_b200 import _lib, problems

HEADER = os.path.join(cases.ROOT, "include", "tinympc_b200.h")

REFERENCE_SYMBOLS = ["setup_solver", "set_x0", "set_x_ref", "set_u_ref", "solve_mpc", "get_states", "get_controls",
                     "cleanup_solver", "codegen", "print_problem_data", "set_cache_terms", "codegen_with_sensitivity",
                     "update_settings", "set_bound_constraints", "set_linear_constraints", "set_cone_constraints"]


def declared_functions():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b([A-Za-z_][A-Za-z0-9_]*)\s*\(", re.sub(r"#.*", "", text))) - {"defined"})


def test_library_exports_every_declared_symbol():
    lib = _lib.load()
    names = declared_functions()
    assert len(names) >= 45
    for n in names:
        assert hasattr(lib, n), f"{n} is declared in include/tinympc_b200.h but not exported"


def test_reference_symbols_present_and_bound():
    lib = _lib.load()
    for n in REFERENCE_SYMBOLS:
        assert hasattr(lib, n)
        assert n in _lib.SIGNATURES
    assert set(_lib.SIGNATURES) <= set(declared_functions())


def test_every_declared_symbol_has_a_prototype():
    assert set(declared_functions()) == set(_lib.SIGNATURES)


def _cuda_available():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


@pytest.mark.skipif(_cuda_available(), reason="only meaningful on a machine without a GPU")
def test_no_cpu_fallback_without_device():
    """The product path must fail loudly when it cannot run on a GPU."""
    import tinympc_julia_b200 as t
    s = problems.cartpole()
    with pytest.raises(t.TinyMPCError):
        t.setup(t.TinyMPCBatchSolver(), s.A, s.B, s.f, s.Q, s.R, s.rho, 4, 1, 10, batch=2)
    with pytest.raises(t.TinyMPCError):
        t.setup(t.TinyMPCSolver(), s.A, s.B, s.f, s.Q, s.R, s.rho, 4, 1, 10)
    assert "no CUDA device" in _lib.last_error() or "CUDA" in _lib.last_error()


def test_calls_before_setup_fail_like_the_reference():
    lib = _lib.load()
    lib.cleanup_solver()
    x = np.zeros(4)
    assert lib.set_x0(x.ctypes.data_as(_lib.c_double_p), 4, 1, 0) == -1          # "Solver not initialized" -> -1
    assert lib.solve_mpc(0) == -1
    r, c = C.c_int(), C.c_int()
    assert lib.get_states(x.ctypes.data_as(_lib.c_double_p), C.byref(r), C.byref(c)) == -1
    assert lib.update_settings(1e-3, 1e-3, 10, 1, 0, 0, 0, 0, 0, 0, 0, 0.1, 10.0, 1, 0) == -1


def test_setup_dimension_mismatch_returns_core_status():
    lib = _lib.load()
    s = problems.cartpole()
    A, B, f, Q, R = [np.asfortranarray(np.asarray(a, dtype=np.float64)) for a in (s.A, s.B, s.f.reshape(-1, 1), s.Q, s.R)]
    P = lambda a: a.ctypes.data_as(_lib.c_double_p)
    # B declared 4x2 while nu = 1: tiny_setup's check_dimension path returns 1 (tiny_api.cpp:53-66)
    assert lib.setup_solver(P(A), 4, 4, P(B), 4, 2, P(f), 4, 1, P(Q), 4, 4, P(R), 1, 1, 1.0, 4, 1, 10, 0) == 1


@pytest.mark.parametrize("make", [problems.cartpole, problems.quadrotor, problems.rocket])
def test_host_lqr_cache_matches_oracle(orc, make):
    lib = _lib.load()
    s = make()
    n, m = s.nx, s.nu
    f64 = lambda a: np.asfortranarray(np.asarray(a, dtype=np.float64))
    A, B, f, Q, R = [f64(a) for a in (s.A, s.B, s.f, s.Q, s.R)]
    out = {k: np.zeros(sh, order="F") for k, sh in dict(Kinf=(m, n), Pinf=(n, n), Quu_inv=(m, m), AmBKt=(n, n), APf=(n,),
                                                         BPf=(m,), Qd=(n,), Rd=(m,)).items()}
    P = lambda a: a.ctypes.data_as(_lib.c_double_p)
    rc = lib.tmpc_compute_cache(P(A), P(B), P(f), P(Q), P(R), s.rho, n, m,
                                *[P(out[k]) for k in ("Kinf", "Pinf", "Quu_inv", "AmBKt", "APf", "BPf", "Qd", "Rd")])
    assert rc == 0
    want = orc.get_cache(cases.to_oracle_problem(dict(cases._base(s, 5))))
    for k in want:
        np.testing.assert_allclose(out[k], want[k], rtol=1e-12, atol=1e-14, err_msg=k)
    np.testing.assert_allclose(out["Qd"], np.diag(s.Q) + s.rho)
    np.testing.assert_allclose(out["Rd"], np.diag(s.R) + s.rho)


def test_host_lqr_cache_rejects_singular():
    lib = _lib.load()
    z = np.zeros((2, 2), order="F")
    P = lambda a: a.ctypes.data_as(_lib.c_double_p)
    f = np.zeros(2)
    # R + rho*I + B'PB = 0 when R = -2*rho... simplest: rho = 0, R = 0, B = 0 -> singular
    none = C.cast(None, _lib.c_double_p)
    rc = lib.tmpc_compute_cache(P(z), P(z), P(f), P(z), P(z), 0.0, 2, 2, *([none] * 8))
    assert rc != 0


def test_noalloc_entry_point_validates_before_touching_the_device():
    """tmpc_solve_noalloc (no counterpart in the reference; operators' request): argument and size checks come before any
    CUDA call, so they can be exercised here.  A working set beyond the static array is refused with the sizes named."""
    lib = _lib.load()
    s = problems.quadrotor()
    n, m, N = 12, 4, 20
    A, B, f, Q, R = (np.asfortranarray(np.asarray(a, dtype=np.float64)) for a in (s.A, s.B, s.f.reshape(n), s.Q, s.R))
    P = lambda a: a.ctypes.data_as(_lib.c_double_p)
    cache = [np.zeros(k) for k in (m * n, n * n, m * m, n * n, n, m, n, m)]
    assert lib.tmpc_compute_cache(P(A), P(B), P(f), P(Q), P(R), float(s.rho), n, m, *[P(c) for c in cache]) == 0
    err = C.create_string_buffer(512)
    fake = C.c_void_p(0x1000)                       # never dereferenced: the call must fail before reaching the device

    def call(B_, max_iter=100, x0=fake):
        return lib.tmpc_solve_noalloc(n, m, N, B_, float(s.rho), P(A), P(B), P(f), *[P(c) for c in cache], x0, None, None, None, None,
                                      None, None, 1e-3, 1e-3, max_iter, 1, fake, fake, fake, fake, None, err, len(err))
    assert call(3073) == 1 and b"too large" in err.value and b"tmpc_batch_create" in err.value
    assert call(1 << 20) == 1 and b"too large" in err.value
    assert call(64, max_iter=0) == 2 and b"max_iter" in err.value
    assert call(64, x0=None) == 1 and b"null device pointer" in err.value
    assert call(0) == 1
