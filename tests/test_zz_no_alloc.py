"""tmpc_solve_noalloc (include/tinympc_b200.h) — the entry point that allocates nothing on the device — GPU only.

No counterpart in the reference: added for callers that cannot call cudaMalloc.
Called through the C ABI with caller-owned device buffers and compared with the existing entry points
(tmpc_batch_create -> set_* -> tmpc_batch_solve on TMPC_KERNEL_THREAD, cold) on two calls in a row: identical status and
iteration counts, states / controls within the tolerance of tests/cases.py (1e-5 rel / 1e-6 abs).  The second call is a
DIFFERENT, smaller batch, so anything the first call left in the shared scratch array would show.
The file name keeps it last in pytest's collection order.
"""
import ctypes as C

import numpy as np
import pytest

import cases
import oracle
import tinympc_julia_b200 as t
from tinympc_julia_b200 import _lib

pytestmark = pytest.mark.gpu

P = lambda a: a.ctypes.data_as(C.POINTER(C.c_double))


def _cache(s):
    """tmpc_compute_cache (host only): Kinf, Pinf, Quu_inv, AmBKt, APf, BPf, Qd, Rd."""
    lib = _lib.load()
    n, m = s.nx, s.nu
    A, B, f, Q, R = (np.asfortranarray(np.asarray(a, dtype=np.float64)) for a in (s.A, s.B.reshape(n, m), s.f.reshape(n), s.Q, s.R))
    out = [np.zeros(k) for k in (m * n, n * n, m * m, n * n, n, m, n, m)]
    assert lib.tmpc_compute_cache(P(A), P(B), P(f), P(Q), P(R), float(s.rho), n, m, *[P(o) for o in out]) == 0
    return (A, B, f), out


def _dev(a):
    """Column-major host array -> device tensor with the same memory layout."""
    import torch
    a = np.asarray(a, dtype=np.float64)
    return torch.from_numpy(np.ascontiguousarray(a.T if a.ndim > 1 else a)).cuda()


def _solve_noalloc(c, stream=None):
    import torch
    lib = _lib.load()
    s, N = c["sys"], c["N"]
    n, m = s.nx, s.nu
    B = c["x0"].shape[1]
    (A, Bm, f), cache = _cache(s)
    x0 = _dev(c["x0"])
    xref = _dev(c["Xref"]) if c["Xref"] is not None else None
    uref = _dev(c["Uref"]) if c["Uref"] is not None else None
    bnd = [_dev(b) for b in c["bounds"]] if c["bounds"] is not None else [None] * 4
    X = torch.full((B, N, n), np.nan, dtype=torch.float64, device="cuda")
    U = torch.full((B, N - 1, m), np.nan, dtype=torch.float64, device="cuda")
    st = torch.full((B,), -1, dtype=torch.int32, device="cuda")
    it = torch.full((B,), -1, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    ptr = lambda x: C.c_void_p(x.data_ptr()) if x is not None else None
    err = C.create_string_buffer(512)
    free0 = torch.cuda.mem_get_info()[0]
    rc = lib.tmpc_solve_noalloc(n, m, N, B, float(s.rho), P(A), P(Bm), P(f), *[P(o) for o in cache],
                                ptr(x0), ptr(xref), ptr(uref), *[ptr(b) for b in bnd],
                                c["abs_pri_tol"], c["abs_dua_tol"], c["max_iter"], int(c["check_termination"]),
                                ptr(X), ptr(U), ptr(st), ptr(it), C.c_void_p(stream.cuda_stream) if stream else None,
                                err, len(err))
    if stream:
        stream.synchronize()
    torch.cuda.synchronize()
    free1 = torch.cuda.mem_get_info()[0]
    return dict(rc=rc, err=err.value.decode(), ret=st.cpu().numpy()[None, :], iters=it.cpu().numpy()[None, :],
                X=X.cpu().numpy().transpose(2, 1, 0), U=U.cpu().numpy().transpose(2, 1, 0), free=(free0, free1))


def _existing(c):
    """The existing entry points: handle-based batched API, thread kernel, cold."""
    out = cases.run_cuda(c, kernel=t.KERNEL_THREAD)
    assert out["solver"].kernel == t.KERNEL_THREAD
    return out


def _quadrotor_with_refs(B, seed, hard):
    c = cases.quadrotor_batch(B, hard=hard, seed=seed, max_iter=40)
    N = c["N"]
    rng = np.random.default_rng(seed)
    c["Xref"] = np.asfortranarray(0.05 * rng.standard_normal((12, N)))
    c["Uref"] = np.asfortranarray(0.01 * rng.standard_normal((4, N - 1)))
    c["ref_mode"] = oracle.REF_SHARED
    return c


def test_two_calls_in_a_row_match_the_existing_entry_point():
    import torch
    stream = torch.cuda.Stream()
    first = _quadrotor_with_refs(700, seed=5, hard=True)          # ragged B (ld = 704), references + box
    second = cases.cartpole_batch(333, seed=8)                    # another shape, smaller: lands on the first call's leftovers
    third = _quadrotor_with_refs(257, seed=6, hard=False)
    for k, (c, s) in enumerate(((first, None), (second, stream), (third, stream))):
        got = _solve_noalloc(c, stream=s)
        assert got["rc"] == 0, got["err"]
        want = _existing(c)
        cases.compare(got, want, what=f"tmpc_solve_noalloc call {k}")
        assert np.isfinite(got["X"]).all() and np.isfinite(got["U"]).all()
        if k > 0:       # nothing was allocated on the device by the call (the first one may load the module)
            assert got["free"][1] >= got["free"][0], got["free"]
    # and the same batch twice gives the same bits
    again = _solve_noalloc(third, stream=stream)
    again2 = _solve_noalloc(third)
    assert np.array_equal(again["X"], again2["X"]) and np.array_equal(again["iters"], again2["iters"])


def test_run_time_shape_and_unconstrained():
    """A shape that is not compiled into the thread kernel (constants go through the head of the scratch array), twice with
    different constants; then no bounds and no references at all."""
    for seed in (7, 9):
        c = cases.random_system(5, 2, 8, 150, seed=seed, with_linear=False)
        got = _solve_noalloc(c)
        assert got["rc"] == 0, got["err"]
        cases.compare(got, _existing(c), what=f"run-time shape, seed {seed}")
    c = cases.cartpole_batch(100, seed=3)
    c["bounds"] = None
    got = _solve_noalloc(c)
    assert got["rc"] == 0, got["err"]
    cases.compare(got, _existing(c), what="unconstrained")


def test_sizes_beyond_the_static_array_are_rejected():
    c = cases.quadrotor_batch(3073)          # ld = 3104 > 3072 instances of 12 / 4 / 20
    got = _solve_noalloc(c)
    assert got["rc"] == 1 and "too large" in got["err"] and "tmpc_batch_create" in got["err"]
    assert (got["ret"] == -1).all()          # nothing ran
    c = cases.quadrotor_batch(3072)          # the largest that fits
    got = _solve_noalloc(c)
    assert got["rc"] == 0, got["err"]
    cases.compare(got, _existing(c), what="largest batch")
    # argument checks
    c = cases.cartpole_batch(8)
    c["max_iter"] = 0
    assert _solve_noalloc(c)["rc"] == 2
