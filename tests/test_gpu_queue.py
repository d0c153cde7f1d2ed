"""Queue order, stream ordering and the output selector of the batched host path — GPU only.

The on-chip kernels are persistent and pull instances from a ticket queue; `tmpc_batch_set_order` changes WHICH
instance a ticket names (longest-first by default, queue_order.cu), never what is computed for it: results must be
identical under every order.  Reference behaviour being protected: iteration counts are data-dependent
(admm.cpp:312-381), instances are independent.
"""
import ctypes as C

import numpy as np
import pytest

import cases
import tinympc_julia_b200 as t
from tinympc_julia_b200 import _lib

pytestmark = pytest.mark.gpu

WAVE = 148 * 64


def _solve_cold(sol, x0):
    t.set_x0(sol, x0)
    sol.reset()
    rc = t.solve(sol)
    assert rc in (0, 1)
    st, it = sol.status()
    X, U = t.get_solution(sol)
    return st, it, X, U


def test_results_do_not_depend_on_queue_order(orc):
    """natural / longest-first / a random permutation: identical status, iteration counts and (bitwise) solutions;
    a sample is checked against the oracle; the automatic order really lists the long instances first."""
    B = 2 * WAVE + 1234
    c = cases.quadrotor_batch(B, hard=False, seed=4)
    sol = cases.make_cuda_solver(c)
    assert sol.kernel == t.KERNEL_TMEM
    sol.set_order(0)
    st0, it0, X0, U0 = _solve_cold(sol, c["x0"])
    assert sol.get_order() is None
    sol.set_order(1)
    st1, it1, X1, U1 = _solve_cold(sol, c["x0"])
    order = sol.get_order()
    assert order is not None and np.array_equal(np.sort(order), np.arange(B))
    # every instance above 12 iterations is among the top 4 of its block of 256, i.e. within the first 4 rank rows of the
    # queue (checked on the reference for seeds 3..10); everything behind the head section (the TOP = 64 highest-ranked of
    # every 256) keeps its natural order
    TOP = 64
    nblk = (B + 255) // 256
    pos = np.empty(B, dtype=np.int64)
    pos[order] = np.arange(B)
    long_ones = np.nonzero(it1 > 12)[0]
    assert long_ones.size == 0 or pos[long_ones].max() < 4 * nblk, (long_ones, pos[long_ones])
    tail = order[TOP * (nblk - 1) + min(TOP, B - 256 * (nblk - 1)):]
    assert np.all(np.diff(tail) > 0)
    rng = np.random.default_rng(1)
    sol.set_order(2, rng.permutation(B).astype(np.int32))
    st2, it2, X2, U2 = _solve_cold(sol, c["x0"])
    for st, it, X, U in ((st1, it1, X1, U1), (st2, it2, X2, U2)):
        assert np.array_equal(st0, st) and np.array_equal(it0, it)
        assert np.array_equal(X0, X) and np.array_equal(U0, U)
    # warm start: the automatic order is built from the previous iteration counts
    sol.set_order(1)
    x0b = np.asfortranarray(c["x0"] * 1.3)
    t.set_x0(sol, x0b)
    assert t.solve(sol) in (0, 1)
    stw, itw = sol.status()
    Xw, Uw = t.get_solution(sol)
    order_w = sol.get_order()
    assert order_w is not None and np.array_equal(np.sort(order_w), np.arange(B))
    for k in (0, 1, nblk - 2):                     # the previous solve's longest instance of every block leads the queue
        assert it2[order_w[k]] == it2[256 * k:256 * (k + 1)].max()
    ref = cases.make_cuda_solver(c)
    ref.set_order(0)
    _solve_cold(ref, c["x0"])
    t.set_x0(ref, x0b)
    assert t.solve(ref) in (0, 1)
    str_, itr = ref.status()
    Xr, Ur = t.get_solution(ref)
    assert np.array_equal(stw, str_) and np.array_equal(itw, itr) and np.array_equal(Xw, Xr) and np.array_equal(Uw, Ur)
    # oracle on a sample of the cold solve
    idx = np.concatenate([np.arange(0, B, 97), long_ones])[:400]
    cs = dict(c)
    cs["x0"] = np.asfortranarray(c["x0"][:, idx])
    want = cases.run_checker(cs, orc, nthreads=8)
    got = {"ret": st0[None, idx], "iters": it0[None, idx], "X": X0[..., idx], "U": U0[..., idx]}
    cases.compare(got, want, what="ordered solve vs oracle sample")


def test_set_order_validates():
    c = cases.quadrotor_batch(64)
    sol = cases.make_cuda_solver(c)
    lib = _lib.load()
    bad = np.zeros(64, dtype=np.int32)
    assert lib.tmpc_batch_set_order(sol._h, 2, bad.ctypes.data_as(C.POINTER(C.c_int))) == 1      # not a permutation
    assert lib.tmpc_batch_set_order(sol._h, 2, None) == 1
    assert lib.tmpc_batch_set_order(sol._h, 7, None) == 1
    perm = np.arange(63, -1, -1, dtype=np.int32)
    assert lib.tmpc_batch_set_order(sol._h, 2, perm.ctypes.data_as(C.POINTER(C.c_int))) == 0
    want = cases.load_golden("quadrotor_easy_64")
    cases.compare(cases.run_cuda(c, sol=sol), want, what="user order on a small batch")


def test_caller_stream_is_ordered_against_getters(orc):
    """a solve enqueued on a CALLER stream, then getters / synchronize on the handle without any caller-side
    synchronisation: the handle waits for the caller stream's work (event), results equal the blocking path.  Also the
    first cone / half-space solve on a caller stream (the duals are zeroed on that stream)."""
    import torch
    stream = torch.cuda.Stream()
    # the third case is a tiny batch: results in mapped host memory, x0 read by the latency kernel from its staging buffer
    for c in (cases.quadrotor_batch(WAVE + 500, hard=True, max_iter=40), cases.rocket(300, N=50),
              cases.quadrotor_batch(20, hard=True, max_iter=100)):
        want = cases.run_cuda(c)                                    # blocking path
        sol = cases.make_cuda_solver(c)
        if c["Xref"] is not None:
            t.set_x_ref(sol, c["Xref"])
        if c["Uref"] is not None:
            t.set_u_ref(sol, c["Uref"])
        t.set_x0(sol, c["x0"])
        sol.solve_async(stream.cuda_stream)
        st, it = sol.status()                                       # no torch / stream synchronisation in between
        X, U = t.get_solution(sol)
        assert np.array_equal(st, want["ret"][0]) and np.array_equal(it, want["iters"][0])
        assert np.array_equal(X, want["X"]) and np.array_equal(U, want["U"])
        # and a setter after an asynchronous solve must not overtake it
        sol.reset()
        t.set_x0(sol, c["x0"])
        sol.solve_async(stream.cuda_stream)
        t.set_x0(sol, np.asfortranarray(c["x0"] * 0.0))             # would corrupt the running solve if it raced
        sol.synchronize()
        st2, it2 = sol.status()
        assert np.array_equal(it2, want["iters"][0])


def test_solve_host_output_selector():
    """tmpc_batch_solve_host_select: u0 (first control of every instance) alone, with controls, with everything."""
    B = 3 * WAVE + 11
    c = cases.quadrotor_batch(B, hard=False, seed=6)
    ref = cases.make_cuda_solver(c)
    st0, it0, X0, U0 = _solve_cold(ref, c["x0"])
    sol = cases.make_cuda_solver(c)
    r, s, st, it = sol.solve_host(c["x0"], cold=True, want_states=False, want_controls=False, want_u0=True)
    assert np.array_equal(st, st0) and np.array_equal(it, it0)
    assert s.states is None and s.controls.shape == (4, B) and np.array_equal(s.controls, U0[:, 0, :])
    r, s, st, it = sol.solve_host(c["x0"], cold=True, want_states=True, want_controls=True, want_u0=True)
    assert np.array_equal(s.states, X0) and np.array_equal(s.controls, U0) and np.array_equal(it, it0)
    r, s, st, it = sol.solve_host(c["x0"], cold=True, chunks=5, want_states=False, want_controls=False, want_u0=True)
    assert np.array_equal(s.controls, U0[:, 0, :]) and np.array_equal(it, it0)
    # warm-started closed-loop shape on page-locked buffers: the kernel reads x0 from the mapped host buffer, the queue
    # order comes from the previous iteration counts, one kernel writes u0 / status / iterations home; the device copy
    # of x0 is refreshed beside the solve (a later plain solve must see it)
    x1 = np.asfortranarray(c["x0"] * 0.8)
    t.set_x0(ref, x1)
    assert t.solve(ref) in (0, 1)
    st1, it1 = ref.status()
    X1, U1 = t.get_solution(ref)
    r, s, st, it = sol.solve_host(x1, cold=False, want_states=False, want_controls=False, want_u0=True)
    assert np.array_equal(st, st1) and np.array_equal(it, it1) and np.array_equal(s.controls, U1[:, 0, :])
    assert t.solve(ref) in (0, 1) and t.solve(sol) in (0, 1)          # third solve, same (device-resident) x0 on both
    st2, it2 = ref.status()
    st3, it3 = sol.status()
    assert np.array_equal(it2, it3) and np.array_equal(st2, st3)
    assert np.array_equal(t.get_solution(ref).controls, t.get_solution(sol).controls)


def test_rejected_settings():
    """max_iter <= 0 is rejected like check_termination = 0 (the reference then returns stale workspace contents,
    admm.cpp:384-388); set_cache_terms on a per-instance batch is rejected instead of silently ignored."""
    from tinympc_julia_b200 import problems
    c = cases.quadrotor_batch(8)
    sol = cases.make_cuda_solver(c)
    with pytest.raises(t.TinyMPCError):
        t.update_settings(sol, max_iter=0)
    assert "max_iter" in _lib.last_error()
    base = problems.quadrotor()
    A = np.asfortranarray(np.repeat(base.A[:, :, None], 8, axis=2))
    Bm = np.asfortranarray(np.repeat(base.B[:, :, None], 8, axis=2))
    t.set_instance_dynamics(sol, A, Bm, None, base.Q, base.R)
    k = sol.cache_terms()
    with pytest.raises(t.TinyMPCError):
        t.set_cache_terms(sol, k["Kinf"], k["Pinf"], k["Quu_inv"], k["AmBKt"])
    assert "per-instance" in _lib.last_error()
