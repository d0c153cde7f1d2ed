"""Host-side mirror of the Julia package's interface for the solve path.

Julia is not installed on the build or GPU machines, so this module is the executable counterpart
of ``src/TinyMPC.jl`` (reference): same function names, argument meaning, defaults and error
behaviour, each function performing exactly the ``ccall`` the Julia function performs, against the
same C symbols.  ``julia/TinyMPC.jl`` in this repository is the Julia-language twin (the reference
file plus the batched methods).

Two solver types:

* :class:`TinyMPCSolver` — the reference's single-instance solver (process-wide singleton in the
  C library, reference src/bindings.cpp:15).  Functions: setup, set_x0, set_x_ref, set_u_ref,
  set_bound_constraints, set_linear_constraints, set_cone_constraints, set_equality_constraints,
  update_settings, set_cache_terms, solve, get_solution, print_problem_data
  (reference src/TinyMPC.jl:55-287).
* :class:`TinyMPCBatchSolver` — B independent instances sharing dynamics and cost; the same
  functions accept it, with arrays carrying a trailing batch dimension.

All matrices cross the boundary column-major (``Matrix{Float64}``), so arrays are converted with
``np.asfortranarray`` before the call.
"""
from __future__ import annotations

import ctypes as C
from typing import NamedTuple, Optional

import numpy as np

from . import _lib

_P, _IP = _lib.c_double_p, _lib.c_int_p


class TinyMPCError(RuntimeError):
    """Raised where the Julia package calls ``error(...)``."""


class Solution(NamedTuple):
    states: np.ndarray
    controls: np.ndarray


def _mat(a, shape=None) -> np.ndarray:
    arr = np.asarray(a, dtype=np.float64)
    if shape is not None:
        arr = arr.reshape(shape, order="F") if arr.ndim != len(shape) else arr
    return np.asfortranarray(arr)


def _dp(arr: np.ndarray):
    return arr.ctypes.data_as(_P)


def _ivec(a) -> np.ndarray:
    return np.ascontiguousarray(np.asarray(a, dtype=np.int32).reshape(-1))


def pinned_empty(shape, dtype=np.float64) -> np.ndarray:
    """Fortran-ordered array in page-locked host memory (tmpc_host_alloc): the buffers of ``solve_host`` must be pinned
    for the chunks' copies to overlap the solves.  Freed when the array (and every view of it) is collected."""
    import weakref
    lib = _lib.load()
    dt = np.dtype(dtype)
    n = int(np.prod(shape)) if np.ndim(shape) else int(shape)
    ptr = lib.tmpc_host_alloc(C.c_size_t(max(n, 1) * dt.itemsize))
    if not ptr:
        raise TinyMPCError("tmpc_host_alloc failed (no CUDA device?)")
    raw = (C.c_char * (max(n, 1) * dt.itemsize)).from_address(ptr)
    arr = np.frombuffer(raw, dtype=dt, count=n).reshape(shape, order="F")
    weakref.finalize(raw, lib.tmpc_host_free, C.c_void_p(ptr))
    return arr


class TinyMPCSolver:
    """Mirror of the Julia ``struct TinyMPCSolver`` (reference src/TinyMPC.jl:35-49)."""

    def __init__(self):
        self.nx = 0
        self.nu = 0
        self.N = 0
        self.rho = 0.0
        self.is_setup = False
        self.A = np.zeros((0, 0))
        self.B = np.zeros((0, 0))
        self.Q = np.zeros((0, 0))
        self.R = np.zeros((0, 0))


class TinyMPCBatchSolver:
    """B problem instances with shared A, B, f, Q, R, rho; device-resident state (opaque handle)."""

    def __init__(self):
        self.nx = 0
        self.nu = 0
        self.N = 0
        self.B = 0
        self.rho = 0.0
        self.is_setup = False
        self._h = C.c_void_p(None)

    def __del__(self):
        try:
            if self._h:
                _lib.load().tmpc_batch_destroy(self._h)
                self._h = C.c_void_p(None)
        except Exception:
            pass

    # -- extras that have no single-solver counterpart
    def reset(self):
        """Cold start: clear duals, slacks and linear cost terms (state right after ``setup``)."""
        _chk(_lib.load().tmpc_batch_reset(self._h), "Failed to reset batch")

    def set_precision(self, bits: int):
        """Arithmetic type of the solve: 64 (``typedef double tinytype``, the reference's default) or 32 (its float
        build, tinympc/TinyMPC/src/tinympc/types.hpp:15).  Call right after ``setup`` / ``reset``."""
        _chk(_lib.load().tmpc_batch_set_precision(self._h, int(bits)), "Failed to set precision")

    @property
    def precision(self) -> int:
        return int(_lib.load().tmpc_batch_get_precision(self._h))

    def set_kernel(self, kernel: int):
        _chk(_lib.load().tmpc_batch_set_kernel(self._h, int(kernel)), "Failed to select kernel")

    @property
    def kernel(self) -> int:
        return int(_lib.load().tmpc_batch_get_kernel(self._h))

    def status(self):
        """(status[B], iters[B]): per instance what ``solve`` would have returned, and its iteration count."""
        st = np.zeros(self.B, dtype=np.int32)
        it = np.zeros(self.B, dtype=np.int32)
        _chk(_lib.load().tmpc_batch_get_status(self._h, st.ctypes.data_as(_IP), it.ctypes.data_as(_IP)), "Failed to get status")
        return st, it

    def cache_terms(self) -> dict:
        n, m = self.nx, self.nu
        out = {"Kinf": np.zeros((m, n), order="F"), "Pinf": np.zeros((n, n), order="F"),
               "Quu_inv": np.zeros((m, m), order="F"), "AmBKt": np.zeros((n, n), order="F"),
               "APf": np.zeros(n), "BPf": np.zeros(m)}
        _chk(_lib.load().tmpc_batch_get_cache_terms(self._h, *[_dp(out[k]) for k in ("Kinf", "Pinf", "Quu_inv", "AmBKt", "APf", "BPf")]),
             "Failed to get cache terms")
        return out

    @property
    def launch_count(self) -> int:
        return int(_lib.load().tmpc_batch_launch_count(self._h))

    def device_ptr(self, which: int) -> int:
        return int(_lib.load().tmpc_batch_device_ptr(self._h, which) or 0)

    # device-pointer variants (integers = CUDA device addresses, e.g. torch.Tensor.data_ptr())
    def bind_x0_device(self, ptr: int):
        _chk(_lib.load().tmpc_batch_bind_x0_device(self._h, C.c_void_p(ptr)), "bind_x0_device failed")

    def bind_x_ref_device(self, ptr: int, batch: int):
        _chk(_lib.load().tmpc_batch_bind_x_ref_device(self._h, C.c_void_p(ptr), int(batch)), "bind_x_ref_device failed")

    def bind_u_ref_device(self, ptr: int, batch: int):
        _chk(_lib.load().tmpc_batch_bind_u_ref_device(self._h, C.c_void_p(ptr), int(batch)), "bind_u_ref_device failed")

    def solve_async(self, stream: int = 0):
        _chk(_lib.load().tmpc_batch_solve_async(self._h, C.c_void_p(stream)), "solve_async failed")

    def synchronize(self):
        _chk(_lib.load().tmpc_batch_synchronize(self._h), "synchronize failed")

    def set_order(self, mode: int = 1, perm=None):
        """Queue order of the on-chip kernels: 0 natural, 1 longest-first (default), 2 the given permutation of 0..B-1."""
        p = None
        if perm is not None:
            pa = _ivec(perm)
            p = pa.ctypes.data_as(_IP)
        _chk(_lib.load().tmpc_batch_set_order(self._h, int(mode), p), "set_order failed")

    def get_order(self):
        """The permutation the last solve used, or None when it ran in natural order."""
        out = np.zeros(self.B, dtype=np.int32)
        rc = _lib.load().tmpc_batch_get_order(self._h, out.ctypes.data_as(_IP))
        return out if rc == 0 else None

    def solve_host(self, x0=None, *, cold: bool = False, chunks: int = 0, want_states: bool = True,
                   want_controls: bool = True, want_u0: bool = False):
        """set_x0 (+ reset) + solve + get_solution + status in one call, pipelined over instance chunks
        (tmpc_batch_solve_host_select).  Output buffers are page-locked (reused between calls), so the chunks' copies
        overlap the solves.  Returns (ret, Solution, status[B], iters[B]); ret as ``solve``; with ``want_u0`` the
        Solution's controls are the first control of every instance (nu x B) when ``want_controls`` is off."""
        n, m, N, B = self.nx, self.nu, self.N, self.B
        buf = self.__dict__.setdefault("_pinned", {})

        def pin(key, shape, dtype=np.float64):
            if key not in buf:
                buf[key] = pinned_empty(shape, dtype)
            return buf[key]

        x0p = None
        if x0 is not None:
            x0a = pin("x0", (n, B))
            x0a[...] = _mat(x0, (n, B))
            x0p = _dp(x0a)
        X = pin("X", (n, N, B)) if want_states else None
        U = pin("U", (m, N - 1, B)) if want_controls else None
        u0 = pin("u0", (m, B)) if want_u0 else None
        st, it = pin("st", (B,), np.int32), pin("it", (B,), np.int32)
        ret = _lib.load().tmpc_batch_solve_host_select(self._h, x0p, int(bool(cold)), _dp(X) if X is not None else None,
                                                       _dp(U) if U is not None else None, _dp(u0) if u0 is not None else None,
                                                       st.ctypes.data_as(_IP), it.ctypes.data_as(_IP), int(chunks))
        if ret not in (0, 1):
            raise TinyMPCError(f"solve_host failed with status {ret}: {_lib.last_error()}")
        controls = U if U is not None else u0
        return ret, Solution(states=None if X is None else X.copy(order="F"),
                             controls=None if controls is None else controls.copy(order="F")), st.copy(), it.copy()

    def rollout_async(self, nsteps: int, x_ref_long_ptr: int = 0, status_out_ptr: int = 0, iters_out_ptr: int = 0, stream: int = 0):
        _chk(_lib.load().tmpc_batch_rollout_async(self._h, int(nsteps), C.c_void_p(x_ref_long_ptr), C.c_void_p(status_out_ptr),
                                                  C.c_void_p(iters_out_ptr), C.c_void_p(stream)), "rollout failed")


def _chk(status: int, msg: str):
    if status != 0:
        detail = _lib.last_error()
        raise TinyMPCError(f"{msg}" + (f" ({detail})" if detail else ""))


# ------------------------------------------------------------------------------------------ setup

def setup(solver, A, B, f, Q, R, rho: float, nx: int, nu: int, N: int, *, batch: Optional[int] = None,
          device: int = -1, verbose: bool = False, abs_pri_tol: float = 1e-3, abs_dua_tol: float = 1e-3,
          max_iter: int = 100, check_termination: bool = True, adaptive_rho: bool = False,
          adaptive_rho_min: float = 0.1, adaptive_rho_max: float = 10.0, adaptive_rho_clipping: bool = True) -> int:
    """reference src/TinyMPC.jl:55-113.  For a :class:`TinyMPCBatchSolver` pass ``batch=B``."""
    lib = _lib.load()
    A_, B_ = _mat(A), _mat(np.asarray(B, dtype=np.float64).reshape(nx, -1) if np.ndim(B) < 2 else B)
    f_ = _mat(np.asarray(f, dtype=np.float64).reshape(-1, 1))
    Q_, R_ = _mat(Q), _mat(np.asarray(R, dtype=np.float64).reshape(nu, -1) if np.ndim(R) < 2 else R)
    solver.nx, solver.nu, solver.N, solver.rho = int(nx), int(nu), int(N), float(rho)
    if isinstance(solver, TinyMPCBatchSolver):
        if batch is None or batch < 1:
            raise TinyMPCError("setup of a batch solver needs batch=B >= 1")
        if A_.shape != (nx, nx) or B_.shape != (nx, nu) or f_.shape != (nx, 1) or Q_.shape != (nx, nx) or R_.shape != (nu, nu):
            raise TinyMPCError("Setup failed with status: 1")
        if solver._h:
            lib.tmpc_batch_destroy(solver._h)
        h = C.c_void_p(None)
        status = lib.tmpc_batch_create(C.byref(h), _dp(A_), _dp(B_), _dp(f_), _dp(Q_), _dp(R_), float(rho), nx, nu, N,
                                       int(batch), int(device), 1 if verbose else 0)
        if status != 0:
            raise TinyMPCError(f"Setup failed with status: {status} ({_lib.last_error()})")
        solver._h, solver.B = h, int(batch)
        solver.A_, solver.B_, solver.Q_, solver.R_ = A_.copy(), B_.copy(), Q_.copy(), R_.copy()
    else:
        solver.A, solver.B, solver.Q, solver.R = A_.copy(), B_.copy(), Q_.copy(), R_.copy()
        status = lib.setup_solver(_dp(A_), *A_.shape, _dp(B_), *B_.shape, _dp(f_), *f_.shape, _dp(Q_), *Q_.shape,
                                  _dp(R_), *R_.shape, float(rho), nx, nu, N, 1 if verbose else 0)
        if status != 0:
            raise TinyMPCError(f"Setup failed with status: {status}")
    solver.is_setup = True
    # the Julia setup immediately pushes its own defaults and clears every en_* flag (:89-104)
    update_settings(solver, abs_pri_tol=abs_pri_tol, abs_dua_tol=abs_dua_tol, max_iter=max_iter,
                    check_termination=check_termination, en_state_bound=False, en_input_bound=False,
                    en_state_soc=False, en_input_soc=False, en_state_linear=False, en_input_linear=False,
                    adaptive_rho=adaptive_rho, adaptive_rho_min=adaptive_rho_min, adaptive_rho_max=adaptive_rho_max,
                    adaptive_rho_enable_clipping=adaptive_rho_clipping, verbose=verbose)
    return 0


def update_settings(solver, *, abs_pri_tol: float = 1e-3, abs_dua_tol: float = 1e-3, max_iter: int = 100,
                    check_termination: bool = True, en_state_bound: bool = False, en_input_bound: bool = False,
                    en_state_soc: bool = False, en_input_soc: bool = False, en_state_linear: bool = False,
                    en_input_linear: bool = False, adaptive_rho: bool = False, adaptive_rho_min: float = 0.1,
                    adaptive_rho_max: float = 10.0, adaptive_rho_enable_clipping: bool = True, verbose: bool = False) -> int:
    """reference src/TinyMPC.jl:181-211 — NOT incremental: every keyword not passed goes back to its default."""
    lib = _lib.load()
    args = (float(abs_pri_tol), float(abs_dua_tol), int(max_iter), 1 if check_termination else 0,
            int(bool(en_state_bound)), int(bool(en_input_bound)), int(bool(en_state_soc)), int(bool(en_input_soc)),
            int(bool(en_state_linear)), int(bool(en_input_linear)), int(bool(adaptive_rho)), float(adaptive_rho_min),
            float(adaptive_rho_max), int(bool(adaptive_rho_enable_clipping)))
    if isinstance(solver, TinyMPCBatchSolver):
        status = lib.tmpc_batch_update_settings(solver._h, *args)
    else:
        status = lib.update_settings(*args, 1 if verbose else 0)
    if status != 0:
        raise TinyMPCError("Failed to update settings")
    return status


# ------------------------------------------------------------------------------------------ per-solve inputs

def _need_setup(solver):
    if not solver.is_setup:
        raise TinyMPCError("Solver not setup")


def set_x0(solver, x0, *, verbose: bool = False) -> int:
    """reference src/TinyMPC.jl:115-123.  Batch solver: ``x0`` is nx x B."""
    _need_setup(solver)
    lib = _lib.load()
    if isinstance(solver, TinyMPCBatchSolver):
        x = _mat(np.asarray(x0, dtype=np.float64).reshape(solver.nx, -1, order="F") if np.ndim(x0) < 2 else x0)
        status = lib.tmpc_batch_set_x0(solver._h, _dp(x), x.shape[0], x.shape[1])
    else:
        x = _mat(np.asarray(x0, dtype=np.float64).reshape(-1, 1))
        status = lib.set_x0(_dp(x), x.shape[0], x.shape[1], 1 if verbose else 0)
    if status != 0:
        raise TinyMPCError("Failed to set initial state")
    return status


def _set_ref(solver, ref, single, batched, what, verbose):
    _need_setup(solver)
    lib = _lib.load()
    r = _mat(ref)
    if isinstance(solver, TinyMPCBatchSolver):
        nb = 1 if r.ndim == 2 else r.shape[2]
        status = getattr(lib, batched)(solver._h, _dp(r), r.shape[0], r.shape[1], nb)
    else:
        if r.ndim != 2:
            raise TinyMPCError(f"Failed to set {what}")
        status = getattr(lib, single)(_dp(r), r.shape[0], r.shape[1], 1 if verbose else 0)
    if status != 0:
        raise TinyMPCError(f"Failed to set {what}")
    return status


def set_x_ref(solver, x_ref, *, verbose: bool = False) -> int:
    """reference src/TinyMPC.jl:125-132.  Batch solver: nx x N (shared) or nx x N x B."""
    return _set_ref(solver, x_ref, "set_x_ref", "tmpc_batch_set_x_ref", "state reference", verbose)


def set_u_ref(solver, u_ref, *, verbose: bool = False) -> int:
    """reference src/TinyMPC.jl:134-141.  Batch solver: nu x (N-1) (shared) or nu x (N-1) x B."""
    return _set_ref(solver, u_ref, "set_u_ref", "tmpc_batch_set_u_ref", "input reference", verbose)


def set_bound_constraints(solver, x_min, x_max, u_min, u_max, *, verbose: bool = False) -> int:
    """reference src/TinyMPC.jl:214-226 (enables both bound flags).  Batch solver: shared or trailing B."""
    lib = _lib.load()
    xm, xM, um, uM = _mat(x_min), _mat(x_max), _mat(u_min), _mat(u_max)
    if isinstance(solver, TinyMPCBatchSolver):
        nb = 1 if xm.ndim == 2 else xm.shape[2]
        exp_x = (solver.nx, solver.N) + (() if nb == 1 and xm.ndim == 2 else (nb,))
        exp_u = (solver.nu, solver.N - 1) + (() if nb == 1 and xm.ndim == 2 else (nb,))
        if xm.shape != exp_x or xM.shape != exp_x or um.shape != exp_u or uM.shape != exp_u:
            raise TinyMPCError("Failed to set bound constraints")
        status = lib.tmpc_batch_set_bound_constraints(solver._h, _dp(xm), _dp(xM), _dp(um), _dp(uM), nb)
    else:
        status = lib.set_bound_constraints(_dp(xm), *xm.shape, _dp(xM), *xM.shape, _dp(um), *um.shape, _dp(uM), *uM.shape,
                                           1 if verbose else 0)
    if status != 0:
        raise TinyMPCError("Failed to set bound constraints")
    return status


def set_linear_constraints(solver, Alin_x, blin_x, Alin_u, blin_u, *, verbose: bool = False) -> int:
    """reference src/TinyMPC.jl:228-241."""
    lib = _lib.load()
    nx, nu = solver.nx, solver.nu
    Ax = _mat(np.asarray(Alin_x, dtype=np.float64).reshape(-1, nx) if np.size(Alin_x) else np.zeros((0, nx)))
    Au = _mat(np.asarray(Alin_u, dtype=np.float64).reshape(-1, nu) if np.size(Alin_u) else np.zeros((0, nu)))
    bx = np.ascontiguousarray(np.asarray(blin_x, dtype=np.float64).reshape(-1))
    bu = np.ascontiguousarray(np.asarray(blin_u, dtype=np.float64).reshape(-1))
    if isinstance(solver, TinyMPCBatchSolver):
        if bx.size != Ax.shape[0] or bu.size != Au.shape[0]:
            raise TinyMPCError("Failed to set linear constraints")
        status = lib.tmpc_batch_set_linear_constraints(solver._h, _dp(Ax), Ax.shape[0], _dp(bx), _dp(Au), Au.shape[0], _dp(bu))
    else:
        status = lib.set_linear_constraints(_dp(Ax), *Ax.shape, _dp(bx), bx.size, _dp(Au), *Au.shape, _dp(bu), bu.size,
                                            1 if verbose else 0)
    if status != 0:
        raise TinyMPCError("Failed to set linear constraints")
    return status


def set_cone_constraints(solver, Acu, qcu, cu, Acx, qcx, cx, *, verbose: bool = False) -> int:
    """reference src/TinyMPC.jl:243-256 — argument order (input cones, state cones) as in Julia.

    The reference stores the first triple in the core's STATE-cone slots and the second in its
    INPUT-cone slots while deriving the enable flags from the caller's naming;
    both solver types reproduce that."""
    lib = _lib.load()
    Acu_, qcu_, Acx_, qcx_ = _ivec(Acu), _ivec(qcu), _ivec(Acx), _ivec(qcx)
    cu_ = np.ascontiguousarray(np.asarray(cu, dtype=np.float64).reshape(-1))
    cx_ = np.ascontiguousarray(np.asarray(cx, dtype=np.float64).reshape(-1))
    ip = lambda a: a.ctypes.data_as(_IP)
    if isinstance(solver, TinyMPCBatchSolver):
        if not (Acu_.size == qcu_.size == cu_.size and Acx_.size == qcx_.size == cx_.size):
            raise TinyMPCError("Failed to set cone constraints")
        status = lib.tmpc_batch_set_cone_constraints_jl(solver._h, ip(Acu_), ip(qcu_), _dp(cu_), Acu_.size,
                                                        ip(Acx_), ip(qcx_), _dp(cx_), Acx_.size)
    else:
        status = lib.set_cone_constraints(ip(Acu_), Acu_.size, ip(qcu_), qcu_.size, _dp(cu_), cu_.size,
                                          ip(Acx_), Acx_.size, ip(qcx_), qcx_.size, _dp(cx_), cx_.size, 1 if verbose else 0)
    if status != 0:
        raise TinyMPCError("Failed to set cone constraints")
    return status


def set_equality_constraints(solver, Aeq_x, beq_x, Aeq_u=None, beq_u=None) -> int:
    """reference src/TinyMPC.jl:258-267: two inequalities per equality."""
    Aeq_x = np.asarray(Aeq_x, dtype=np.float64).reshape(-1, solver.nx)
    beq_x = np.asarray(beq_x, dtype=np.float64).reshape(-1)
    Aeq_u = np.zeros((0, solver.nu)) if Aeq_u is None else np.asarray(Aeq_u, dtype=np.float64).reshape(-1, solver.nu)
    beq_u = np.zeros(0) if beq_u is None else np.asarray(beq_u, dtype=np.float64).reshape(-1)
    return set_linear_constraints(solver, np.vstack([Aeq_x, -Aeq_x]), np.concatenate([beq_x, -beq_x]),
                                  np.vstack([Aeq_u, -Aeq_u]), np.concatenate([beq_u, -beq_u]))


def set_cache_terms(solver, Kinf, Pinf, Quu_inv, AmBKt, *, verbose: bool = False) -> int:
    """reference src/TinyMPC.jl:276-290: replaces Kinf, Pinf, Quu_inv, AmBKt only."""
    _need_setup(solver)
    lib = _lib.load()
    K, P, Qi, Am = _mat(Kinf), _mat(Pinf), _mat(Quu_inv), _mat(AmBKt)
    if isinstance(solver, TinyMPCBatchSolver):
        n, m = solver.nx, solver.nu
        if K.shape != (m, n) or P.shape != (n, n) or Qi.shape != (m, m) or Am.shape != (n, n):
            raise TinyMPCError("Failed to set cache terms")
        status = lib.tmpc_batch_set_cache_terms(solver._h, _dp(K), _dp(P), _dp(Qi), _dp(Am))
    else:
        status = lib.set_cache_terms(_dp(K), *K.shape, _dp(P), *P.shape, _dp(Qi), *Qi.shape, _dp(Am), *Am.shape, 1 if verbose else 0)
    if status != 0:
        raise TinyMPCError("Failed to set cache terms")
    return status


# ------------------------------------------------------------------------------------------ solve / results

def set_instance_dynamics(solver, A, B, f, Q, R) -> int:
    """Batched solver only: per-instance dynamics.  A (nx, nx, B), B (nx, nu, B), f (nx, B) or None, Q (nx, nx[, B]),
    R (nu, nu[, B]); the LQR cache of every instance is computed on the device (tmpc_batch_set_instance_dynamics)."""
    _need_setup(solver)
    if not isinstance(solver, TinyMPCBatchSolver):
        raise TinyMPCError("set_instance_dynamics needs a batched solver")
    n, m, Bn = solver.nx, solver.nu, solver.B
    Aa = np.asfortranarray(np.asarray(A, dtype=np.float64).reshape((n, n, Bn), order="F"))
    Ba = np.asfortranarray(np.asarray(B, dtype=np.float64).reshape((n, m, Bn), order="F"))
    fa = None if f is None else np.asfortranarray(np.asarray(f, dtype=np.float64).reshape((n, Bn), order="F"))
    Qa = np.asfortranarray(np.asarray(Q, dtype=np.float64))
    Ra = np.asfortranarray(np.asarray(R, dtype=np.float64))
    qb = Bn if Qa.ndim == 3 else 1
    if (Ra.ndim == 3) != (Qa.ndim == 3):
        raise TinyMPCError("Q and R must both be shared or both per instance")
    _chk(_lib.load().tmpc_batch_set_instance_dynamics(solver._h, _dp(Aa), _dp(Ba), _dp(fa) if fa is not None else None,
                                                      _dp(Qa), _dp(Ra), qb), "Failed to set per-instance dynamics")
    return 0


def solve(solver, *, verbose: bool = False) -> int:
    """reference src/TinyMPC.jl:143-148: returns the status WITHOUT raising (0 converged, 1 max_iter).
    Batch solver: 0 when every instance converged, 1 when at least one stopped at max_iter."""
    _need_setup(solver)
    lib = _lib.load()
    if isinstance(solver, TinyMPCBatchSolver):
        return int(lib.tmpc_batch_solve(solver._h))
    return int(lib.solve_mpc(1 if verbose else 0))


def get_solution(solver) -> Solution:
    """reference src/TinyMPC.jl:150-177.  Batch solver: states nx x N x B, controls nu x (N-1) x B."""
    _need_setup(solver)
    lib = _lib.load()
    nx, nu, N = solver.nx, solver.nu, solver.N
    if isinstance(solver, TinyMPCBatchSolver):
        X = np.zeros((nx, N, solver.B), order="F")
        U = np.zeros((nu, N - 1, solver.B), order="F")
        if lib.tmpc_batch_get_states(solver._h, _dp(X)) != 0 or lib.tmpc_batch_get_controls(solver._h, _dp(U)) != 0:
            raise TinyMPCError("Failed to get solution")
        return Solution(X, U)
    sb, cb = np.zeros(nx * N), np.zeros(nu * (N - 1))
    sr, sc, cr, cc = C.c_int(), C.c_int(), C.c_int(), C.c_int()
    s1 = lib.get_states(_dp(sb), C.byref(sr), C.byref(sc))
    s2 = lib.get_controls(_dp(cb), C.byref(cr), C.byref(cc))
    if s1 != 0 or s2 != 0:
        raise TinyMPCError("Failed to get solution")
    return Solution(sb[: sr.value * sc.value].reshape((sr.value, sc.value), order="F"),
                    cb[: cr.value * cc.value].reshape((cr.value, cc.value), order="F"))


def get_solve_info(solver):
    """(iter, solved) of the last solve — the reference only prints these (src/bindings.cpp:236-237)."""
    if isinstance(solver, TinyMPCBatchSolver):
        return solver.status()
    it, so = C.c_int(), C.c_int()
    if _lib.load().get_solve_info(C.byref(it), C.byref(so)) != 0:
        raise TinyMPCError("Solver not setup")
    return it.value, so.value


def print_problem_data(solver, *, verbose: bool = False) -> int:
    _need_setup(solver)
    if isinstance(solver, TinyMPCBatchSolver):
        st, it = solver.status()
        print(f"=== TinyMPC batch: B={solver.B} nx={solver.nx} nu={solver.nu} N={solver.N} rho={solver.rho} "
              f"solved={int((st == 0).sum())}/{solver.B} iters[min/med/max]={it.min()}/{int(np.median(it))}/{it.max()}")
        return 0
    return int(_lib.load().print_problem_data(1 if verbose else 0))


def _solve_lqr(A, B, Q, R, rho_val: float):
    """reference src/TinyMPC.jl:328-351 (private helper of the Julia package): fixed-point Riccati iteration with the
    rho-regularised costs, <= 5000 sweeps, stop when ||K - K_prev|| < 1e-10; returns K, P, C1 = Quu^-1, C2 = (A - B K)'."""
    nx, nu = A.shape[0], B.shape[1]
    Q_rho = Q + rho_val * np.eye(nx)
    R_rho = R + rho_val * np.eye(nu)
    P = Q_rho.copy()
    K = np.zeros((nu, nx))
    for it in range(5000):
        K_prev = K.copy()
        K = np.linalg.solve(R_rho + B.T @ P @ B + 1e-8 * np.eye(nu), B.T @ P @ A)
        P = Q_rho + A.T @ P @ (A - B @ K)
        if it > 0 and np.linalg.norm(K - K_prev) < 1e-10:
            break
    C1 = np.linalg.inv(R_rho + B.T @ P @ B)
    C2 = (A - B @ K).T
    return K, P, C1, C2


def lqr_sensitivity(A, B, Q, R, rho: float, h: float = 1e-6):
    """Forward differences of (K, P, C1, C2) in rho, step 1e-6: the arithmetic of compute_sensitivity_autograd
    (reference src/TinyMPC.jl:301-323) on plain matrices.  Returns (dK, dP, dC1, dC2), column-major float64."""
    A, B, Q, R = (np.asarray(m, dtype=np.float64) for m in (A, B, Q, R))
    B = B.reshape(A.shape[0], -1)
    a0 = _solve_lqr(A, B, Q, R, float(rho))
    a1 = _solve_lqr(A, B, Q, R, float(rho) + h)
    return tuple(np.asfortranarray((y - x) / h) for x, y in zip(a0, a1))


def compute_sensitivity_autograd(solver):
    """reference src/TinyMPC.jl:301-323 — numerical differentiation, the name notwithstanding."""
    _need_setup(solver)
    if isinstance(solver, TinyMPCBatchSolver):
        return lqr_sensitivity(solver.A_, solver.B_, solver.Q_, solver.R_, solver.rho)
    return lqr_sensitivity(solver.A, solver.B, solver.Q, solver.R, solver.rho)


def set_sensitivity(solver, dK, dP, dC1, dC2, *, verbose: bool = False) -> int:
    """Loads the sensitivity matrices of the adaptive-rho branch into the live solver: the half of the reference's
    codegen_with_sensitivity (codegen.cpp:90-97) that the solve path needs, without the code generation."""
    _need_setup(solver)
    lib = _lib.load()
    nx, nu = solver.nx, solver.nu
    mats = [_mat(dK, (nu, nx)), _mat(dP, (nx, nx)), _mat(dC1, (nu, nu)), _mat(dC2, (nx, nx))]
    if isinstance(solver, TinyMPCBatchSolver):
        status = lib.tmpc_batch_set_sensitivity(solver._h, *[_dp(m) for m in mats])
    else:
        args = []
        for m in mats:
            args += [_dp(m), m.shape[0], m.shape[1]]
        status = lib.set_sensitivity_matrices(*args, 1 if verbose else 0)
    if status != 0:
        raise TinyMPCError(f"Failed to set sensitivity matrices ({_lib.last_error()})")
    return 0


def get_adapted_cache(solver):
    """rho [B], Kinf [nu, nx, B], Pinf [nx, nx, B] every instance of a batch currently solves with (adaptive rho)."""
    _need_setup(solver)
    nx, nu, B = solver.nx, solver.nu, solver.B
    rho = np.zeros(B)
    K = np.zeros((nu, nx, B), order="F")
    P = np.zeros((nx, nx, B), order="F")
    _chk(_lib.load().tmpc_batch_get_adapted_cache(solver._h, _dp(rho), _dp(K), _dp(P)), "get_adapted_cache")
    return rho, K, P


def codegen_with_sensitivity(solver, output_dir: str, dK, dP, dC1, dC2, *, verbose: bool = False) -> int:
    """reference src/TinyMPC.jl:374-395.  The library stores the matrices when adaptive_rho is on (as the reference
    does before it writes any file) and then reports that code generation is not part of it: raises like a failed
    codegen, with the matrices in place."""
    _need_setup(solver)
    if isinstance(solver, TinyMPCBatchSolver):
        raise TinyMPCError("Code generation with sensitivity failed (batch solvers: use set_sensitivity)")
    nx, nu = solver.nx, solver.nu
    args = []
    for m in (_mat(dK, (nu, nx)), _mat(dP, (nx, nx)), _mat(dC1, (nu, nu)), _mat(dC2, (nx, nx))):
        args += [_dp(m), m.shape[0], m.shape[1]]
    status = _lib.load().codegen_with_sensitivity(str(output_dir).encode(), *args, 1 if verbose else 0)
    if status != 0:
        raise TinyMPCError("Code generation with sensitivity failed")
    return status


def codegen(solver, output_dir: str, *, verbose: bool = False) -> int:
    """reference src/TinyMPC.jl:357-366.  Not part of the CUDA library: raises like a failed codegen."""
    _need_setup(solver)
    status = _lib.load().codegen(str(output_dir).encode(), 1 if verbose else 0)
    if status != 0:
        raise TinyMPCError("Code generation failed")
    return status


def cleanup():
    try:
        _lib.load().cleanup_solver()
    except Exception:
        pass


