"""TEST INFRASTRUCTURE — ctypes front end for the two checkers under oracle/.

* ``liboracle.so``                 — plain-C restatement (oracle/tinympc_oracle.c), prefix ``orc_``
* ``_ref/libtinympc_refbatch.so``  — the unmodified reference core + our batch driver
                                     (oracle/ref_driver.cpp), prefix ``ref_``
* ``_ref/libtinympc_jl.so``        — the reference's own C ABI (src/bindings.cpp), i.e. exactly
                                     the library Julia dlopens (src/TinyMPC.jl:12-14)

Only tests/, ``__graft_entry__.smoke()`` and bench.py's CPU-baseline legs may import this
package.  Nothing under ``tinympc_julia_b200/`` does.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from dataclasses import dataclass, field
from typing import Optional, Sequence

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_SO = os.path.join(HERE, "liboracle.so")
REFBATCH32_SO = os.path.join(HERE, "_ref", "libtinympc_refbatch32.so")
REFBATCH_SO = os.path.join(HERE, "_ref", "libtinympc_refbatch.so")
REFJL_SO = os.path.join(HERE, "_ref", "libtinympc_jl.so")
REFERENCE_ROOT = "/root/reference"

REF_ZERO, REF_SHARED, REF_PER_INSTANCE, REF_SLIDING = 0, 1, 2, 3

_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int)


class _CProblem(C.Structure):
    """Mirror of ``tmpc_problem`` in oracle/batch_problem.h (field order matters)."""
    _fields_ = [
        ("nx", C.c_int), ("nu", C.c_int), ("N", C.c_int), ("rho", C.c_double),
        ("A", _dp), ("B", _dp), ("f", _dp), ("Q", _dp), ("R", _dp),
        ("abs_pri_tol", C.c_double), ("abs_dua_tol", C.c_double),
        ("max_iter", C.c_int), ("check_termination", C.c_int),
        ("en_state_bound", C.c_int), ("en_input_bound", C.c_int),
        ("en_state_soc", C.c_int), ("en_input_soc", C.c_int),
        ("en_state_linear", C.c_int), ("en_input_linear", C.c_int),
        ("x_min", _dp), ("x_max", _dp), ("u_min", _dp), ("u_max", _dp),
        ("bounds_per_instance", C.c_int),
        ("num_state_cones", C.c_int), ("Acx", _ip), ("qcx", _ip), ("cx", _dp),
        ("num_input_cones", C.c_int), ("Acu", _ip), ("qcu", _ip), ("cu", _dp),
        ("num_state_linear", C.c_int), ("Alin_x", _dp), ("blin_x", _dp),
        ("num_input_linear", C.c_int), ("Alin_u", _dp), ("blin_u", _dp),
        ("Kinf", _dp), ("Pinf", _dp), ("Quu_inv", _dp), ("AmBKt", _dp),
        ("adaptive_rho", C.c_int), ("adaptive_rho_min", C.c_double), ("adaptive_rho_max", C.c_double),
        ("adaptive_rho_enable_clipping", C.c_int),
        ("dKinf_drho", _dp), ("dPinf_drho", _dp), ("dC1_drho", _dp), ("dC2_drho", _dp),
        ("rho_final", _dp), ("Kinf_final", _dp), ("Pinf_final", _dp),
    ]


def _f(a) -> np.ndarray:
    """float64, Fortran (column-major) contiguous — Julia's Matrix{Float64} layout."""
    return np.asfortranarray(np.asarray(a, dtype=np.float64))


@dataclass
class Problem:
    """Core-level problem description (what TinySolver holds after the C-ABI glue ran)."""
    A: np.ndarray
    B: np.ndarray
    Q: np.ndarray
    R: np.ndarray
    rho: float
    N: int
    f: Optional[np.ndarray] = None
    abs_pri_tol: float = 1e-3
    abs_dua_tol: float = 1e-3
    max_iter: int = 100
    check_termination: int = 1
    en_state_bound: int = 0
    en_input_bound: int = 0
    en_state_soc: int = 0
    en_input_soc: int = 0
    en_state_linear: int = 0
    en_input_linear: int = 0
    x_min: Optional[np.ndarray] = None
    x_max: Optional[np.ndarray] = None
    u_min: Optional[np.ndarray] = None
    u_max: Optional[np.ndarray] = None
    bounds_per_instance: bool = False
    Acx: Sequence[int] = ()
    qcx: Sequence[int] = ()
    cx: Sequence[float] = ()
    Acu: Sequence[int] = ()
    qcu: Sequence[int] = ()
    cu: Sequence[float] = ()
    Alin_x: Optional[np.ndarray] = None
    blin_x: Optional[np.ndarray] = None
    Alin_u: Optional[np.ndarray] = None
    blin_u: Optional[np.ndarray] = None
    cache: Optional[dict] = None  # {"Kinf","Pinf","Quu_inv","AmBKt"} override (set_cache_terms semantics)
    adaptive_rho: int = 0            # types.hpp:76-79; defaults as tiny_api.cpp:367-370
    adaptive_rho_min: float = 1.0
    adaptive_rho_max: float = 100.0
    adaptive_rho_enable_clipping: int = 1
    sensitivity: Optional[dict] = None   # {"dK","dP","dC1","dC2"}: what codegen_with_sensitivity stores (codegen.cpp:91-96)
    _keep: list = field(default_factory=list, repr=False)

    @property
    def nx(self) -> int:
        return int(np.asarray(self.A).shape[0])

    @property
    def nu(self) -> int:
        return int(np.asarray(self.B).reshape(self.nx, -1).shape[1])

    def with_bounds(self, x_min, x_max, u_min, u_max, per_instance=False) -> "Problem":
        """Same effect as set_bound_constraints through the C ABI (bindings.cpp:400-404)."""
        self.x_min, self.x_max, self.u_min, self.u_max = x_min, x_max, u_min, u_max
        self.bounds_per_instance = per_instance
        self.en_state_bound = self.en_input_bound = 1
        return self

    def to_c(self) -> _CProblem:
        nx, nu, N = self.nx, self.nu, self.N
        keep = self._keep
        keep.clear()

        def dp(a, shape=None):
            if a is None:
                return None
            arr = _f(a) if shape is None else _f(np.asarray(a, dtype=np.float64).reshape(shape))
            keep.append(arr)
            return arr.ctypes.data_as(_dp)

        def ip(a):
            arr = np.ascontiguousarray(np.asarray(a, dtype=np.int32))
            keep.append(arr)
            return arr.ctypes.data_as(_ip)

        c = _CProblem()
        c.nx, c.nu, c.N, c.rho = nx, nu, N, float(self.rho)
        c.A, c.B = dp(self.A, (nx, nx)), dp(np.asarray(self.B, dtype=np.float64).reshape(nx, nu))
        c.f = dp(np.zeros(nx) if self.f is None else np.asarray(self.f, dtype=np.float64).reshape(nx))
        c.Q, c.R = dp(self.Q, (nx, nx)), dp(self.R, (nu, nu))
        c.abs_pri_tol, c.abs_dua_tol = float(self.abs_pri_tol), float(self.abs_dua_tol)
        c.max_iter, c.check_termination = int(self.max_iter), int(self.check_termination)
        for k in ("en_state_bound", "en_input_bound", "en_state_soc", "en_input_soc", "en_state_linear", "en_input_linear"):
            setattr(c, k, int(getattr(self, k)))
        c.x_min, c.x_max, c.u_min, c.u_max = dp(self.x_min), dp(self.x_max), dp(self.u_min), dp(self.u_max)
        c.bounds_per_instance = int(bool(self.bounds_per_instance))
        c.num_state_cones, c.Acx, c.qcx, c.cx = len(self.Acx), ip(self.Acx), ip(self.qcx), dp(np.asarray(self.cx, dtype=np.float64))
        c.num_input_cones, c.Acu, c.qcu, c.cu = len(self.Acu), ip(self.Acu), ip(self.qcu), dp(np.asarray(self.cu, dtype=np.float64))
        nlx = 0 if self.Alin_x is None else int(np.asarray(self.Alin_x).reshape(-1, nx).shape[0])
        nlu = 0 if self.Alin_u is None else int(np.asarray(self.Alin_u).reshape(-1, nu).shape[0])
        c.num_state_linear, c.num_input_linear = nlx, nlu
        c.Alin_x = dp(np.zeros((0, nx)) if nlx == 0 else np.asarray(self.Alin_x, dtype=np.float64).reshape(nlx, nx))
        c.blin_x = dp(np.zeros(0) if nlx == 0 else self.blin_x)
        c.Alin_u = dp(np.zeros((0, nu)) if nlu == 0 else np.asarray(self.Alin_u, dtype=np.float64).reshape(nlu, nu))
        c.blin_u = dp(np.zeros(0) if nlu == 0 else self.blin_u)
        if self.cache is not None:
            c.Kinf, c.Pinf = dp(self.cache["Kinf"], (nu, nx)), dp(self.cache["Pinf"], (nx, nx))
            c.Quu_inv, c.AmBKt = dp(self.cache["Quu_inv"], (nu, nu)), dp(self.cache["AmBKt"], (nx, nx))
        c.adaptive_rho = int(self.adaptive_rho)
        c.adaptive_rho_min, c.adaptive_rho_max = float(self.adaptive_rho_min), float(self.adaptive_rho_max)
        c.adaptive_rho_enable_clipping = int(self.adaptive_rho_enable_clipping)
        if self.sensitivity is not None:
            sv = self.sensitivity
            c.dKinf_drho, c.dPinf_drho = dp(sv["dK"], (nu, nx)), dp(sv["dP"], (nx, nx))
            c.dC1_drho, c.dC2_drho = dp(sv["dC1"], (nu, nu)), dp(sv["dC2"], (nx, nx))
        return c


class BatchLib:
    """One of the two checkers, addressed through the common batch contract."""

    def __init__(self, path: str, prefix: str, kind: str):
        self.path, self.prefix, self.kind = path, prefix, kind
        self.lib = C.CDLL(path)
        self._run = getattr(self.lib, prefix + "_batch_run")
        self._run.restype = C.c_int
        self._run.argtypes = [C.POINTER(_CProblem), C.c_int, C.c_int, _dp, _dp, _dp, C.c_int, C.c_int,
                              _ip, _ip, _dp, _dp, _dp, _dp]
        self._cache = getattr(self.lib, prefix + "_get_cache")
        self._cache.restype = C.c_int
        self._cache.argtypes = [C.POINTER(_CProblem)] + [_dp] * 6

    def get_cache(self, prob: Problem) -> dict:
        nx, nu = prob.nx, prob.nu
        out = {"Kinf": np.zeros((nu, nx), order="F"), "Pinf": np.zeros((nx, nx), order="F"),
               "Quu_inv": np.zeros((nu, nu), order="F"), "AmBKt": np.zeros((nx, nx), order="F"),
               "APf": np.zeros(nx), "BPf": np.zeros(nu)}
        c = prob.to_c()
        rc = self._cache(C.byref(c), *[out[k].ctypes.data_as(_dp) for k in ("Kinf", "Pinf", "Quu_inv", "AmBKt", "APf", "BPf")])
        if rc != 0:
            raise RuntimeError(f"{self.prefix}_get_cache failed with status {rc}")
        return out

    def batch_run(self, prob: Problem, x0, Xref=None, Uref=None, ref_mode=REF_ZERO, nsteps=1, nthreads=1) -> dict:
        """x0: (nx, B).  Returns dict(ret, iters [nsteps,B], X [nx,N,B], U [nu,N-1,B], x_final [nx,B], seconds)."""
        nx, nu, N = prob.nx, prob.nu, prob.N
        x0 = _f(np.asarray(x0, dtype=np.float64).reshape(nx, -1))
        B = x0.shape[1]
        ret = np.zeros((nsteps, B), dtype=np.int32)
        iters = np.zeros((nsteps, B), dtype=np.int32)
        X = np.zeros((nx, N, B), order="F")
        U = np.zeros((nu, N - 1, B), order="F")
        xf = np.zeros((nx, B), order="F")
        sec = C.c_double(0.0)
        xr = None if Xref is None else _f(Xref)
        ur = None if Uref is None else _f(Uref)
        c = prob.to_c()
        adapt = {}
        if prob.adaptive_rho:
            adapt = {"rho": np.zeros(B), "Kinf": np.zeros((nu, nx, B), order="F"), "Pinf": np.zeros((nx, nx, B), order="F")}
            c.rho_final, c.Kinf_final, c.Pinf_final = (adapt[k].ctypes.data_as(_dp) for k in ("rho", "Kinf", "Pinf"))
        rc = self._run(C.byref(c), B, int(nsteps), x0.ctypes.data_as(_dp),
                       None if xr is None else xr.ctypes.data_as(_dp), None if ur is None else ur.ctypes.data_as(_dp),
                       int(ref_mode), int(nthreads), ret.ctypes.data_as(_ip), iters.ctypes.data_as(_ip),
                       X.ctypes.data_as(_dp), U.ctypes.data_as(_dp), xf.ctypes.data_as(_dp), C.byref(sec))
        if rc != 0:
            raise RuntimeError(f"{self.prefix}_batch_run rejected the problem (status {rc})")
        out = {"ret": ret, "iters": iters, "X": X, "U": U, "x_final": xf, "seconds": sec.value}
        if adapt:
            out["adapted"] = adapt
        return out


def build(ref: bool = True, quiet: bool = True) -> None:
    """Compile the checkers (``make -C oracle``).  The reference part needs /root/reference and is
    skipped when that tree is absent (e.g. on the GPU box, which uses the prebuilt oracle/_ref)."""
    targets = ["oracle"]
    if ref and os.path.isdir(os.path.join(REFERENCE_ROOT, "tinympc", "TinyMPC", "src")):
        targets += ["ref", "ref32"]
    subprocess.run(["make", "-C", HERE, "-j8"] + targets, check=True,
                   stdout=subprocess.DEVNULL if quiet else None)


def load_oracle() -> BatchLib:
    if not os.path.exists(ORACLE_SO):
        build(ref=False)
    return BatchLib(ORACLE_SO, "orc", "port")


def load_ref() -> Optional[BatchLib]:
    """The reference-backed checker, or None when oracle/_ref has not been built."""
    if not os.path.exists(REFBATCH_SO):
        return None
    return BatchLib(REFBATCH_SO, "ref", "reference")


def load_ref32() -> Optional[BatchLib]:
    """The FLOAT build of the reference (`typedef float tinytype`; oracle/Makefile target ref32), same batch
    contract with doubles at the boundary (cast where src/bindings.cpp casts), or None when it has not been built."""
    if not os.path.exists(REFBATCH32_SO):
        return None
    return BatchLib(REFBATCH32_SO, "ref32", "reference-f32")
