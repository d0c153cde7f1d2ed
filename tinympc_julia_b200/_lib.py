"""ctypes binding of the C-ABI library (include/tinympc_b200.h).

The library is the product: there is no Python or CPU implementation of the solve path behind
this module.  If ``lib/libtinympc_jl.so`` is missing the import of the solver classes fails loudly
(``build()`` compiles it with nvcc for sm_100a).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "lib", "libtinympc_jl.so")   # same file name the Julia package dlopens
CSRC = os.path.join(HERE, "csrc")

c_double_p = C.POINTER(C.c_double)
c_int_p = C.POINTER(C.c_int)

# name -> (restype, argtypes); mirrors include/tinympc_b200.h declaration by declaration
_D, _I, _P, _IP, _V = C.c_double, C.c_int, c_double_p, c_int_p, C.c_void_p
_MAT = [_P, _I, _I]
SIGNATURES = {
    # PART 1: the reference's sixteen symbols (src/bindings.cpp)
    "setup_solver": (_I, _MAT * 5 + [_D, _I, _I, _I, _I]),
    "set_x0": (_I, _MAT + [_I]),
    "set_x_ref": (_I, _MAT + [_I]),
    "set_u_ref": (_I, _MAT + [_I]),
    "solve_mpc": (_I, [_I]),
    "get_states": (_I, [_P, _IP, _IP]),
    "get_controls": (_I, [_P, _IP, _IP]),
    "cleanup_solver": (None, []),
    "codegen": (_I, [C.c_char_p, _I]),
    "codegen_with_sensitivity": (_I, [C.c_char_p] + _MAT * 4 + [_I]),
    "set_sensitivity_matrices": (_I, _MAT * 4 + [_I]),
    "print_problem_data": (_I, [_I]),
    "set_cache_terms": (_I, _MAT * 4 + [_I]),
    "update_settings": (_I, [_D, _D, _I, _I, _I, _I, _I, _I, _I, _I, _I, _D, _D, _I, _I]),
    "set_bound_constraints": (_I, _MAT * 4 + [_I]),
    "set_linear_constraints": (_I, _MAT + [_P, _I] + _MAT + [_P, _I, _I]),
    "set_cone_constraints": (_I, [_IP, _I, _IP, _I, _P, _I, _IP, _I, _IP, _I, _P, _I, _I]),
    "get_solve_info": (_I, [_IP, _IP]),
    # PART 2: batched variants
    "tmpc_compute_cache": (_I, [_P, _P, _P, _P, _P, _D, _I, _I] + [_P] * 8),
    "tmpc_compute_cache_f32": (_I, [_P, _P, _P, _P, _P, _D, _I, _I] + [_P] * 8),
    "tmpc_batch_create": (_I, [C.POINTER(_V), _P, _P, _P, _P, _P, _D, _I, _I, _I, _I, _I, _I]),
    "tmpc_batch_destroy": (_I, [_V]),
    "tmpc_batch_update_settings": (_I, [_V, _D, _D, _I, _I, _I, _I, _I, _I, _I, _I, _I, _D, _D, _I]),
    "tmpc_batch_set_precision": (_I, [_V, _I]),
    "tmpc_batch_get_precision": (_I, [_V]),
    "tmpc_batch_set_kernel": (_I, [_V, _I]),
    "tmpc_batch_get_kernel": (_I, [_V]),
    "tmpc_batch_set_order": (_I, [_V, _I, _IP]),
    "tmpc_batch_get_order": (_I, [_V, _IP]),
    "tmpc_batch_set_x0": (_I, [_V, _P, _I, _I]),
    "tmpc_batch_set_x_ref": (_I, [_V, _P, _I, _I, _I]),
    "tmpc_batch_set_u_ref": (_I, [_V, _P, _I, _I, _I]),
    "tmpc_batch_set_bound_constraints": (_I, [_V, _P, _P, _P, _P, _I]),
    "tmpc_batch_set_linear_constraints": (_I, [_V, _P, _I, _P, _P, _I, _P]),
    "tmpc_batch_set_cone_constraints": (_I, [_V, _IP, _IP, _P, _I, _IP, _IP, _P, _I]),
    "tmpc_batch_set_cone_constraints_jl": (_I, [_V, _IP, _IP, _P, _I, _IP, _IP, _P, _I]),
    "tmpc_batch_get_settings": (_I, [_V, _P, _IP]),
    "tmpc_batch_set_cache_terms": (_I, [_V, _P, _P, _P, _P]),
    "tmpc_batch_get_cache_terms": (_I, [_V, _P, _P, _P, _P, _P, _P]),
    "tmpc_batch_set_sensitivity": (_I, [_V, _P, _P, _P, _P]),
    "tmpc_batch_get_adapted_cache": (_I, [_V, _P, _P, _P]),
    "tmpc_batch_reset": (_I, [_V]),
    # no counterpart in the reference: device pointers are passed as integers
    "tmpc_solve_noalloc": (_I, [_I, _I, _I, _I, _D] + [_P] * 11 + [_V] * 7 + [_D, _D, _I, _I] + [_V] * 5 + [C.c_char_p, C.c_size_t]),
    "tmpc_batch_solve": (_I, [_V]),
    "tmpc_batch_solve_async": (_I, [_V, _V]),
    "tmpc_batch_set_instance_dynamics": (_I, [_V, _P, _P, _P, _P, _P, _I]),
    "tmpc_batch_get_instance_cache": (_I, [_V, _I, _P, _P, _P, _P, _P, _P]),
    "tmpc_batch_set_rollout_fused": (_I, [_V, _I]),
    "tmpc_batch_solve_host": (_I, [_V, _P, _I, _P, _P, _IP, _IP, _I]),
    "tmpc_batch_solve_host_select": (_I, [_V, _P, _I, _P, _P, _P, _IP, _IP, _I]),
    "tmpc_batch_synchronize": (_I, [_V]),
    "tmpc_batch_get_status": (_I, [_V, _IP, _IP]),
    "tmpc_batch_get_states": (_I, [_V, _P]),
    "tmpc_batch_get_controls": (_I, [_V, _P]),
    "tmpc_batch_bind_x0_device": (_I, [_V, _V]),
    "tmpc_batch_bind_x_ref_device": (_I, [_V, _V, _I]),
    "tmpc_batch_bind_u_ref_device": (_I, [_V, _V, _I]),
    "tmpc_batch_device_ptr": (_V, [_V, _I]),
    "tmpc_batch_rollout_async": (_I, [_V, _I, _V, _V, _V, _V]),
    "tmpc_batch_get_x0": (_I, [_V, _V, _I]),
    "tmpc_batch_dims": (_I, [_V, _IP, _IP, _IP, _IP]),
    "tmpc_batch_launch_count": (C.c_longlong, [_V]),
    "tmpc_host_alloc": (_V, [C.c_size_t]),
    "tmpc_host_free": (None, [_V]),
    "tmpc_measure_fma_peak": (_I, [_I, _I, _I, _P]),
    "tmpc_last_error": (C.c_char_p, []),
    "tmpc_version": (C.c_char_p, []),
}

_lib = None


def build(verbose: bool = False) -> str:
    """Compile the CUDA library in-tree (nvcc, sm_100a).  Returns the path of the .so."""
    subprocess.run(["make", "-C", CSRC, "-j8"], check=True, stdout=None if verbose else subprocess.DEVNULL)
    return LIB_PATH


def load() -> C.CDLL:
    """dlopen the library and attach the prototypes.  Raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: the CUDA library is the only implementation of the solve path "
            "(no CPU fallback). Build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            f"or `make -C {CSRC}`.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError here = a declared symbol is not exported
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def last_error() -> str:
    return load().tmpc_last_error().decode()
