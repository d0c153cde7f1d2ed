"""tinympc_julia_b200 — Blackwell-native batched TinyMPC ADMM solve path.

The product is the C-ABI CUDA library ``lib/libtinympc_jl.so`` (include/tinympc_b200.h); this
package is the thin host-side mirror of the Julia interface that binds it (solver.py) plus the
problem fixtures of the named configurations (problems.py).
"""
from . import problems  # noqa: F401
from .solver import (  # noqa: F401
    Solution, TinyMPCBatchSolver, TinyMPCError, TinyMPCSolver, cleanup, codegen, get_solution, get_solve_info,
    print_problem_data, set_bound_constraints, set_cache_terms, set_cone_constraints, set_equality_constraints,
    set_instance_dynamics, set_linear_constraints, set_u_ref, set_x0, set_x_ref, setup, solve, update_settings,
    codegen_with_sensitivity, compute_sensitivity_autograd, get_adapted_cache, lqr_sensitivity, set_sensitivity,
)

KERNEL_AUTO, KERNEL_THREAD, KERNEL_SUBWARP, KERNEL_TMEM, KERNEL_CONE, KERNEL_WARP = 0, 1, 2, 3, 4, 5
