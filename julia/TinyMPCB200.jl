# TinyMPCB200.jl — batched methods for the B200 library, to be loaded next to the reference package.
#
# The reference module (src/TinyMPC.jl) keeps working unchanged on lib/libtinympc_jl.so built from this repository
# (same sixteen symbols, INTEGRATION.md §1).  This file adds the batched solver type and one thin method per
# `tmpc_batch_*` entry point of include/tinympc_b200.h, written in the style of the reference's own wrappers
# (reference src/TinyMPC.jl:55-287): Julia values in, `ccall`, `error()` on a non-zero status.
#
# Not executed on the build machines (no Julia there); tinympc_julia_b200/solver.py performs the same calls and is
# what the tests drive.
module TinyMPCB200

using Libdl

export TinyMPCBatchSolver, setup_batch, set_x0, set_x_ref, set_u_ref, set_bound_constraints, set_linear_constraints,
       set_cone_constraints, set_cache_terms, update_settings, reset!, solve, get_solution, get_status,
       set_kernel, set_precision, set_order, rollout!, solve_host, solve_host_u0, set_instance_dynamics, destroy!,
       set_sensitivity, get_adapted_cache

const lib_path = joinpath(dirname(@__DIR__), "lib", "libtinympc_jl." * Libdl.dlext)   # as reference src/TinyMPC.jl:12-13

const KERNEL_AUTO, KERNEL_THREAD, KERNEL_SUBWARP, KERNEL_TMEM, KERNEL_CONE, KERNEL_WARP = Int32(0), Int32(1), Int32(2), Int32(3), Int32(4), Int32(5)

mutable struct TinyMPCBatchSolver
    handle::Ptr{Cvoid}
    nx::Int
    nu::Int
    N::Int
    B::Int
    rho::Float64
end

last_error() = unsafe_string(ccall((:tmpc_last_error, lib_path), Cstring, ()))
check(status, msg) = status == 0 ? status : error("$msg ($(last_error()))")
f64(a) = Array{Float64}(a)

"B instances sharing A, B, f, Q, R, rho; same meaning as `setup` (reference src/TinyMPC.jl:55-113)."
function setup_batch(A, Bm, f, Q, R, rho::Real, nx::Integer, nu::Integer, N::Integer, B::Integer;
                     device::Integer = -1, verbose::Bool = false, kwargs...)
    h = Ref{Ptr{Cvoid}}(C_NULL)
    status = ccall((:tmpc_batch_create, lib_path), Int32,
                   (Ref{Ptr{Cvoid}}, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}, Ptr{Float64},
                    Float64, Int32, Int32, Int32, Int32, Int32, Int32),
                   h, f64(A), f64(reshape(Bm, nx, nu)), f64(vec(f)), f64(Q), f64(reshape(R, nu, nu)),
                   Float64(rho), nx, nu, N, B, device, verbose ? 1 : 0)
    status == 0 || error("Setup failed with status: $status ($(last_error()))")
    s = TinyMPCBatchSolver(h[], nx, nu, N, B, Float64(rho))
    finalizer(destroy!, s)
    update_settings(s; kwargs...)          # the reference's setup pushes its defaults and clears every en_* flag (:89-104)
    return s
end

function destroy!(s::TinyMPCBatchSolver)
    if s.handle != C_NULL
        ccall((:tmpc_batch_destroy, lib_path), Int32, (Ptr{Cvoid},), s.handle)
        s.handle = C_NULL
    end
    return nothing
end

"NOT incremental, like the reference (src/TinyMPC.jl:181-211): keywords not passed return to their defaults."
function update_settings(s::TinyMPCBatchSolver; abs_pri_tol = 1e-3, abs_dua_tol = 1e-3, max_iter = 100,
                         check_termination = true, en_state_bound = false, en_input_bound = false,
                         en_state_soc = false, en_input_soc = false, en_state_linear = false, en_input_linear = false,
                         adaptive_rho = false, adaptive_rho_min = 0.1, adaptive_rho_max = 10.0,
                         adaptive_rho_enable_clipping = true)
    check(ccall((:tmpc_batch_update_settings, lib_path), Int32,
                (Ptr{Cvoid}, Float64, Float64, Int32, Int32, Int32, Int32, Int32, Int32, Int32, Int32, Int32, Float64, Float64, Int32),
                s.handle, abs_pri_tol, abs_dua_tol, max_iter, check_termination ? 1 : 0, en_state_bound ? 1 : 0,
                en_input_bound ? 1 : 0, en_state_soc ? 1 : 0, en_input_soc ? 1 : 0, en_state_linear ? 1 : 0,
                en_input_linear ? 1 : 0, adaptive_rho ? 1 : 0, adaptive_rho_min, adaptive_rho_max,
                adaptive_rho_enable_clipping ? 1 : 0), "Failed to update settings")
end

"x0: nx x B"
set_x0(s::TinyMPCBatchSolver, x0) =
    check(ccall((:tmpc_batch_set_x0, lib_path), Int32, (Ptr{Cvoid}, Ptr{Float64}, Int32, Int32),
                s.handle, f64(x0), size(x0, 1), size(x0, 2)), "Failed to set initial state")

batchdim(a) = ndims(a) == 3 ? size(a, 3) : 1

"x_ref: nx x N (shared) or nx x N x B"
set_x_ref(s::TinyMPCBatchSolver, x_ref) =
    check(ccall((:tmpc_batch_set_x_ref, lib_path), Int32, (Ptr{Cvoid}, Ptr{Float64}, Int32, Int32, Int32),
                s.handle, f64(x_ref), size(x_ref, 1), size(x_ref, 2), batchdim(x_ref)), "Failed to set state reference")

"u_ref: nu x (N-1) (shared) or nu x (N-1) x B"
set_u_ref(s::TinyMPCBatchSolver, u_ref) =
    check(ccall((:tmpc_batch_set_u_ref, lib_path), Int32, (Ptr{Cvoid}, Ptr{Float64}, Int32, Int32, Int32),
                s.handle, f64(u_ref), size(u_ref, 1), size(u_ref, 2), batchdim(u_ref)), "Failed to set input reference")

"bounds shared (matrices) or per instance (3-D arrays); enables both bound flags (reference src/bindings.cpp:400-404)"
set_bound_constraints(s::TinyMPCBatchSolver, x_min, x_max, u_min, u_max) =
    check(ccall((:tmpc_batch_set_bound_constraints, lib_path), Int32,
                (Ptr{Cvoid}, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}, Int32),
                s.handle, f64(x_min), f64(x_max), f64(u_min), f64(u_max), batchdim(x_min)), "Failed to set bound constraints")

set_linear_constraints(s::TinyMPCBatchSolver, Alin_x, blin_x, Alin_u, blin_u) =
    check(ccall((:tmpc_batch_set_linear_constraints, lib_path), Int32,
                (Ptr{Cvoid}, Ptr{Float64}, Int32, Ptr{Float64}, Ptr{Float64}, Int32, Ptr{Float64}),
                s.handle, f64(Alin_x), length(blin_x), f64(blin_x), f64(Alin_u), length(blin_u), f64(blin_u)),
          "Failed to set linear constraints")

"Argument order and (crossed) behaviour of the reference's set_cone_constraints (src/TinyMPC.jl:243-257, SURVEY.md Q1)."
set_cone_constraints(s::TinyMPCBatchSolver, Acu, qcu, cu, Acx, qcx, cx) =
    check(ccall((:tmpc_batch_set_cone_constraints_jl, lib_path), Int32,
                (Ptr{Cvoid}, Ptr{Int32}, Ptr{Int32}, Ptr{Float64}, Int32, Ptr{Int32}, Ptr{Int32}, Ptr{Float64}, Int32),
                s.handle, Vector{Int32}(Acu), Vector{Int32}(qcu), f64(cu), length(Acu),
                Vector{Int32}(Acx), Vector{Int32}(qcx), f64(cx), length(Acx)), "Failed to set cone constraints")

set_cache_terms(s::TinyMPCBatchSolver, Kinf, Pinf, Quu_inv, AmBKt) =
    check(ccall((:tmpc_batch_set_cache_terms, lib_path), Int32,
                (Ptr{Cvoid}, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}),
                s.handle, f64(Kinf), f64(Pinf), f64(Quu_inv), f64(AmBKt)), "Failed to set cache terms")

"cold start: back to the workspace `setup` leaves"
reset!(s::TinyMPCBatchSolver) = check(ccall((:tmpc_batch_reset, lib_path), Int32, (Ptr{Cvoid},), s.handle), "Failed to reset")

"arithmetic type of the solve: 64 (`typedef double tinytype`) or 32 (the reference's float build); right after `setup` / `reset!`"
set_precision(s::TinyMPCBatchSolver, bits::Integer) =
    check(ccall((:tmpc_batch_set_precision, lib_path), Int32, (Ptr{Cvoid}, Int32), s.handle, bits), "Failed to set precision")

set_kernel(s::TinyMPCBatchSolver, k::Integer) =
    check(ccall((:tmpc_batch_set_kernel, lib_path), Int32, (Ptr{Cvoid}, Int32), s.handle, k), "Failed to select kernel")

"0 when every instance converged, 1 when at least one stopped at max_iter (never raises, like the reference's solve)"
solve(s::TinyMPCBatchSolver) = ccall((:tmpc_batch_solve, lib_path), Int32, (Ptr{Cvoid},), s.handle)

"(status[B], iterations[B]); status = what `solve` returns for that instance alone"
function get_status(s::TinyMPCBatchSolver)
    st = Vector{Int32}(undef, s.B); it = Vector{Int32}(undef, s.B)
    check(ccall((:tmpc_batch_get_status, lib_path), Int32, (Ptr{Cvoid}, Ptr{Int32}, Ptr{Int32}), s.handle, st, it), "Failed to get status")
    return st, it
end

function get_solution(s::TinyMPCBatchSolver)
    X = Array{Float64}(undef, s.nx, s.N, s.B)
    U = Array{Float64}(undef, s.nu, s.N - 1, s.B)
    check(ccall((:tmpc_batch_get_states, lib_path), Int32, (Ptr{Cvoid}, Ptr{Float64}), s.handle, X), "Failed to get states")
    check(ccall((:tmpc_batch_get_controls, lib_path), Int32, (Ptr{Cvoid}, Ptr{Float64}), s.handle, U), "Failed to get controls")
    return (states = X, controls = U)
end

"""
Per-instance dynamics: A (nx, nx, B), Bdyn (nx, nu, B), f (nx, B), Q (nx, nx[, B]), R (nu, nu[, B]); the LQR cache of every
instance is computed on the device (tmpc_batch_set_instance_dynamics).
"""
function set_instance_dynamics(s::TinyMPCBatchSolver, A::Array{Float64,3}, Bdyn::Array{Float64,3}, f::Matrix{Float64},
                               Q::Array{Float64}, R::Array{Float64})
    qb = ndims(Q) == 3 ? s.B : 1
    check(ccall((:tmpc_batch_set_instance_dynamics, lib_path), Int32,
                (Ptr{Cvoid}, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}, Int32),
                s.handle, A, Bdyn, f, Q, R, Int32(qb)), "Failed to set per-instance dynamics")
end

"""
One call on host arrays: set_x0 (+ cold start) + solve + get_solution + get_status, the batch cut into chunks so that
the solution of one chunk crosses PCIe while the next chunk solves (tmpc_batch_solve_host).
Returns (ret, (states, controls), status, iterations); `ret` as `solve`.
"""
function solve_host(s::TinyMPCBatchSolver, x0::Matrix{Float64}; cold::Bool = false, chunks::Integer = 0)
    X = Array{Float64}(undef, s.nx, s.N, s.B)
    U = Array{Float64}(undef, s.nu, s.N - 1, s.B)
    st = Vector{Int32}(undef, s.B); it = Vector{Int32}(undef, s.B)
    ret = ccall((:tmpc_batch_solve_host, lib_path), Int32,
                (Ptr{Cvoid}, Ptr{Float64}, Int32, Ptr{Float64}, Ptr{Float64}, Ptr{Int32}, Ptr{Int32}, Int32),
                s.handle, x0, Int32(cold), X, U, st, it, Int32(chunks))
    ret in (0, 1) || error("solve_host failed with status $ret")
    return ret, (states = X, controls = U), st, it
end

"""
The same with only the first control of every instance (nu x B) and the status / iteration counts coming home: what an
MPC loop reads from a solve (reference examples/cartpole_example_mpc.jl:41-42).  tmpc_batch_solve_host_select.
"""
function solve_host_u0(s::TinyMPCBatchSolver, x0::Matrix{Float64}; cold::Bool = false)
    u0 = Array{Float64}(undef, s.nu, s.B)
    st = Vector{Int32}(undef, s.B); it = Vector{Int32}(undef, s.B)
    ret = ccall((:tmpc_batch_solve_host_select, lib_path), Int32,
                (Ptr{Cvoid}, Ptr{Float64}, Int32, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}, Ptr{Int32}, Ptr{Int32}, Int32),
                s.handle, x0, Int32(cold), C_NULL, C_NULL, u0, st, it, Int32(0))
    ret in (0, 1) || error("solve_host_u0 failed with status $ret")
    return ret, u0, st, it
end

"""
Adaptive rho (reference admm.cpp:331-357): loads the sensitivity matrices `compute_sensitivity_autograd` returns
(reference src/TinyMPC.jl:301-323) into the batch - the part of the reference's `codegen_with_sensitivity`
(codegen.cpp:90-97) that the solve path needs.  With `update_settings(s; adaptive_rho = true, ...)` every instance then
adapts its own rho / Kinf / Pinf (tmpc_batch_set_sensitivity).  Single solver: `set_sensitivity_matrices` with the
argument list of `codegen_with_sensitivity` minus the directory.
"""
function set_sensitivity(s::TinyMPCBatchSolver, dK::Matrix{Float64}, dP::Matrix{Float64}, dC1::Matrix{Float64},
                         dC2::Matrix{Float64})
    size(dK) == (s.nu, s.nx) && size(dP) == (s.nx, s.nx) && size(dC1) == (s.nu, s.nu) && size(dC2) == (s.nx, s.nx) ||
        error("sensitivity matrices must be dK nu x nx, dP nx x nx, dC1 nu x nu, dC2 nx x nx")
    check(ccall((:tmpc_batch_set_sensitivity, lib_path), Int32,
                (Ptr{Cvoid}, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}), s.handle, dK, dP, dC1, dC2),
          "Failed to set sensitivity matrices")
end

"rho (B), Kinf (nu x nx x B), Pinf (nx x nx x B) every instance currently solves with (tmpc_batch_get_adapted_cache)."
function get_adapted_cache(s::TinyMPCBatchSolver)
    rho = Vector{Float64}(undef, s.B)
    K = Array{Float64}(undef, s.nu, s.nx, s.B); P = Array{Float64}(undef, s.nx, s.nx, s.B)
    check(ccall((:tmpc_batch_get_adapted_cache, lib_path), Int32, (Ptr{Cvoid}, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}),
                s.handle, rho, K, P), "Failed to read the adapted cache")
    return rho, K, P
end

"""
Queue order of the persistent kernels: 0 natural, 1 longest-first (default), 2 `perm` (0-based permutation of 0..B-1).
Results never depend on it (tmpc_batch_set_order).
"""
set_order(s::TinyMPCBatchSolver, mode::Integer; perm::Vector{Int32} = Int32[]) =
    check(ccall((:tmpc_batch_set_order, lib_path), Int32, (Ptr{Cvoid}, Int32, Ptr{Int32}),
                s.handle, mode, isempty(perm) ? C_NULL : pointer(perm)), "Failed to set the queue order")

"""
Closed loop on the device: `nsteps` x (solve, apply controls[:, 1], x <- A x + B u + f), references sliding by one
column per step when `x_ref_long` (a device pointer to nx x (N+nsteps-1) x B) is given
(pattern of reference examples/cartpole_example_mpc.jl:35-51).
"""
function rollout!(s::TinyMPCBatchSolver, nsteps::Integer; x_ref_long::Ptr{Cvoid} = C_NULL,
                  status_out::Ptr{Cvoid} = C_NULL, iters_out::Ptr{Cvoid} = C_NULL, stream::Ptr{Cvoid} = C_NULL)
    check(ccall((:tmpc_batch_rollout_async, lib_path), Int32,
                (Ptr{Cvoid}, Int32, Ptr{Cvoid}, Ptr{Cvoid}, Ptr{Cvoid}, Ptr{Cvoid}),
                s.handle, nsteps, x_ref_long, status_out, iters_out, stream), "Rollout failed")
    check(ccall((:tmpc_batch_synchronize, lib_path), Int32, (Ptr{Cvoid},), s.handle), "Synchronize failed")
end

end # module
