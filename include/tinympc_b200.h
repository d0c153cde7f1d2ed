/*
 * tinympc_b200 — C ABI of the Blackwell-native batched TinyMPC ADMM solver.
 *
 * The library is built as  tinympc_julia_b200/lib/libtinympc_jl.so  — the file name and
 * location the Julia package dlopens (reference: src/TinyMPC.jl:12-14).
 *
 * PART 1 re-exports, symbol for symbol and signature for signature, the sixteen `extern "C"`
 * entry points of the reference glue (reference: src/bindings.cpp) so that the unmodified
 * `src/TinyMPC.jl` drives the CUDA solver through its existing `ccall`s.  They operate on one
 * process-wide solver (reference: src/bindings.cpp:15) which here is a batch of B = 1 living
 * on the current CUDA device (box-constrained quadrotor / cartpole shapes: on the latency kernel,
 * TMPC_KERNEL_WARP, a block of four warps for the one instance).
 *
 * PART 2 adds the batched, handle-based variants (B independent problem instances that share
 * dynamics, cost and cache) and device-pointer variants for callers that keep their data in HBM.
 *
 * Conventions (identical to the reference):
 *   - every matrix is column-major double (Julia `Matrix{Float64}`), passed as pointer + rows/cols;
 *   - inputs are copied during the call, outputs are caller-allocated;
 *   - return 0 = success, non-zero = failure; C++ exceptions never cross the boundary (-1);
 *   - calls are synchronous unless the name ends in `_async`.
 * Batched arrays carry the instance index as the slowest dimension (Julia `Array{Float64,3}` of
 * size (nx, N, B)), i.e. each instance's matrix is one contiguous block.
 *
 * There is no CPU fallback: every solve runs the CUDA kernels of this library; when no CUDA
 * device is usable the calls fail with TMPC_ERR_CUDA.
 */
#ifndef TINYMPC_B200_H
#define TINYMPC_B200_H
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ------------------------------------------------------------------------------------------
 * PART 1 — drop-in symbols of the reference library
 * ------------------------------------------------------------------------------------------ */

/* replaces src/bindings.cpp:21-73 (called from src/TinyMPC.jl:77-83) */
int setup_solver(double* A_data, int A_rows, int A_cols,
                 double* B_data, int B_rows, int B_cols,
                 double* fdyn_data, int fdyn_rows, int fdyn_cols,
                 double* Q_data, int Q_rows, int Q_cols,
                 double* R_data, int R_rows, int R_cols,
                 double rho, int nx, int nu, int N, int verbose);
/* replaces src/bindings.cpp:75-96 (src/TinyMPC.jl:119-120) */
int set_x0(double* x0_data, int x0_rows, int x0_cols, int verbose);
/* replaces src/bindings.cpp:98-119 (src/TinyMPC.jl:128-129) */
int set_x_ref(double* x_ref_data, int x_ref_rows, int x_ref_cols, int verbose);
/* replaces src/bindings.cpp:121-142 (src/TinyMPC.jl:137-138) */
int set_u_ref(double* u_ref_data, int u_ref_rows, int u_ref_cols, int verbose);
/* replaces src/bindings.cpp:144-159 (src/TinyMPC.jl:146): 0 converged, 1 max_iter reached, -1 error */
int solve_mpc(int verbose);
/* replaces src/bindings.cpp:161-181 / :183-203 (src/TinyMPC.jl:165-168) */
int get_states(double* states_buffer, int* rows, int* cols);
int get_controls(double* controls_buffer, int* rows, int* cols);
/* replaces src/bindings.cpp:205-208 (src/TinyMPC.jl:428-429) */
void cleanup_solver(void);
/* replaces src/bindings.cpp:210-225 and :298-331 — code generation for MCUs is outside the
 * accelerated path; the symbols exist and return a non-zero status.  codegen_with_sensitivity keeps the half of the
 * reference call that belongs to the solve path: with adaptive_rho on it stores dK, dP, dC1, dC2 in the live solver
 * (tinympc/TinyMPC/src/tinympc/codegen.cpp:90-97: the only loader the Julia package has, tests/test_sensitivity.jl:40-53)
 * before it reports that no code was generated. */
int codegen(const char* output_dir, int verbose);
int codegen_with_sensitivity(const char* output_dir,
                             double* dK_data, int dK_rows, int dK_cols,
                             double* dP_data, int dP_rows, int dP_cols,
                             double* dC1_data, int dC1_rows, int dC1_cols,
                             double* dC2_data, int dC2_rows, int dC2_cols, int verbose);
/* NEW (no counterpart in src/bindings.cpp): the loader alone, returns 0.  dK nu x nx, dP nx x nx, dC1 nu x nu, dC2 nx x nx,
 * column-major (what compute_sensitivity_autograd returns, src/TinyMPC.jl:301-323). */
int set_sensitivity_matrices(double* dK_data, int dK_rows, int dK_cols,
                             double* dP_data, int dP_rows, int dP_cols,
                             double* dC1_data, int dC1_rows, int dC1_cols,
                             double* dC2_data, int dC2_rows, int dC2_cols, int verbose);
/* replaces src/bindings.cpp:228-259 (src/TinyMPC.jl:273) */
int print_problem_data(int verbose);
/* replaces src/bindings.cpp:262-293 (src/TinyMPC.jl:281-286): overrides Kinf, Pinf, Quu_inv, AmBKt only */
int set_cache_terms(double* Kinf_data, int Kinf_rows, int Kinf_cols,
                    double* Pinf_data, int Pinf_rows, int Pinf_cols,
                    double* Quu_inv_data, int Quu_inv_rows, int Quu_inv_cols,
                    double* AmBKt_data, int AmBKt_rows, int AmBKt_cols, int verbose);
/* replaces src/bindings.cpp:336-376 (src/TinyMPC.jl:199-207): overwrites ALL settings.
 * Deviations: check_termination == 0 is rejected (the reference divides by it, admm.cpp:255);
 * adaptive_rho != 0 is rejected. */
int update_settings(double abs_pri_tol, double abs_dua_tol, int max_iter, int check_termination,
                    int en_state_bound, int en_input_bound, int en_state_soc, int en_input_soc,
                    int en_state_linear, int en_input_linear,
                    int adaptive_rho, double adaptive_rho_min, double adaptive_rho_max,
                    int adaptive_rho_enable_clipping, int verbose);
/* replaces src/bindings.cpp:378-411 (src/TinyMPC.jl:218-223): enables both bound flags on success */
int set_bound_constraints(double* x_min_data, int x_min_rows, int x_min_cols,
                          double* x_max_data, int x_max_rows, int x_max_cols,
                          double* u_min_data, int u_min_rows, int u_min_cols,
                          double* u_max_data, int u_max_rows, int u_max_cols, int verbose);
/* replaces src/bindings.cpp:414-450 (src/TinyMPC.jl:233-238) */
int set_linear_constraints(double* Alin_x_data, int Alin_x_rows, int Alin_x_cols,
                           double* blin_x_data, int blin_x_len,
                           double* Alin_u_data, int Alin_u_rows, int Alin_u_cols,
                           double* blin_u_data, int blin_u_len, int verbose);
/* replaces src/bindings.cpp:453-490 (src/TinyMPC.jl:248-253).  The reference stores the FIRST
 * triple in the state-cone slots and the SECOND in the input-cone slots while deriving the
 * enable flags the other way round; that behaviour is reproduced. */
int set_cone_constraints(int* Acu_data, int Acu_len, int* qcu_data, int qcu_len, double* cu_data, int cu_len,
                         int* Acx_data, int Acx_len, int* qcx_data, int qcx_len, double* cx_data, int cx_len,
                         int verbose);

/* extra accessors on the process-wide solver (the reference exposes iter/solved only through
 * print_problem_data, src/bindings.cpp:236-237) */
int get_solve_info(int* iter, int* solved);

/* ------------------------------------------------------------------------------------------
 * PART 2 — batched variants
 * ------------------------------------------------------------------------------------------ */

typedef struct tmpc_batch tmpc_batch; /* opaque */

enum {
    TMPC_OK = 0,
    TMPC_ERR_DIM = 1,       /* dimension mismatch / null pointer (the reference core's `1`) */
    TMPC_ERR_STATE = 2,     /* call order / unsupported setting */
    TMPC_ERR_CUDA = 3,      /* CUDA runtime failure (message on stderr) */
    TMPC_ERR_EXCEPTION = -1
};

/* kernel selection */
enum {
    TMPC_KERNEL_AUTO = 0,
    TMPC_KERNEL_THREAD = 1,   /* one thread per instance, state streamed through L2/HBM each iteration; any nx, nu */
    TMPC_KERNEL_SUBWARP = 2,  /* one instance per 4-lane sub-warp, state resident in shared memory (box constraints) */
    TMPC_KERNEL_TMEM = 3,     /* 4-lane sub-warp, state rows in Tensor Memory + input rows in shared memory, 8 warps/SM,
                                 per-instance refill (box constraints with shared bounds; nx=12 nu=4 with N <= 42, nx=4 nu=1) */
    TMPC_KERNEL_CONE = 4,     /* box + second-order cones + half-spaces on chip: 3 instances per warp, recursions with a lane
                                 per matrix row, projections with a lane per knot point (nx=6 nu=3, N <= 50) */
    TMPC_KERNEL_WARP = 5      /* a block per instance, built for latency (recursions on one warp, the rest of an iteration over all knot
                                 points at once on four; one warp per instance beyond 296 instances): what AUTO picks for
                                 B <= 2048 (nx=4: 4096), the single solve behind the reference's own symbols first of all;
                                 box constraints, nx=12 nu=4 / nx=4 nu=1, double build, shared dynamics */
};

/* Host-only: the infinite-horizon LQR cache the solve loop reads (replaces tiny_precompute_and_set_cache,
 * tiny_api.cpp:244-318, fed as tiny_setup feeds it, :107-108,:126).  Column-major outputs, any may be NULL:
 * Kinf nu x nx, Pinf nx x nx, Quu_inv nu x nu, AmBKt nx x nx, APf nx, BPf nu, Qd = diag(Q)+rho, Rd = diag(R)+rho. */
int tmpc_compute_cache(const double* A, const double* Bdyn, const double* f, const double* Q, const double* R, double rho,
                       int nx, int nu, double* Kinf, double* Pinf, double* Quu_inv, double* AmBKt, double* APf,
                       double* BPf, double* Qd, double* Rd);

/* The same for the reference's float build (`typedef float tinytype`): inputs rounded to float as in
 * src/bindings.cpp:50-55, every operation in float; float values returned in doubles. */
int tmpc_compute_cache_f32(const double* A, const double* Bdyn, const double* f, const double* Q, const double* R, double rho,
                           int nx, int nu, double* Kinf, double* Pinf, double* Quu_inv, double* AmBKt, double* APf,
                           double* BPf, double* Qd, double* Rd);

/* Create a batch of B instances sharing A, B, f, Q, R, rho (same meaning as setup_solver) on
 * CUDA device `device` (-1 = current).  All instances start from the zero workspace tiny_setup
 * produces (reference: tiny_api.cpp:68-105); settings start at the Julia defaults
 * (src/TinyMPC.jl:57-61: tol 1e-3, max_iter 100, check_termination 1, all en_* = 0). */
int tmpc_batch_create(tmpc_batch** out,
                      const double* A, const double* Bdyn, const double* f, const double* Q, const double* R,
                      double rho, int nx, int nu, int N, int B, int device, int verbose);
int tmpc_batch_destroy(tmpc_batch* h);

/* Overwrites all settings (reference: src/bindings.cpp:336-376).  Rejected with TMPC_ERR_STATE: check_termination = 0 (the
 * reference computes iter % 0) and max_iter <= 0.  adaptive_rho != 0 is accepted as a setting (tests/test_settings.jl:63-73
 * only sets it up); solves with the flag on need tmpc_batch_set_sensitivity (below). */
int tmpc_batch_update_settings(tmpc_batch* h, double abs_pri_tol, double abs_dua_tol, int max_iter,
                               int check_termination, int en_state_bound, int en_input_bound,
                               int en_state_soc, int en_input_soc, int en_state_linear, int en_input_linear,
                               int adaptive_rho, double adaptive_rho_min, double adaptive_rho_max,
                               int adaptive_rho_enable_clipping);
/* Arithmetic type of the solve = the reference's `tinytype` (tinympc/TinyMPC/src/tinympc/types.hpp:15: `typedef double
 * tinytype`, float being its other build).  bits = 32: the setup is redone as the float build does it (problem data and
 * rho rounded to float as in src/bindings.cpp:50-55, cache computed in float) and later solves run the float
 * instantiation of the kernel, inputs rounded on load as in bindings.cpp:82,105,128 and results widened as in :172,194;
 * the ABI stays double.  Only on a fresh / reset batch with shared dynamics; the float build runs the float instantiation
 * of the TMEM kernel where the shape is compiled (else the shared-memory sub-warp kernel, else the thread kernel); the cone
 * kernel, per-instance dynamics and the device closed loop are 64-bit. */
int tmpc_batch_set_precision(tmpc_batch* h, int bits);   /* 32 or 64 (default) */
int tmpc_batch_get_precision(tmpc_batch* h);
int tmpc_batch_set_kernel(tmpc_batch* h, int kernel);   /* TMPC_KERNEL_*; AUTO picks WARP for small batches, else TMEM, then SUBWARP, then CONE, where they apply, else THREAD */
int tmpc_batch_get_kernel(tmpc_batch* h);               /* kernel the next solve will run (resolved AUTO) */

/* Queue order of the on-chip (persistent) kernels.  Iteration counts are data-dependent (admm.cpp:312-381), so an
 * instance that needs 25 iterations and is pulled in the last wave holds one sub-warp while the other SMs idle.
 * AUTO (default): instances expected to run longest go first - cold start: by the saturation of the unconstrained LQR
 * input, max_j |(Kinf x0)_j| / min(|u_min_j|, |u_max_j|); warm start: by the iteration count of the previous solve.
 * USER: `perm` (host, B entries, a permutation of 0..B-1) is the order.  Results never depend on the order. */
enum { TMPC_ORDER_NATURAL = 0, TMPC_ORDER_AUTO = 1, TMPC_ORDER_USER = 2 };
int tmpc_batch_set_order(tmpc_batch* h, int mode, const int* perm);
int tmpc_batch_get_order(tmpc_batch* h, int* perm);   /* the permutation the last solve used (TMPC_ERR_STATE: natural) */

/* Host-pointer setters.  `batch` = 1: one matrix shared by all instances; `batch` = B: one per instance. */
int tmpc_batch_set_x0(tmpc_batch* h, const double* x0, int rows, int batch);                 /* nx x B */
int tmpc_batch_set_x_ref(tmpc_batch* h, const double* x_ref, int rows, int cols, int batch); /* nx x N [x B] */
int tmpc_batch_set_u_ref(tmpc_batch* h, const double* u_ref, int rows, int cols, int batch); /* nu x (N-1) [x B] */
int tmpc_batch_set_bound_constraints(tmpc_batch* h, const double* x_min, const double* x_max,
                                     const double* u_min, const double* u_max, int batch);
int tmpc_batch_set_linear_constraints(tmpc_batch* h, const double* Alin_x, int nlx, const double* blin_x,
                                      const double* Alin_u, int nlu, const double* blin_u);
/* first triple -> state-cone slots, second triple -> input-cone slots (core order, tiny_api.cpp:166-168);
 * enable flags are set per non-empty triple of the SAME side.  Cone size must be 3 (admm.cpp:53). */
int tmpc_batch_set_cone_constraints(tmpc_batch* h, const int* Acx, const int* qcx, const double* cx, int ncx,
                                    const int* Acu, const int* qcu, const double* cu, int ncu);
/* Julia argument order (input-cone triple first, as src/TinyMPC.jl:243-256 passes them) WITH the
 * reference's crossed behaviour: the first triple lands in the state-cone slots,
 * the second in the input-cone slots, en_state_soc follows the second triple's length and
 * en_input_soc the first's.  This is what set_cone_constraints (PART 1) does. */
int tmpc_batch_set_cone_constraints_jl(tmpc_batch* h, const int* Acu, const int* qcu, const double* cu, int ncu_len,
                                       const int* Acx, const int* qcx, const double* cx, int ncx_len);
/* tols[2] = {abs_pri_tol, abs_dua_tol}; ints[8] = {max_iter, check_termination, en_state_bound,
 * en_input_bound, en_state_soc, en_input_soc, en_state_linear, en_input_linear} */
int tmpc_batch_get_settings(tmpc_batch* h, double* tols, int* ints);
/* Per-instance dynamics: B different systems in one batch.  The setup the reference runs once per
 * solver (tiny_setup + tiny_precompute_and_set_cache, tiny_api.cpp:21-137, :244-318) runs on the device, one thread per
 * instance, operation by operation as the host path does (the Riccati stop rule makes the cache only 1e-5-converged, so
 * iteration-count parity needs the same arithmetic).  A nx x nx x B, Bdyn nx x nu x B, f nx x B or NULL, Q nx x nx [x B],
 * R nu x nu [x B] (qr_batch = 1 shared / B per instance); host pointers, Julia layout; rho as given to create.  Later
 * solves (tmpc_batch_solve*, tmpc_batch_solve_host, tmpc_batch_rollout_async with each instance's own plant) use the
 * TMEM kernel reading one packed operand block per instance where its shapes apply (nx = 12 / nu = 4, nx = 4 / nu = 1,
 * box constraints), else the thread kernel with coalesced per-instance operands; nx <= 12, nu <= 4. */
int tmpc_batch_set_instance_dynamics(tmpc_batch* h, const double* A, const double* Bdyn, const double* f, const double* Q,
                                     const double* R, int qr_batch);
int tmpc_batch_get_instance_cache(tmpc_batch* h, int b, double* Kinf, double* Pinf, double* Quu_inv, double* AmBKt,
                                  double* APf, double* BPf);
int tmpc_batch_set_cache_terms(tmpc_batch* h, const double* Kinf, const double* Pinf,
                               const double* Quu_inv, const double* AmBKt);
int tmpc_batch_get_cache_terms(tmpc_batch* h, double* Kinf, double* Pinf, double* Quu_inv, double* AmBKt,
                               double* APf, double* BPf);

/* Adaptive rho (reference: solve, tinympc/TinyMPC/src/tinympc/admm.cpp:331-357; benchmark_rho_adaptation,
 * rho_benchmark.cpp:213-249).  With adaptive_rho on (tmpc_batch_update_settings) every instance re-estimates rho after
 * iterations 6, 11, 16, ... from the reference's residual ratio and applies the first-order update
 * Kinf += d_rho * dKinf_drho, Pinf += d_rho * dPinf_drho to ITS OWN copy of the cache; like the reference's cache, the
 * adapted rho / Kinf / Pinf stay with the instance over later solves (also after the flag is cleared) until
 * tmpc_batch_reset.  The sensitivity matrices are shared (column-major dKinf nu x nx, dPinf nx x nx, dC1 nu x nu,
 * dC2 nx x nx - the reference stores them with codegen_with_sensitivity, codegen.cpp:90-97); a solve with the flag on
 * and no matrices returns TMPC_ERR_STATE (the reference crashes there).  64-bit build, shared dynamics, thread kernel
 * (TMPC_KERNEL_AUTO selects it). */
int tmpc_batch_set_sensitivity(tmpc_batch* h, const double* dKinf_drho, const double* dPinf_drho, const double* dC1_drho,
                               const double* dC2_drho);
/* The cache every instance currently solves with: rho [B], Kinf nu x nx x B, Pinf nx x nx x B (any may be NULL). */
int tmpc_batch_get_adapted_cache(tmpc_batch* h, double* rho, double* Kinf, double* Pinf);

/* Back to the zero workspace (cold start): duals, slacks, linear cost terms and solution cleared, every instance back
 * on the cache the setup computed; x0, references, constraints and settings kept. */
int tmpc_batch_reset(tmpc_batch* h);

/* Solve all instances (warm-started from each instance's own carried state).  Returns 0 when every
 * instance converged, 1 when at least one stopped at max_iter, <0 / >1 on error. */
int tmpc_batch_solve(tmpc_batch* h);
int tmpc_batch_solve_async(tmpc_batch* h, void* cuda_stream);   /* enqueue only; no host sync */
/* set_x0 (+ reset when cold != 0) + solve + get_states / get_controls / get_status in ONE call on host buffers, the
 * batch cut into `chunks` (0 = default) so that the solution of one chunk crosses PCIe while the next chunk solves.
 * x0 nx x B or NULL (keep the current one); X nx x N x B, U nu x (N-1) x B, status, iters: B each, any may be NULL.
 * Pinned host buffers make the copies truly asynchronous.  Returns like tmpc_batch_solve (0 / 1 / 100 + error). */
int tmpc_batch_solve_host(tmpc_batch* h, const double* x0, int cold, double* X, double* U, int* status, int* iters,
                          int chunks);
/* The same with an output selector: u0 (nu x B) receives the FIRST control of every instance, which is all a closed
 * loop consumes from a solve (examples/cartpole_example_mpc.jl:41-42: `u = solution.controls[:, 1]`).  Any of X, U, u0,
 * status, iters may be NULL; with X = U = NULL the device-to-host traffic is nu + 1 words per instance instead of the
 * 2.5 KB trajectories. */
int tmpc_batch_solve_host_select(tmpc_batch* h, const double* x0, int cold, double* X, double* U, double* u0,
                                 int* status, int* iters, int chunks);
int tmpc_batch_synchronize(tmpc_batch* h);

/* Per-instance results.  status[b] = what solve_mpc would return for that instance (0 / 1). */
int tmpc_batch_get_status(tmpc_batch* h, int* status, int* iters);
int tmpc_batch_get_states(tmpc_batch* h, double* buf);     /* nx x N x B */
int tmpc_batch_get_controls(tmpc_batch* h, double* buf);   /* nu x (N-1) x B */

/* Device-pointer variants: the pointers must be device memory on the batch's device, laid out as
 * above, and must stay valid until replaced (they are NOT copied).  NULL restores the internal buffer. */
int tmpc_batch_bind_x0_device(tmpc_batch* h, const double* d_x0);
int tmpc_batch_bind_x_ref_device(tmpc_batch* h, const double* d_x_ref, int batch);
int tmpc_batch_bind_u_ref_device(tmpc_batch* h, const double* d_u_ref, int batch);
/* device addresses of the result buffers: 0 states (nx x N x B), 1 controls, 2 status (int), 3 iters (int), 4 x0 */
void* tmpc_batch_device_ptr(tmpc_batch* h, int which);

/* Closed-loop MPC on the device (reference pattern: examples/cartpole_example_mpc.jl:35-51):
 * for k in 0..nsteps-1:  [x_ref <- columns k..k+N-1 of x_ref_long];  solve;  u = controls(:,0);
 *                        x0 <- A x0 + B u + f.
 * x_ref_long: device pointer, nx x (N+nsteps-1) x B, or NULL to keep the current references.
 * iters_out / status_out: optional device int arrays of nsteps x B (step-major). */
int tmpc_batch_rollout_async(tmpc_batch* h, int nsteps, const double* d_x_ref_long,
                             int* d_status_out, int* d_iters_out, void* cuda_stream);
/* 1 (default): where the TMEM kernel applies, tmpc_batch_rollout_async is ONE launch that keeps each instance's ADMM
 * state on chip across its MPC steps; 0: one solve + record + plant-step launch per step (any kernel). */
int tmpc_batch_set_rollout_fused(tmpc_batch* h, int fused);

/* Current x0 of every instance (after a rollout: the plant state reached), nx x B.  `device_dst` != 0: dst is a device
 * pointer on the batch's device (device-to-device copy), else a host buffer. */
int tmpc_batch_get_x0(tmpc_batch* h, void* dst, int device_dst);

/* Page-locked host memory for the buffers of tmpc_batch_solve_host*: with pageable buffers every copy of the call
 * blocks the host and the chunks cannot overlap.  NULL when no CUDA device is usable. */
void* tmpc_host_alloc(size_t bytes);
void tmpc_host_free(void* p);

/* One batched cold solve that allocates nothing on the device.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  It is the thread-per-instance kernel with its working set (carried ADMM state + feed-forward scratch,
 * (4 nx N + 5 nu (N-1) + nx) doubles per instance, instances rounded up to a multiple of 32) in a file-scope
 * `static __device__` array of TMPC_NOALLOC_SCRATCH_BYTES, which every GPU that loads the library holds until the process
 * exits.  A batch whose working set does not fit returns TMPC_ERR_DIM with the sizes in the message (quadrotor 12/4/20:
 * B <= 3,072; cartpole 4/1/20: B <= 9,984): use tmpc_batch_create / tmpc_batch_solve for those.
 *   host pointers   A nx x nx, Bdyn nx x nu, f nx, and Kinf .. Rd exactly as tmpc_compute_cache returns them (call it once);
 *   device pointers d_x0 nx x B; d_x_ref nx x N and d_u_ref nu x (N-1) shared by all instances, or NULL (zeros);
 *                   d_x_min / d_x_max nx x N and d_u_min / d_u_max nu x (N-1) shared, or NULL pairs (no box on that side);
 *                   outputs d_states nx x N x B, d_controls nu x (N-1) x B, d_status B, d_iters B - all on the CURRENT device;
 *   cuda_stream     the cudaStream_t everything is enqueued on (NULL = default stream); the call returns without waiting.
 * Semantics = tmpc_batch_reset + tmpc_batch_solve on TMPC_KERNEL_THREAD with box constraints (results are identical).
 * The calls it makes: cudaGetSymbolAddress (once per device), cudaMemsetAsync of the working set, one kernel launch; for
 * shapes other than 12/4, 4/1, 6/3 also a cudaMemcpyAsync of the constants when they changed.  No cudaMalloc / cudaFree.
 * The array is SHARED by every call and keeps the last call's contents (each call clears what it uses): calls that overlap
 * on the GPU - other streams without ordering, other host threads - race.  errbuf (may be NULL) receives the message of a
 * non-zero return. */
#define TMPC_NOALLOC_SCRATCH_BYTES 33554432   /* 32 MiB */
int tmpc_solve_noalloc(int nx, int nu, int N, int B, double rho, const double* A, const double* Bdyn, const double* f,
                       const double* Kinf, const double* Pinf, const double* Quu_inv, const double* AmBKt,
                       const double* APf, const double* BPf, const double* Qd, const double* Rd,
                       const double* d_x0, const double* d_x_ref, const double* d_u_ref,
                       const double* d_x_min, const double* d_x_max, const double* d_u_min, const double* d_u_max,
                       double abs_pri_tol, double abs_dua_tol, int max_iter, int check_termination,
                       double* d_states, double* d_controls, int* d_status, int* d_iters, void* cuda_stream,
                       char* errbuf, size_t errbuf_len);

/* Introspection */
int tmpc_batch_dims(tmpc_batch* h, int* nx, int* nu, int* N, int* B);
long long tmpc_batch_launch_count(tmpc_batch* h);   /* kernels launched by this handle so far */
/* Measurement aid (not on the solve path): best-of-`reps` FMA throughput of register-resident,
 * dependency-free chains on `device` (-1 = current), in TFLOP/s (2 flop per FMA).  fp64 != 0: DFMA
 * (the reference's tinytype, types.hpp:15); fp64 == 0: FFMA.  The roofline denominator of bench.py. */
int tmpc_measure_fma_peak(int device, int fp64, int reps, double* tflops);
const char* tmpc_last_error(void);
const char* tmpc_version(void);

#ifdef __cplusplus
}
#endif
#endif /* TINYMPC_B200_H */
