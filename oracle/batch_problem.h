/*
 * TEST INFRASTRUCTURE — not part of the product path.
 *
 * Flat description of one TinyMPC problem plus a batch of instances, shared by
 *   - oracle/tinympc_oracle.c   (plain-C restatement of the reference ADMM loop), and
 *   - oracle/ref_driver.cpp     (thin driver that links the UNMODIFIED reference core
 *                                from /root/reference and loops it over a batch).
 * Both export the same two entry points (prefix `orc_` / `ref_`), so tests can swap one
 * for the other and compare the CUDA library against either.
 *
 * Fields are at the level of the reference core (`TinySolver` after the C-ABI glue has
 * run), i.e. cones are already in the slots the core stores them in
 * (reference: tinympc/TinyMPC/src/tinympc/tiny_api.cpp:166-198, types.hpp:86-187).
 * All matrices are column-major, as Julia passes them (reference: src/bindings.cpp:40-44).
 */
#ifndef TINYMPC_ORACLE_BATCH_PROBLEM_H
#define TINYMPC_ORACLE_BATCH_PROBLEM_H

#ifdef __cplusplus
extern "C" {
#endif

typedef struct {
    int nx, nu, N;
    double rho;
    const double *A;      /* nx x nx */
    const double *B;      /* nx x nu */
    const double *f;      /* nx      */
    const double *Q;      /* nx x nx (only used as the reference uses it: full matrix into Riccati, diagonal+rho into the loop) */
    const double *R;      /* nu x nu */

    /* TinySettings (types.hpp:63-80) */
    double abs_pri_tol, abs_dua_tol;
    int max_iter, check_termination;
    int en_state_bound, en_input_bound;
    int en_state_soc, en_input_soc;
    int en_state_linear, en_input_linear;

    /* box bounds; NULL => the workspace keeps whatever tiny_setup left (empty matrices in the
       reference; the matching en_* flag must then be 0).  bounds_per_instance=1: arrays carry a
       trailing batch dimension (nx x N x B, nu x (N-1) x B). */
    const double *x_min, *x_max, *u_min, *u_max;
    int bounds_per_instance;

    /* second-order cones as stored in TinyWorkspace (state slots / input slots) */
    int num_state_cones; const int *Acx; const int *qcx; const double *cx;
    int num_input_cones; const int *Acu; const int *qcu; const double *cu;

    /* linear inequalities a_k . z <= b_k */
    int num_state_linear; const double *Alin_x; /* num_state_linear x nx, column-major */ const double *blin_x;
    int num_input_linear; const double *Alin_u; /* num_input_linear x nu */                const double *blin_u;

    /* optional cache override (set_cache_terms semantics, bindings.cpp:277-280): replaces
       Kinf (nu x nx), Pinf (nx x nx), Quu_inv (nu x nu), AmBKt (nx x nx) only. NULL => computed. */
    const double *Kinf, *Pinf, *Quu_inv, *AmBKt;

    /* adaptive rho (admm.cpp:282-357, rho_benchmark.cpp:14-249): the four settings (types.hpp:76-79) and the sensitivity
       matrices as `codegen_with_sensitivity` leaves them in the LIVE cache (codegen.cpp:91-96: the only loader of
       arbitrary-shape matrices; tests/test_sensitivity.jl:40-53 is the call sequence).  adaptive_rho != 0 needs all
       four (dKinf nu x nx, dPinf nx x nx, dC1 nu x nu, dC2 nx x nx); without them the reference adds 0 x 0 matrices
       and crashes.  The adapted cache (rho, Kinf, Pinf, C1, C2) is solver state: it carries over the steps of ONE
       instance and every instance starts from the cache tiny_setup computed. */
    int adaptive_rho; double adaptive_rho_min, adaptive_rho_max; int adaptive_rho_enable_clipping;
    const double *dKinf_drho, *dPinf_drho, *dC1_drho, *dC2_drho;
    /* optional outputs of an adaptive run (NULL = not wanted): cache->rho [B], Kinf [nu*nx*B], Pinf [nx*nx*B] after the
       instance's last step */
    double *rho_final, *Kinf_final, *Pinf_final;
} tmpc_problem;

/* reference modes for batch_run */
enum { TMPC_REF_ZERO = 0,        /* Xref = Uref = 0 (never set)                                  */
       TMPC_REF_SHARED = 1,      /* Xref nx x N, Uref nu x (N-1), same for every instance         */
       TMPC_REF_PER_INSTANCE = 2,/* Xref nx x N x B, Uref nu x (N-1) x B, fixed over steps        */
       TMPC_REF_SLIDING = 3      /* Xref nx x (N+nsteps-1) x B; step k uses columns k..k+N-1;
                                    Uref as in mode 1 when non-NULL, else zero                   */ };

/*
 * Run `nsteps` closed-loop MPC steps for each of B independent instances, each instance starting
 * from the zero workspace tiny_setup produces (tiny_api.cpp:68-105) and warm-starting from its own
 * previous step (nothing is reset between steps; examples/cartpole_example_mpc.jl:35-51):
 *     set_x0(x); [set references]; solve; u = solution.u(:,0); x <- A x + B u + f
 * nsteps = 1 is a single cold-start solve per instance.
 *
 * Outputs (any may be NULL): ret[nsteps*B] solve() return value (0 converged / 1 max_iter), step-major
 * (ret[k*B + b]); iters[nsteps*B]; Xsol nx x N x B and Usol nu x (N-1) x B of the LAST step;
 * x_final nx x B plant state after the last step; seconds = wall time of the instance loop.
 * Returns 0, or a non-zero code when the problem description is rejected.
 */
#define TMPC_DECLARE_BATCH_API(prefix)                                                             \
    int prefix##_batch_run(const tmpc_problem *P, int B, int nsteps, const double *x0,             \
                           const double *Xref, const double *Uref, int ref_mode, int nthreads,     \
                           int *ret, int *iters, double *Xsol, double *Usol, double *x_final,      \
                           double *seconds);                                                       \
    /* Riccati cache as the solve loop reads it (tiny_api.cpp:244-318) */                          \
    int prefix##_get_cache(const tmpc_problem *P, double *Kinf, double *Pinf, double *Quu_inv,     \
                           double *AmBKt, double *APf, double *BPf);

#ifdef __cplusplus
}
#endif
#endif
