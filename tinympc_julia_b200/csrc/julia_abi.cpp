// PART 1 of include/tinympc_b200.h: the sixteen symbols the Julia package binds with `ccall`
// (reference: src/bindings.cpp, called from src/TinyMPC.jl), re-exported on top of the batched
// CUDA solver with B = 1.  Same names, argument meaning, return codes and flag side effects, so the
// unmodified src/TinyMPC.jl works against this library.
//
// Like the reference (src/bindings.cpp:15) there is one solver per process; a second setup_solver
// replaces the first; the calls are not thread-safe.
#include <cmath>
#include <cstdio>
#include <cstring>
#include <vector>

#include "../../include/tinympc_b200.h"

namespace {

tmpc_batch* g_solver = nullptr;
int g_nx = 0, g_nu = 0, g_N = 0;
int g_last_status = 0, g_last_iter = 0, g_solved = 0;
double g_rho = 0.0;
struct Settings {
    double abs_pri_tol = 1e-3, abs_dua_tol = 1e-3;
    int max_iter = 1000, check_termination = 1;               // C++-side defaults, tiny_api_constants.hpp:5-14
    int en_state_bound = 1, en_input_bound = 1, en_state_soc = 0, en_input_soc = 0, en_state_linear = 0, en_input_linear = 0;
    int adaptive_rho = 0; double adaptive_rho_min = 1.0, adaptive_rho_max = 100.0; int adaptive_rho_enable_clipping = 1;   // tiny_api.cpp:367-370
} g_set;

int fail(const char* fn, const char* why) {
    std::fprintf(stderr, "%s failed: %s\n", fn, why);      // reference prints "<fn> failed: <what>" on stderr and returns -1
    return -1;
}

int push_settings() {
    return tmpc_batch_update_settings(g_solver, g_set.abs_pri_tol, g_set.abs_dua_tol, g_set.max_iter,
                                      g_set.check_termination, g_set.en_state_bound, g_set.en_input_bound,
                                      g_set.en_state_soc, g_set.en_input_soc, g_set.en_state_linear,
                                      g_set.en_input_linear, g_set.adaptive_rho, g_set.adaptive_rho_min,
                                      g_set.adaptive_rho_max, g_set.adaptive_rho_enable_clipping);
}

}  // namespace

extern "C" {

int setup_solver(double* A_data, int A_rows, int A_cols, double* B_data, int B_rows, int B_cols,
                 double* fdyn_data, int fdyn_rows, int fdyn_cols, double* Q_data, int Q_rows, int Q_cols,
                 double* R_data, int R_rows, int R_cols, double rho, int nx, int nu, int N, int verbose) {
    if (verbose) std::printf("Setting up TinyMPC solver with nx=%d, nu=%d, N=%d, rho=%g\n", nx, nu, N, rho);
    // dimension checks of tiny_setup (tiny_api.cpp:53-66): status 1, previous solver left in place
    if (A_rows != nx || A_cols != nx || B_rows != nx || B_cols != nu || fdyn_rows != nx || fdyn_cols != 1 ||
        Q_rows != nx || Q_cols != nx || R_rows != nu || R_cols != nu) {
        std::printf("setup_solver: matrix dimensions do not match nx=%d, nu=%d\n", nx, nu);
        return 1;
    }
    if (!A_data || !B_data || !fdyn_data || !Q_data || !R_data) return fail("setup_solver", "null matrix");
    tmpc_batch* fresh = nullptr;
    const int rc = tmpc_batch_create(&fresh, A_data, B_data, fdyn_data, Q_data, R_data, rho, nx, nu, N, 1, -1, verbose);
    if (rc != 0) {
        if (rc == TMPC_ERR_DIM) return 1;
        cleanup_solver();                                   // reference resets the singleton when setup throws (:70)
        return fail("setup_solver", tmpc_last_error());
    }
    if (g_solver) tmpc_batch_destroy(g_solver);
    g_solver = fresh;
    g_nx = nx; g_nu = nu; g_N = N; g_rho = rho;
    g_last_status = 0; g_last_iter = 0; g_solved = 0;
    g_set = Settings();
    if (push_settings() != 0) return fail("setup_solver", tmpc_last_error());
    if (verbose) std::printf("TinyMPC solver setup completed successfully\n");
    return 0;
}

int set_x0(double* x0_data, int x0_rows, int x0_cols, int verbose) {
    if (!g_solver) return fail("set_x0", "Solver not initialized");
    if (!x0_data || x0_rows * x0_cols != g_nx) return fail("set_x0", "x0 is not the correct length");
    if (tmpc_batch_set_x0(g_solver, x0_data, g_nx, 1) != 0) return fail("set_x0", tmpc_last_error());
    if (verbose) std::printf("Initial state set\n");
    return 0;
}

int set_x_ref(double* x_ref_data, int x_ref_rows, int x_ref_cols, int verbose) {
    if (!g_solver) return fail("set_x_ref", "Solver not initialized");
    if (tmpc_batch_set_x_ref(g_solver, x_ref_data, x_ref_rows, x_ref_cols, 1) != 0) return fail("set_x_ref", tmpc_last_error());
    if (verbose) std::printf("State reference set\n");
    return 0;
}

int set_u_ref(double* u_ref_data, int u_ref_rows, int u_ref_cols, int verbose) {
    if (!g_solver) return fail("set_u_ref", "Solver not initialized");
    if (tmpc_batch_set_u_ref(g_solver, u_ref_data, u_ref_rows, u_ref_cols, 1) != 0) return fail("set_u_ref", tmpc_last_error());
    if (verbose) std::printf("Input reference set\n");
    return 0;
}

int solve_mpc(int verbose) {
    if (!g_solver) return fail("solve_mpc", "Solver not initialized");
    const int rc = tmpc_batch_solve(g_solver);
    if (rc != 0 && rc != 1) return fail("solve_mpc", tmpc_last_error());
    int st = 0, it = 0;
    if (tmpc_batch_get_status(g_solver, &st, &it) != 0) return fail("solve_mpc", tmpc_last_error());
    g_last_status = st; g_last_iter = it; g_solved = (st == 0);
    if (g_solved) std::printf("Solver converged in %d iterations\n", it);   // reference admm.cpp:373
    if (verbose) std::printf("Solve completed with status: %d\n", st);
    return st;
}

int get_states(double* states_buffer, int* rows, int* cols) {
    if (!g_solver || !states_buffer || !rows || !cols) return -1;
    *rows = g_nx; *cols = g_N;
    return tmpc_batch_get_states(g_solver, states_buffer) == 0 ? 0 : -1;
}

int get_controls(double* controls_buffer, int* rows, int* cols) {
    if (!g_solver || !controls_buffer || !rows || !cols) return -1;
    *rows = g_nu; *cols = g_N - 1;
    return tmpc_batch_get_controls(g_solver, controls_buffer) == 0 ? 0 : -1;
}

void cleanup_solver(void) {
    if (g_solver) tmpc_batch_destroy(g_solver);
    g_solver = nullptr;
    g_nx = g_nu = g_N = 0;
}

int codegen(const char* output_dir, int verbose) {
    (void)output_dir; (void)verbose;
    if (!g_solver) return fail("codegen", "Solver not initialized");
    return fail("codegen", "embedded code generation is not part of the CUDA library (use the reference library for codegen)");
}

namespace {
int store_sensitivity(const char* fn, double* dK_data, int dK_rows, int dK_cols, double* dP_data, int dP_rows, int dP_cols,
                      double* dC1_data, int dC1_rows, int dC1_cols, double* dC2_data, int dC2_rows, int dC2_cols) {
    if (!dK_data || !dP_data || !dC1_data || !dC2_data) return fail(fn, "null sensitivity matrix");
    if (dK_rows != g_nu || dK_cols != g_nx || dP_rows != g_nx || dP_cols != g_nx || dC1_rows != g_nu || dC1_cols != g_nu ||
        dC2_rows != g_nx || dC2_cols != g_nx)
        return fail(fn, "sensitivity matrices must be dK nu x nx, dP nx x nx, dC1 nu x nu, dC2 nx x nx");
    if (tmpc_batch_set_sensitivity(g_solver, dK_data, dP_data, dC1_data, dC2_data) != 0) return fail(fn, tmpc_last_error());
    return 0;
}
}  // namespace

// The reference does two things here (codegen.cpp:82-100): with adaptive_rho on it first stores the four matrices in
// the LIVE solver's cache - the only way the Julia package can load them (tests/test_sensitivity.jl:40-53) - and then
// writes the MCU sources.  The first half is kept (later solve_mpc calls adapt rho with these matrices); the second is
// outside this library, so the call still returns non-zero.
int codegen_with_sensitivity(const char* output_dir, double* dK_data, int dK_rows, int dK_cols, double* dP_data,
                             int dP_rows, int dP_cols, double* dC1_data, int dC1_rows, int dC1_cols, double* dC2_data,
                             int dC2_rows, int dC2_cols, int verbose) {
    (void)output_dir; (void)verbose;
    if (!g_solver) return fail("codegen_with_sensitivity", "Solver not initialized");
    if (g_set.adaptive_rho) {     // "Only store sensitivity matrices if adaptive rho is enabled", codegen.cpp:90-97
        const int rc = store_sensitivity("codegen_with_sensitivity", dK_data, dK_rows, dK_cols, dP_data, dP_rows, dP_cols,
                                         dC1_data, dC1_rows, dC1_cols, dC2_data, dC2_rows, dC2_cols);
        if (rc != 0) return rc;
    }
    return fail("codegen_with_sensitivity",
                g_set.adaptive_rho ? "the sensitivity matrices are stored; embedded code generation is not part of the CUDA library"
                                   : "embedded code generation is not part of the CUDA library");
}

// The loader on its own (no reference counterpart: there it only exists inside codegen_with_sensitivity).  Same
// argument list minus the output directory; stores regardless of the adaptive_rho flag; returns 0.
int set_sensitivity_matrices(double* dK_data, int dK_rows, int dK_cols, double* dP_data, int dP_rows, int dP_cols,
                             double* dC1_data, int dC1_rows, int dC1_cols, double* dC2_data, int dC2_rows, int dC2_cols,
                             int verbose) {
    if (!g_solver) return fail("set_sensitivity_matrices", "Solver not initialized");
    const int rc = store_sensitivity("set_sensitivity_matrices", dK_data, dK_rows, dK_cols, dP_data, dP_rows, dP_cols,
                                     dC1_data, dC1_rows, dC1_cols, dC2_data, dC2_rows, dC2_cols);
    if (rc == 0 && verbose) std::printf("Sensitivity matrices stored\n");
    return rc;
}

int print_problem_data(int verbose) {
    if (!g_solver) return fail("print_problem_data", "Solver not initialized");
    std::printf("=== TinyMPC Problem Data ===\n");
    std::printf("Solution: iter=%d, solved=%d\n", g_last_iter, g_solved);
    std::printf("Cache: rho=%g\n", g_rho);
    std::printf("Settings: max_iter=%d, abs_pri_tol=%g, abs_dua_tol=%g\n", g_set.max_iter, g_set.abs_pri_tol, g_set.abs_dua_tol);
    std::printf("Problem: nx=%d, nu=%d\n", g_nx, g_nu);
    if (verbose) {
        std::vector<double> X(static_cast<size_t>(g_nx) * g_N), U(static_cast<size_t>(g_nu) * (g_N - 1));
        std::vector<double> K(static_cast<size_t>(g_nu) * g_nx), P(static_cast<size_t>(g_nx) * g_nx);
        tmpc_batch_get_states(g_solver, X.data());
        tmpc_batch_get_controls(g_solver, U.data());
        tmpc_batch_get_cache_terms(g_solver, K.data(), P.data(), nullptr, nullptr, nullptr, nullptr);
        auto dump = [](const char* name, const std::vector<double>& M, int r, int c) {
            std::printf("%s:\n", name);
            for (int i = 0; i < r; ++i) { for (int j = 0; j < c; ++j) std::printf("%s%g", j ? " " : "", M[static_cast<size_t>(j) * r + i]); std::printf("\n"); }
        };
        dump("States x", X, g_nx, g_N); dump("Controls u", U, g_nu, g_N - 1);
        dump("Cache Kinf", K, g_nu, g_nx); dump("Cache Pinf", P, g_nx, g_nx);
    }
    std::fflush(stdout);
    return 0;
}

int set_cache_terms(double* Kinf_data, int Kinf_rows, int Kinf_cols, double* Pinf_data, int Pinf_rows, int Pinf_cols,
                    double* Quu_inv_data, int Quu_inv_rows, int Quu_inv_cols, double* AmBKt_data, int AmBKt_rows,
                    int AmBKt_cols, int verbose) {
    if (!g_solver) return fail("set_cache_terms", "Solver not initialized");
    if (Kinf_rows != g_nu || Kinf_cols != g_nx || Pinf_rows != g_nx || Pinf_cols != g_nx || Quu_inv_rows != g_nu ||
        Quu_inv_cols != g_nu || AmBKt_rows != g_nx || AmBKt_cols != g_nx)
        return fail("set_cache_terms", "cache matrices have the wrong shape");
    if (tmpc_batch_set_cache_terms(g_solver, Kinf_data, Pinf_data, Quu_inv_data, AmBKt_data) != 0)
        return fail("set_cache_terms", tmpc_last_error());
    if (verbose) {
        double nk = 0, np = 0;
        for (int i = 0; i < g_nu * g_nx; ++i) nk += Kinf_data[i] * Kinf_data[i];
        for (int i = 0; i < g_nx * g_nx; ++i) np += Pinf_data[i] * Pinf_data[i];
        std::printf("Cache terms set - Kinf norm: %g, Pinf norm: %g\n", std::sqrt(nk), std::sqrt(np));
    }
    return 0;
}

int update_settings(double abs_pri_tol, double abs_dua_tol, int max_iter, int check_termination, int en_state_bound,
                    int en_input_bound, int en_state_soc, int en_input_soc, int en_state_linear, int en_input_linear,
                    int adaptive_rho, double adaptive_rho_min, double adaptive_rho_max,
                    int adaptive_rho_enable_clipping, int verbose) {
    if (!g_solver) return fail("update_settings", "Solver not initialized");
    const int rc = tmpc_batch_update_settings(g_solver, abs_pri_tol, abs_dua_tol, max_iter, check_termination,
                                              en_state_bound, en_input_bound, en_state_soc, en_input_soc,
                                              en_state_linear, en_input_linear, adaptive_rho, adaptive_rho_min,
                                              adaptive_rho_max, adaptive_rho_enable_clipping);
    if (rc != 0) return fail("update_settings", tmpc_last_error());
    g_set.abs_pri_tol = abs_pri_tol; g_set.abs_dua_tol = abs_dua_tol; g_set.max_iter = max_iter;
    g_set.check_termination = check_termination;
    g_set.en_state_bound = en_state_bound; g_set.en_input_bound = en_input_bound;
    g_set.en_state_soc = en_state_soc; g_set.en_input_soc = en_input_soc;
    g_set.en_state_linear = en_state_linear; g_set.en_input_linear = en_input_linear;
    g_set.adaptive_rho = adaptive_rho; g_set.adaptive_rho_min = adaptive_rho_min; g_set.adaptive_rho_max = adaptive_rho_max;
    g_set.adaptive_rho_enable_clipping = adaptive_rho_enable_clipping;
    if (verbose) std::printf("Updated settings with status: 0\n");
    return 0;
}

int set_bound_constraints(double* x_min_data, int x_min_rows, int x_min_cols, double* x_max_data, int x_max_rows,
                          int x_max_cols, double* u_min_data, int u_min_rows, int u_min_cols, double* u_max_data,
                          int u_max_rows, int u_max_cols, int verbose) {
    if (!g_solver) return fail("set_bound_constraints", "Solver not initialized");
    // The reference computes a dimension status and ignores it (tiny_api.cpp:148-163),
    // then reads out of bounds during the solve; a mismatch is an error here.
    if (x_min_rows != g_nx || x_min_cols != g_N || x_max_rows != g_nx || x_max_cols != g_N || u_min_rows != g_nu ||
        u_min_cols != g_N - 1 || u_max_rows != g_nu || u_max_cols != g_N - 1)
        return fail("set_bound_constraints", "bound matrices must be nx x N and nu x (N-1)");
    if (tmpc_batch_set_bound_constraints(g_solver, x_min_data, x_max_data, u_min_data, u_max_data, 1) != 0)
        return fail("set_bound_constraints", tmpc_last_error());
    g_set.en_state_bound = 1; g_set.en_input_bound = 1;
    if (verbose) std::printf("set_bound_constraints status: 0\n");
    return 0;
}

int set_linear_constraints(double* Alin_x_data, int Alin_x_rows, int Alin_x_cols, double* blin_x_data, int blin_x_len,
                           double* Alin_u_data, int Alin_u_rows, int Alin_u_cols, double* blin_u_data, int blin_u_len,
                           int verbose) {
    if (!g_solver) return fail("set_linear_constraints", "Solver not initialized");
    // tiny_set_linear_constraints (tiny_api.cpp:201-242): status 1 on inconsistent shapes
    if (Alin_x_rows > 0 && (Alin_x_cols != g_nx || blin_x_len != Alin_x_rows)) return 1;
    if (Alin_u_rows > 0 && (Alin_u_cols != g_nu || blin_u_len != Alin_u_rows)) return 1;
    if (tmpc_batch_set_linear_constraints(g_solver, Alin_x_data, Alin_x_rows > 0 ? Alin_x_rows : 0, blin_x_data,
                                          Alin_u_data, Alin_u_rows > 0 ? Alin_u_rows : 0, blin_u_data) != 0)
        return fail("set_linear_constraints", tmpc_last_error());
    if (Alin_x_rows > 0 && blin_x_len > 0) g_set.en_state_linear = 1;
    if (Alin_u_rows > 0 && blin_u_len > 0) g_set.en_input_linear = 1;
    if (verbose) std::printf("set_linear_constraints status: 0\n");
    return 0;
}

int set_cone_constraints(int* Acu_data, int Acu_len, int* qcu_data, int qcu_len, double* cu_data, int cu_len,
                         int* Acx_data, int Acx_len, int* qcx_data, int qcx_len, double* cx_data, int cx_len,
                         int verbose) {
    if (!g_solver) return fail("set_cone_constraints", "Solver not initialized");
    // tiny_set_cone_constraints size checks (tiny_api.cpp:176-183)
    if (qcu_len != Acu_len || cu_len != Acu_len || qcx_len != Acx_len || cx_len != Acx_len) return 1;
    // crossed slots and flag rule of the reference live in the batched call
    if (tmpc_batch_set_cone_constraints_jl(g_solver, Acu_data, qcu_data, cu_data, Acu_len, Acx_data, qcx_data, cx_data,
                                           Acx_len) != 0)
        return fail("set_cone_constraints", tmpc_last_error());
    if (Acx_len > 0) g_set.en_state_soc = 1;
    if (Acu_len > 0) g_set.en_input_soc = 1;
    if (verbose) std::printf("set_cone_constraints status: 0\n");
    return 0;
}

int get_solve_info(int* iter, int* solved) {
    if (!g_solver) return -1;
    if (iter) *iter = g_last_iter;
    if (solved) *solved = g_solved;
    return 0;
}

}  // extern "C"
