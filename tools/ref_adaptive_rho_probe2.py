"""Probe (this container only): the compiled reference with adaptive_rho = 1 AND the sensitivity matrices loaded the way
tests/test_sensitivity.jl:40-53 loads them (codegen_with_sensitivity stores them in the live cache, codegen.cpp:91-96)."""
import ctypes as C, numpy as np, sys, os, shutil
sys.path.insert(0, '/root/repo')
from tinympc_julia_b200 import problems
which = sys.argv[1]
L = C.CDLL('/root/repo/oracle/_ref/libtinympc_jl.so')
s = problems.cartpole() if which == 'cartpole' else problems.quadrotor()
N = 20
def P(a): return np.asfortranarray(a, dtype=np.float64)
A, B, f, Q, R = P(s.A), P(s.B), P(np.asarray(s.f).reshape(-1, 1)), P(s.Q), P(s.R)
dp = C.POINTER(C.c_double)
def p(a): return a.ctypes.data_as(dp)
nx, nu = s.nx, s.nu

def solve_lqr(A, B, Q, R, rho):
    Qr = Q + rho * np.eye(nx); Rr = R + rho * np.eye(nu)
    Pm = Qr.copy(); K = np.zeros((nu, nx))
    for it in range(5000):
        Kp = K.copy()
        K = np.linalg.solve(Rr + B.T @ Pm @ B + 1e-8 * np.eye(nu), B.T @ Pm @ A)
        Pm = Qr + A.T @ Pm @ (A - B @ K)
        if it > 0 and np.linalg.norm(K - Kp) < 1e-10: break
    return K, Pm, np.linalg.inv(Rr + B.T @ Pm @ B), (A - B @ K).T

h = 1e-6
a0 = solve_lqr(A, B, Q, R, s.rho); a1 = solve_lqr(A, B, Q, R, s.rho + h)
dK, dP, dC1, dC2 = [P((y - x) / h) for x, y in zip(a0, a1)]
print('norms', [float(np.linalg.norm(m)) for m in (dK, dP, dC1, dC2)])
rc = L.setup_solver(p(A), nx, nx, p(B), nx, nu, p(f), nx, 1, p(Q), nx, nx, p(R), nu, nu, C.c_double(s.rho), nx, nu, N, 0)
print('setup', rc, flush=True)
L.update_settings.argtypes = [C.c_double, C.c_double] + [C.c_int] * 9 + [C.c_double, C.c_double, C.c_int, C.c_int]
rc = L.update_settings(1e-3, 1e-3, 100, 1, 0, 0, 0, 0, 0, 0, 1, 0.1, 10.0, 1, 0)
print('update_settings(adaptive_rho=1)', rc, flush=True)
out = '/root/repo/oracle/_ref/scratch/cg'
shutil.rmtree(out, ignore_errors=True)
rc = L.codegen_with_sensitivity(out.encode(), p(dK), nu, nx, p(dP), nx, nx, p(dC1), nu, nu, p(dC2), nx, nx, 0)
print('codegen_with_sensitivity', rc, flush=True)
umin = P(np.full((nu, N - 1), -0.5 if which != 'cartpole' else -5.0)); umax = P(-umin)
xmin = P(np.full((nx, N), -1e17)); xmax = P(-xmin)
print('bounds', L.set_bound_constraints(p(xmin), nx, N, p(xmax), nx, N, p(umin), nu, N - 1, p(umax), nu, N - 1, 0))
for trial in range(3):
    x0 = P(np.full((nx, 1), (float(sys.argv[2]) if len(sys.argv) > 2 else 0.3) * (1 + trial)))
    print('set_x0', L.set_x0(p(x0), nx, 1, 0), flush=True)
    print('solve', L.solve_mpc(0), flush=True)
    U = np.zeros((nu, N - 1), order='F'); r_ = C.c_int(0); c_ = C.c_int(0)
    print('get', L.get_controls(p(U), C.byref(r_), C.byref(c_)), U[:, 0])
