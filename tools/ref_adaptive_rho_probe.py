"""Probe (this container only): the compiled reference with adaptive_rho = 1 through its own Julia-facing symbols."""
import ctypes as C, numpy as np, sys
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
rc = L.setup_solver(p(A), nx, nx, p(B), nx, nu, p(f), nx, 1, p(Q), nx, nx, p(R), nu, nu, C.c_double(s.rho), nx, nu, N, 0)
print('setup', rc, flush=True)
L.update_settings.argtypes = [C.c_double, C.c_double] + [C.c_int] * 9 + [C.c_double, C.c_double, C.c_int, C.c_int]
rc = L.update_settings(1e-3, 1e-3, 100, 1, 0, 0, 0, 0, 0, 0, 1, 0.1, 10.0, 1, 0)
print('update_settings(adaptive_rho=1)', rc, flush=True)
x0 = P(np.full((nx, 1), float(sys.argv[2]) if len(sys.argv) > 2 else 0.3))
print('set_x0', L.set_x0(p(x0), nx, 1, 0), flush=True)
print('solve', L.solve_mpc(0), flush=True)
