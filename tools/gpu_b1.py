import sys, time, os, ctypes as C
sys.path.insert(0, "/root/repo/tests"); sys.path.insert(0, "/root/repo")
import numpy as np, torch
import cases
import tinympc_julia_b200 as t
from tinympc_julia_b200 import _lib, problems
lib = _lib.load()
s = problems.quadrotor(); N = 20
sol = t.TinyMPCSolver()
t.setup(sol, s.A, s.B, s.f, s.Q, s.R, s.rho, 12, 4, N)
t.set_bound_constraints(sol, *s.bounds(N))
rng = np.random.default_rng(0)
P = lambda a: a.ctypes.data_as(C.POINTER(C.c_double))
U = np.zeros((4, N - 1), order="F"); r = C.c_int(); c = C.c_int()
ts = {"set_x0": [], "solve": [], "get": []}
for k in range(400):
    x0 = np.asfortranarray(rng.uniform(-0.1, 0.1, (12, 1)))
    a = time.perf_counter(); lib.set_x0(P(x0), 12, 1, 0); b = time.perf_counter(); lib.solve_mpc(0); d = time.perf_counter(); lib.get_controls(P(U), C.byref(r), C.byref(c)); e = time.perf_counter()
    if k >= 100:
        ts["set_x0"].append(b - a); ts["solve"].append(d - b); ts["get"].append(e - d)
print({k: round(float(np.median(v)) * 1e6, 2) for k, v in ts.items()})
# kernel alone: batched handle with B = 1 on a torch stream with events
c1 = cases.quadrotor_batch(1)
bs = cases.make_cuda_solver(c1)
t.set_x0(bs, c1["x0"])
st = torch.cuda.Stream(); ev = []
for k in range(60):
    x0 = np.asfortranarray(rng.uniform(-0.1, 0.1, (12, 1))); t.set_x0(bs, x0)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(st):
        e0.record(st); bs.solve_async(st.cuda_stream); e1.record(st)
    torch.cuda.synchronize(); ev.append(e0.elapsed_time(e1) * 1e3)
stt, it = bs.status()
print("device us (memset + kernel) median", round(float(np.median(ev[10:])), 2), "iters", it, "kernel", bs.kernel)
for kern in (1, 2):
    bs.set_kernel(kern); ev = []
    for k in range(40):
        x0 = np.asfortranarray(rng.uniform(-0.1, 0.1, (12, 1))); t.set_x0(bs, x0)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(st):
            e0.record(st); bs.solve_async(st.cuda_stream); e1.record(st)
        torch.cuda.synchronize(); ev.append(e0.elapsed_time(e1) * 1e3)
    print("kernel", kern, "device us", round(float(np.median(ev[10:])), 2), bs.status()[1])
