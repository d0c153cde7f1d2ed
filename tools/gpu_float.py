"""Float build (`typedef float tinytype`): parity against the reference compiled with the float
typedef, iteration-flip rates, and throughput of the float instantiation of the thread kernel.

Prints one JSON object.  Sections:
  golden     - every tests/golden/f32_*.npz case: status / iteration mismatches and the worst tolerance excess
  flip_*     - config 3 (quadrotor, B instances, easy and hard x0): iteration counts of
                 the float GPU solve  vs the float reference on the host (same cache installed)   -> parity at scale
                 the float GPU solve  vs the double GPU solve                                     -> what the precision costs
  time_*     - thread kernel, float vs double, same batch
"""
import json, os, sys, time
ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import cases, oracle
import tinympc_julia_b200 as t

out = {}
ref32 = oracle.load_ref32()

def with_cache(c, g):
    s = c["sys"]; n, m = s.nx, s.nu
    c["cache"] = {"Kinf": g["cache_Kinf"].reshape(m, n, order="F"), "Pinf": g["cache_Pinf"].reshape(n, n, order="F"),
                  "Quu_inv": g["cache_Quu_inv"].reshape(m, m, order="F"), "AmBKt": g["cache_AmBKt"].reshape(n, n, order="F")}
    return c

gold = {}
for name in cases.GOLDEN_F32_CASES:
    g = cases.load_golden("f32_" + name)
    c = with_cache(cases.GOLDEN_CASES[name](), g)
    got = cases.run_cuda(c, kernel=t.KERNEL_THREAD, precision=32)
    same = (got["iters"][-1] == g["iters"][-1]) & (got["ret"][-1] == g["ret"][-1])
    ex = 0.0
    for k in ("X", "U"):
        a, w = got[k][..., same], g[k][..., same]
        if a.size: ex = max(ex, float((np.abs(a - w) - (cases.ATOL_F32 + cases.RTOL_F32 * np.abs(w))).max()))
    gold[name] = {"B": int(same.size), "iter_mismatch": int((~same).sum()), "tol_excess": ex,
                  "max_abs_dU": float(np.abs(got["U"][..., same] - g["U"][..., same]).max()) if same.any() else None}
out["golden"] = gold

B = int(os.environ.get("QB", 65536))
for hard in (False, True):
    tag = "hard" if hard else "easy"
    c = cases.quadrotor_batch(B, hard=hard)
    r = {}
    if ref32 is not None:
        cache = ref32.get_cache(cases.to_oracle_problem(c))
        c["cache"] = {k: cache[k] for k in ("Kinf", "Pinf", "Quu_inv", "AmBKt")}
        nth = os.cpu_count() or 1
        want = cases.run_checker(c, ref32, nthreads=nth)
        r["ref32_host_seconds"] = want["seconds"]; r["ref32_threads"] = nth
    g32 = cases.run_cuda(c, kernel=t.KERNEL_THREAD, precision=32)
    g32s = cases.run_cuda(c, kernel=t.KERNEL_SUBWARP, precision=32)
    r["float_subwarp_equals_float_thread"] = bool(all(np.array_equal(g32[k], g32s[k]) for k in ("ret", "iters", "X", "U")))
    if ref32 is not None:
        same = (g32["iters"][-1] == want["iters"][-1]) & (g32["ret"][-1] == want["ret"][-1])
        r["gpu32_vs_ref32_iter_mismatch"] = int((~same).sum())
        r["gpu32_vs_ref32_flip_rate"] = float((~same).mean())
        r["gpu32_vs_ref32_max_iter_delta"] = int(np.abs(g32["iters"][-1] - want["iters"][-1]).max())
        ex = 0.0
        for k in ("X", "U"):
            a, w = g32[k][..., same], want[k][..., same]
            ex = max(ex, float((np.abs(a - w) - (cases.ATOL_F32 + cases.RTOL_F32 * np.abs(w))).max()))
        r["gpu32_vs_ref32_tol_excess"] = ex
        r["gpu32_vs_ref32_max_abs_dU"] = float(np.abs(g32["U"][..., same] - want["U"][..., same]).max())
    # float vs double on the GPU: the double solve uses ITS OWN (double) cache, as a user switching precision would
    c64 = cases.quadrotor_batch(B, hard=hard)
    g64 = cases.run_cuda(c64, kernel=t.KERNEL_AUTO)
    d = g32["iters"][-1] - g64["iters"][-1]
    r["gpu32_vs_gpu64_iter_differs"] = float((d != 0).mean())
    r["gpu32_vs_gpu64_status_differs"] = float((g32["ret"][-1] != g64["ret"][-1]).mean())
    r["gpu32_vs_gpu64_iter_delta_hist"] = {int(k): int(v) for k, v in zip(*np.unique(np.clip(d, -3, 3), return_counts=True))}
    r["mean_iters_32"] = float(g32["iters"].mean()); r["mean_iters_64"] = float(g64["iters"].mean())
    r["gpu32_vs_gpu64_max_abs_dU"] = float(np.abs(g32["U"] - g64["U"])[..., d == 0].max())
    # thread and sub-warp kernels, float vs double
    for bits, kern, kname in ((32, t.KERNEL_THREAD, "thread"), (64, t.KERNEL_THREAD, "thread"),
                              (32, t.KERNEL_SUBWARP, "subwarp"), (64, t.KERNEL_SUBWARP, "subwarp"), (64, t.KERNEL_TMEM, "tmem")):
        sol = cases.make_cuda_solver(cases.quadrotor_batch(B, hard=hard), kernel=kern, precision=bits)
        t.set_x0(sol, np.asfortranarray(c["x0"]))
        best = 1e9
        for rep in range(4):
            sol.reset()
            t0 = time.perf_counter(); sol.solve_async(0); sol.synchronize(); best = min(best, time.perf_counter() - t0)
        st, it = sol.status()
        r[f"{kname}_kernel_{bits}_ms"] = best * 1e3
        r[f"{kname}_kernel_{bits}_solves_per_s"] = B / best
        r[f"{kname}_kernel_{bits}_admm_iters_per_s"] = float(it.sum()) / best
        del sol
    out["flip_" + tag] = r
# ---- configs 2 (cartpole) and 4 (rocket: box + both cones + half-spaces, thread kernel): float vs double, flips vs the float reference
for tag, make in (("cartpole", lambda: cases.cartpole_batch(B)), ("rocket", lambda: cases.rocket(B, N=50))):
    r = {}
    c = make()
    want = None
    if ref32 is not None:
        cache = ref32.get_cache(cases.to_oracle_problem(c))
        c["cache"] = {k: cache[k] for k in ("Kinf", "Pinf", "Quu_inv", "AmBKt")}
        want = cases.run_checker(c, ref32, nthreads=os.cpu_count() or 1)
        r["ref32_host_seconds"] = want["seconds"]
    for bits in (32, 64):
        cc = c if bits == 32 else make()
        sol = cases.make_cuda_solver(cc, kernel=t.KERNEL_AUTO, precision=bits)
        t.set_x0(sol, np.asfortranarray(cc["x0"]))
        if cc["Xref"] is not None: t.set_x_ref(sol, cc["Xref"])
        if cc["Uref"] is not None: t.set_u_ref(sol, cc["Uref"])
        best = 1e9
        for rep in range(3):
            sol.reset()
            t0 = time.perf_counter(); sol.solve_async(0); sol.synchronize(); best = min(best, time.perf_counter() - t0)
        st, it = sol.status()
        r[f"kernel_{bits}"] = sol.kernel
        r[f"ms_{bits}"] = best * 1e3
        r[f"solves_per_s_{bits}"] = B / best
        r[f"admm_iters_per_s_{bits}"] = float(it.sum()) / best
        if bits == 32 and want is not None:
            same = (it == want["iters"][-1]) & (st == want["ret"][-1])
            r["gpu32_vs_ref32_flip_rate"] = float((~same).mean())
            r["gpu32_vs_ref32_max_iter_delta"] = int(np.abs(it - want["iters"][-1]).max())
        del sol
    out["config_" + tag] = r
print(json.dumps(out))
