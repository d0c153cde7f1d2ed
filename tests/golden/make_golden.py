"""Generate tests/golden/*.npz from the COMPILED, UNMODIFIED reference (oracle/_ref).

Run in the build container (needs /root/reference to have been compiled by `make -C oracle ref`):
    python tests/golden/make_golden.py
Each fixture holds the reference's outputs (solve return value, iteration count, solution states and
controls, final plant state) for one seeded case of tests/cases.py:GOLDEN_CASES, plus a digest of the
inputs so a test can tell when a case definition drifted from its fixture.  Two cases are additionally
run through the reference's own C ABI (oracle/_ref/libtinympc_jl.so, exactly what Julia dlopens) to
check that the batch driver and the ccall path agree.
"""
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import cases  # noqa: E402
import oracle  # noqa: E402


def digest(c):
    h = hashlib.sha256()
    for k in ("x0", "Xref", "Uref"):
        if c[k] is not None:
            h.update(np.ascontiguousarray(np.asfortranarray(c[k]).T).tobytes())
    s = c["sys"]
    for a in (s.A, s.B, s.f, s.Q, s.R):
        h.update(np.ascontiguousarray(a).tobytes())
    h.update(repr((c["N"], c["max_iter"], c["abs_pri_tol"], c["abs_dua_tol"], c["nsteps"], c["ref_mode"])).encode())
    if c.get("adaptive"):
        ad = c["adaptive"]
        h.update(repr((ad["rho_min"], ad["rho_max"], ad["clip"])).encode())
    return h.hexdigest()


def main():
    ref = oracle.load_ref()
    if ref is None:
        raise SystemExit("oracle/_ref is not built: run `make -C oracle ref` where /root/reference exists")
    only = set(sys.argv[1:])        # optional: names of the fixtures to (re)generate; default all
    for name, make in cases.GOLDEN_CASES.items():
        if only and name not in only:
            continue
        c = make()
        out = cases.run_checker(c, ref)
        extra = {}
        if "adapted" in out:        # adaptive rho: the cache every instance ends with
            extra = {"adapted_rho": out["adapted"]["rho"], "adapted_Kinf": out["adapted"]["Kinf"]}
        np.savez_compressed(os.path.join(HERE, name + ".npz"), ret=out["ret"], iters=out["iters"], X=out["X"], U=out["U"],
                            x_final=out["x_final"], digest=np.array(digest(c)), **extra)
        print(f"{name:32s} B={c['x0'].shape[1]:4d} steps={c['nsteps']:3d} iters[min/max]={out['iters'].min()}/{out['iters'].max()} "
              f"solved={(out['ret'] == 0).mean():.2f}")
    # ---- float build of the reference (oracle/Makefile target ref32)
    ref32 = oracle.load_ref32()
    if ref32 is None:
        raise SystemExit("oracle/_ref/libtinympc_refbatch32.so is not built: run `make -C oracle ref32`")
    for name in cases.GOLDEN_F32_CASES:
        if only and ("f32_" + name) not in only:
            continue
        c = cases.GOLDEN_CASES[name]()
        out = cases.run_checker(c, ref32)
        cache = ref32.get_cache(cases.to_oracle_problem(c))
        np.savez_compressed(os.path.join(HERE, "f32_" + name + ".npz"), ret=out["ret"], iters=out["iters"], X=out["X"],
                            U=out["U"], x_final=out["x_final"], digest=np.array(digest(c)),
                            **{"cache_" + k: v for k, v in cache.items()})
        print(f"f32 {name:28s} B={c['x0'].shape[1]:4d} iters[min/max]={out['iters'].min()}/{out['iters'].max()} "
              f"solved={(out['ret'] == 0).mean():.2f}")


if __name__ == "__main__":
    main()
