#!/usr/bin/env python
"""bench.py — throughput of the batched ADMM solve path (BASELINE.json metric) on N GPUs of one node.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload W]      # this library (CUDA, sm_100a)
    python bench.py --impl reference [--gpus N] [--steps K] [--workload W]  # the reference's own CPU solver

Workloads (config.workload; SURVEY.md §8d, BASELINE.json configs):
  quadrotor_easy  (default, the headline = configs[2]) batched quadrotor hover nx=12 nu=4 N=20, box bounds shared, shared
                  Kinf/Pinf cache, rho=5, tol 1e-3, max_iter=100, B = 65,536 per GPU, x0 ~ U(+-0.1)^12, cold start
  quadrotor_hard  the same with x0 ~ U(+-0.5)^12 (about half of the instances run into max_iter)
  cartpole        configs[1]: nx=4 nu=1 N=20, x1 in +-2, u in +-5, x0 ~ U(+-1)^4 o [1, .5, .2, .5], B = 65,536
  rocket          configs[3]: nx=6 nu=3 N=50, box + both second-order cones (Julia argument order) + half-spaces
  rollout         configs[4]: closed loop, quadrotor, 200 warm-started MPC steps per instance, per-instance sliding
                  line references, B = 65,536 per GPU; one "step" of the bench = one 200-step rollout of the batch
Inputs come from numpy default_rng(seed); timed step k of rank r uses seed  base + 8 r + (k mod 8)  (eight input sets
resident in HBM), so a line is an average over eight seeds and no rank repeats another's inputs.

A "step" is one pass of the hot path over one batch: reset to the zero workspace + one batched solve (or rollout).
  value          QP solves/s, whole job, inputs resident in HBM, device time (CUDA events on the launch stream, max over ranks)
  admm_iters_per_s  ADMM iterations/s over the same region (sum of per-instance iteration counts)
  e2e            the same through the C ABI with HOST buffers (pinned): x0 up, solve, solution + status down, wall clock
                 (tmpc_batch_solve_host); e2e.u0_only: only the first control + status come home (what a closed loop reads)
  roofline       FP64 CUDA-core FMA roofline of the solve kernel (the binding limit, north_star): achieved =
                 sum(iterations) x F(nx,nu,N) flop / kernel time; peak = DFMA peak measured live on this GPU
                 (tmpc_measure_fma_peak; MEASURED_PEAKS.json has no CUDA-core figure), peak_nominal = SMs x 64 x 2 x clock
                 beside it; frac_useful_fma counts the mat-vec FMAs only (2M).  traffic = DRAM bytes per launch read from
                 the ncu --set full summary committed under profiles/ (ncu_full_<workload>_summary.csv).
  per_rank       kernel ms per step (min / median / max) and the largest iteration count of every rank
  cpu_baseline   the compiled reference core (oracle/_ref, kind "reference") or the C restatement (kind "port") on all
                 host cores, a bounded sample of the same instances — N=1, rank 0 only — and parity_full_batch against it
  single_solve   p50 latency of set_x0 -> solve_mpc -> get_controls at B=1 through the reference's own symbols
Only the `cpu_baseline` / `--impl reference` legs touch oracle/ (and tests/cases.py); the product path has no CPU fallback.
"""
from __future__ import annotations

import argparse
import csv
import ctypes as C
import json
import os
import statistics
import sys
import threading
import time

import numpy as np

# The contract is ONE JSON line on stdout.  Libraries write there too (NCCL prints its version banner on stdout during
# init), so everything else is sent to stderr: fd 1 is duplicated for the final line and then pointed at fd 2.
_JSON_OUT = os.fdopen(os.dup(1), "w")
os.dup2(2, 1)


def emit(line: dict) -> None:
    _JSON_OUT.write(json.dumps(line) + "\n")
    _JSON_OUT.flush()


ROOT = os.path.dirname(os.path.abspath(__file__))
E2E_CHUNKS = int(os.environ.get("TMPC_E2E_CHUNKS", "0"))   # 0 = the library default; sweep aid for tmpc_batch_solve_host
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

METRIC = "qp_solves_per_sec"
UNIT = "solves/s"
L2_FLUSH_BYTES = 256 << 20
NSEEDS = 8


# ------------------------------------------------------------------------------------------ workloads

class Workload:
    """One of BASELINE.json's configurations: the system, its constraints and the seeded per-instance inputs."""

    def __init__(self, key, batch, nsteps):
        from tinympc_julia_b200 import problems
        self.key, self.B, self.problems = key, batch, problems
        self.nsteps = 1
        self.tol_pri = self.tol_dua = 1e-3
        self.max_iter = 100
        self.cones = self.linear = None
        self.Xref = self.Uref = None
        if key in ("quadrotor_easy", "quadrotor_hard", "rollout"):
            self.sys, self.N, self.seed0 = problems.quadrotor(), 20, 3
            self.hard = key == "quadrotor_hard"
            if key == "rollout":
                self.nsteps, self.seed0 = nsteps, 5
        elif key == "cartpole":
            self.sys, self.N, self.seed0 = problems.cartpole(), 20, 2
        elif key == "rocket":
            self.sys, self.N, self.seed0 = problems.rocket(), 50, 4
            e = self.sys.extra
            self.cones = dict(Acu=e["Acu"], qcu=e["qcu"], cu=e["cu"], Acx=e["Acx"], qcx=e["qcx"], cx=e["cx"])
            self.linear = problems.rocket_linear_constraints()
            self.tol_pri = 2e-3
            _, self.Xref, self.Uref = problems.rocket_batch(1, self.N, self.seed0)
        else:
            raise SystemExit(f"unknown workload {key}")
        self.nx, self.nu = self.sys.nx, self.sys.nu
        self.flops = dict(problems.flops_per_iteration(self.nx, self.nu, self.N))
        self.flops["box_only_F"] = self.flops["F"]
        if self.cones is not None:
            # SURVEY.md 8d, rocket row: + ~30 flop per cone per column, 2 (len + 1) per half-space row per column, 5 flop per
            # element per extra slack / dual pair (cone and half-space pairs on both sides)
            n, m, N = self.nx, self.nu, self.N
            ncx, ncu = len(self.cones["Acu"]), len(self.cones["Acx"])          # Julia order: first triple -> state slots
            nlx, nlu = (len(self.linear[1]), len(self.linear[3])) if self.linear is not None else (0, 0)
            extra = 30 * (ncx * N + ncu * (N - 1)) + 2 * (n + 1) * nlx * N + 2 * (m + 1) * nlu * (N - 1)
            pairs = (1 if ncx else 0) + (1 if nlx else 0), (1 if ncu else 0) + (1 if nlu else 0)
            extra += 5 * (pairs[0] * n * N + pairs[1] * m * (N - 1))
            self.flops["F"] += extra

    def seed(self, rank, k):
        return self.seed0 + NSEEDS * rank + (k % NSEEDS)

    def x0(self, seed):
        p = self.problems
        if self.key in ("quadrotor_easy", "quadrotor_hard"):
            return p.quadrotor_batch_x0(self.B, self.hard, seed)
        if self.key == "cartpole":
            return p.cartpole_batch_x0(self.B, seed)
        if self.key == "rocket":
            return p.rocket_batch(self.B, self.N, seed)[0]
        return p.quadrotor_line_refs(self.B, self.N, self.nsteps, seed)[0]

    def xref_long(self, seed):
        return self.problems.quadrotor_line_refs(self.B, self.N, self.nsteps, seed)[1]

    def describe(self, precision=64):
        B = self.B
        prec = "" if precision == 64 else ", float build (tinytype = float)"
        if self.key in ("quadrotor_easy", "quadrotor_hard"):
            tag = "hard x0~U(+-0.5)" if self.hard else "easy x0~U(+-0.1)"
            return f"quadrotor hover nx=12 nu=4 N=20, box bounds shared, shared cache, B={B}/GPU, cold start, {tag}{prec}"
        if self.key == "cartpole":
            return f"cartpole nx=4 nu=1 N=20, box bounds shared (x1 in +-2, u in +-5), shared cache, B={B}/GPU, cold start{prec}"
        if self.key == "rocket":
            return (f"rocket landing nx=6 nu=3 N=50, box + state/input second-order cones (Julia argument order) + half-spaces, "
                    f"shared references, B={B}/GPU, cold start{prec}")
        return (f"closed loop: quadrotor nx=12 nu=4 N=20, {self.nsteps} warm-started MPC steps per instance, per-instance sliding "
                f"line references, B={B}/GPU, plant simulated on the owning GPU{prec}")

    def make_solver(self, device, precision=64, kernel=0):
        """The CUDA solver configured through the Julia-mirror API, in the order a Julia user would."""
        import tinympc_julia_b200 as t
        s = self.sys
        sol = t.TinyMPCBatchSolver()
        t.setup(sol, s.A, s.B, s.f, s.Q, s.R, s.rho, self.nx, self.nu, self.N, batch=self.B, device=device,
                abs_pri_tol=self.tol_pri, abs_dua_tol=self.tol_dua, max_iter=self.max_iter)
        if precision == 32:
            sol.set_precision(32)       # float build: setup redone in float, solves compute in float
        t.set_bound_constraints(sol, *s.bounds(self.N))
        if self.cones is not None:
            k = self.cones
            t.set_cone_constraints(sol, k["Acu"], k["qcu"], k["cu"], k["Acx"], k["qcx"], k["cx"])
        if self.linear is not None:
            t.set_linear_constraints(sol, *self.linear)
        if self.Xref is not None:
            t.set_x_ref(sol, self.Xref)
            t.set_u_ref(sol, self.Uref)
        if kernel:
            sol.set_kernel(kernel)
        return sol

    # ---- CPU legs only (tests/cases.py + oracle/)
    def cpu_case(self, seed, n=None):
        import cases
        B = self.B if n is None else n
        if self.key in ("quadrotor_easy", "quadrotor_hard"):
            c = cases.quadrotor_batch(self.B, hard=self.hard, seed=seed)
        elif self.key == "cartpole":
            c = cases.cartpole_batch(self.B, seed=seed)
        elif self.key == "rocket":
            c = cases.rocket(self.B, N=self.N, seed=seed)
        else:
            c = cases.quadrotor_closed_loop(self.B, self.nsteps, seed=seed)
        if B != self.B:     # the first n instances of the same seeded batch
            c["x0"] = np.asfortranarray(c["x0"][:, :B])
            if c["Xref"] is not None and c["Xref"].ndim == 3:
                c["Xref"] = np.asfortranarray(c["Xref"][:, :, :B])
        return c

    def cpu_sample_size(self):
        """Instances of one CPU pass, sized for a few seconds on the box's host cores."""
        return {"rocket": 8192, "rollout": 4096}.get(self.key, self.B)


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return json.load(f), "measured"
    except Exception:
        return {"hbm_gbs": 6650.0}, "fallback"


def ncu_traffic(workload, precision, kernel_name_part):
    """DRAM bytes per launch of the solve kernel from the committed ncu --set full summary of this workload
    (profiles/ncu_full_<workload>[_f32]_summary.csv, written by tools/ncu_summary.py from the capture at HEAD)."""
    path = os.path.join(ROOT, "profiles", f"ncu_full_{workload}{'_f32' if precision == 32 else ''}_summary.csv")
    try:
        rows = {r[0]: r for r in csv.reader(open(path)) if r}
        if kernel_name_part not in rows["Kernel Name"][2]:
            return None, None
        scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
        tot = 0.0
        for k in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
            vals = [float(v) for v in rows[k][2:] if v]
            tot += scale[rows[k][1]] * sum(vals) / len(vals)
        return tot, os.path.relpath(path, ROOT)
    except Exception:
        return None, None


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons of one GPU through NVML while the timed region runs."""

    def __init__(self, torch_index: int, period_s: float = 0.004):
        super().__init__(daemon=True)
        self.period, self.samples, self.reasons = period_s, [], set()
        self.max_mhz, self.ok = None, False
        self._halt = threading.Event()
        try:
            import pynvml
            import torch
            pynvml.nvmlInit()
            self.nv = pynvml
            try:
                uuid = "GPU-" + str(torch.cuda.get_device_properties(torch_index).uuid)
                self.h = pynvml.nvmlDeviceGetHandleByUUID(uuid.encode() if hasattr(uuid, "encode") else uuid)
            except Exception:
                vis = os.environ.get("CUDA_VISIBLE_DEVICES")
                idx = int(vis.split(",")[torch_index]) if vis and vis.split(",")[torch_index].isdigit() else torch_index
                self.h = pynvml.nvmlDeviceGetHandleByIndex(idx)
            self.max_mhz = int(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            self.ok = True
        except Exception as e:  # pragma: no cover - depends on the box
            self.err = repr(e)

    NAMES = {0x1: "gpu_idle", 0x2: "applications_clocks_setting", 0x4: "sw_power_cap", 0x8: "hw_slowdown",
             0x10: "sync_boost", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown", 0x80: "hw_power_brake_slowdown",
             0x100: "display_clock_setting"}

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        while not self._halt.is_set():
            try:
                self.samples.append(int(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
                try:
                    r = int(nv.nvmlDeviceGetCurrentClocksEventReasons(self.h))
                except Exception:
                    r = int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
                for bit, name in self.NAMES.items():
                    if r & bit and name != "gpu_idle":
                        self.reasons.add(name)
            except Exception:
                pass
            self._halt.wait(self.period)

    def stop(self):
        self._halt.set()
        if self.is_alive():
            self.join(timeout=2)
        med = int(statistics.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(self.samples)}


def silence_stdout():
    """The reference prints one line per converged solve (admm.cpp:373): send fd 1 to /dev/null meanwhile."""
    sys.stdout.flush()
    saved = os.dup(1)
    devnull = os.open(os.devnull, os.O_WRONLY)
    os.dup2(devnull, 1)
    os.close(devnull)
    return saved


def restore_stdout(saved):
    os.dup2(saved, 1)
    os.close(saved)


def config_dict(wl, args):
    """Identical for both arms (the driver compares the two `config` dicts); what differs per arm is in `impl_config`."""
    return {"workload": wl.describe(args.precision), "nx": wl.nx, "nu": wl.nu, "N": wl.N, "batch_per_gpu": wl.B,
            "mpc_steps_per_instance": wl.nsteps,
            "seeds": f"numpy default_rng({wl.seed0} + 8 rank + (step mod 8)): eight input sets per rank"}


def impl_config(kernel, world, impl):
    if impl == "b200":
        return {"kernel": kernel, "l2": "flushed between timed steps (256 MiB write, outside the event pair)",
                "timing": "CUDA events on the launch stream per step, summed, max over ranks",
                "parallelism": f"instances sharded over {world} GPU(s), no collective inside the solve",
                "inputs": "eight input sets per rank resident in HBM before the timed region"}
    return {"kernel": kernel, "l2": "n/a (CPU)",
            "timing": "wall clock of the instance loop (std::chrono inside the driver), all host threads",
            "parallelism": "instances looped over the host threads, rank 0 only; n_gpus is nominal",
            "inputs": "a bounded sample of the step's seeded batch (cpu_baseline.sample)"}


# ------------------------------------------------------------------------------------------ CPU legs (oracle/)

def load_cpu_checker(precision=64):
    import oracle
    if precision == 32:
        r32 = oracle.load_ref32()
        if r32 is not None:
            return r32
    ref = oracle.load_ref()
    if ref is not None:
        return ref
    return oracle.load_oracle()


def cpu_run(lib, wl, seed, n, nthreads):
    """One CPU pass over the first n instances of the seeded batch.  Returns (seconds, total iterations, out)."""
    import cases
    c = wl.cpu_case(seed, n)
    out = cases.run_checker(c, lib, nthreads=nthreads)
    return out["seconds"], int(out["iters"].sum()), out


def cpu_single_solve_p50(n=300):
    """p50 of set_x0 -> solve_mpc -> get_controls on the compiled reference's own C ABI (what Julia ccalls)."""
    import oracle
    from tinympc_julia_b200 import _lib, problems
    if not os.path.exists(oracle.REFJL_SO):
        return None
    lib = C.CDLL(oracle.REFJL_SO)
    for name, (res, args) in _lib.SIGNATURES.items():
        if hasattr(lib, name) and not name.startswith("tmpc_") and name != "get_solve_info":
            fn = getattr(lib, name)
            fn.restype, fn.argtypes = res, args
    return _single_solve_loop(lib, n, problems)


def _single_solve_loop(lib, n, problems):
    NX, NU, NH = 12, 4, 20
    s = problems.quadrotor()
    P = lambda a: a.ctypes.data_as(C.POINTER(C.c_double))
    f64 = lambda a: np.asfortranarray(np.asarray(a, dtype=np.float64))
    A, Bm, f, Q, R = f64(s.A), f64(s.B), f64(s.f.reshape(-1, 1)), f64(s.Q), f64(s.R)
    saved = silence_stdout()
    try:
        rc = lib.setup_solver(P(A), NX, NX, P(Bm), NX, NU, P(f), NX, 1, P(Q), NX, NX, P(R), NU, NU, float(s.rho), NX, NU, NH, 0)
        if rc != 0:
            return None
        lib.update_settings(1e-3, 1e-3, 100, 1, 0, 0, 0, 0, 0, 0, 0, 0.1, 10.0, 1, 0)
        xm, xM, um, uM = s.bounds(NH)
        lib.set_bound_constraints(P(xm), NX, NH, P(xM), NX, NH, P(um), NU, NH - 1, P(uM), NU, NH - 1, 0)
        x0s = problems.quadrotor_batch_x0(n + 20, hard=False, seed=3)
        ubuf = np.zeros((NU, NH - 1), order="F")
        r_, c_ = C.c_int(), C.c_int()
        ts = []
        for k in range(n + 20):
            x = np.ascontiguousarray(x0s[:, k])
            t0 = time.perf_counter()
            lib.set_x0(P(x), NX, 1, 0)
            lib.solve_mpc(0)
            lib.get_controls(P(ubuf), C.byref(r_), C.byref(c_))
            t1 = time.perf_counter()
            if k >= 20:
                ts.append((t1 - t0) * 1e6)
        lib.cleanup_solver()
    finally:
        restore_stdout(saved)
    ts.sort()
    return {"p50_us": round(ts[len(ts) // 2], 2), "p99_us": round(ts[int(len(ts) * 0.99)], 2), "n": len(ts),
            "note": "warm-started sequence of random x0 (the reference keeps duals between solves)"}


def run_reference(args, rank, world):
    """--impl reference: the reference's own CPU implementation on all host cores, same config and metric; a step is a
    bounded sample of the workload (cpu_sample_size instances of the step's seed)."""
    if rank != 0:
        return
    wl = Workload(args.workload, args.batch, args.rollout_steps)
    lib = load_cpu_checker(args.precision)
    cores = os.cpu_count() or 1
    n = wl.cpu_sample_size()
    for k in range(args.warmup):
        cpu_run(lib, wl, wl.seed(0, k), min(n, max(cores * 16, 256)), cores)
    secs, iters = [], 0
    for k in range(args.steps):
        s, it, _o = cpu_run(lib, wl, wl.seed(0, k), n, cores)
        secs.append(s)
        iters += it
    total = sum(secs)
    value = n * wl.nsteps * args.steps / total
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64" if args.precision == 64 else "f32", "data": "synthetic",
        "admm_iters_per_s": iters / total,
        "config": config_dict(wl, args),
        "impl_config": impl_config("reference CPU solver (admm.cpp) compiled here" if lib.kind.startswith("reference") else "C restatement of admm.cpp", world, "reference"),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": lib.kind,
                         "sample": f"the first {n} of the {wl.B} instances of the step's seed"
                                   + (f", {wl.nsteps} closed-loop steps each" if wl.nsteps > 1 else "") + f", {args.steps} steps"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


# ------------------------------------------------------------------------------------------ GPU arm

def bind_to_gpu_numa(torch, local_rank):
    """Multi-GPU runs: pin this rank's threads to the CPUs local to its GPU (sysfs local_cpulist of the PCI device) so
    that its pinned host buffers are allocated on that NUMA node; 8 ranks sharing one node's memory do not scale on the
    host <-> device leg.  Best effort: returns a short description or None."""
    try:
        p = torch.cuda.get_device_properties(local_rank)
        bus = f"{p.pci_domain_id:04x}:{p.pci_bus_id:02x}:{p.pci_device_id:02x}.0"
        base = f"/sys/bus/pci/devices/{bus}"
        cpus = set()
        for part in open(f"{base}/local_cpulist").read().strip().split(","):
            lo, _, hi = part.partition("-")
            cpus.update(range(int(lo), int(hi or lo) + 1))
        cpus &= os.sched_getaffinity(0)
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        node = open(f"{base}/numa_node").read().strip()
        return {"pci": bus, "numa_node": int(node), "cpus": len(cpus)}
    except Exception:
        return None


def run_gpu(args, rank, local_rank, world):
    import torch
    import torch.distributed as dist
    import tinympc_julia_b200 as t
    from tinympc_julia_b200 import _lib, problems, sharding

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the solve path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    numa = None
    if world > 1:
        numa = bind_to_gpu_numa(torch, local_rank)      # before any pinned allocation
        dist.init_process_group("nccl", device_id=dev)
    lib = _lib.load()
    wl = Workload(args.workload, args.batch, args.rollout_steps)
    B, K, W = wl.B, args.steps, max(args.warmup, 3)
    NX, NU, NH, NS = wl.nx, wl.nu, wl.N, wl.nsteps
    rollout = wl.key == "rollout"
    F = wl.flops["F"]

    sol = wl.make_solver(local_rank, args.precision, args.kernel)
    if args.order >= 0:
        sol.set_order(args.order)
    nseeds = min(NSEEDS, K + W) if not rollout else 1       # the rollout keeps ONE input set (1.4 GB of references per set)
    seeds = [wl.seed(rank, k) for k in range(nseeds)]
    x0_host = [wl.x0(s) for s in seeds]                                         # (nx, B) Fortran == (B, nx) C
    x0_pinned = [torch.from_numpy(np.ascontiguousarray(x.T)).pin_memory() for x in x0_host]
    x0_dev = [x.to(dev) for x in x0_pinned]
    xl_pinned = xl_dev = st_out = it_out = None
    if rollout:
        xl = wl.xref_long(seeds[0])                                             # (nx, N + NS - 1, B)
        xl_pinned = torch.from_numpy(np.ascontiguousarray(np.asfortranarray(xl).transpose(2, 1, 0))).pin_memory()
        del xl
        xl_dev = xl_pinned.to(dev)
        st_out = torch.full((NS, B), -1, dtype=torch.int32, device=dev)
        it_out = torch.full((NS, B), -1, dtype=torch.int32, device=dev)
    flush = torch.empty(L2_FLUSH_BYTES, dtype=torch.uint8, device=dev)
    # a real (non-default) stream: the library treats a NULL stream argument as "use the handle's own stream",
    # which CUDA events recorded on torch's default stream would not see
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)
    sptr = stream.cuda_stream
    assert sptr != 0

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step(k):
        sol.bind_x0_device(x0_dev[k % nseeds].data_ptr())
        sol.reset()                 # cold start: logical reset, the kernel rebuilds all carried state
        if rollout:                 # the plant state is advanced in place on an internal copy of x0
            sol.rollout_async(NS, xl_dev.data_ptr(), st_out.data_ptr(), it_out.data_ptr(), sptr)
        else:
            sol.solve_async(sptr)

    # iteration counts per input set (and the results kept for the parity leg)
    iters_seed, maxit_seed, solved_seed, keep = [], [], [], {}
    for j in range(nseeds):
        step(j)
        torch.cuda.synchronize()
        if rollout:
            iters_seed.append(int(it_out.sum(dtype=torch.int64).item()))
            maxit_seed.append(int(it_out.max().item()))
            solved_seed.append(int((st_out == 0).sum().item()))
        else:
            st, it = sol.status()
            iters_seed.append(int(it.sum())); maxit_seed.append(int(it.max())); solved_seed.append(int((st == 0).sum()))
            if world == 1 and not args.no_cpu:
                keep[j] = (st.copy(), it.copy())
    for k in range(W):
        flush.zero_()
        step(k)
    torch.cuda.synchronize()
    l0 = sol.launch_count

    sampler = ClockSampler(local_rank)
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
    barrier()
    sampler.start()
    w0 = time.perf_counter()
    for k in range(K):
        flush.zero_()               # L2 flush between timed steps (untimed: outside the event pair)
        ev[k][0].record(stream)
        step(k)
        ev[k][1].record(stream)
    barrier()
    w1 = time.perf_counter()
    clocks = sampler.stop()
    step_ms = [a.elapsed_time(b) for a, b in ev]
    launches = sol.launch_count - l0
    iters_timed = sum(iters_seed[k % nseeds] for k in range(K))
    solved_timed = sum(solved_seed[k % nseeds] for k in range(K))

    # multi-GPU plumbing (tinympc_julia_b200/sharding.py): device time = max over ranks; counts = sum over ranks; the one
    # collective of the path gathers the per-instance (status, iterations) over NCCL, outside the solve
    tot_ms_max = sharding.max_over_ranks(sum(step_ms), dev)
    all_iters, all_solved, all_launches = sharding.sum_over_ranks([iters_timed, solved_timed, launches], dev)
    if rollout:     # [nsteps][B] status / iteration counts of every rank (north_star: the only collective of the closed loop)
        if world > 1:
            gat = torch.empty((world, 2, NS, B), dtype=torch.int32, device=dev)
            dist.all_gather_into_tensor(gat, torch.stack([st_out, it_out]))
            assert int(gat[:, 1].sum(dtype=torch.int64).item()) * 1.0 == sharding.sum_over_ranks([iters_seed[(K - 1) % nseeds]], dev)[0]
            gathered = int(gat.numel())
            del gat
        else:
            gathered = int(2 * st_out.numel())
    else:
        st_last, it_last = sol.status()
        st_all, it_all = sharding.gather_status(st_last, it_last, dev)
        assert st_all.size == B * world
        gathered = int(2 * st_all.size)
    total_s = tot_ms_max * 1e-3
    value = B * NS * world * K / total_s
    iters_per_s = all_iters / total_s
    my = {"rank": rank, "kernel_ms_min": min(step_ms), "kernel_ms_median": statistics.median(step_ms), "kernel_ms_max": max(step_ms),
          "kernel_ms_mean": sum(step_ms) / K, "max_iterations": max(maxit_seed), "seeds": seeds}
    if world > 1:
        per_rank = [None] * world
        dist.all_gather_object(per_rank, my)
    else:
        per_rank = [my]

    # ---- live FMA peak + roofline of the solve kernel (one solve launch per step; per-rank numbers, rank 0 reported)
    pk = C.c_double(0.0)
    rc = lib.tmpc_measure_fma_peak(local_rank, 1, 5, C.byref(pk))
    pk32 = C.c_double(0.0)
    lib.tmpc_measure_fma_peak(local_rank, 0, 5, C.byref(pk32))
    f64 = args.precision == 64
    peak = (pk.value if rc == 0 else None) if f64 else pk32.value
    nsm = torch.cuda.get_device_properties(local_rank).multi_processor_count
    clk = (clocks.get("sm_mhz") or clocks.get("sm_max_mhz") or 1965) * 1e6
    peak_nominal = nsm * (64 if f64 else 128) * 2 * clk / 1e12
    my_s = sum(step_ms) * 1e-3
    kern_ms = sum(step_ms) / K
    achieved_tf = iters_timed * F / my_s / 1e12
    useful_tf = iters_timed * 2 * wl.flops["M"] / my_s / 1e12
    peaks_json, peaks_kind = peaks()
    nxN, nuN = NX * NH, NU * (NH - 1)
    if rollout:
        bytes_in = 8 * NX + 8 * NX * (NH + NS - 1)                # x0 + the long per-instance reference, read once
        bytes_out = 8 * (nxN + nuN) + 8 * NS + 8 * NX             # last solution + per-step status/iterations + plant state
    else:
        bytes_in = 8 * NX                                         # x0; bounds / references are shared (staged once per CTA)
        bytes_out = 8 * (nxN + nuN) + 8                           # solution + status/iterations
    bytes_state = 8 * (3 * nxN - NX + 3 * nuN + NX)               # carried state written for the next warm start: g, v, q, y, z, r, p_N
    if wl.cones is not None:
        bytes_state += 8 * 2 * (nxN + nuN)                        # + cone and half-space duals
    hbm_bytes = B * (bytes_in + bytes_out + bytes_state)
    hbm_gbs = hbm_bytes / (kern_ms * 1e-3) / 1e9

    kname = {1: "thread", 2: "subwarp", 3: "tmem", 4: "cone"}.get(sol.kernel, str(sol.kernel))
    kernel_symbol = {1: "admm_thread_kernel", 2: "admm_subwarp_kernel", 3: "admm_tmem_kernel", 4: "admm_cone_kernel"}.get(sol.kernel, "admm")
    traffic, traffic_src = ncu_traffic(wl.key, args.precision, kernel_symbol)
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": tot_ms_max / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64" if f64 else "f32", "data": "synthetic",
        "admm_iters_per_s": iters_per_s,
        "mean_iters_per_solve": all_iters / (B * NS * world * K), "solved_fraction": all_solved / (B * NS * world * K),
        "config": config_dict(wl, args),
        "impl_config": impl_config(kname, world, "b200"),
        "gpu_launches": int(all_launches),
        "wall_ms_per_step_incl_flush": 1e3 * (w1 - w0) / K,
        "clocks": clocks,
        "per_rank": per_rank,
        "collective": {"op": "all_gather of int32 (status, iterations)" + (" [nsteps][B]" if rollout else " [B]") + " per rank, outside the timed region",
                       "backend": "nccl" if world > 1 else "none (1 rank)", "elements": gathered},
        "roofline": {"bound": "fp64_fma" if f64 else "fp32_fma", "achieved": achieved_tf, "peak": peak, "unit": "TFLOP/s",
                     "frac": (achieved_tf / peak) if peak else None,
                     "peak_nominal": peak_nominal, "frac_nominal": achieved_tf / peak_nominal,
                     "peak_nominal_note": f"{nsm} SMs x {64 if f64 else 128} FMA/clk x 2 x {clk / 1e6:.0f} MHz (median SM clock sampled under load)",
                     "achieved_useful_fma": useful_tf, "frac_useful_fma": (useful_tf / peak) if peak else None,
                     "traffic": traffic, "traffic_source": traffic_src,
                     "flop_per_iter": F, "flop_per_iter_box_part": wl.flops["box_only_F"], "fma_per_iter": wl.flops["M"], "kernel_ms": kern_ms,
                     "peak_source": "FMA chains measured live on this GPU by tmpc_measure_fma_peak (MEASURED_PEAKS.json has no CUDA-core peak)",
                     "fp32_fma_peak_tflops": pk32.value},
        "roofline_hbm": {"bound": "hbm", "achieved": hbm_gbs, "peak": peaks_json.get("hbm_gbs"), "unit": "GB/s",
                         "frac": hbm_gbs / peaks_json.get("hbm_gbs", 6650.0), "peak_source": peaks_kind,
                         "algorithmic_bytes": hbm_bytes, "traffic": traffic,
                         "bytes_per_instance": {"in": bytes_in, "out": bytes_out, "carried_state_out": bytes_state}},
    }

    # ---- e2e through the C ABI with host buffers (every rank, same K steps)
    dp = lambda tt: C.cast(tt.data_ptr(), C.POINTER(C.c_double))
    ip = lambda tt: C.cast(tt.data_ptr(), C.POINTER(C.c_int))
    h = sol._h
    lib.tmpc_batch_bind_x0_device(h, None)
    if rollout:
        stp = torch.empty((NS, B), dtype=torch.int32).pin_memory()
        itp = torch.empty((NS, B), dtype=torch.int32).pin_memory()
        xfp = torch.empty((B, NX), dtype=torch.float64).pin_memory()
        x0_in = torch.empty_like(x0_dev[0])

        def e2e_step():
            # host buffers in: x0 and the per-instance reference trajectories; out: per-step status / iterations, final plant state
            x0_in.copy_(x0_pinned[0], non_blocking=True)
            xl_dev.copy_(xl_pinned, non_blocking=True)
            sol.bind_x0_device(x0_in.data_ptr())
            sol.reset()
            sol.rollout_async(NS, xl_dev.data_ptr(), st_out.data_ptr(), it_out.data_ptr(), sptr)
            stp.copy_(st_out, non_blocking=True); itp.copy_(it_out, non_blocking=True)
            assert lib.tmpc_batch_get_x0(h, C.cast(xfp.data_ptr(), C.c_void_p), 0) == 0     # synchronises the handle
            torch.cuda.synchronize()

        KE = max(2, K // 4)
        e2e_step()
        barrier()
        e0 = time.perf_counter()
        for _ in range(KE):
            e2e_step()
        barrier()
        e2e_s = sharding.max_over_ranks(time.perf_counter() - e0, dev)
        assert int(itp.sum(dtype=torch.int64).item()) == iters_seed[0], "e2e rollout solved a different problem"
        line["e2e"] = {"value": B * NS * world * KE / e2e_s, "unit": UNIT, "steps": KE,
                       "h2d_bytes_per_step": int(x0_pinned[0].numel() * 8 + xl_pinned.numel() * 8),
                       "d2h_bytes_per_step": int(stp.numel() * 4 + itp.numel() * 4 + xfp.numel() * 8),
                       "ms_per_step": 1e3 * e2e_s / KE,
                       "path": "pinned host x0 + reference trajectories up, tmpc_batch_rollout_async (one fused launch), per-step status / iterations + final plant state down"}
    else:
        X_host = torch.empty((B, NH, NX), dtype=torch.float64).pin_memory()
        U_host = torch.empty((B, NH - 1, NU), dtype=torch.float64).pin_memory()
        u0_host = torch.empty((B, NU), dtype=torch.float64).pin_memory()
        st_host = torch.empty(B, dtype=torch.int32).pin_memory()
        it_host = torch.empty(B, dtype=torch.int32).pin_memory()

        def e2e_step(k):
            # one call: x0 up, solve cut into instance chunks, each chunk's solution crossing PCIe while the next one solves
            assert lib.tmpc_batch_solve_host(h, dp(x0_pinned[k % nseeds]), 1, dp(X_host), dp(U_host), ip(st_host), ip(it_host), E2E_CHUNKS) in (0, 1)

        def e2e_step_u0(k):
            # the same with only the first control + status coming home (what an MPC loop reads, cartpole_example_mpc.jl:41-42)
            assert lib.tmpc_batch_solve_host_select(h, dp(x0_pinned[k % nseeds]), 1, None, None, dp(u0_host), ip(st_host), ip(it_host), E2E_CHUNKS) in (0, 1)

        def e2e_step_calls(k):
            # the same through the five reference-style calls (no overlap): reported next to the pipelined number
            assert lib.tmpc_batch_set_x0(h, dp(x0_pinned[k % nseeds]), NX, B) == 0
            assert lib.tmpc_batch_reset(h) == 0
            assert lib.tmpc_batch_solve(h) in (0, 1)
            assert lib.tmpc_batch_get_states(h, dp(X_host)) == 0
            assert lib.tmpc_batch_get_controls(h, dp(U_host)) == 0
            assert lib.tmpc_batch_get_status(h, ip(st_host), ip(it_host)) == 0

        def timed(fn, n):
            for k in range(2):
                fn(k)
            barrier()
            c0 = time.perf_counter()
            for k in range(n):
                fn(k)
            barrier()
            return sharding.max_over_ranks(time.perf_counter() - c0, dev)

        KC = max(3, K // 5)
        calls_s = timed(e2e_step_calls, KC) / KC
        e2e_step_calls(0)
        X_chk = X_host[:64].clone(); U_chk = U_host[:64].clone()
        u0_s = timed(e2e_step_u0, K) / K
        e2e_step_u0(0)
        assert torch.equal(u0_host[:64], U_chk[:, 0, :]), "u0-only e2e path returned a different first control"
        e2e_s = timed(e2e_step, K)
        e2e_step(0)
        assert int(it_host.sum()) == iters_seed[0], "e2e path solved a different problem than the device-resident path"
        assert torch.equal(X_chk, X_host[:64]) and torch.equal(U_chk, U_host[:64]), "pipelined e2e path returned a different solution"
        line["e2e"] = {"value": B * world * K / e2e_s, "unit": UNIT,
                       "h2d_bytes_per_step": int(x0_pinned[0].numel() * 8),
                       "d2h_bytes_per_step": int(X_host.numel() * 8 + U_host.numel() * 8 + 2 * 4 * B),
                       "ms_per_step": 1e3 * e2e_s / K,
                       "path": "tmpc_batch_solve_host on pinned host buffers (x0 chunks up on their own stream, chunked solve overlapped with the D2H of the previous chunk's solution)",
                       "unpipelined_ms_per_step": 1e3 * calls_s,
                       "unpipelined_path": "tmpc_batch_set_x0 / reset / solve / get_states / get_controls / get_status",
                       "u0_only": {"value": B * world / u0_s, "unit": UNIT, "ms_per_step": 1e3 * u0_s,
                                   "d2h_bytes_per_step": int(u0_host.numel() * 8 + 2 * 4 * B),
                                   "path": "tmpc_batch_solve_host_select(X = U = NULL, u0): the first control of every instance + status / iterations come home",
                                   "vs_device_value": (B * world / u0_s) / value}}
    if numa is not None:
        line["e2e"]["host_binding_rank0"] = numa

    # ---- CPU baseline + parity against it + single-solve latency: N=1, rank 0 only
    if world == 1 and rank == 0 and not args.no_cpu:
        cpu = load_cpu_checker(args.precision)
        cores = os.cpu_count() or 1
        n = wl.cpu_sample_size()
        secs, cpu_iters, out = cpu_run(cpu, wl, seeds[0], n, cores)
        passes, first_secs = 1, secs
        if secs < 2.5:      # a short pass (config 2 / 3: a fraction of a second): the first one warms the thread pool up, then ~10 s of passes
            tot_s, tot_it, passes = 0.0, 0, 0
            while tot_s < 10.0 and passes < 200:
                s_k, it_k, _ = cpu_run(cpu, wl, seeds[0], n, cores)
                tot_s += s_k; tot_it += it_k; passes += 1
            secs, cpu_iters = tot_s / passes, tot_it / passes
        line["cpu_baseline"] = {"value": n * NS / secs, "unit": UNIT, "cores": cores, "kind": cpu.kind,
                                "admm_iters_per_s": cpu_iters / secs, "seconds": secs * passes,
                                "sample": f"the first {n} of the {B} instances of seed {seeds[0]}"
                                          + (f", {NS} closed-loop steps each" if NS > 1 else "")
                                          + (f", mean of {passes} passes after a warm-up pass ({first_secs:.3f} s)" if passes > 1 or first_secs < 2.5 else ", one pass")
                                          + f", {cores} host threads"}
        par = {"instances_checked": 0, "seeds_checked": [], "iter_mismatches": 0, "status_mismatches": 0, "max_abs_dU": 0.0}
        if rollout:
            st_g, it_g = st_out[:, :n].cpu().numpy(), it_out[:, :n].cpu().numpy()     # left by the last e2e rollout (same inputs)
            par.update(instances_checked=n, seeds_checked=[seeds[0]], steps_checked=NS,
                       iter_mismatches=int((out["iters"] != it_g).sum()), status_mismatches=int((out["ret"] != st_g).sum()),
                       max_abs_dx_final=float(np.abs(out["x_final"] - xfp[:n].numpy().T).max()))
            par.pop("max_abs_dU")
        else:
            par_seeds = list(range(nseeds)) if n == B and args.precision == 64 and wl.key != "rocket" else [0]
            for j in par_seeds:
                if j > 0:
                    _s, _i, out = cpu_run(cpu, wl, seeds[j], n, cores)
                e2e_step(j)                                                   # this seed's GPU solution into U_host
                st_g, it_g = keep[j]
                par["instances_checked"] += n
                par["seeds_checked"].append(seeds[j])
                par["iter_mismatches"] += int((out["iters"][0] != it_g[:n]).sum())
                par["status_mismatches"] += int((out["ret"][0] != st_g[:n]).sum())
                par["max_abs_dU"] = max(par["max_abs_dU"], float(np.abs(out["U"] - U_host[:n].numpy().transpose(2, 1, 0)).max()))
        line["parity_full_batch"] = par
        if wl.key == "quadrotor_easy":
            del sol
            gpu_single = _single_solve_loop(lib, 300, problems)
            cpu_single = cpu_single_solve_p50(300)
            line["single_solve"] = {"gpu": gpu_single, "cpu_reference": cpu_single,
                                    "path": "set_x0 -> solve_mpc -> get_controls through the reference's own symbols, B=1"}

    if rank == 0:
        emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=None)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=65536)
    ap.add_argument("--workload", default="quadrotor_easy", choices=["quadrotor_easy", "quadrotor_hard", "cartpole", "rocket", "rollout"])
    ap.add_argument("--rollout-steps", type=int, default=200, help="MPC steps per instance of the rollout workload")
    ap.add_argument("--kernel", type=int, default=0, help="0 auto, 1 thread-per-instance, 2 sub-warp (shared memory), 3 sub-warp (TMEM + shared memory), 4 sub-warp with cones / half-spaces")
    ap.add_argument("--order", type=int, default=-1, help="queue order of the on-chip kernels: -1 library default (longest-first), 0 natural, 1 longest-first")
    ap.add_argument("--precision", type=int, default=64, choices=[64, 32],
                    help="arithmetic type of the solve = the reference's tinytype: 64 double (default, the headline), 32 its float build")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline / parity / single_solve legs")
    args = ap.parse_args()
    if args.steps is None:      # defaults that finish within minutes: the slow workloads take fewer steps
        args.steps = {"rollout": 8, "rocket": 10}.get(args.workload, 48) if args.impl == "b200" else \
                     {"rollout": 2, "rocket": 2}.get(args.workload, 8)
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    if world != args.gpus and world > 1:
        args.gpus = world
    run_gpu(args, rank, local_rank, world)


if __name__ == "__main__":
    main()
