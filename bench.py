#!/usr/bin/env python
"""bench.py — throughput of the batched ADMM solve path (BASELINE.json metric) on N GPUs of one node.

    python bench.py [--gpus N] [--steps K] [--warmup W]            # this library (CUDA, sm_100a)
    python bench.py --impl reference [--gpus N] [--steps K] ...    # the reference's own CPU solver

Workload (config.workload): BASELINE.json configs[2] — batched quadrotor hover, nx=12 nu=4 N=20, input bounds
u in +-0.5 and state bounds x in +-5 shared by all instances, shared Kinf/Pinf cache, rho=5, tol 1e-3,
max_iter=100, B = 65,536 instances per GPU, x0 ~ U(+-0.1)^12 from numpy default_rng(3 + rank) ("easy",
SURVEY.md §8d; `--workload quadrotor_hard` draws U(+-0.5)^12), Xref = Uref = 0, COLD start for every step.

A "step" is one batched solve of the B instances of a rank (reset to the zero workspace + solve).
  value          QP solves/s, whole job, x0 resident in HBM, device time (CUDA events, max over ranks)
  admm_iters_per_s  ADMM iterations/s over the same region (sum of per-instance iteration counts)
  e2e            the same solves through the C ABI with HOST buffers: tmpc_batch_set_x0 (pinned host -> device),
                 tmpc_batch_solve, tmpc_batch_get_states + get_controls + get_status (device -> pinned host),
                 wall clock around the synchronous calls
  roofline       FP64 CUDA-core FMA roofline of the solve kernel (the binding limit, north_star): achieved =
                 sum(iterations) x F(nx,nu,N) flop / kernel time; peak = DFMA peak measured live on this GPU
                 (tmpc_measure_fma_peak) since MEASURED_PEAKS.json has no CUDA-core figure.  roofline_hbm gives
                 the once-per-solve HBM streams against MEASURED_PEAKS.json hbm_gbs.
  cpu_baseline   the compiled reference core (oracle/_ref, kind "reference") or the C restatement
                 (kind "port") on all host cores, same instances — N=1, rank 0 only
  single_solve   p50 latency of set_x0 -> solve_mpc -> get_controls at B=1 through the reference's own symbols,
                 GPU library next to the compiled reference
Only the `cpu_baseline` / `--impl reference` legs touch oracle/; the product path has no CPU fallback.
"""
from __future__ import annotations

import argparse
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
NX, NU, NH = 12, 4, 20
L2_FLUSH_BYTES = 256 << 20


def flops_per_iter(nx=NX, nu=NU, N=NH):
    from tinympc_julia_b200 import problems
    return problems.flops_per_iteration(nx, nu, N)


def make_inputs(B, rank, hard):
    from tinympc_julia_b200 import problems
    return problems.quadrotor_batch_x0(B, hard=hard, seed=3 + rank)


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return json.load(f), "measured"
    except Exception:
        return {"hbm_gbs": 6650.0}, "fallback"


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


# ------------------------------------------------------------------------------------------ CPU legs (oracle/)

def cpu_problem(hard):
    import cases
    c = cases.quadrotor_batch(8, hard=hard)
    return cases.to_oracle_problem(c)


def load_cpu_checker():
    import oracle
    ref = oracle.load_ref()
    if ref is not None:
        return ref
    return oracle.load_oracle()


def cpu_batch_run(lib, x0, hard, nthreads):
    import oracle
    out = lib.batch_run(cpu_problem(hard), x0, None, None, oracle.REF_ZERO, 1, nthreads)
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
    """--impl reference: the reference's own CPU implementation on all host cores, same config and metric."""
    if rank != 0:
        return
    lib = load_cpu_checker()
    cores = os.cpu_count() or 1
    hard = args.workload == "quadrotor_hard"
    B = args.batch
    x0 = make_inputs(B, 0, hard)
    for _ in range(args.warmup):
        cpu_batch_run(lib, x0[:, : max(cores * 64, 1024)], hard, cores)
    secs, iters = [], 0
    for _ in range(args.steps):
        s, it, _o = cpu_batch_run(lib, x0, hard, cores)
        secs.append(s)
        iters += it
    total = sum(secs)
    value = B * args.steps / total
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "admm_iters_per_s": iters / total,
        "config": {"workload": workload_name(args), "nx": NX, "nu": NU, "N": NH, "batch_per_step": B,
                   "note": "CPU solver: one step = the same B instances looped over all host threads; n_gpus is nominal"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": lib.kind,
                         "sample": f"all {B} instances of the workload per step, {args.steps} steps"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


def workload_name(args):
    tag = "hard x0~U(+-0.5)" if args.workload == "quadrotor_hard" else "easy x0~U(+-0.1)"
    prec = "" if getattr(args, "precision", 64) == 64 else ", float build (tinytype = float)"
    return f"quadrotor hover nx=12 nu=4 N=20, box bounds shared, shared cache, B={args.batch}/GPU, cold start, {tag}{prec}"


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
    from tinympc_julia_b200 import _lib, problems

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the solve path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    numa = None
    if world > 1:
        numa = bind_to_gpu_numa(torch, local_rank)      # before any pinned allocation
        dist.init_process_group("nccl", device_id=dev)
    lib = _lib.load()
    hard = args.workload == "quadrotor_hard"
    B, K, W = args.batch, args.steps, args.warmup
    s = problems.quadrotor()
    F = flops_per_iter()["F"]

    sol = t.TinyMPCBatchSolver()
    t.setup(sol, s.A, s.B, s.f, s.Q, s.R, s.rho, NX, NU, NH, batch=B, device=local_rank)
    if args.precision == 32:
        sol.set_precision(32)       # float build: setup redone in float, solves compute in float
    t.set_bound_constraints(sol, *s.bounds(NH))
    if args.kernel:
        sol.set_kernel(args.kernel)
    x0_host = make_inputs(B, rank, hard)                                       # (12, B) Fortran == (B, 12) C
    x0_pinned = torch.from_numpy(np.ascontiguousarray(x0_host.T)).pin_memory()  # instance-major, == Julia nx x B
    x0_dev = x0_pinned.to(dev)
    sol.bind_x0_device(x0_dev.data_ptr())
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

    def step():
        sol.reset()                 # cold start: logical reset, the kernel rebuilds all carried state
        sol.solve_async(sptr)

    for _ in range(max(W, 3)):
        flush.zero_()
        step()
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
        step()
        ev[k][1].record(stream)
    barrier()
    w1 = time.perf_counter()
    clocks = sampler.stop()
    step_ms = [a.elapsed_time(b) for a, b in ev]
    launches = sol.launch_count - l0
    st, it = sol.status()
    iters_step = int(it.sum())

    # multi-GPU plumbing (tinympc_julia_b200/sharding.py): device time = max over ranks; counts = sum over ranks; the one
    # collective of the path gathers the per-instance (status, iterations) over NCCL, outside the solve
    from tinympc_julia_b200 import sharding
    tot_ms_max = sharding.max_over_ranks(sum(step_ms), dev)
    all_iters_step, all_solved, all_launches = sharding.sum_over_ranks([iters_step, int((st == 0).sum()), launches], dev)
    st_all, it_all = sharding.gather_status(st, it, dev)
    assert st_all.size == B * world and int(it_all.sum()) == int(all_iters_step)
    total_s = tot_ms_max * 1e-3
    value = B * world * K / total_s
    iters_per_s = all_iters_step * K / total_s

    # ---- live FMA peak + roofline of the solve kernel (one launch per step; per-rank numbers, rank 0 reported)
    pk = C.c_double(0.0)
    rc = lib.tmpc_measure_fma_peak(local_rank, 1, 5, C.byref(pk))
    fma_peak = pk.value if rc == 0 else None
    pk32 = C.c_double(0.0)
    lib.tmpc_measure_fma_peak(local_rank, 0, 5, C.byref(pk32))
    my_s = sum(step_ms) * 1e-3
    kern_ms = sum(step_ms) / K
    achieved_tf = iters_step * K * F / my_s / 1e12
    peaks_json, peaks_kind = peaks()
    nxN, nuN = NX * NH, NU * (NH - 1)
    bytes_in = 8 * NX                                            # x0; bounds / references are shared (staged once per CTA)
    bytes_out = 8 * (nxN + nuN) + 8                              # solution + status/iterations
    bytes_state = 8 * (3 * nxN - NX + 3 * nuN + NX)              # carried state written for the next warm start: g, v, q, y, z, r, p_N
    hbm_bytes = B * (bytes_in + bytes_out + bytes_state)
    hbm_gbs = hbm_bytes / (kern_ms * 1e-3) / 1e9

    kname = {1: "thread", 2: "subwarp", 3: "tmem", 4: "tmem2"}.get(sol.kernel, str(sol.kernel))
    traffic = None
    try:
        rec = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get(kname + ("_f32" if args.precision == 32 else ""))
        if rec and rec.get("workload") == args.workload and rec.get("B") == B:
            traffic = rec["dram_read_bytes"] + rec["dram_write_bytes"]
    except Exception:
        pass
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": max(W, 3),
        "ms_per_step": tot_ms_max / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64" if args.precision == 64 else "f32", "data": "synthetic",
        "admm_iters_per_s": iters_per_s,
        "mean_iters_per_solve": all_iters_step / (B * world), "solved_fraction": all_solved / (B * world),
        "config": {"workload": workload_name(args), "nx": NX, "nu": NU, "N": NH, "batch_per_gpu": B,
                   "kernel": kname,
                   "l2": "flushed between timed steps (256 MiB write, outside the event pair)",
                   "timing": "CUDA events on the launch stream per step, summed, max over ranks",
                   "parallelism": f"instances sharded over {world} GPU(s), no collective inside the solve"},
        "gpu_launches": int(all_launches),
        "wall_ms_per_step_incl_flush": 1e3 * (w1 - w0) / K,
        "clocks": clocks,
        "roofline": {"bound": "fp64_fma" if args.precision == 64 else "fp32_fma", "achieved": achieved_tf,
                     "peak": fma_peak if args.precision == 64 else pk32.value, "unit": "TFLOP/s",
                     "frac": ((achieved_tf / fma_peak) if fma_peak else None) if args.precision == 64
                             else ((achieved_tf / pk32.value) if pk32.value else None), "traffic": traffic,
                     "traffic_note": "DRAM bytes per launch from the ncu --set full capture recorded in profiles/traffic.json (null when none matches this kernel/workload)",
                     "flop_per_iter": F, "kernel_ms": kern_ms,
                     "peak_source": "DFMA chains measured live on this GPU by tmpc_measure_fma_peak (MEASURED_PEAKS.json has no CUDA-core peak)",
                     "fp32_fma_peak_tflops": pk32.value},
        "roofline_hbm": {"bound": "hbm", "achieved": hbm_gbs, "peak": peaks_json.get("hbm_gbs"), "unit": "GB/s",
                         "frac": hbm_gbs / peaks_json.get("hbm_gbs", 6650.0), "peak_source": peaks_kind,
                         "algorithmic_bytes": hbm_bytes, "traffic": traffic,
                         "bytes_per_instance": {"in": bytes_in, "out": bytes_out, "carried_state_out": bytes_state}},
    }

    # ---- e2e through the C ABI with host buffers (every rank, same K steps)
    X_host = torch.empty((B, NH, NX), dtype=torch.float64).pin_memory()
    U_host = torch.empty((B, NH - 1, NU), dtype=torch.float64).pin_memory()
    st_host = torch.empty(B, dtype=torch.int32).pin_memory()
    it_host = torch.empty(B, dtype=torch.int32).pin_memory()
    dp = lambda tt: C.cast(tt.data_ptr(), C.POINTER(C.c_double))
    ip = lambda tt: C.cast(tt.data_ptr(), C.POINTER(C.c_int))
    h = sol._h
    lib.tmpc_batch_bind_x0_device(h, None)

    def e2e_step():
        # one call: x0 up, solve cut into instance chunks, each chunk's solution crossing PCIe while the next one solves
        assert lib.tmpc_batch_solve_host(h, dp(x0_pinned), 1, dp(X_host), dp(U_host), ip(st_host), ip(it_host), E2E_CHUNKS) in (0, 1)

    def e2e_step_calls():
        # the same through the five reference-style calls (no overlap): reported next to the pipelined number
        assert lib.tmpc_batch_set_x0(h, dp(x0_pinned), NX, B) == 0
        assert lib.tmpc_batch_reset(h) == 0
        assert lib.tmpc_batch_solve(h) in (0, 1)
        assert lib.tmpc_batch_get_states(h, dp(X_host)) == 0
        assert lib.tmpc_batch_get_controls(h, dp(U_host)) == 0
        assert lib.tmpc_batch_get_status(h, ip(st_host), ip(it_host)) == 0

    for _ in range(2):
        e2e_step_calls()
    barrier()
    c0 = time.perf_counter()
    for _ in range(max(3, K // 5)):
        e2e_step_calls()
    barrier()
    calls_s = sharding.max_over_ranks((time.perf_counter() - c0) / max(3, K // 5), dev)
    X_chk = X_host[:64].clone(); U_chk = U_host[:64].clone()
    for _ in range(3):
        e2e_step()
    barrier()
    e0 = time.perf_counter()
    for _ in range(K):
        e2e_step()
    barrier()
    e1 = time.perf_counter()
    e2e_s = sharding.max_over_ranks(e1 - e0, dev)
    assert int(it_host.sum()) == iters_step, "e2e path solved a different problem than the device-resident path"
    line["e2e"] = {"value": B * world * K / e2e_s, "unit": UNIT,
                   "h2d_bytes_per_step": int(x0_pinned.numel() * 8),
                   "d2h_bytes_per_step": int(X_host.numel() * 8 + U_host.numel() * 8 + 2 * 4 * B),
                   "ms_per_step": 1e3 * e2e_s / K,
                   "path": "tmpc_batch_solve_host on pinned host buffers (x0 up, chunked solve overlapped with the D2H of the previous chunk's solution)",
                   "unpipelined_ms_per_step": 1e3 * calls_s,
                   "unpipelined_path": "tmpc_batch_set_x0 / reset / solve / get_states / get_controls / get_status"}
    if numa is not None:
        line["e2e"]["host_binding_rank0"] = numa
    assert torch.equal(X_chk, X_host[:64]) and torch.equal(U_chk, U_host[:64]), "pipelined e2e path returned a different solution"

    # ---- CPU baseline + single-solve latency: N=1, rank 0 only
    if world == 1 and rank == 0 and not args.no_cpu:
        cpu = load_cpu_checker()
        if args.precision == 32:    # the float build is checked against (and timed beside) the reference compiled with the float typedef
            import oracle
            cpu32 = oracle.load_ref32()
            if cpu32 is not None:
                cpu = cpu32
        cores = os.cpu_count() or 1
        secs, cpu_iters, out = cpu_batch_run(cpu, x0_host, hard, cores)
        line["cpu_baseline"] = {"value": B / secs, "unit": UNIT, "cores": cores, "kind": cpu.kind,
                                "admm_iters_per_s": cpu_iters / secs, "seconds": secs,
                                "sample": f"all {B} instances of the workload, one pass, {cores} host threads"}
        # parity at full size against the same CPU run (status + iteration count per instance)
        line["parity_full_batch"] = {"iter_mismatches": int((out["iters"][0] != it).sum()),
                                     "status_mismatches": int((out["ret"][0] != st).sum()),
                                     "max_abs_dU": float(np.abs(out["U"] - U_host.numpy().transpose(2, 1, 0)).max())}
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
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=65536)
    ap.add_argument("--workload", default="quadrotor_easy", choices=["quadrotor_easy", "quadrotor_hard"])
    ap.add_argument("--kernel", type=int, default=0, help="0 auto, 1 thread-per-instance, 2 sub-warp (shared memory), 3 sub-warp (TMEM + shared memory), 4 the same with two instances per lane")
    ap.add_argument("--precision", type=int, default=64, choices=[64, 32],
                    help="arithmetic type of the solve = the reference's tinytype: 64 double (default, the headline), 32 its float build")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline / single_solve legs")
    args = ap.parse_args()
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
