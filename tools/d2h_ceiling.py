"""Host-link ceiling of the e2e path: plain pinned cudaMemcpyAsync device -> host of one solution-sized
buffer per rank (166 MB = quadrotor X + U at B = 65,536), all ranks at once, no solver involved.

    python tools/d2h_ceiling.py                                             # 1 rank
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 tools/d2h_ceiling.py

Per rank: best and median GB/s of `reps` copies, for three kinds of pinned host memory — torch's pin_memory()
(cudaHostAlloc default), tmpc_host_alloc (cudaHostAllocPortable) after first-touch on the rank's NUMA-local CPUs, and
the same WITHOUT the NUMA binding.  Rank 0 prints one JSON line with the aggregate (sum of per-rank bytes / max time)."""
import ctypes as C
import json
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tinympc_julia_b200 import _lib  # noqa: E402

BYTES = 8 * 65536 * (12 * 20 + 4 * 19)      # 165,675,008
REPS = 12


def bind_numa(local_rank):
    try:
        p = torch.cuda.get_device_properties(local_rank)
        bus = f"{p.pci_domain_id:04x}:{p.pci_bus_id:02x}:{p.pci_device_id:02x}.0"
        cpus = set()
        for part in open(f"/sys/bus/pci/devices/{bus}/local_cpulist").read().strip().split(","):
            lo, _, hi = part.partition("-")
            cpus.update(range(int(lo), int(hi or lo) + 1))
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
        return {"pci": bus, "numa_node": int(open(f"/sys/bus/pci/devices/{bus}/numa_node").read()), "cpus": len(cpus)}
    except Exception as e:
        return {"error": repr(e)}


def main():
    rank = int(os.environ.get("RANK", "0")); local = int(os.environ.get("LOCAL_RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    lib = _lib.load()
    src = torch.empty(BYTES, dtype=torch.uint8, device=dev).random_(0, 255)
    stream = torch.cuda.Stream(device=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def measure(dst_tensor):
        ts = []
        for r in range(REPS + 2):
            barrier()
            t0 = time.perf_counter()
            with torch.cuda.stream(stream):
                dst_tensor.copy_(src, non_blocking=True)
            stream.synchronize()
            t1 = time.perf_counter()
            barrier()
            if r >= 2:
                ts.append(t1 - t0)
        return ts

    out = {}
    # (a) no NUMA binding, cudaHostAllocPortable
    p0 = lib.tmpc_host_alloc(C.c_size_t(BYTES))
    a0 = torch.frombuffer((C.c_char * BYTES).from_address(p0), dtype=torch.uint8)
    a0.zero_()
    out["portable_unbound"] = measure(a0)
    numa = bind_numa(local)
    # (b) torch pinned after binding
    a1 = torch.empty(BYTES, dtype=torch.uint8).pin_memory()
    a1.zero_()
    out["torch_pinned_bound"] = measure(a1)
    # (c) portable, first touch on the bound CPUs
    p2 = lib.tmpc_host_alloc(C.c_size_t(BYTES))
    a2 = torch.frombuffer((C.c_char * BYTES).from_address(p2), dtype=torch.uint8)
    a2.zero_()
    out["portable_bound"] = measure(a2)
    res = {k: {"best_gbs": BYTES / min(v) / 1e9, "median_gbs": BYTES / float(np.median(v)) / 1e9, "max_s": max(v), "median_s": float(np.median(v))} for k, v in out.items()}
    allres = [None] * world
    if world > 1:
        dist.all_gather_object(allres, {"rank": rank, "numa": numa, "res": res})
    else:
        allres = [{"rank": rank, "numa": numa, "res": res}]
    if rank == 0:
        agg = {}
        for k in res:
            slowest = max(r["res"][k]["median_s"] for r in allres)
            agg[k] = {"aggregate_gbs_median": world * BYTES / slowest / 1e9,
                      "per_rank_median_gbs": [round(r["res"][k]["median_gbs"], 2) for r in allres]}
        print(json.dumps({"what": "pinned cudaMemcpyAsync D2H, one 166 MB copy per rank, all ranks at once", "ranks": world,
                          "bytes_per_rank": BYTES, "reps": REPS, "aggregate": agg, "numa_rank0": allres[0]["numa"],
                          "host_cpus": os.cpu_count()}))
    lib.tmpc_host_free(C.c_void_p(p0)); lib.tmpc_host_free(C.c_void_p(p2))
    if world > 1:
        dist.barrier(); dist.destroy_process_group()


if __name__ == "__main__":
    main()
