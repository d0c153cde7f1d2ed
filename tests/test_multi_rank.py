"""N > 1 path on CPU: two gloo ranks shard a batch, each solves its shard (with the oracle standing in for the GPU
kernel — this is a test), and the one collective of the path gathers (status, iterations).  The gathered result must
equal the single-process run: sharding is by contiguous blocks and nothing else is exchanged."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import cases
from tinympc_julia_b200 import sharding


def test_shard_range_partitions_exactly():
    for total in (0, 1, 7, 64, 65536, 65537):
        for world in (1, 2, 3, 8):
            blocks = [sharding.shard_range(total, r, world) for r in range(world)]
            assert blocks[0][0] == 0 and blocks[-1][1] == total
            assert all(blocks[r][1] == blocks[r + 1][0] for r in range(world - 1))
            sizes = [hi - lo for lo, hi in blocks]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        sharding.shard_range(10, 2, 2)


def test_gather_without_process_group_is_identity():
    st, it = np.array([0, 1, 0], dtype=np.int32), np.array([5, 100, 7], dtype=np.int32)
    a, b = sharding.gather_status(st, it)
    assert np.array_equal(a, st) and np.array_equal(b, it)
    assert sharding.max_over_ranks(1.5) == 1.5


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, B, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import oracle
        orc = oracle.load_oracle()
        c = cases.quadrotor_batch(B, hard=True, max_iter=30)
        lo, hi = sharding.shard_range(B, rank, world)
        shard = dict(c)
        shard["x0"] = np.asfortranarray(c["x0"][:, lo:hi])
        out = cases.run_checker(shard, orc)
        st, it = sharding.gather_status(out["ret"][0], out["iters"][0])
        tmax = sharding.max_over_ranks(float(rank + 1))
        tot = sharding.sum_over_ranks([hi - lo, int(out["iters"].sum())])
        np.savez(os.path.join(out_dir, f"rank{rank}.npz"), st=st, it=it, tmax=tmax, tot=np.array(tot))
    finally:
        dist.destroy_process_group()


def test_two_ranks_gather_equals_single_process(tmp_path, orc):
    B, world = 37, 2           # odd on purpose: shards of 19 and 18
    mp.spawn(_worker, args=(world, _free_port(), B, str(tmp_path)), nprocs=world, join=True)
    want = cases.run_checker(cases.quadrotor_batch(B, hard=True, max_iter=30), orc)
    for rank in range(world):
        z = np.load(tmp_path / f"rank{rank}.npz")
        assert np.array_equal(z["st"], want["ret"][0])
        assert np.array_equal(z["it"], want["iters"][0])
        assert float(z["tmax"]) == float(world)
        assert z["tot"].tolist() == [float(B), float(want["iters"].sum())]
