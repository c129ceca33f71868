"""Multi-GPU plumbing on CPU: two gloo ranks shard a batch of robots, solve their shards (with the CPU
oracle standing in for the GPU solver -- the sharding logic does not care who solves), gather the
first-step forces and must reproduce the single-process result bit for bit."""
import os
import socket
import sys

import numpy as np
import pytest


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, B, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    import torch
    import torch.distributed as dist

    from oracle import bindings as ob
    from qrw_b200 import fleet, workloads as wl

    dist.init_process_group("gloo", rank=rank, world_size=world)
    w = wl.MpcWorkload(B, 3, gaits=("trot", "walk"), ticks=2)
    xr, fs = w.inputs(0)
    lo, hi = fleet.shard_bounds(B, rank, world)
    o = ob.OracleMPC(hi - lo)
    o.run(0, fleet.shard(xr, rank, world), fleet.shard(fs, rank, world), 2)
    f0 = torch.from_numpy(fleet.first_step_forces(o.get_latest_result()).copy())
    full = fleet.gather_rows(f0, B)
    q.put((rank, full.numpy()))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("B", [6, 7])  # even and ragged shards
def test_two_rank_shard_and_gather(oracle, B):
    import torch.multiprocessing as mp

    from qrw_b200 import fleet, workloads as wl

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, B, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=300) for _ in range(2))
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    w = wl.MpcWorkload(B, 3, gaits=("trot", "walk"), ticks=2)
    xr, fs = w.inputs(0)
    o = oracle.OracleMPC(B)
    o.run(0, xr, fs, 2)
    want = fleet.first_step_forces(o.get_latest_result())
    assert np.array_equal(got[0], want) and np.array_equal(got[1], want)


def test_shard_bounds_cover_without_overlap():
    from qrw_b200 import fleet

    for total in (0, 1, 7, 4096, 1048576):
        for world in (1, 2, 4, 8):
            b = [fleet.shard_bounds(total, r, world) for r in range(world)]
            assert b[0][0] == 0 and b[-1][1] == total
            assert all(b[i][1] == b[i + 1][0] for i in range(world - 1))
            sizes = [hi - lo for lo, hi in b]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        fleet.shard_bounds(4, 2, 2)
