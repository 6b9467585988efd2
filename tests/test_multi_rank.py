"""N > 1 path on CPU: batch sharding + the statistics all_gather with torch.distributed (gloo, world_size 2).
The per-problem work needs a GPU and is covered by tests/test_gpu_parity.py::test_batch_order_and_sharding_invariance;
here the host logic that bench.py runs under torchrun is exercised end to end with per-rank fake solver outputs."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from mpc_motion_planner_b200 import sharding


def test_shard_ranges_partition_the_batch():
    for n in (0, 1, 7, 1000, 100003):
        for world in (1, 2, 3, 4, 8):
            cuts = [sharding.shard_range(r, world, n) for r in range(world)]
            assert cuts[0][0] == 0 and cuts[-1][1] == n
            for a, b in zip(cuts, cuts[1:]):
                assert a[1] == b[0]
            sizes = [hi - lo for lo, hi in cuts]
            assert max(sizes) - min(sizes) <= 1 and sizes == sorted(sizes, reverse=True)


def _fake_outputs(n_global, seed=3):
    rng = np.random.default_rng(seed)
    return {"T": rng.uniform(0.5, 2.5, n_global), "status": rng.integers(1, 3, n_global), "qp": rng.integers(25, 1401, n_global),
            "sqp": rng.integers(1, 3, n_global)}


def _worker(rank, world, port, n_global, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        g = _fake_outputs(n_global)
        lo, hi = sharding.shard_range(rank, world, n_global)
        loc = sharding.local_stats(10.0 + rank, 0.5 + rank, 1e9 * (rank + 1), g["status"][lo:hi], g["qp"][lo:hi], g["sqp"][lo:hi], g["T"][lo:hi])
        allr = sharding.gather_stats(loc, dist)
        q.put((rank, allr))
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(120)
def test_stats_gather_world_size_2_matches_host_reduction():
    world, n_global = 2, 1001
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_global, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=100) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert np.array_equal(got[0], got[1])                 # every rank sees the same gathered matrix
    red = sharding.reduce_stats(got[0])
    g = _fake_outputs(n_global)
    assert red["n_ocp"] == n_global
    assert red["solved"] == int((g["status"] == 1).sum())
    assert red["qp_iters"] == float(g["qp"].sum()) and red["sqp_iters"] == float(g["sqp"].sum())
    assert abs(red["mean_T"] - g["T"].mean()) < 1e-12 and red["min_T"] == g["T"].min() and red["max_T"] == g["T"].max()
    assert red["ms_total_max"] == 11.0 and red["e2e_s_max"] == 1.5      # times: max over ranks


def test_single_rank_gather_is_identity():
    loc = sharding.local_stats(1.0, 2.0, 3.0, [1, 2, 1], [10, 20, 30], [2, 2, 1], [1.0, 2.0, 3.0])
    allr = sharding.gather_stats(loc)
    assert allr.shape == (1, len(sharding.STAT_FIELDS)) and np.array_equal(allr[0], loc)
