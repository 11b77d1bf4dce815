"""CPU, world_size 2, gloo: the N>1 host logic (shard ranges, batch-summary all-reduce, result
gather) of mim_solvers_b200/sharding.py.  Each rank runs the CPU oracle (test
infrastructure) on its shard as a stand-in for its GPU engine; the gathered result must be
the single-process result bit for bit (no data-path collective: SURVEY 8e)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from mim_solvers_b200.sharding import allreduce_max, allreduce_summary, gather_field, shard_range


def test_shard_range_covers_batch():
    for B in (1, 7, 8, 4096, 16384):
        for world in (1, 2, 3, 4, 8):
            r = [shard_range(B, k, world) for k in range(world)]
            assert r[0][0] == 0 and r[-1][1] == B
            assert all(r[i][1] == r[i + 1][0] for i in range(world - 1))
            assert max(b - a for a, b in r) == -(-B // world)
    with pytest.raises(ValueError):
        shard_range(8, 2, 2)


def _worker(rank, world, port, B, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from mim_solvers_b200.problems import make_random_batch
        from oracle import oracle
        full = make_random_batch(B, 6, 4, 2)
        lo, hi = shard_range(B, rank, world)
        o = oracle.Oracle(full.subset(np.arange(lo, hi)), nthreads=1)
        o.sqp_direction()
        kkt = o.scalar("KKT")
        local = dict(max_kkt=float(kkt.max()), sum_qp_iters=int(rank + 1), max_qp_iters=int(10 * (rank + 1)),
                     n_converged=int((kkt <= 1e-6).sum()), n_failed=int((o.status != 0).sum()))
        glob = allreduce_summary(local)
        t = allreduce_max(1.0 + rank)
        dx = gather_field(torch.from_numpy(o.field("dx").copy()), B, dst=0)
        if rank == 0:
            torch.save(dict(glob=glob, t=t, dx=dx), out)
    finally:
        dist.destroy_process_group()


def test_two_rank_summary_and_gather(tmp_path, oracle_lib):
    from mim_solvers_b200.problems import make_random_batch
    B, world = 5, 2           # ragged: shards of 3 and 2
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    out = str(tmp_path / "r0.pt")
    mp.spawn(_worker, args=(world, port, B, out), nprocs=world, join=True)
    res = torch.load(out, weights_only=False)
    o = oracle_lib.Oracle(make_random_batch(B, 6, 4, 2), nthreads=1)
    o.sqp_direction()
    assert np.array_equal(res["dx"].numpy(), o.field("dx"))
    assert res["glob"]["max_kkt"] == float(o.scalar("KKT").max())
    assert res["glob"]["sum_qp_iters"] == 3 and res["glob"]["max_qp_iters"] == 20
    assert res["glob"]["n_failed"] == 0
    assert res["t"] == 2.0
