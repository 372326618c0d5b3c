"""CPU, world_size 2, gloo: the exchange plumbing of the gene-sharded path
(scran_b200/distributed.py) - effects from gene shards to comparison shards, MIN-combine of the
min-ranks, gather of the result columns - checked against the single-process answer."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import rlevel
from scran_b200.distributed import INT32_MAX, combine_minrank, exchange_effects, gather_tables, shard_ranges


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, ngenes, Q, K, result):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        rng = np.random.default_rng(5)
        eff = rng.normal(size=(3, Q, ngenes))
        eff[0, 1, ::7] = np.nan
        tables = rng.normal(size=(K, 19, ngenes))
        granges, per = shard_ranges(ngenes, world)
        g0, g1 = granges[rank]
        local = torch.zeros((3, Q, per), dtype=torch.float64)
        local[:, :, :g1 - g0] = torch.from_numpy(eff[:, :, g0:g1])
        full, (q0, q1) = exchange_effects(local, g1 - g0, ngenes, Q, dist, world, rank)
        assert np.array_equal(full[:, q0:q1, :].numpy(), eff[:, q0:q1, :], equal_nan=True)
        # every rank ranks its own columns over all genes; tables: column q belongs to table q % K
        mr = torch.full((K, 3, ngenes), INT32_MAX, dtype=torch.int32)
        for e in range(3):
            for q in range(q0, q1):
                r = rlevel.compute_min_rank(full[e, q].numpy()[:, None])
                mr[q % K, e] = torch.minimum(mr[q % K, e], torch.from_numpy(r.astype(np.int32)))
        combine_minrank(mr, dist, world)
        out_local = torch.zeros((K, 19, per), dtype=torch.float64)
        out_local[:, :, :g1 - g0] = torch.from_numpy(tables[:, :, g0:g1])
        out = gather_tables(out_local, ngenes, dist, world)
        assert np.array_equal(out.numpy(), tables)
        if rank == 0:
            want = np.full((K, 3, ngenes), INT32_MAX, dtype=np.int64)
            for e in range(3):
                for k in range(K):
                    cols = [q for q in range(Q) if q % K == k]
                    want[k, e] = rlevel.compute_min_rank(eff[e][cols].T)
            result.put(bool(np.array_equal(mr.numpy().astype(np.int64), want)))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("ngenes,Q,K", [(101, 7, 3), (64, 2, 2)])
def test_sharded_exchange_world2(ngenes, Q, K):
    ctxm = mp.get_context("spawn")
    result = ctxm.Queue()
    port = _free_port()
    procs = [ctxm.Process(target=_worker, args=(r, 2, port, ngenes, Q, K, result)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert result.get(timeout=10) is True


def test_shard_ranges_cover_everything():
    for n, w in ((33538, 8), (10, 4), (3, 8), (0, 2)):
        ranges, per = shard_ranges(n, w)
        assert len(ranges) == w and ranges[0][0] == 0 and ranges[-1][1] == n
        assert all(hi - lo <= per for lo, hi in ranges)
        assert sum(hi - lo for lo, hi in ranges) == n
