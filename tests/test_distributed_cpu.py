"""CPU, world_size 2 and 3, gloo: the exchange step of the gene-sharded path. The block positions (which
comparison columns go to which rank, where a received block lands in the staging buffer) come from
scranb200_exchange_plan - the very function csrc/dist.cu drives its ncclSend / ncclRecv calls with - and the
row ranges from scranb200_balance_rows; the transport here is gloo on CPU tensors. Every rank must end up
with its comparison columns over ALL genes, the MIN-combination of the per-table ranks must equal the
single-process min-rank, and the gathered result rows must equal the single-process table."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import rlevel
from scran_b200._lib import check, lib, ptr
from scran_b200.distributed import INT32_MAX, balanced_row_starts, shard_ranges


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _plan(Q, world, rank, starts):
    q = np.zeros(world + 1, np.int64)
    so, sl, ro, rl = (np.zeros(world, np.int64) for _ in range(4))
    check(lib().scranb200_exchange_plan(Q, world, rank, ptr(starts), ptr(q), ptr(so), ptr(sl), ptr(ro), ptr(rl)))
    return q, so, sl, ro, rl


def _worker(rank, world, port, ngenes, Q, K, result):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        rng = np.random.default_rng(5)
        eff = rng.normal(size=(3, Q, ngenes))
        eff[0, 1, ::7] = np.nan
        tables = rng.normal(size=(K, 19, ngenes))
        weights = rng.random(ngenes) + 0.1
        weights[::9] *= 12.0                       # a few heavy genes: the ranges come out uneven
        starts = balanced_row_starts(weights, world)
        assert starts[0] == 0 and starts[-1] == ngenes and (np.diff(starts) >= 0).all()
        g0, g1 = int(starts[rank]), int(starts[rank + 1])
        n = g1 - g0
        local = np.ascontiguousarray(eff[:, :, g0:g1])          # [3][Q][n]: this rank's gene shard
        q, so, sl, ro, rl = _plan(Q, world, rank, starts)
        nq = int(q[rank + 1] - q[rank])
        # "send": every rank publishes its blocks; gloo has no all-to-all, so all blocks are gathered and each rank
        # keeps the ones addressed to it
        blocks = [[local[e].ravel()[so[s]:so[s] + sl[s]].copy() for e in range(3)] for s in range(world)]
        everyone = [None] * world
        dist.all_gather_object(everyone, blocks)
        staging = np.full(3 * nq * ngenes, np.nan)
        for s in range(world):
            for e in range(3):
                blk = everyone[s][rank][e]
                assert len(blk) == rl[s]
                staging[ro[s] + e * rl[s]: ro[s] + (e + 1) * rl[s]] = blk
        cols = np.full((3, nq, ngenes), np.nan)                  # what dist_unpack builds
        for s in range(world):
            ns = int(starts[s + 1] - starts[s])
            cols[:, :, starts[s]:starts[s + 1]] = staging[ro[s]: ro[s] + 3 * nq * ns].reshape(3, nq, ns)
        assert np.array_equal(cols, eff[:, q[rank]:q[rank + 1], :], equal_nan=True)
        # every rank ranks its own columns over all genes; tables: column q belongs to table q % K
        mr = torch.full((K, 3, ngenes), INT32_MAX, dtype=torch.int32)
        for e in range(3):
            for j in range(nq):
                qq = int(q[rank]) + j
                r = rlevel.compute_min_rank(cols[e, j][:, None])
                mr[qq % K, e] = torch.minimum(mr[qq % K, e], torch.from_numpy(r.astype(np.int32)))
        dist.all_reduce(mr, op=dist.ReduceOp.MIN)
        parts = [None] * world
        dist.all_gather_object(parts, np.ascontiguousarray(tables[:, :, g0:g1]))
        out = np.concatenate(parts, axis=2)                       # rows land at [starts[s], starts[s+1])
        assert np.array_equal(out, tables)
        if rank == 0:
            want = np.full((K, 3, ngenes), INT32_MAX, dtype=np.int64)
            for e in range(3):
                for k in range(K):
                    cs = [c for c in range(Q) if c % K == k]
                    if cs:
                        want[k, e] = rlevel.compute_min_rank(eff[e][cs].T)
            result.put(bool(np.array_equal(mr.numpy().astype(np.int64), want)))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,ngenes,Q,K", [(2, 101, 7, 3), (2, 64, 2, 2), (3, 90, 10, 4), (3, 40, 2, 2)])
def test_sharded_exchange_plan_over_gloo(world, ngenes, Q, K):
    ctxm = mp.get_context("spawn")
    result = ctxm.Queue()
    port = _free_port()
    procs = [ctxm.Process(target=_worker, args=(r, world, port, ngenes, Q, K, result)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(180)
        assert p.exitcode == 0
    assert result.get(timeout=10) is True


def test_row_ranges_cover_everything():
    for n, w in ((33538, 8), (10, 4), (3, 8), (0, 2)):
        ranges, per = shard_ranges(n, w)
        assert len(ranges) == w and ranges[0][0] == 0 and ranges[-1][1] == n
        assert sum(hi - lo for lo, hi in ranges) == n
        starts = balanced_row_starts(np.ones(n), w)
        assert starts[0] == 0 and starts[-1] == n and (np.diff(starts) >= 0).all()
    heavy = np.ones(1000)
    heavy[:10] = 100.0
    starts = balanced_row_starts(heavy, 4)
    sums = [heavy[starts[k]:starts[k + 1]].sum() for k in range(4)]
    assert max(sums) / min(sums) < 1.25
