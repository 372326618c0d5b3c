"""Gene-sharded scoreMarkers across the GPUs of one box: one process per GPU (torchrun),
torch.distributed for the plumbing.

The reference's only parallelism is the same decomposition: beachmat::rowBlockApply cuts the
genes into row blocks, one worker per block, results rbind-ed (R/scoreMarkers.R:234-248), and the
min-rank is computed after the blocks are joined because it ranks across ALL genes (:251-256).
Here rank r owns genes [g0_r, g1_r) for all cells. Everything up to the block-averaged effects and
the mean/min/median/max summaries is local. The one exchange step: the effect matrix moves from
gene shards to comparison shards (all-to-all over NVLink), every rank ranks its own comparison
columns over all genes, the per-table minima are combined with an all-reduce(MIN), and the
result columns are gathered on rank 0.

The three collective helpers below work on CPU tensors with the gloo backend too, which is how
tests/test_distributed_cpu.py checks the plumbing without a GPU.
"""
import numpy as np

INT32_MAX = np.iinfo(np.int32).max


def shard_ranges(n, world):
    """Contiguous, equally sized (padded) shards: [(lo, hi)] per rank and the padded size."""
    per = (n + world - 1) // world
    return [(min(n, r * per), min(n, (r + 1) * per)) for r in range(world)], per


def exchange_effects(eff_local, nrows_local, ngenes, Q, dist, world, rank):
    """eff_local: [3, Q, per] (this rank's genes in the first nrows_local columns of the last axis).
    Returns (eff_cols [3, Q, ngenes] with this rank's comparison columns filled, (q0, q1))."""
    import torch
    qranges, qper = shard_ranges(Q, world)
    per = eff_local.shape[2]
    send = []
    for s in range(world):
        q0, q1 = qranges[s]
        buf = torch.zeros((3, qper, per), dtype=eff_local.dtype, device=eff_local.device)
        if q1 > q0:
            buf[:, :q1 - q0, :] = eff_local[:, q0:q1, :]
        send.append(buf)
    recv = [torch.empty_like(send[0]) for _ in range(world)]
    if world == 1:
        recv = send
    elif dist.get_backend() == "nccl":
        dist.all_to_all(recv, send)
    else:  # gloo has no all-to-all: gather every rank's send list and keep the slices addressed to us
        mine = torch.stack(send, 0)
        everyone = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(everyone, mine)
        recv = [everyone[r][rank] for r in range(world)]
    q0, q1 = qranges[rank]
    granges, _ = shard_ranges(ngenes, world)
    full = torch.zeros((3, Q, ngenes), dtype=eff_local.dtype, device=eff_local.device)
    for r in range(world):
        g0, g1 = granges[r]
        if g1 > g0 and q1 > q0:
            full[:, q0:q1, g0:g1] = recv[r][:, :q1 - q0, :g1 - g0]
    return full, (q0, q1)


def combine_minrank(minrank, dist, world):
    """minrank: int32 [K, 3, ngenes], INT32_MAX where this rank saw no column. In-place MIN."""
    if world > 1:
        dist.all_reduce(minrank, op=dist.ReduceOp.MIN)
    return minrank


def gather_tables(out_local, ngenes, dist, world):
    """out_local: [K, 19, per] -> [K, 19, ngenes] on every rank (rows of rank r at [g0_r, g1_r))."""
    import torch
    if world == 1:
        return out_local[:, :, :ngenes].contiguous()
    parts = [torch.empty_like(out_local) for _ in range(world)]
    dist.all_gather(parts, out_local.contiguous())
    granges, _ = shard_ranges(ngenes, world)
    full = torch.empty((out_local.shape[0], out_local.shape[1], ngenes), dtype=out_local.dtype, device=out_local.device)
    for r in range(world):
        g0, g1 = granges[r]
        full[:, :, g0:g1] = parts[r][:, :, :g1 - g0]
    return full


class ShardedScorer:
    """scoreMarkers over a gene shard per rank, all numerics in libscranb200 on this rank's GPU."""

    def __init__(self, ctx, plan, ngenes_total, rank=0, world=1, dist=None):
        import torch
        self.torch = torch
        self.ctx, self.plan = ctx, plan
        self.ngenes, self.rank, self.world, self.dist = int(ngenes_total), rank, world, dist
        self.granges, self.per = shard_ranges(self.ngenes, world)
        self.g0, self.g1 = self.granges[rank]
        dev = torch.device("cuda", ctx.device)
        K, Q = max(1, plan.K), max(1, plan.Q)
        self.eff_local = torch.zeros((3, Q, self.per), dtype=torch.float64, device=dev)
        self.out_local = torch.zeros((K, 19, self.per), dtype=torch.float64, device=dev)
        self.minrank = torch.empty((K, 3, self.ngenes), dtype=torch.int32, device=dev)
        torch.cuda.synchronize(dev)

    def step(self, mat):
        """One full pass; returns the [K, 19, ngenes] result tensor (complete on every rank)."""
        from ._lib import check, lib, ptr
        torch, plan, ctx = self.torch, self.plan, self.ctx
        L = lib()
        check(L.scranb200_score_effects_dev(ctx.handle, mat.handle, plan.handle, self.per, 0, ptr(self.eff_local),
                                            ptr(self.out_local)))
        self.last_timing = ctx.last_timing()
        if self.world > 1:
            full, (q0, q1) = exchange_effects(self.eff_local, self.g1 - self.g0, self.ngenes, plan.Q, self.dist,
                                              self.world, self.rank)
            torch.cuda.current_stream().synchronize()
            eff, ld = full, self.ngenes
        else:
            eff, ld, q0, q1 = self.eff_local, self.per, 0, plan.Q
        check(L.scranb200_fill_i32(ctx.handle, ptr(self.minrank), self.minrank.numel(), INT32_MAX))
        check(L.scranb200_score_minrank_dev(ctx.handle, plan.handle, self.ngenes, ld, ptr(eff), q0, q1,
                                            ptr(self.minrank)))
        self.last_timing["minrank"] = ctx.last_timing()["minrank"]
        combine_minrank(self.minrank, self.dist, self.world)
        out = gather_tables(self.out_local, self.ngenes, self.dist, self.world)
        torch.cuda.current_stream().synchronize()
        check(L.scranb200_score_finish_dev(ctx.handle, plan.handle, self.ngenes, self.ngenes, ptr(self.minrank),
                                           ptr(out)))
        return out
