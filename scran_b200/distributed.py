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

On the GPU the exchange runs inside the library (csrc/dist.cu, NCCL on the context's stream;
ShardedScorer below only passes the row ranges). The three collective helpers in this file restate
that exchange plan on torch tensors; they work with the gloo backend on CPU tensors, which is how
tests/test_distributed_cpu.py checks the plan (who sends which block where, how the min-ranks
combine, how the rows are gathered) without a GPU.
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


def balanced_row_starts(weights, world):
    """Contiguous row ranges of equal total weight (scranb200_balance_rows): int64[world + 1]."""
    from ._lib import check, lib, ptr
    w = np.ascontiguousarray(weights, np.float64)
    starts = np.zeros(world + 1, np.int64)
    check(lib().scranb200_balance_rows(ptr(w), len(w), int(world), ptr(starts)))
    return starts


def init_comm(ctx, rank=0, world=1, dist=None):
    """Gives the context its NCCL communicator. The 128-byte unique id is made on rank 0 and travels
    over the process group the caller already has (any backend)."""
    from ._lib import check, lib, ptr
    if getattr(ctx, "_comm", None) == (world, rank):
        return  # this context already has the communicator of this world
    ident = np.zeros(128, np.uint8)
    if world > 1:
        import torch
        if rank == 0:
            check(lib().scranb200_comm_unique_id(ptr(ident)))
        t = torch.from_numpy(ident)
        if dist.get_backend() == "nccl":
            t = t.cuda()
            dist.broadcast(t, 0)
            ident = t.cpu().numpy()
        else:
            dist.broadcast(t, 0)
    check(lib().scranb200_ctx_comm_init(ctx.handle, int(world), int(rank), ptr(ident)))
    ctx._comm = (world, rank)


class ShardedScorer:
    """scoreMarkers over one gene shard per rank: scranb200_score_markers_sharded, i.e. all numerics AND
    the exchange (NCCL) inside libscranb200 on this rank's GPU and stream. row_starts: int64[world + 1]
    (default: equally many genes per rank)."""

    def __init__(self, ctx, plan, ngenes_total, rank=0, world=1, dist=None, row_starts=None, root=0):
        self.ctx, self.plan = ctx, plan
        self.ngenes, self.rank, self.world, self.root = int(ngenes_total), rank, world, root
        if row_starts is None:
            granges, _ = shard_ranges(self.ngenes, world)
            row_starts = [g[0] for g in granges] + [self.ngenes]
        self.row_starts = np.ascontiguousarray(row_starts, np.int64)
        self.g0, self.g1 = int(self.row_starts[rank]), int(self.row_starts[rank + 1])
        init_comm(ctx, rank, world, dist)

    def step(self, mat, out=None):
        """One full pass. out (optional): host array [K, 19, ngenes] filled on the ranks that receive the
        result (root, or all if root < 0). Returns the device address of the result table (0 elsewhere)."""
        import ctypes as C

        from ._lib import check, lib, ptr
        d_out = C.c_void_p()
        check(lib().scranb200_score_markers_sharded(self.ctx.handle, mat.handle, self.plan.handle, self.ngenes,
                                                    ptr(self.row_starts), int(self.root), ptr(out), C.byref(d_out)))
        self.last_timing = self.ctx.last_timing()
        return d_out.value or 0
