#!/usr/bin/env python
"""bench.py - scoreMarkers throughput on B200 (BASELINE.json metric: gene·pair scores/sec).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload config3] [--impl ours|reference]

One "step" = one full scoreMarkers pass (transpose -> blocked stats -> sort + U count -> effects
-> summaries -> min-rank, plus the exchange when N > 1) over one synthetic matrix of the named
config. For N > 1 launch under torchrun (one rank per GPU); the genes of the SAME matrix are
sharded over the ranks (strong scaling), which is BASELINE.json's "gene-sharded over 8xB200".

Prints ONE JSON line on rank 0. `value` = gene·pairs / s with inputs resident in HBM; `e2e` = the
same through the host-buffer C-ABI call (H2D of the CSC arrays and D2H of the result tables
inside the timed region); `roofline` = the dominant kernel phase against the measured HBM peak;
`cpu_baseline` = the CPU oracle (the reference's own object code when oracle/_ref is present)
timed on a bounded gene sample of the same workload on this box's host cores.

--impl reference times that CPU path alone with all host threads (R + BiocParallel cannot run:
R is not installed on either machine).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: genes, cells, groups, batches, density, seed (1000 + config id)
    "config1": dict(ngenes=10000, ncells=3000, G=10, B=1, density=0.63, seed=1001),
    "config2": dict(ngenes=33538, ncells=100000, G=20, B=4, density=0.07, seed=1002),
    "config3": dict(ngenes=33538, ncells=1300000, G=50, B=8, density=0.07, seed=1003),
    # config 4: pairwiseTTests(lfc = 0.5) + modelGeneVar's blocked statistics on the 1.3M-cell matrix of config 3
    "config4": dict(ngenes=33538, ncells=1300000, G=50, B=8, density=0.07, seed=1003),
    "config5": dict(ngenes=20000, ncells=500000, G=200, B=1, density=0.07, seed=1005),
    "tiny": dict(ngenes=2000, ncells=20000, G=8, B=2, density=0.07, seed=1999),
    # one eighth of config 3's genes over all its cells: the per-GPU share of the 8-GPU run, for profiling
    "config3s": dict(ngenes=4192, ncells=1300000, G=50, B=8, density=0.07, seed=1003),
}
PHASES = ["transpose", "stats", "sort", "count", "effects", "summaries", "minrank"]


def labels_for(w):
    """Uniform group and batch labels with every combination non-empty."""
    rng = np.random.default_rng(w["seed"])
    groups = rng.integers(0, w["G"], w["ncells"]).astype(np.int32)
    block = rng.integers(0, w["B"], w["ncells"]).astype(np.int32)
    k = 0
    for g in range(w["G"]):
        for b in range(w["B"]):
            groups[k], block[k] = g, b
            k += 1
    return groups, (block if w["B"] > 1 else None)


def measured_peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        return float(json.load(open(path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (profiling recipe's line)."""

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + q,
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc:
            self.proc.terminate()
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx.append(float(r[1]))
            except (ValueError, IndexError):
                continue
            for name, v in zip(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"], r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        busy = [s for s in sm if s > 0]
        return {"sm_mhz": float(np.median(busy)) if busy else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def algorithmic_bytes(w, nnz, P, K):
    """what one pass must move at least, in bytes."""
    E = 3 * w["ngenes"] * P * 8
    S = w["ngenes"] * K * 19 * 8
    return nnz * 12 + (w["ncells"] + 1) * 8 + w["ncells"] * 4 + 2 * E + S


def phase_bytes(w, nnz, plan):
    """Algorithmic bytes per kernel phase."""
    ng, C, PB, Q, K = w["ngenes"], plan.C, plan.npairs, plan.Q, plan.K
    return {
        "transpose": nnz * (4 + 12) + nnz * 8,          # count pass reads row indices, scatter reads 12 B, writes 8 B
        "stats": nnz * 8 + 3 * ng * C * 8,
        "sort": nnz * 8 + nnz * 2,                      # values in, 2-byte label stream out
        "count": nnz * 2 + ng * PB * 8,                 # label stream in, U per block-level pair out
        "effects": ng * PB * 8 + 3 * ng * C * 8 + 3 * ng * Q * 8,
        "summaries": 3 * ng * Q * 8 + ng * K * 16 * 8,
        "minrank": 3 * ng * Q * 8 + ng * K * 3 * 4 * 2,
    }


def cpu_sample(w, groups, block, ctx, nsample, threads):
    """Times the CPU oracle on `nsample` genes x all cells of the same synthetic workload."""
    from concurrent.futures import ThreadPoolExecutor

    import scipy.sparse as sp

    import scran_b200
    from oracle import native, rlevel
    lib = native.load("best")
    g0 = min(1000, max(0, w["ngenes"] - nsample))
    m = scran_b200.DeviceMatrix.synthetic(ctx, w["ngenes"], g0, g0 + nsample, w["ncells"], groups, w["G"], w["density"],
                                          w["seed"])
    p, i, x = m.download()
    m.free()
    full = sp.csc_matrix((x, i, p), shape=(nsample, w["ncells"]))
    st = rlevel.ScoreSetup(groups.tolist(), None if block is None else block.tolist(), None, 0.0)
    parts = np.array_split(np.arange(nsample), threads)
    rows = full.tocsr()
    pieces = [rows[idx].tocsc() for idx in parts if len(idx)]

    def work(piece):  # what one rowBlockApply worker does (R/scoreMarkers.R:351-366)
        lib.blocked_stats(piece, st.comb_id.astype(np.int32), len(st.combos))
        lib.overlap_paired(piece, st.left + 1, st.right + 1, st.involved, 0.0)

    t0 = time.perf_counter()
    with ThreadPoolExecutor(max_workers=threads) as ex:
        list(ex.map(work, pieces))
    dt = time.perf_counter() - t0
    return dt, lib.kind, nsample


def run_reference(args, w):
    """--impl reference: the reference's CPU path on the host cores, bounded sample per step."""
    import scran_b200
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    groups, block = labels_for(w)
    ctx = scran_b200.Context(int(os.environ.get("LOCAL_RANK", "0")))
    threads = os.cpu_count() or 1
    P = w["G"] * (w["G"] - 1) // 2
    per_gene = {"config3": 0.09, "config5": 0.12, "config2": 0.006}.get(args.workload, 0.001)
    nsample = int(max(threads, min(w["ngenes"], threads * max(1, int(6.0 / per_gene)))))
    nsample = min(nsample, 4096)
    times = []
    for s in range(args.warmup + args.steps):
        dt, kind, n = cpu_sample(w, groups, block, ctx, nsample, threads)
        if s >= args.warmup:
            times.append(dt)
    t = float(np.mean(times))
    value = nsample * P / t
    line = {
        "impl": "reference", "metric": "scoreMarkers gene·pair scores/sec", "value": value, "unit": "gene·pairs/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": t * 1e3,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": args.workload, **{k: w[k] for k in ("ngenes", "ncells", "G", "B", "density")}},
        "cpu_baseline": {"value": value, "unit": "gene·pairs/s", "cores": threads,
                         "kind": "reference" if kind == "ref" else "port",
                         "sample": "%d genes x all %d cells per step, native kernels only (blocked stats + U for every "
                                   "block-level pair), %d threads" % (nsample, w["ncells"], threads)},
        "e2e": {"value": value, "unit": "gene·pairs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "R + BiocParallel unavailable on this box; this is the reference's own C++ compiled unmodified "
                "(oracle/_ref) when kind=reference, else the C port",
    }
    print(json.dumps(line))


def run_config4(args, w):
    """BASELINE.json config 4: compute_blocked_stats_none on the 1.3M-cell matrix with the 400 block-major ids of
    pairwiseTTests (R/pairwiseTTests.R:242) and with modelGeneVar's 8 block ids (R/modelGeneVar.R:148-157), plus the
    t-test layer with lfc = 0.5. Metric: mean / variance cells per second = ngenes x nblocks /
    device time of the blocked statistics (transpose + stats), with their achieved GB/s against
    nnz 12 + ncells 4 + 2 ngenes nblocks 8 bytes. N > 1: genes sharded, no exchange (gather by rows only)."""
    import ctypes as C

    import torch

    import scran_b200
    from scran_b200._lib import check, lib, ptr
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    L = lib()
    groups, block = labels_for(w)
    G, B, ng, nc = w["G"], w["B"], w["ngenes"], w["ncells"]
    ids400 = np.ascontiguousarray(groups + block * G, np.int32)
    ids8 = np.ascontiguousarray(block, np.int32)
    starts = shard_starts(w, world)
    g0, g1 = int(starts[rank]), int(starts[rank + 1])
    nrow = g1 - g0
    if args.impl == "reference":
        if rank == 0:
            run_reference_config4(args, w, groups, ids400, ids8)
        return
    ctx = scran_b200.Context(local)
    mat = scran_b200.DeviceMatrix.synthetic(ctx, ng, g0, g1, nc, groups, G, w["density"], w["seed"])
    nnz_t = torch.tensor([mat.nnz], dtype=torch.int64, device="cuda")
    if world > 1:
        dist.all_reduce(nnz_t)
    nnz = int(nnz_t.item())
    out400 = [torch.empty((G * B, max(1, nrow)), dtype=torch.float64).pin_memory() for _ in range(2)]
    out8 = [torch.empty((B, max(1, nrow)), dtype=torch.float64).pin_memory() for _ in range(2)]
    npair = G * (G - 1)
    outt = [torch.empty((npair, max(1, nrow)), dtype=torch.float64).pin_memory() for _ in range(3)]
    gcodes = np.ascontiguousarray(groups, np.int32)

    def step():
        ms = {}
        check(L.scranb200_blocked_stats(ctx.handle, mat.handle, ptr(ids400), G * B, ptr(out400[0]), ptr(out400[1]), None))
        lt = ctx.last_timing()
        ms["stats400"], ms["k1_400"], ms["k2_400"] = lt["total"], lt["transpose"], lt["stats"]
        check(L.scranb200_blocked_stats(ctx.handle, mat.handle, ptr(ids8), B, ptr(out8[0]), ptr(out8[1]), None))
        lt = ctx.last_timing()
        ms["stats8"], ms["k1_8"], ms["k2_8"] = lt["total"], lt["transpose"], lt["stats"]
        t0 = time.perf_counter()
        check(L.scranb200_pairwise_t_tests(ctx.handle, mat.handle, ptr(gcodes), G, ptr(ids8), B, 0.5, 0, 0, ptr(outt[0]),
                                           ptr(outt[1]), ptr(outt[2])))
        ms["ttests_call"] = (time.perf_counter() - t0) * 1e3
        return ms

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    for _ in range(max(args.warmup, 1)):
        step()
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    launches0 = ctx.launches
    acc = {}
    t0 = time.perf_counter()
    for _ in range(args.steps):
        for k, v in step().items():
            acc[k] = acc.get(k, 0.0) + v / args.steps
    barrier()
    wall = (time.perf_counter() - t0) / args.steps
    clocks = sampler.stop() if rank == 0 else None
    launches = ctx.launches - launches0
    dev = torch.tensor([acc["stats400"] + acc["stats8"], acc["stats400"], acc["stats8"], acc["ttests_call"], wall * 1e3],
                       dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(dev, op=dist.ReduceOp.MAX)
    ms_stats, ms400, ms8, ms_tt, ms_wall = [float(v) for v in dev.tolist()]
    # end to end: upload from pinned host memory + the two blocked-statistics calls
    p_h, i_h, x_h = mat.download()
    mat.free()
    pin = [torch.from_numpy(a).pin_memory() for a in (p_h, i_h, x_h)]
    del p_h, i_h, x_h

    def e2e_step():
        h = C.c_void_p()
        check(L.scranb200_upload_csc(ctx.handle, nrow, nc, ptr(pin[0]), ptr(pin[1]), ptr(pin[2]), C.byref(h)))
        try:
            check(L.scranb200_blocked_stats(ctx.handle, h, ptr(ids400), G * B, ptr(out400[0]), ptr(out400[1]), None))
            check(L.scranb200_blocked_stats(ctx.handle, h, ptr(ids8), B, ptr(out8[0]), ptr(out8[1]), None))
        finally:
            L.scranb200_mat_free(h)
    e2e_step()
    barrier()
    t0 = time.perf_counter()
    e2e_step()
    barrier()
    t_e = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t_e, op=dist.ReduceOp.MAX)
    if rank != 0:
        dist.destroy_process_group()
        return
    peak, peak_src = measured_peak()
    cells = ng * (G * B + B)
    bytes400 = nnz * 12 + nc * 4 + 2 * ng * G * B * 8
    bytes8 = nnz * 12 + nc * 4 + 2 * ng * B * 8
    gbs = (bytes400 + bytes8) / world / (ms_stats * 1e-3) / 1e9
    cpu = None
    if not args.no_cpu and world == 1:
        cpu = cpu_sample_config4(w, groups, ids400, ids8, ctx, args.cpu_genes or 512, 1)
    line = {
        "metric": "compute_blocked_stats mean/variance cells/sec (1.3M cells; 400 + 8 blocks)", "value": cells / (ms_stats * 1e-3),
        "unit": "gene·block cells/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_stats,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "config4", "ngenes": ng, "ncells": nc, "G": G, "B": B, "density": w["density"], "nnz": nnz,
                   "blocks": [G * B, B], "t_tests": "lfc=0.5, direction any, %d ordered pairs" % npair,
                   "sharding": "genes x%d, ranges of equal expected nnz" % world,
                   "l2": "inputs (%.1f GB) larger than L2" % (nnz * 12 / 1e9)},
        "e2e": {"value": cells / float(t_e.item()), "unit": "gene·block cells/s", "h2d_bytes_per_step": int(nnz * 12 + (nc + 1) * 8),
                "d2h_bytes_per_step": int(2 * ng * (G * B + B) * 8), "ms_per_step": float(t_e.item()) * 1e3, "host_memory": "pinned",
                "call": "scranb200_upload_csc + 2 x scranb200_blocked_stats (host buffers)"},
        "gpu_launches": int(launches), "clocks": clocks,
        "roofline": {"bound": "hbm", "kernel": "transpose + stats (both calls)", "achieved": round(gbs, 1), "peak": peak, "unit": "GB/s",
                     "frac": round(gbs / peak, 4), "traffic": None, "peak_source": peak_src},
        "kernels": {"stats_400_blocks": {"ms": round(ms400, 3), "gbs": round(bytes400 / world / (ms400 * 1e-3) / 1e9, 1),
                                         "transpose_ms": round(acc["k1_400"], 3), "stats_ms": round(acc["k2_400"], 3)},
                    "stats_8_blocks": {"ms": round(ms8, 3), "gbs": round(bytes8 / world / (ms8 * 1e-3) / 1e9, 1),
                                       "transpose_ms": round(acc["k1_8"], 3), "stats_ms": round(acc["k2_8"], 3)},
                    "pairwise_t_tests_call": {"ms": round(ms_tt, 3),
                                              "note": "host-buffer call: blocked statistics + t-test layer + log BH on the device, "
                                                      "three [%d x ngenes] result arrays copied to the host" % npair}},
        "wall_ms_per_step_all_three_calls": round(ms_wall, 2), "cpu_baseline": cpu,
    }
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def cpu_sample_config4(w, groups, ids400, ids8, ctx, nsample, threads):
    """The reference's compute_blocked_stats_none (object code when oracle/_ref is there) on `nsample` genes x all cells."""
    from concurrent.futures import ThreadPoolExecutor

    import scipy.sparse as sp

    import scran_b200
    from oracle import native
    lib = native.load("best")
    m = scran_b200.DeviceMatrix.synthetic(ctx, w["ngenes"], 1000, 1000 + nsample, w["ncells"], groups, w["G"], w["density"], w["seed"])
    p, i, x = m.download()
    m.free()
    rows = sp.csc_matrix((x, i, p), shape=(nsample, w["ncells"])).tocsr()
    pieces = [rows[idx].tocsc() for idx in np.array_split(np.arange(nsample), threads) if len(idx)]

    def work(piece):
        lib.blocked_stats(piece, ids400, w["G"] * w["B"])
        lib.blocked_stats(piece, ids8, w["B"])
    t0 = time.perf_counter()
    with ThreadPoolExecutor(max_workers=threads) as ex:
        list(ex.map(work, pieces))
    dt = time.perf_counter() - t0
    cells = nsample * (w["G"] * w["B"] + w["B"])
    return {"value": cells / dt, "unit": "gene·block cells/s", "cores": threads, "kind": "reference" if lib.kind == "ref" else "port",
            "sample": "%d genes x all %d cells, compute_blocked_stats_none with 400 and with 8 blocks, %.1f s" % (nsample, w["ncells"], dt)}


def run_reference_config4(args, w, groups, ids400, ids8):
    import scran_b200
    ctx = scran_b200.Context(int(os.environ.get("LOCAL_RANK", "0")))
    threads = os.cpu_count() or 1
    vals = []
    for s in range(args.warmup + args.steps):
        r = cpu_sample_config4(w, groups, ids400, ids8, ctx, 256 * threads, threads)
        if s >= args.warmup:
            vals.append(r)
    value = float(np.mean([r["value"] for r in vals]))
    print(json.dumps({
        "impl": "reference", "metric": "compute_blocked_stats mean/variance cells/sec (1.3M cells; 400 + 8 blocks)", "value": value,
        "unit": "gene·block cells/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "config4", **{k: w[k] for k in ("ngenes", "ncells", "G", "B", "density")}},
        "cpu_baseline": {**vals[-1], "value": value},
        "e2e": {"value": value, "unit": "gene·block cells/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


def shard_starts(w, world):
    """Gene ranges of equal expected nnz (housekeeping genes are ten times heavier
    than the average, so equal gene counts leave the ranks unbalanced): int64[world + 1]."""
    from scran_b200._lib import check, lib, ptr
    from scran_b200.distributed import balanced_row_starts
    wts = np.zeros(w["ngenes"], np.float64)
    check(lib().scranb200_synth_row_weights(w["ngenes"], w["ncells"], float(w["density"]), int(w["seed"]), ptr(wts)))
    return balanced_row_starts(wts, world)


def multi_gpu_check(scran_b200, ctx, rank, world, dist):
    """Before anything is timed at N > 1: a config-2-sized matrix (33,538 x 100,000, 20 groups, 4 batches)
    is scored sharded over all ranks and, on rank 0, whole through the fused single-GPU call; the two
    [K][19][ngenes] tables must be bit-identical (the reference's multi-worker == serial check,
    tests/testthat/test-pairwise-wilcox.R:287-296). Exits non-zero otherwise."""
    from scran_b200.distributed import ShardedScorer
    w = dict(WORKLOADS["config2"])
    groups, block = labels_for(w)
    starts = shard_starts(w, world)
    plan = scran_b200.ScorePlan(groups.tolist(), block.tolist(), None, 0.0, ctx=ctx)
    scorer = ShardedScorer(ctx, plan, w["ngenes"], rank, world, dist, row_starts=starts, root=0)
    part = scran_b200.DeviceMatrix.synthetic(ctx, w["ngenes"], int(starts[rank]), int(starts[rank + 1]), w["ncells"], groups,
                                             w["G"], w["density"], w["seed"])
    got = np.zeros((plan.K, 19, w["ngenes"]), np.float64) if rank == 0 else None
    scorer.step(part, got)
    part.free()
    note = None
    if rank == 0:
        from scran_b200._lib import check, lib, ptr
        whole = scran_b200.DeviceMatrix.synthetic(ctx, w["ngenes"], 0, w["ngenes"], w["ncells"], groups, w["G"], w["density"],
                                                  w["seed"])
        want = np.zeros_like(got)
        check(lib().scranb200_score_markers(ctx.handle, whole.handle, plan.handle, ptr(want), None))
        whole.free()
        same = np.array_equal(got, want, equal_nan=True)
        note = "bit-identical" if same else "MISMATCH"
        if not same:
            bad = np.argwhere(~((got == want) | (np.isnan(got) & np.isnan(want))))
            print(json.dumps({"multi_gpu_check": "MISMATCH", "n_gpus": world, "first_bad": bad[:5].tolist()}), flush=True)
    flag = [note]
    dist.broadcast_object_list(flag, 0)
    plan.free()
    if flag[0] != "bit-identical":
        dist.destroy_process_group()
        sys.exit(3)
    return "bit-identical: config2 (33538 x 100000, 20 groups, 4 batches) sharded over %d ranks == fused single-GPU call" % world


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default=os.environ.get("SCRANB200_WORKLOAD", "config3"))
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--cpu-genes", type=int, default=0, help="genes in the CPU baseline sample (0 = auto)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-pageable", action="store_true", help="skip the end-to-end step from pageable host memory")
    args = ap.parse_args()
    w = dict(WORKLOADS[args.workload])
    args.warmup = max(args.warmup, 0)

    # stdout carries the JSON line and nothing else: whatever libraries print on file descriptor 1 while the
    # job runs (NCCL's version banner, for one) goes to stderr; Python's own prints still reach the real stdout
    sys.stdout.flush()
    sys.stdout = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)

    if args.workload == "config4":
        run_config4(args, w)
        return
    if args.impl == "reference":
        run_reference(args, w)
        return

    import torch

    import scran_b200
    from scran_b200._lib import check, lib, ptr
    from scran_b200.distributed import ShardedScorer

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    torch.cuda.set_device(local)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    ctx = scran_b200.Context(local)
    groups, block = labels_for(w)
    row_starts = shard_starts(w, world)
    g0, g1 = int(row_starts[rank]), int(row_starts[rank + 1])
    check_note = multi_gpu_check(scran_b200, ctx, rank, world, dist) if world > 1 else None
    t_gen = time.perf_counter()
    mat = scran_b200.DeviceMatrix.synthetic(ctx, w["ngenes"], g0, g1, w["ncells"], groups, w["G"], w["density"], w["seed"])
    t_gen = time.perf_counter() - t_gen
    nnz_local = mat.nnz
    nnz_t = torch.tensor([nnz_local], dtype=torch.int64, device="cuda")
    nnz_all = [torch.zeros_like(nnz_t) for _ in range(world)]
    if world > 1:
        dist.all_gather(nnz_all, nnz_t)
    else:
        nnz_all = [nnz_t]
    nnz_ranks = [int(t.item()) for t in nnz_all]
    nnz = sum(nnz_ranks)

    plan = scran_b200.ScorePlan(groups.tolist(), None if block is None else block.tolist(), None, 0.0, ctx=ctx)
    scorer = ShardedScorer(ctx, plan, w["ngenes"], rank, world, dist, row_starts=row_starts, root=0)
    P = w["G"] * (w["G"] - 1) // 2
    gene_pairs = w["ngenes"] * P
    ext = torch.cuda.ExternalStream(lib().scranb200_ctx_stream(ctx.handle), device=torch.device("cuda", local))

    for _ in range(args.warmup):
        scorer.step(mat)
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    launches0 = ctx.launches
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    phase_ms = {k: 0.0 for k in PHASES}
    ev0.record(ext)
    for _ in range(args.steps):
        scorer.step(mat)
        for k in PHASES:
            phase_ms[k] += scorer.last_timing.get(k, 0.0)
    ev1.record(ext)
    barrier()
    launches = ctx.launches - launches0
    ms_total = ev0.elapsed_time(ev1)
    t_ms = torch.tensor([ms_total], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t_ms, op=dist.ReduceOp.MAX)
    ms_step = float(t_ms.item()) / args.steps
    if world == 1:
        # a timed region shorter than nvidia-smi's sampling period (config 1: a few milliseconds) would leave no
        # clock sample at all: the same steps keep running, untimed, until one has arrived (at most a second)
        t_wait = time.perf_counter()
        while len(sampler.rows) < 2 and time.perf_counter() - t_wait < 1.0:
            scorer.step(mat)
    clocks = sampler.stop() if rank == 0 else None
    if clocks is not None and ms_total < 200.0:
        clocks["note"] = "timed region shorter than two sampling periods: sampled while the same steps kept running"
    for k in PHASES:
        phase_ms[k] /= args.steps
    value = gene_pairs / (ms_step * 1e-3)

    # ---- end to end through the host-buffer C-ABI calls --------------------------------------
    # N = 1: scranb200_upload_csc + scranb200_score_markers. N > 1: every rank uploads its row block and
    # all ranks run scranb200_score_markers_sharded; the complete [K][19][ngenes] table (min-ranks over
    # ALL genes) lands in the root's host buffer. Timed with the host clock around barriers.
    e2e = None
    if not args.no_e2e:
        p_h, i_h, x_h = mat.download()
        mat.free()
        torch.cuda.empty_cache()
        pin = [torch.from_numpy(a).pin_memory() for a in (p_h, i_h, x_h)]
        del p_h, i_h, x_h
        nrow = g1 - g0
        out_rows = w["ngenes"] if (world > 1 and rank == 0) else nrow
        out_h = torch.empty((max(1, plan.K), 19, max(1, out_rows)), dtype=torch.float64).pin_memory()
        L = lib()

        def e2e_step():
            import ctypes as C
            h = C.c_void_p()
            check(L.scranb200_upload_csc(ctx.handle, nrow, w["ncells"], ptr(pin[0]), ptr(pin[1]), ptr(pin[2]), C.byref(h)))
            try:
                if world == 1:
                    check(L.scranb200_score_markers(ctx.handle, h, plan.handle, ptr(out_h), None))
                else:
                    check(L.scranb200_score_markers_sharded(ctx.handle, h, plan.handle, w["ngenes"], ptr(scorer.row_starts), 0,
                                                            ptr(out_h) if rank == 0 else None, None))
            finally:
                L.scranb200_mat_free(h)

        e2e_step()
        barrier()
        t0 = time.perf_counter()
        n_e2e = max(1, min(args.steps, 2))
        for _ in range(n_e2e):
            e2e_step()
        barrier()
        dt = (time.perf_counter() - t0) / n_e2e
        t_e = torch.tensor([dt], dtype=torch.float64, device="cuda")
        h2d_t = torch.tensor([sum(int(a.numel() * a.element_size()) for a in pin)], dtype=torch.int64, device="cuda")
        if world > 1:
            dist.all_reduce(t_e, op=dist.ReduceOp.MAX)
            dist.all_reduce(h2d_t)
        d2h = int(max(1, plan.K) * 19 * w["ngenes"] * 8)
        pageable = None
        if world == 1 and not args.no_pageable:
            # the same call from ordinary (pageable) host memory, which is what an R process hands over
            src = [a.numpy().copy() for a in pin]
            pin_keep, pin[:] = list(pin), [torch.from_numpy(a) for a in src]
            e2e_step()  # (the first call from pageable memory also allocates the library's pinned staging buffers)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            e2e_step()
            torch.cuda.synchronize()
            dtp = time.perf_counter() - t0
            pin[:] = pin_keep
            del src
            pageable = {"value": gene_pairs / dtp, "unit": "gene·pairs/s", "ms_per_step": dtp * 1e3, "host_memory": "pageable"}
        e2e = {"value": gene_pairs / float(t_e.item()), "unit": "gene·pairs/s", "h2d_bytes_per_step": int(h2d_t.item()),
               "d2h_bytes_per_step": d2h, "ms_per_step": float(t_e.item()) * 1e3, "host_memory": "pinned",
               "pageable": pageable,
               "call": "scranb200_upload_csc + scranb200_score_markers (host buffers)" if world == 1 else
                       "per rank scranb200_upload_csc of its row block + scranb200_score_markers_sharded; complete table "
                       "(min-ranks over all genes) to the root's host buffer"}

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    peak, peak_src = measured_peak()
    pb = phase_bytes(w, nnz, plan)
    dom = max(PHASES, key=lambda k: phase_ms[k])
    kernels = {}
    for k in PHASES:
        ms = phase_ms[k]
        gbs = (pb[k] / world) / (ms * 1e-3) / 1e9 if ms > 0 else None
        kernels[k] = {"ms": round(ms, 4), "alg_bytes": int(pb[k] / world),
                      "gbs": None if gbs is None else round(gbs, 1), "frac": None if gbs is None else round(gbs / peak, 4)}
    # DRAM bytes of the dominant phase per step, from the committed ncu capture of this very workload
    # (profiles/r0X_dram_<workload>_n<N>.json; re-captured whenever that kernel changes); null when there is none
    traffic = traffic_src = None
    try:
        prof = os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles")
        for rnd in ("r02", "r01"):  # the newest capture of this workload
            path = os.path.join(prof, "%s_dram_%s_n%d.json" % (rnd, args.workload, world))
            if os.path.exists(path):
                cap = json.load(open(path))
                traffic = int(cap["phase_bytes"][dom])
                traffic_src = cap["source"]
                break
    except (OSError, KeyError, ValueError):
        pass
    roofline = {"bound": "hbm", "kernel": dom, "achieved": kernels[dom]["gbs"], "peak": peak, "unit": "GB/s",
                "frac": kernels[dom]["frac"], "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
                "pipeline_alg_bytes": int(algorithmic_bytes(w, nnz, P, plan.K)),
                "pipeline_frac": round(algorithmic_bytes(w, nnz, P, plan.K) / (ms_step * 1e-3) / 1e9 / (peak * world), 4)}

    cpu = None
    if not args.no_cpu and world == 1:
        threads = os.cpu_count() or 1
        per_gene = {"config3": 0.09, "config5": 0.12, "config2": 0.006}.get(args.workload, 0.001)
        nsample = args.cpu_genes or int(max(4, min(w["ngenes"], 15.0 / per_gene)))
        nsample = min(nsample, 1024)
        dt, kind, n = cpu_sample(w, groups, block, ctx, nsample, 1)
        cpu = {"value": n * P / dt, "unit": "gene·pairs/s", "cores": 1, "kind": "reference" if kind == "ref" else "port",
               "sample": "%d genes x all %d cells, native kernels only (blocked stats + U for %d block-level pairs), "
                         "%.1f s" % (n, w["ncells"], plan.npairs, dt),
               "host_cores_available": threads}

    line = {
        "metric": "scoreMarkers gene·pair scores/sec", "value": value, "unit": "gene·pairs/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": args.workload, **{k: w[k] for k in ("ngenes", "ncells", "G", "B", "density")},
                   "nnz": nnz, "pairs": P, "block_level_pairs": plan.npairs,
                   "sharding": "genes x%d, ranges of equal expected nnz" % world, "nnz_per_rank": nnz_ranks,
                   "l2": "inputs (%.1f GB) larger than L2" % (nnz * 12 / 1e9)},
        "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks, "roofline": roofline, "kernels": kernels,
        "cpu_baseline": cpu, "synth_seconds": round(t_gen, 2),
    }
    if check_note:
        line["multi_gpu_check"] = check_note
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
