import sys, numpy as np
sys.path.insert(0, "/root/repo")
import scran_b200 as sb
from bench import WORKLOADS, labels_for
w = WORKLOADS["config5"]
groups, block = labels_for(w)
ctx = sb.Context(0)
plan = sb.ScorePlan(groups.tolist(), None, None, 0.0, ctx=ctx)
for g0, g1 in ((0, 2500), (8, 1000), (1100, 2000), (5000, 7500), (2100, 4600)):
    m = sb.DeviceMatrix.synthetic(ctx, w["ngenes"], g0, g1, w["ncells"], groups, w["G"], w["density"], w["seed"])
    for _ in range(2):
        sb.score_markers(m, groups, plan=plan)
    t = ctx.last_timing()
    print(g0, g1, "nnz", m.nnz, {k: round(v, 2) for k, v in t.items()}, flush=True)
    m.free()
