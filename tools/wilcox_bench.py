"""pairwiseWilcox and pairwiseTTests end to end (statistics + p-value layer + log BH + the copy of the three
result arrays to the host) on a device-resident synthetic matrix of config 2's shape (33,538 x 100,000,
20 groups, 4 batches). Prints one JSON line."""
import json
import sys
import time

import numpy as np

sys.path.insert(0, ".")
import scran_b200 as sb  # noqa: E402

ngenes, ncells, G, B = 33538, 100000, 20, 4
rng = np.random.default_rng(5)
groups = rng.integers(0, G, ncells).astype(np.int32)
block = rng.integers(0, B, ncells).astype(np.int32)
ctx = sb.Context()
m = sb.DeviceMatrix.synthetic(ctx, ngenes, 0, ngenes, ncells, groups, G, 0.07, 1002)
res = {}
for name, fn in (("wilcox", sb.pairwise_wilcox), ("t_tests", sb.pairwise_t_tests)):
    fn(m, groups, block=block)  # warm-up
    t0 = time.perf_counter()
    for _ in range(3):
        out = fn(m, groups, block=block)
    res[name + "_s"] = (time.perf_counter() - t0) / 3
    res[name + "_min_log_p"] = float(min(np.nanmin(q["log_p"]) for q in out))
    res[name + "_gene_pairs_per_s"] = ngenes * G * (G - 1) / 2 / res[name + "_s"]
res["workload"] = "%d x %d, %d groups, %d batches, %d ordered pairs" % (ngenes, ncells, G, B, G * (G - 1))
print(json.dumps(res))
