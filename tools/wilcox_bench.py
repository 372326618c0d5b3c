"""pairwiseWilcox end to end on a device-resident synthetic matrix of config 2's shape (33,538 x 100,000,
20 groups, 4 batches): the p-value layer on the device against the same layer in numpy / scipy on the
host (both on top of the device U and tie terms). Prints one JSON line."""
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
for name, host_layer, reps in (("device_layer", False, 3), ("host_layer", True, 1)):
    sb.pairwise_wilcox(m, groups, block=block, host_layer=host_layer) if not host_layer else None  # warm-up (device)
    t0 = time.perf_counter()
    for _ in range(reps):
        out = sb.pairwise_wilcox(m, groups, block=block, host_layer=host_layer)
    res[name + "_s"] = (time.perf_counter() - t0) / reps
    res[name + "_min_log_p"] = float(min(np.nanmin(q["log_p"]) for q in out))
res["workload"] = "pairwiseWilcox, %d x %d, %d groups, %d batches, %d ordered pairs" % (ngenes, ncells, G, B, G * (G - 1))
res["gene_pairs_per_s_device_layer"] = ngenes * G * (G - 1) / 2 / res["device_layer_s"]
print(json.dumps(res))
