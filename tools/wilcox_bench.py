# pairwiseWilcox-style timing on config 2 (20 groups, 4 batches): overlap_exprs per block level with ties
import sys, time, numpy as np
sys.path.insert(0, '.')
import scran_b200, bench
w = dict(bench.WORKLOADS['config2']); ctx = scran_b200.default_context()
groups, block = bench.labels_for(w)
m = scran_b200.DeviceMatrix.synthetic(ctx, w['ngenes'], 0, w['ngenes'], w['ncells'], groups, w['G'], w['density'], w['seed'])
from scran_b200._lib import lib, check, ptr
import ctypes as C
for rep in range(2):
    t0 = time.perf_counter(); tk = 0.0
    for b in range(w['B']):
        chosen = np.flatnonzero(block == b)
        by = [chosen[groups[chosen] == g].astype(np.int32) for g in range(w['G'])]
        gptr = np.zeros(w['G'] + 1, np.int64); gptr[1:] = np.cumsum([len(v) for v in by]); cells = np.concatenate(by)
        U = np.zeros((w['G'], w['G'], w['ngenes'])); T = np.zeros_like(U)
        check(lib().scranb200_overlap_exprs(ctx.handle, m.handle, w['G'], ptr(gptr), ptr(cells), 0.0, ptr(U), ptr(T)))
        tk += ctx.last_timing()['total']
    print("pairwiseWilcox native part, config2: wall %.1f ms, device %.1f ms, last call phases %s" % ((time.perf_counter() - t0) * 1e3, tk, ctx.last_timing()))
