"""Generates tests/golden/*.npz: small seeded inputs together with what the REFERENCE'S OWN
OBJECT CODE (oracle/_ref/liboracle_ref.so = /root/reference/src/overlap_exprs.cpp and
compute_blocked_stats.cpp compiled unmodified, see oracle/Makefile) returns for them.

Run in the build container (it needs /root/reference):   python tests/golden/make_golden.py
The fixtures then pin the C port (tests/test_oracle.py) and the CUDA path (tests/test_gpu_golden.py)
on machines where /root/reference does not exist.
"""
import os
import sys

import numpy as np
import scipy.sparse as sp

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, os.path.dirname(HERE))

from helpers import all_pairs, make_matrix  # noqa: E402
from oracle import native  # noqa: E402


def main():
    native.build(with_ref=True)
    ref = native.load("ref")
    cases = [
        dict(name="g4_lfc0", ngenes=40, ncells=61, G=4, lfc=0.0, seed=1),
        dict(name="g3_lfc05", ngenes=30, ncells=47, G=3, lfc=0.5, seed=2),
        dict(name="g5_lfc1_int", ngenes=25, ncells=80, G=5, lfc=1.0, seed=3),
        dict(name="g2_neg", ngenes=20, ncells=33, G=2, lfc=-0.75, seed=4),
    ]
    for c in cases:
        x = make_matrix(c["ngenes"], c["ncells"], seed=c["seed"])
        rng = np.random.default_rng(100 + c["seed"])
        g = rng.integers(0, c["G"], c["ncells"]).astype(np.int32)
        block = rng.integers(0, 3, c["ncells"]).astype(np.int32)
        block[::11] = native.NA_INT
        groups = [np.flatnonzero(g == k).astype(np.int32) for k in range(c["G"])]
        left, right = all_pairs(c["G"])
        m = sp.csc_matrix(x)
        mean, var = ref.blocked_stats(m, block, 4)   # block 3 is empty
        dmean, dvar = ref.blocked_stats(x, block, 4)
        u = ref.overlap_paired(m, left, right, groups, c["lfc"])
        U, T = ref.overlap(m, groups, c["lfc"])
        np.savez_compressed(os.path.join(HERE, c["name"] + ".npz"), x=x, g=g, block=block, lfc=c["lfc"], left=left,
                            right=right, mean=mean, var=var, dense_mean=dmean, dense_var=dvar, u_paired=u, U=U, T=T)
        print("wrote", c["name"])


if __name__ == "__main__":
    main()
