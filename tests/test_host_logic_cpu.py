"""CPU: the C++ host bookkeeping behind scoreMarkers (scran_b200/csrc/score_plan.cu) against the
independent numpy restatement of R/scoreMarkers.R:191-217, 293-345, 463-521 in oracle/rlevel.py,
and .expand_pairings against R/expandPairings.R (tests/testthat/test-expand-pairings.R:4-94)."""
import ctypes as C

import numpy as np
import pytest

import scran_b200
from oracle import rlevel
from scran_b200._lib import check, lib, ptr


def host_plan(groups, block, pairings, lfc):
    groups = list(groups)
    universe = sorted(set(groups))
    code = {g: k for k, g in enumerate(universe)}
    gc = np.array([code[g] for g in groups], np.int32)
    bc = None
    if block is not None:
        lv = sorted(set(block))
        bc = np.array([lv.index(b) for b in block], np.int32)
    h = C.c_void_p()
    if pairings is None:
        check(lib().scranb200_plan_create(None, len(groups), ptr(gc), len(universe), ptr(bc), None, None, 0, lfc, C.byref(h)))
    else:
        i1, i2 = scran_b200.expand_pairings(pairings, universe)
        dl, dr = np.array(i1, np.int32), np.array(i2, np.int32)
        check(lib().scranb200_plan_create(None, len(groups), ptr(gc), len(universe), ptr(bc), ptr(dl), ptr(dr), len(i1), lfc,
                                          C.byref(h)))
    info = np.zeros(5, np.int64)
    check(lib().scranb200_plan_info(h, ptr(info), 5))
    K, nd, Q, npairs, Cn = [int(v) for v in info]
    a = dict(K=K, nd=nd, Q=Q, npairs=npairs, C=Cn, universe=universe)
    a["comb_group"], a["comb_block"] = np.zeros(Cn, np.int32), np.zeros(Cn, np.int32)
    a["ncells"] = np.zeros(Cn)
    a["left"], a["right"], a["avg_idx"] = (np.zeros(max(1, npairs), np.int32) for _ in range(3))
    a["avg_left"], a["avg_right"] = np.zeros(max(1, Q), np.int32), np.zeros(max(1, Q), np.int32)
    a["avg_ptr"] = np.zeros(Q + 1, np.int32)
    a["q_direct"], a["q_flipped"] = np.zeros(max(1, nd), np.int32), np.zeros(max(1, nd), np.int32)
    a["comb_id"] = np.zeros(len(groups), np.int32)
    check(lib().scranb200_plan_pairs(h, ptr(a["comb_group"]), ptr(a["comb_block"]), ptr(a["ncells"]), ptr(a["left"]),
                                     ptr(a["right"]), ptr(a["avg_left"]), ptr(a["avg_right"]), ptr(a["avg_ptr"]),
                                     ptr(a["avg_idx"]), ptr(a["q_direct"]), ptr(a["q_flipped"]), ptr(a["comb_id"])))
    a["out_groups"] = np.zeros(max(1, K), np.int32)
    a["right_ptr"] = np.zeros(K + 1, np.int64)
    a["right_group"] = np.zeros(max(1, nd), np.int32)
    a["weight"] = np.zeros(max(1, nd))
    check(lib().scranb200_plan_tables(h, ptr(a["out_groups"]), ptr(a["right_ptr"]), ptr(a["right_group"]), ptr(a["weight"])))
    lib().scranb200_plan_destroy(h)
    return a


CASES = [
    dict(G=4, B=None, pairings=None, lfc=0.0),
    dict(G=5, B=3, pairings=None, lfc=0.0),
    dict(G=4, B=3, pairings=None, lfc=0.5),
    dict(G=6, B=None, pairings=[1, 3, 4], lfc=0.0),
    dict(G=5, B=2, pairings=([0, 1], [2, 3, 4]), lfc=0.0),
    dict(G=5, B=2, pairings=np.array([[0, 1], [2, 1], [4, 0], [1, 0]]), lfc=1.0),
    dict(G=4, B=3, pairings=None, lfc=0.0, confound=True),
]


@pytest.mark.parametrize("case", CASES, ids=[str(i) for i in range(len(CASES))])
def test_plan_matches_the_restated_r_bookkeeping(case):
    rng = np.random.default_rng(17)
    n = 400
    groups = rng.integers(0, case["G"], n)
    block = None if case["B"] is None else rng.integers(0, case["B"], n)
    if case.get("confound"):
        block[groups == 3] = 2
        block[(groups == 0) & (block == 2)] = 0
    a = host_plan(groups.tolist(), None if block is None else block.tolist(), case["pairings"], case["lfc"])
    st = rlevel.ScoreSetup(groups.tolist(), None if block is None else block.tolist(), case["pairings"], case["lfc"])
    uni = a["universe"]
    assert a["C"] == len(st.combos)
    assert [uni[g] for g in a["comb_group"]] == [c[0] for c in st.combos]
    assert np.array_equal(a["ncells"], st.ncells)
    kept = a["comb_id"] >= 0
    assert np.array_equal(kept, st.keep)
    assert np.array_equal(a["comb_id"][kept], st.comb_id)
    assert np.array_equal(a["left"][:a["npairs"]], st.left) and np.array_equal(a["right"][:a["npairs"]], st.right)
    assert a["Q"] == len(st.averaged)
    assert [(uni[l], uni[r]) for l, r in zip(a["avg_left"][:a["Q"]], a["avg_right"][:a["Q"]])] == list(st.averaged)
    for q in range(a["Q"]):
        assert np.array_equal(a["avg_idx"][a["avg_ptr"][q]:a["avg_ptr"][q + 1]], st.to_average[q])
    assert [uni[g] for g in a["out_groups"][:a["K"]]] == list(st.desired_indices.keys())
    for k, (g, cur) in enumerate(st.desired_indices.items()):
        lo, hi = a["right_ptr"][k], a["right_ptr"][k + 1]
        assert [uni[r] for r in a["right_group"][lo:hi]] == list(cur["right"])
        assert [(-1 if d is None else d) for d in cur["direct"]] == a["q_direct"][lo:hi].tolist()
        flipped = [(-1 if (f is None or d is not None) else f) for d, f in zip(cur["direct"], cur["flipped"])]
        assert flipped == a["q_flipped"][lo:hi].tolist()


def test_expand_pairings_like_the_reference():
    uni = ["a", "b", "c", "d"]
    for pairings in (None, ["b", "d", "zz"], (["a", "b"], ["b", "c", "d"]), (None, ["c"]),
                     np.array([["a", "b"], ["c", "zz"], ["d", "a"]])):
        assert scran_b200.expand_pairings(pairings, uni) == rlevel.expand_pairings(pairings, uni)
    i1, i2 = scran_b200.expand_pairings(None, uni)
    assert len(i1) == 12 and (i1[0], i2[0]) == (1, 0)   # expand.grid: first factor fastest
    with pytest.raises(ValueError):
        scran_b200.expand_pairings(np.zeros((2, 3)), uni)


def test_plan_rejects_bad_input():
    g = np.array([0, 1, 5], np.int32)
    h = C.c_void_p()
    assert lib().scranb200_plan_create(None, 3, ptr(g), 2, None, None, None, 0, 0.0, C.byref(h)) != 0
    assert "out of range" in lib().scranb200_last_error().decode()
