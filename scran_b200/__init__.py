"""scran_b200: the marker-scoring hot path of scran (scoreMarkers / pairwiseWilcox /
pairwiseTTests and the native routines under them) as hand-written sm_100a CUDA behind a C ABI.

``scran_b200.api`` mirrors the reference's function surface; ``scran_b200._lib`` is the raw
ctypes binding of include/scranb200.h. Nothing here computes on the CPU.
"""
from .api import (COLUMNS, EFFECTS, Context, DeviceMatrix, ScorePlan, combine_markers, compute_blocked_stats_lognorm, compute_blocked_stats_none, compute_blocked_stats_norm,
                  compute_min_rank,
                  default_context, expand_pairings, find_markers_wilcox, overlap_exprs, overlap_exprs_paired, pairwise_t_tests,
                  pairwise_wilcox, score_markers, summary_marker_stats)
from ._lib import NA_INT, ScranB200Error

__all__ = ["COLUMNS", "EFFECTS", "Context", "DeviceMatrix", "ScorePlan", "combine_markers", "find_markers_wilcox", "compute_blocked_stats_none", "compute_blocked_stats_lognorm", "compute_blocked_stats_norm",
           "compute_min_rank", "default_context", "expand_pairings", "overlap_exprs", "overlap_exprs_paired",
           "pairwise_t_tests", "pairwise_wilcox", "score_markers", "summary_marker_stats", "NA_INT", "ScranB200Error"]
