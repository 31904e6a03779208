"""Mirror of the numeric part of node2vec2rank/post_utils.py (:45-128): DCG / nDCG of a ranking against a
set of relevant items and the empirical permutation p-value, computed on the B200 (K12, n2v2r_ndcg: one
CTA per ranking).  The GSEA / enrichment / plotting functions of the reference module (gseapy, plotly,
matplotlib) are out of scope.

Same names, argument order, return values and zero-relevance behaviour as the reference.  Only the
reference's default weighting runs on the device (binary relevance, denominator log2(position + 2), ideal
ranking = the relevant items that occur in the ranking first); other weightings raise NotImplementedError
rather than falling back to the CPU.
"""
from __future__ import annotations

import numpy as np
import pandas as pd
import torch

from . import _native as N


def ndcg_of_rankings(rankings, relevant_mask, want_dcg=False):
    """nDCG (and DCG) of R rankings.  rankings: (R, len) integer item ids in rank order (host or device);
    relevant_mask: (n_items,) bool/uint8.  Returns float64 ndarray (R,) [and DCG (R,)]."""
    dev = N.require_gpu()
    r = torch.as_tensor(np.asarray(rankings) if not isinstance(rankings, torch.Tensor) else rankings)
    if r.dim() == 1:
        r = r[None, :]
    r = r.to(device=dev, dtype=torch.int32).contiguous()
    m = torch.as_tensor(np.asarray(relevant_mask) if not isinstance(relevant_mask, torch.Tensor) else relevant_mask)
    m = (m != 0).to(device=dev, dtype=torch.uint8).contiguous()
    R, length = r.shape
    out = torch.empty(R, dtype=torch.float64, device=dev)
    dcg = torch.empty(R, dtype=torch.float64, device=dev) if want_dcg else None
    N.check(N.load().n2v2r_ndcg(N.stream_ptr(), N.ptr(r), R, length, N.ptr(m), m.numel(), N.ptr(out), N.ptr(dcg)))
    if want_dcg:
        return out.cpu().numpy(), dcg.cpu().numpy()
    return out.cpu().numpy()


def _default_weighting(binary_relevance, denominator, ideal_numerator):
    if not binary_relevance or denominator is not None or ideal_numerator is not None:
        raise NotImplementedError("only the reference's default weighting (binary relevance, log2 discount) runs on the device")


def discounted_cumulative_gain(relevance_vector, denominator=None):
    """post_utils.py:45-50 for a 0/1 relevance vector in rank order: sum (2**rel - 1) / log2(pos + 2)."""
    rel = np.asarray(relevance_vector)
    if denominator is not None or not np.all((rel == 0) | (rel == 1)):
        raise NotImplementedError("only binary relevance with the log2 discount runs on the device")
    return float(ndcg_of_rankings(np.arange(rel.shape[0]), rel, want_dcg=True)[1][0])


def normalized_discounted_cumulative_gain(relevance_vector, binary_relevance=False, denominator=None, ideal_numerator=None,
                                          denom_all_rel=False, all_rel=None):
    """post_utils.py:80-108.  For a 0/1 vector the binary and the graded definitions coincide (2**1 - 1 = 1)."""
    rel = np.asarray(relevance_vector)
    if np.sum(rel) == 0:
        return 0
    if denominator is not None or ideal_numerator is not None or denom_all_rel or not np.all((rel == 0) | (rel == 1)):
        raise NotImplementedError("only binary relevance with the log2 discount runs on the device")
    return float(ndcg_of_rankings(np.arange(rel.shape[0]), rel)[0])


def _ids(universe: pd.Index, items):
    return universe.get_indexer(pd.Index(list(items))).astype(np.int32)          # -1 = unknown item = not relevant


def relevance_and_ndcg(ranked_items, relevant_items, binary_relevance=True, denominator=None, ideal_numerator=None):
    """post_utils.py:111-117: nDCG of one ranking (list of item names, best first); 0 when nothing relevant occurs."""
    _default_weighting(binary_relevance, denominator, ideal_numerator)
    ranked = list(ranked_items)
    universe = pd.Index(ranked)
    ids = np.arange(len(ranked), dtype=np.int32)
    if universe.has_duplicates:                       # the reference scores every position by membership
        rel = set(relevant_items)
        mask = np.fromiter((x in rel for x in ranked), dtype=np.uint8, count=len(ranked))
    else:
        mask = np.zeros(len(ranked), dtype=np.uint8)
        hit = _ids(universe, relevant_items)
        mask[hit[hit >= 0]] = 1
    if mask.sum() == 0:
        return 0
    return float(ndcg_of_rankings(ids, mask)[0])


def parallel_empirical_p_ndcg(ndcg, permutations, relevant_items, denominator=None, ideal_numerator=None):
    """post_utils.py:125-128: fraction of the given permutations whose nDCG exceeds `ndcg` (all permutations in one launch).
    `permutations`: list of rankings (item names), or an (R, len) integer array of item ids with `relevant_items` a mask."""
    _default_weighting(True, denominator, ideal_numerator)
    if isinstance(permutations, (np.ndarray, torch.Tensor)) and getattr(permutations, "ndim", 0) == 2:
        scores = ndcg_of_rankings(permutations, relevant_items)
    else:
        perms = [list(p) for p in permutations]
        universe = pd.Index(perms[0])
        mask = np.zeros(len(universe), dtype=np.uint8)
        hit = _ids(universe, relevant_items)
        mask[hit[hit >= 0]] = 1
        scores = ndcg_of_rankings(np.stack([_ids(universe, p) for p in perms]), mask)
    return (np.asarray(scores) > ndcg).sum() / len(scores)


def borda_ndcg(aggregate_ranks: pd.DataFrame, relevant_items, n_permutations=0, seed=0):
    """Quality gate for a Borda table (N2V2R.aggregate_transform()[key]): nDCG of the nodes ordered by descending Borda
    points (ties by node position, like the rank kernel) against `relevant_items`, and -- when n_permutations > 0 -- the
    empirical p-value over random permutations drawn on the device.  Returns (ndcg, p | None)."""
    from . import device as D
    dev = N.require_gpu()
    names = pd.Index(aggregate_ranks.index)
    pts = torch.from_numpy(aggregate_ranks.iloc[:, 0].to_numpy(dtype=np.float64)[None, :].copy()).to(dev)
    _, pos = D.rank_borda(pts, want_positions=True)                  # position of every node in the descending order
    n = names.shape[0]
    order = torch.empty(n, dtype=torch.int32, device=dev)
    order[pos[0].long()] = torch.arange(n, dtype=torch.int32, device=dev)
    mask = np.zeros(n, dtype=np.uint8)
    hit = _ids(names, relevant_items)
    mask[hit[hit >= 0]] = 1
    value = float(ndcg_of_rankings(order, mask)[0]) if mask.sum() else 0
    if not n_permutations:
        return value, None
    g = torch.Generator(device=dev)
    g.manual_seed(int(seed))
    perms = torch.stack([torch.randperm(n, generator=g, device=dev, dtype=torch.int32) for _ in range(int(n_permutations))])
    return value, float(parallel_empirical_p_ndcg(value, perms, mask))


__all__ = ["ndcg_of_rankings", "discounted_cumulative_gain", "normalized_discounted_cumulative_gain", "relevance_and_ndcg",
           "parallel_empirical_p_ndcg", "borda_ndcg"]
