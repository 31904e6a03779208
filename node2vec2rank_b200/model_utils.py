"""Mirror of node2vec2rank/model_utils.py on the B200 (kernels K9-K11).

Same names, arguments, return types and errors as the reference:
  compute_pairwise_distances  model_utils.py:39-67
  borda_aggregate_parallel    model_utils.py:28-36 (+ _get_ranking :22-25)
  signed_transform_single     model_utils.py:7-19
"""
from __future__ import annotations

import numpy as np
import pandas as pd
import torch

from . import _native as N
from . import device as D


def compute_pairwise_distances(mat1, mat2, distance='cosine'):
    """Row-wise distance between two (n, d) embedding matrices; returns a list of n floats.

    Runs the K9 kernel in float64 on the device (one thread per node).  cosine of a zero row
    is NaN, like scipy.spatial.distance.cosine in scipy >= 1.11."""
    if distance not in N.METRICS:
        raise NotImplementedError("Unsupported metric")
    a = np.ascontiguousarray(np.asarray(mat1, dtype=np.float64))
    b = np.ascontiguousarray(np.asarray(mat2, dtype=np.float64))
    if a.shape != b.shape or a.ndim != 2:
        raise ValueError("mat1 and mat2 must be 2-d arrays of the same shape")
    n, d = a.shape
    if n == 0:
        return []
    dev = N.require_gpu()
    emb = torch.from_numpy(np.stack([a, b])).to(dev)
    scores = D.distances(emb, np.asarray([d], dtype=np.int32),
                         np.asarray([N.METRICS[distance]], dtype=np.int32), N.SEQUENTIAL)
    return scores[0, 0].cpu().numpy().tolist()


def borda_aggregate_parallel(rankings: list):
    """Borda points of every candidate over a list of rankings (lists of names, best first).

    Output like the reference: DataFrame indexed by rankings[0], column 'borda_ranks' (int64),
    value = sum over rankings of (n - position).  The per-ranking O(n^2) list.index of the
    reference becomes a device rank + accumulate (K10/K11) on scores n - position."""
    index = rankings[0]
    n = len(index)
    slot = {node: i for i, node in enumerate(index)}
    if len(slot) != n:
        raise ValueError("duplicate candidates in a ranking")
    scores = np.empty((len(rankings), n), dtype=np.float64)
    for c, ranking in enumerate(rankings):
        if len(ranking) != n:
            raise ValueError(f"{ranking[-1] if ranking else None!r} is not in list")
        ids = np.fromiter((slot[node] for node in ranking), dtype=np.int64, count=n)   # KeyError ~ reference's ValueError
        col = np.full(n, np.nan)
        col[ids] = n - np.arange(n, dtype=np.float64)
        if np.isnan(col).any():
            raise ValueError("a candidate is missing from a ranking")
        scores[c] = col
    dev = N.require_gpu()
    borda, _ = D.rank_borda(torch.from_numpy(scores).to(dev))
    return pd.DataFrame(borda.cpu().numpy(), index=index, columns=['borda_ranks'])


def signed_transform_single(ranks: pd.Series, prior_signed_ranks: pd.Series):
    """Keep the score where the prior is > 0, negate it otherwise; nodes missing from the
    prior are dropped (model_utils.py:7-19).  Host-side: one vectorised select over n values."""
    keep = ranks.index.isin(prior_signed_ranks.index)
    r = ranks[keep]
    prior = prior_signed_ranks.loc[r.index].to_numpy()
    vals = np.where(prior > 0, r.to_numpy(), -r.to_numpy())
    return pd.Series(vals.tolist(), index=list(r.index))


__all__ = ["compute_pairwise_distances", "borda_aggregate_parallel", "signed_transform_single"]
