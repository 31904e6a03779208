"""Host-side mirror of the reference's preprocessing_utils for the hot path.

Same names, arguments and error behaviour as node2vec2rank/preprocessing_utils.py
(network_transform :116-141, bipartite_to_unipartite_projection :16-32, match_networks
:291-306); the arithmetic runs on the B200 (kernels K1a/K1b/K2).
"""
from __future__ import annotations

import numpy as np
import pandas as pd

from . import device as D


def bipartite_to_unipartite_projection(graph, project_unipartite_on="columns"):
    """W^T W ('columns') or W W^T ('rows') -- preprocessing_utils.py:16-32 -- on the tensor cores.
    Returns a float32 ndarray (the reference returns the dtype of its input; its only caller casts
    to float32 right after, preprocessing_utils.py:141)."""
    r, c = np.shape(graph)
    assert r != c, "Graph is not bipartite"
    which = project_unipartite_on.casefold()
    if which not in ("columns", "rows"):
        raise ValueError("Unknown projection type, options are columns or rows")
    out = D.project(D.to_device(graph), which)
    assert out.shape[0] == out.shape[1], "Graph should be square"
    return out.cpu().numpy()


def network_transform_device(network, threshold=None, top_percent_keep=100, binarize=False, absolute=False,
                             project_unipartite_on="columns", storage_planes=None, want_f32=False, storage="auto", block_width=None):
    """Device-resident variant: returns (DeviceLayer, fp32 device tensor | None).  block_width: the solver block the layer
    will be multiplied with, when known (sizes the panels of a row-list layer, device.rows_panel_cols)."""
    x = D.to_device(network)
    return D.preprocess_to_layer(x, threshold=threshold, top_percent_keep=top_percent_keep, binarize=binarize,
                                 absolute=absolute, project_unipartite_on=project_unipartite_on,
                                 storage_planes=storage_planes, want_f32=want_f32, storage=storage, block_width=block_width)


def network_transform(network, threshold=None, top_percent_keep=100, binarize=False, absolute=False,
                      project_unipartite_on="columns"):
    """Drop-in for preprocessing_utils.py:116-141: returns a new float32 ndarray, input untouched."""
    x = D.to_device(network)
    rows, cols = x.shape
    if rows != cols:
        if absolute or threshold is not None:
            x, _, _, _ = D.transform(x, absolute=absolute, threshold=threshold, want_f32=True)
        x = D.project(x, project_unipartite_on)
        absolute, threshold = False, None
    cut = D.percentile_cut(x, 100 - top_percent_keep, absolute, threshold)
    out, _, _, _ = D.transform(x, absolute, threshold, float(cut), binarize, want_f32=True)
    return out.cpu().numpy()


def match_networks(graphs):
    """Common row / column labels of all graphs (preprocessing_utils.py:291-306).

    The reference orders them by Python set iteration (hash order, different per process and
    different for rows and columns).  Here the order is deterministic: the order of
    the first graph's labels.  Results keyed by node name are identical."""
    common_cols = set.intersection(*[set(g.columns.to_list()) for g in graphs])
    common_rows = set.intersection(*[set(g.index.to_list()) for g in graphs])
    cols = [c for c in graphs[0].columns.to_list() if c in common_cols]
    rows = [r for r in graphs[0].index.to_list() if r in common_rows]
    return [g.loc[rows, cols] for g in graphs]


__all__ = ["network_transform", "network_transform_device", "bipartite_to_unipartite_projection", "match_networks"]
