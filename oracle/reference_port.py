"""CPU oracle for the UASE embed-and-rank path.  TEST INFRASTRUCTURE ONLY.

This module is a NumPy/SciPy restatement of the reference's algorithm for the hot
path.  It is *not* part of the product: only ``tests/``, ``__graft_entry__.smoke()``
and ``bench.py``'s ``cpu_baseline`` / ``--impl reference`` legs may import it, and only
as the checker or the timed CPU baseline.  The product (``node2vec2rank_b200``) never
imports anything under ``oracle/`` and has no CPU fallback.

Parity pin
----------
The reference has no tests and its core arithmetic (``spectral_embedding.UASE``) is an
un-vendored, un-pinned pip-git dependency (``environment.yaml:19``,
``git+https://github.com/iggallagher/Spectral-Embedding``).  The pin is therefore
"outputs of the reference itself run in the build container": ``oracle/make_golden.py``
imports the reference's own ``dataloader`` / ``model`` / ``model_utils`` /
``preprocessing_utils`` modules unmodified (behind two ``sys.modules`` stubs: an empty
``plotly`` and a ``spectral_embedding`` exposing the UASE restated below) and writes
golden vectors to ``tests/golden/``.  ``tests/test_oracle_golden.py`` checks every
function in this file against those vectors.  The UASE boundary itself is pinned by the
demo notebook's published recall (0.68 / 0.0) which only the S^1/2 scaling reproduces.

Every function cites the reference file:line it follows (paths relative to the
reference checkout).
"""
from __future__ import annotations

import math
import time

import numpy as np
import scipy.sparse
import scipy.sparse.linalg
import scipy.spatial.distance

# --------------------------------------------------------------------------------------
# preprocessing_utils.py
# --------------------------------------------------------------------------------------


def bipartite_to_unipartite_projection(graph, project_unipartite_on="columns"):
    """preprocessing_utils.py:16-32 -- W^T W ('columns') or W W^T ('rows')."""
    r, c = graph.shape
    assert r != c, "Graph is not bipartite"
    which = project_unipartite_on.casefold()  # AttributeError on None, as the reference
    if which == "columns":
        out = np.matmul(graph.T, graph)
    elif which == "rows":
        out = np.matmul(graph, graph.T)
    else:
        raise ValueError("Unknown projection type, options are columns or rows")
    assert out.shape[0] == out.shape[1], "Graph should be square"
    return out


def network_transform(network, threshold=None, top_percent_keep=100, binarize=False,
                      absolute=False, project_unipartite_on="columns"):
    """preprocessing_utils.py:116-141.

    copy -> abs -> (< threshold -> 0) -> projection if non-square -> percentile over the
    nonzeros (``np.percentile`` linear interpolation) -> (< cut -> 0) -> binarize -> fp32.
    """
    r, c = np.shape(network)
    x = np.array(getattr(network, "values", network), copy=True)   # :119-122
    if absolute:                                                   # :124-125
        x = np.abs(x)
    if threshold is not None:                                      # :127-128
        x[x < threshold] = 0
    if r != c:                                                     # :130-132
        x = bipartite_to_unipartite_projection(x, project_unipartite_on)
    cut_off = np.percentile(x[x.nonzero()], 100 - top_percent_keep)  # :134
    x[x < cut_off] = 0                                             # :135
    if binarize:                                                   # :138-139
        x[x != 0] = 1
    return np.float32(x)                                           # :141


def match_networks_labels(row_label_sets, col_label_sets):
    """preprocessing_utils.py:291-306 restated on labels only.

    The reference keeps ``list(set.intersection(...))`` order (hash order, different per
    process).  The oracle returns the common labels as *sets*; callers compare results
    keyed by node name, never by position.
    """
    return set.intersection(*map(set, row_label_sets)), set.intersection(*map(set, col_label_sets))


# --------------------------------------------------------------------------------------
# spectral_embedding.UASE (third party, restated; call site model.py:54-55)
# --------------------------------------------------------------------------------------


def UASE(As, d, svds_kwargs=None):
    """Unfolded adjacency spectral embedding.

    ``A = hstack(As)`` (n x T*n); top-d singular triplets by ``scipy.sparse.linalg.svds``
    (ARPACK on the Gram operator, tol=0, v0 from the global NumPy RNG -- scipy
    ``_svds.py:428-535``); sigma sorted descending; ``XA = U*sqrt(S)``;
    ``YA[t] = V[n*t:n*(t+1)] * sqrt(S)``.  dtype follows the layers (fp32 in the reference
    because ``network_transform`` returns fp32 -- preprocessing_utils.py:141, model.py:53).
    """
    T = len(As)
    n = As[0].shape[0]
    A = scipy.sparse.hstack(As)
    U, S, Vt = scipy.sparse.linalg.svds(A, d, **(svds_kwargs or {}))
    idx = np.argsort(S)[::-1]
    U, S, V = U[:, idx], S[idx], Vt.T[:, idx]
    rt = np.sqrt(S)
    XA = U * rt
    YA = np.zeros((T, n, d))
    for t in range(T):
        YA[t] = V[n * t:n * (t + 1), :] * rt
    return XA, YA


def uase_dense_fp64(graphs, d):
    """Ground-truth UASE: dense float64 LAPACK SVD of the unfolded matrix.

    Used as the parity *truth* (fp64 ARPACK == fp64 LAPACK to 1e-12, while the
    reference's fp32 ARPACK differs from both by up to 1.9e-3 at bulk cuts).
    Returns (sigma (d,), V (T*n, d), YA (T, n, d)).  Feasible for n*T <= ~8000.
    """
    A = np.hstack([np.asarray(g, dtype=np.float64) for g in graphs])
    n = A.shape[0]
    T = len(graphs)
    # eigen-decomposition of A A^T (n x n) is cheaper than the SVD of n x T*n
    M = A @ A.T
    lam, U = np.linalg.eigh(M)
    idx = np.argsort(lam)[::-1][:d]
    lam, U = lam[idx], U[:, idx]
    S = np.sqrt(np.maximum(lam, 0))
    V = (A.T @ U) / S
    YA = np.stack([V[n * t:n * (t + 1), :] * np.sqrt(S) for t in range(T)])
    return S, V, YA


# --------------------------------------------------------------------------------------
# model_utils.py
# --------------------------------------------------------------------------------------


def compute_pairwise_distances(mat1, mat2, distance="cosine"):
    """model_utils.py:39-67 -- one scipy.spatial.distance call per node (faithful)."""
    if distance == "cosine":
        f = scipy.spatial.distance.cosine
    elif distance == "euclidean":
        f = scipy.spatial.distance.euclidean
    elif distance == "correlation":
        f = scipy.spatial.distance.correlation
    else:
        raise NotImplementedError("Unsupported metric")
    return [f(r1, r2) for r1, r2 in zip(mat1, mat2)]


def compute_pairwise_distances_vec(mat1, mat2, distance="cosine"):
    """Vectorised float64 equivalent of :func:`compute_pairwise_distances`.

    euclidean = ||u-v||_2; cosine = clip(1 - u.v/sqrt(u.u * v.v), 0, 2) (NaN when a row is
    zero, scipy 1.18 ``distance.correlation(centered=False)``); correlation = cosine of the
    mean-centred rows.  Used for the large-n parity cases where n Python calls per column
    are too slow; agrees with the faithful version to ~1e-15.
    """
    u = np.asarray(mat1, dtype=np.float64)
    v = np.asarray(mat2, dtype=np.float64)
    if distance == "euclidean":
        return np.sqrt(np.sum((u - v) ** 2, axis=1))
    if distance == "correlation":
        u = u - u.mean(axis=1, keepdims=True)
        v = v - v.mean(axis=1, keepdims=True)
    elif distance != "cosine":
        raise NotImplementedError("Unsupported metric")
    uv = np.sum(u * v, axis=1)
    uu = np.sum(u * u, axis=1)
    vv = np.sum(v * v, axis=1)
    with np.errstate(divide="ignore", invalid="ignore"):
        d = 1.0 - uv / np.sqrt(uu * vv)
    return np.clip(d, 0.0, 2.0)


def _get_ranking(ranking, index):
    """model_utils.py:22-25 -- O(n^2) list.index (faithful)."""
    n = len(index)
    return [n - ranking.index(node) for node in index]


def borda_aggregate(rankings, faithful=True, n_jobs=-2):
    """model_utils.py:28-36.  rankings: list of lists of node names (best first).

    Returns (index, borda_ranks int64) with index = rankings[0] order, exactly like the
    reference's DataFrame.  ``faithful`` uses the reference's joblib + list.index route
    (what the CPU baseline times); otherwise an O(n) dictionary route with identical output.
    """
    index = rankings[0]
    if faithful:
        from joblib import Parallel, delayed
        res = np.asarray(Parallel(n_jobs=n_jobs)(delayed(_get_ranking)(r, index) for r in rankings))
    else:
        n = len(index)
        res = np.empty((len(rankings), n), dtype=np.int64)
        for c, r in enumerate(rankings):
            pos = {node: i for i, node in enumerate(r)}
            res[c] = [n - pos[node] for node in index]
    return index, np.sum(res, axis=0)


def descending_positions(scores):
    """Position (0 = best) of every node in a descending sort of one score column.

    model.py:171-175 uses ``Series.sort_values(ascending=False)`` whose default
    ``kind='quicksort'`` is unstable for ties and puts NaN last.  The product defines the
    tie order as *stable by node position* (SURVEY 7.3-6); this is that definition on the
    CPU: sort by (-score, position), NaN last.
    """
    s = np.asarray(scores, dtype=np.float64)
    n = s.shape[0]
    nan = np.isnan(s)
    key = np.where(nan, -np.inf, s)
    order = np.lexsort((np.arange(n), -key, nan))     # last key is primary
    pos = np.empty(n, dtype=np.int64)
    pos[order] = np.arange(n)
    return pos


def borda_from_scores(score_table):
    """Borda points for a table (n, C) of scores: sum_c (n - position_c).  model.py:167-185
    + model_utils.py:22-36 with the stable tie rule above.  Output in node order."""
    s = np.asarray(score_table, dtype=np.float64)
    n = s.shape[0]
    out = np.zeros(n, dtype=np.int64)
    for c in range(s.shape[1]):
        out += n - descending_positions(s[:, c])
    return out


# --------------------------------------------------------------------------------------
# model.py
# --------------------------------------------------------------------------------------


def rank_columns(node_embeddings, embed_dimensions, distance_metrics, comp_strategy,
                 vectorised=True):
    """model.py:57-96 (__rank).  Returns dict key -> (column names, (n, C) float64)."""
    emb = np.asarray(node_embeddings)
    T = emb.shape[0]
    dist = compute_pairwise_distances_vec if vectorised else compute_pairwise_distances
    out = {}
    for i in range(T):
        if i == 0 and comp_strategy != "one_vs_rest":            # :59-61
            continue
        key = str(i) if comp_strategy != "one_vs_rest" else str(i + 1)   # :63-66
        names, cols = [], []
        for dim in embed_dimensions:                             # :73
            if comp_strategy == "sequential":                    # :74-76
                one, two = emb[i - 1, :, :dim], emb[i, :, :dim]
            elif comp_strategy == "one_vs_before":               # :78-80
                one, two = np.mean(emb[:i, :, :dim], axis=0), emb[i, :, :dim]
            elif comp_strategy == "one_vs_rest":                 # :82-84
                one = np.mean(emb[np.arange(T) != i, :, :dim], axis=0)
                two = emb[i, :, :dim]
            else:
                raise UnboundLocalError("embed_one")             # reference behaviour
            for metric in distance_metrics:                      # :87
                if metric == "cosine" and dim == 1:              # :88-89
                    continue
                names.append(f"dim-{dim}_distance-{metric}")     # :90
                cols.append(np.asarray(dist(one, two, metric), dtype=np.float64))
        out[key] = (names, np.stack(cols, axis=1) if cols else np.zeros((emb.shape[1], 0)))
    return out


def degree_difference(graphs):
    """model.py:282-311 -- column sums of consecutive layers; keys '1'..'T-1'."""
    out = {}
    for i in range(1, len(graphs)):
        dedi = np.sum(graphs[i - 1], axis=0) - np.sum(graphs[i], axis=0)
        out[str(i)] = (dedi, np.abs(dedi))
    return out


def signed_transform(scores, prior):
    """model_utils.py:7-19 -- keep score where prior > 0, negate otherwise."""
    return np.where(np.asarray(prior) > 0, scores, -np.asarray(scores))


def fit_rank_borda(graphs, embed_dimensions, distance_metrics, comp_strategy,
                   faithful=True, dtype=None, timings=None):
    """The whole timed path of the reference: model.py:98-201 on already-transformed graphs.

    csc_matrix per layer (model.py:53) -> UASE at max(dims) -> distance columns -> per
    column descending sort -> Borda.  ``dtype=np.float64`` gives the fp64 oracle.
    Returns (node_embeddings, {key: (names, scores)}, {key: borda int64 in node order}).
    """
    t0 = time.perf_counter()
    gs = [g if dtype is None else np.asarray(g, dtype=dtype) for g in graphs]
    sparse = [scipy.sparse.csc_matrix(g) for g in gs]
    _, emb = UASE(sparse, max(embed_dimensions))
    t1 = time.perf_counter()
    ranks = rank_columns(emb, embed_dimensions, distance_metrics, comp_strategy,
                         vectorised=not faithful)
    t2 = time.perf_counter()
    n = emb.shape[1]
    borda = {}
    for key, (names, scores) in ranks.items():
        if faithful:
            import pandas as pd
            nodes = list(range(n))
            lists = []
            for c in range(scores.shape[1]):
                s = pd.Series(scores[:, c], index=nodes)
                s.sort_values(ascending=False, inplace=True)      # model.py:174
                lists.append(s.index.to_list())
            index, pts = borda_aggregate(lists, faithful=True)
            arr = np.empty(n, dtype=np.int64)
            arr[np.asarray(index)] = pts                           # .loc[node_names] :185
            borda[key] = arr
        else:
            borda[key] = borda_from_scores(scores)
    t3 = time.perf_counter()
    if timings is not None:
        timings.update(fit=t1 - t0, distances=t2 - t1, borda=t3 - t2, total=t3 - t0)
    return emb, ranks, borda


def stability_borda(graphs, embed_dimensions, distance_metrics, comp_strategy, seeds, dtype=None):
    """README.md:54,173 -- "running it multiple times and aggregating": one fit_transform_rank
    (model.py:98-147) per seed (model.py:36-38 seeds np.random, which is svds' start vector),
    all runs' columns concatenated per comparison (seeds outer), then aggregate_transform's Borda
    (model.py:149-201) over the union.  Returns ({key: (names, scores (n, S*C))}, {key: borda})."""
    union = {}
    for seed in seeds:
        np.random.seed(seed)
        _, ranks, _ = fit_rank_borda(graphs, embed_dimensions, distance_metrics, comp_strategy,
                                     faithful=False, dtype=dtype)
        for key, (names, scores) in ranks.items():
            u = union.setdefault(key, ([], []))
            u[0].extend(f"seed-{seed}_{c}" for c in names)
            u[1].append(scores)
    tables = {key: (names, np.concatenate(blocks, axis=1)) for key, (names, blocks) in union.items()}
    return tables, {key: borda_from_scores(t[1]) for key, t in tables.items()}


# --------------------------------------------------------------------------------------
# dataloader.py (edge-list branch)
# --------------------------------------------------------------------------------------


def read_weighted_edgelist_dense(text, delimiter=" "):
    """dataloader.py:100-104 restated without networkx: nx.read_weighted_edgelist(path, delimiter, nodetype=str) followed by
    nx.to_numpy_array.  Lines are cut at '#', split on `delimiter`, need exactly (u, v, weight); the graph is undirected
    (a later line of the same edge, in either direction, overwrites the weight); nodes are numbered in order of first
    appearance.  Returns (node labels, dense float64 adjacency)."""
    nodes, index, weights = [], {}, {}
    for line in text.splitlines():
        p = line.find("#")
        if p >= 0:
            line = line[:p]
        if not line:
            continue
        s = line.rstrip("\n").split(delimiter)
        if len(s) < 2:
            continue
        u, v, d = s[0], s[1], s[2:]
        if len(d) != 1:
            raise IndexError(f"Edge data {d} and data_keys (('weight', float),) are not the same length")
        for x in (u, v):
            if x not in index:
                index[x] = len(nodes)
                nodes.append(x)
        a, b = index[u], index[v]
        weights[(min(a, b), max(a, b))] = float(d[0])
    A = np.zeros((len(nodes), len(nodes)), dtype=np.float64)
    for (a, b), w in weights.items():
        A[a, b] = A[b, a] = w
    return nodes, A


# --------------------------------------------------------------------------------------
# post_utils.py (numeric part only)
# --------------------------------------------------------------------------------------


def discounted_cumulative_gain(relevance_vector, denominator=None):
    """post_utils.py:45-50."""
    rel = np.array(relevance_vector)
    if denominator is None:
        denominator = np.log2(np.arange(rel.shape[0]) + 2)
    return np.sum(np.divide(2 ** rel - 1, denominator))


def relevance_and_ndcg(ranked_items, relevant_items):
    """post_utils.py:111-117 with the defaults of :80-108 (binary relevance: the ideal ranking has the
    sum(relevance) relevant items first); 0 when no ranked item is relevant."""
    relevant = set(relevant_items)
    rel = np.array([1 if x in relevant else 0 for x in ranked_items])
    if rel.sum() == 0:
        return 0
    denominator = np.log2(np.arange(len(rel)) + 2)
    ideal = np.pad(np.ones(int(rel.sum())), (0, len(rel) - int(rel.sum())), mode="constant")
    return discounted_cumulative_gain(rel, denominator) / np.sum(np.divide(ideal, denominator))


def empirical_p_ndcg(ndcg, permutations, relevant_items):
    """post_utils.py:125-128 (the joblib loop, serially)."""
    results = [relevance_and_ndcg(p, relevant_items) for p in permutations]
    return (np.array(results) > ndcg).sum() / len(permutations)


# --------------------------------------------------------------------------------------
# parity measures
# --------------------------------------------------------------------------------------


def sin_principal_angle(V_ref, V_test):
    """sin of the largest principal angle between span(V_ref) and span(V_test):
    spectral norm of (I - Qr Qr^T) Qt with orthonormalised bases (float64)."""
    Qr, _ = np.linalg.qr(np.asarray(V_ref, dtype=np.float64))
    Qt, _ = np.linalg.qr(np.asarray(V_test, dtype=np.float64))
    R = Qt - Qr @ (Qr.T @ Qt)
    return float(np.linalg.norm(R, 2))


def unfolded_right_basis(YA, d):
    """Stack the per-layer embeddings (T, n, k) into the (T*n, d) right-singular basis
    (column scaling by sqrt(sigma) does not change the span)."""
    YA = np.asarray(YA, dtype=np.float64)
    return YA[:, :, :d].reshape(-1, d)


def percentile_cut(values, top_percent_keep):
    """preprocessing_utils.py:134 on a flat array of the nonzeros."""
    return np.percentile(values, 100 - top_percent_keep)


__all__ = [n for n in dir() if not n.startswith("_")]
