"""GPU parity of the BENCHMARK shapes in miniature, through bench.py's own code path.

bench.py's timed step is build_workload_layers() (K0 generate -> K1 load pass with the workload's transform) followed by
hot_path() (UASE at k = 24 with l = 48, tol 1e-5, the gap-aware stop at d = 4..24 -> distances -> rank + Borda).  Here the
same two functions run the C3 / C2 / C5 workloads at sizes the fp64 oracle finishes in seconds and every stage is compared
with the oracle (oracle/reference_port.py: dense float64 SVD, the reference's distance / rank / Borda restatements):

  * embeddings: sin(largest principal angle) vs the fp64 truth at EVERY swept d <= max(1e-4, 2 x the angle the reference's
    own fp32 ARPACK path (port.UASE on csc float32 layers) reaches on the same graphs);
  * scores: |ours - fp64| <= 1e-4 |fp64| + 1e-5 median, or at least as close as the reference's fp32 run (x2);
  * Borda: bit-exact when the rank kernel is fed the oracle's scores; end to end Spearman > 0.9999 against the fp64 table;
  * the solver's own record: converged status, per-d angle bounds, and the cross-solve angle bench.py reports.
"""
import os
import sys

import numpy as np
import pytest
import scipy.sparse
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from oracle import reference_port as port  # noqa: E402

ANGLE_TOL = 1e-4
SCORE_RTOL = 1e-4


def _spearman(a, b):
    ra = np.argsort(np.argsort(a, kind="stable"), kind="stable").astype(np.float64)
    rb = np.argsort(np.argsort(b, kind="stable"), kind="stable").astype(np.float64)
    return float(np.corrcoef(ra, rb)[0, 1])


_ORACLE = {}        # (workload, n) -> oracle results: both storages of one shape are checked against the same truth


def _oracle(key, graphs, k, seed):
    if key not in _ORACLE:
        S, V, YA = port.uase_dense_fp64(graphs, k)
        np.random.seed(seed)
        _, ref32 = port.UASE([scipy.sparse.csc_matrix(g) for g in graphs], k)
        _ORACLE.clear()
        _ORACLE[key] = (S, YA, ref32)
    return _ORACLE[key]


def _log(name, payload):
    d = os.environ.get("N2V2R_TEST_LOG")
    if d:
        import json
        os.makedirs(d, exist_ok=True)
        with open(os.path.join(d, name + ".json"), "w") as f:
            json.dump(payload, f, indent=1, default=float)


def _run_shape(gpu, workload, n, storage, l=48):
    import bench
    from node2vec2rank_b200 import _native as N
    from node2vec2rank_b200 import device as D
    w = dict(bench.WORKLOADS[workload])
    T = w["T"]
    layers, f32s, _ = bench.build_workload_layers(w, n, range(T), want_f32=True, storage=storage)
    graphs = [f.cpu().numpy() for f in f32s]
    for L, g in zip(layers, graphs):
        assert L.storage == storage and np.array_equal(L.numpy(), g)           # the stored layer IS the transformed graph
    info, emb, scores, bordas = bench.hot_path(layers, w, l)
    cross = bench.cross_solve_check(layers, w, l, emb, bordas, T, 0, None)
    # ---- oracle: fp64 truth and the reference's own fp32 path on the same transformed graphs ----
    k = max(bench.DIMS)
    S, YA, ref32 = _oracle((workload, n), graphs, k, bench.SOLVER_SEED)
    e = emb.cpu().numpy().astype(np.float64)
    assert np.allclose(info["sigma"], S, rtol=5e-6), (info["sigma"], S)
    worst, table = 0.0, {}
    for d in bench.DIMS:
        truth = port.unfolded_right_basis(YA, d)
        ours = port.sin_principal_angle(truth, port.unfolded_right_basis(e, d))
        ref = port.sin_principal_angle(truth, port.unfolded_right_basis(ref32, d))
        table[d] = dict(ours=ours, reference_fp32=ref, bound=info["angle_bounds"][d], cross=cross["sin_angle_per_d"][str(d)])
        worst = max(worst, ours)
    _log(f"bench_shape_{workload}_{n}_{storage}", dict(per_d=table, iterations=info["iterations"], status=info["converged_status"],
                                                        angle_bound=info["angle_bound"], cross=cross))
    for d in bench.DIMS:
        assert table[d]["ours"] <= max(ANGLE_TOL, 2 * table[d]["reference_fp32"]), (workload, storage, d, table[d], info["converged_status"])
    # the solver's own bound is an upper bound of the true angle (up to the fp32 noise of the residuals)
    assert info["converged"] in (1, 2) and info["status"] == N.OK
    assert worst <= max(2 * info["angle_bound"], ANGLE_TOL)
    assert set(info["angle_bounds"]) == set(bench.DIMS)
    # two independent solves agree to the tolerance at every swept d
    assert cross["max_sin_angle"] <= max(2 * ANGLE_TOL, 4 * worst), cross
    # ---- scores ----
    r64 = port.rank_columns(YA, bench.DIMS, bench.METRICS, w["strategy"])
    r32 = port.rank_columns(ref32, bench.DIMS, bench.METRICS, w["strategy"])
    got = scores.cpu().numpy()                                  # (C, cols, n)
    assert got.shape[0] == len(r64)
    for c, key in enumerate(sorted(r64, key=int)):
        ref64, ref_fp32 = r64[key][1], r32[key][1]
        atol = 1e-5 * np.median(ref64, axis=0)
        err = np.abs(got[c].T - ref64)
        referr = np.abs(ref_fp32 - ref64)
        bound = np.maximum(SCORE_RTOL * np.abs(ref64) + atol, 2 * referr.max(axis=0))
        assert np.all(err <= bound), (workload, key, float((err / bound).max()))
        # Borda: bit-exact on identical scores, and end to end against the fp64 table
        fed = D.rank_borda(torch.from_numpy(np.ascontiguousarray(ref64.T)).to(gpu))[0].cpu().numpy()
        truth_borda = port.borda_from_scores(ref64)
        assert np.array_equal(fed, truth_borda)
        ours_borda = bordas[c].cpu().numpy()
        assert ours_borda.sum() == ref64.shape[1] * n * (n + 1) // 2
        assert _spearman(ours_borda, truth_borda) > 0.9999, (workload, key, _spearman(ours_borda, truth_borda))
    return info, cross


@pytest.mark.parametrize("storage", ["rows", "tiles", "planes"])
def test_c3_in_miniature(gpu, storage):
    """BASELINE configs[2] (the configuration the metric is quoted on) at n = 6000: 8 weighted layers, K = 32, top 10 %
    kept (2.6 % dense, 16-bit weights -> 2 planes), one_vs_rest, d = 4..24, euclidean + cosine, Borda; the three layer
    storages a sparse layer can have (row lists are what the full-size benchmark runs on)."""
    info, cross = _run_shape(gpu, "c3", 6000, storage)
    assert info["iterations"] <= 80


def test_c2_in_miniature(gpu):
    """BASELINE configs[1] at n = 5000: 2 binarised dense layers, sequential."""
    _run_shape(gpu, "c2", 5000, "planes")


def test_c5_in_miniature(gpu):
    """BASELINE configs[4] at n = 3000 with 4 of the 16 seeds: every seed's run against the fp64 truth through the same
    functions, then Borda across runs against the oracle's stability_borda restatement fed the fp64 scores."""
    import bench
    from node2vec2rank_b200 import device as D
    w = dict(bench.WORKLOADS["c5"])
    n, T, seeds = 3000, w["T"], [0, 1, 2, 3]
    layers, f32s, _ = bench.build_workload_layers(w, n, range(T), want_f32=True, storage="planes")
    graphs = [f.cpu().numpy() for f in f32s]
    S, V, YA = port.uase_dense_fp64(graphs, max(bench.DIMS))
    np.random.seed(0)
    _, ref32 = port.UASE([scipy.sparse.csc_matrix(g) for g in graphs], max(bench.DIMS))
    r64 = port.rank_columns(YA, bench.DIMS, bench.METRICS, w["strategy"])
    per_seed = []
    for seed in seeds:
        info, emb, scores, _ = bench.hot_path(layers, w, 48, seed=seed)
        e = emb.cpu().numpy().astype(np.float64)
        for d in bench.DIMS:
            truth = port.unfolded_right_basis(YA, d)
            ours = port.sin_principal_angle(truth, port.unfolded_right_basis(e, d))
            ref = port.sin_principal_angle(truth, port.unfolded_right_basis(ref32, d))
            assert ours <= max(ANGLE_TOL, 2 * ref), (seed, d, ours, ref)
        per_seed.append(scores)
    sc = torch.stack(per_seed)                                   # (S, C, cols, n)
    Sn, C, cols, _ = sc.shape
    borda = D.rank_borda(sc.permute(1, 0, 2, 3).reshape(C, Sn * cols, n))[0].cpu().numpy()
    for c, key in enumerate(sorted(r64, key=int)):
        union64 = np.concatenate([r64[key][1]] * Sn, axis=1)     # in exact arithmetic every seed gives the same scores
        truth = port.borda_from_scores(union64)
        assert borda[c].sum() == Sn * cols * n * (n + 1) // 2
        assert _spearman(borda[c], truth) > 0.9999
