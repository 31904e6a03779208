"""GPU parity of the ranking-quality read-outs (K12, post_utils mirror) against the golden vectors of the
reference's post_utils.py:45-128 and the oracle restatement."""
import json
import os

import numpy as np
import pandas as pd
import pytest

pytestmark = pytest.mark.gpu

from oracle import reference_port as port

RTOL = 1e-12          # fp64 sums in a different (fixed) order than np.sum's pairwise order


def _cases(golden_dir):
    for c in json.load(open(os.path.join(golden_dir, "ndcg.json")))["cases"]:
        c["ranked"] = [f"g{i}" for i in c["ranked"]]
        c["permutations"] = [[f"g{i}" for i in p] for p in c["permutations"]]
        yield c


def test_ndcg_mirror_vs_reference_golden(gpu, golden_dir):
    from node2vec2rank_b200 import post_utils as pu
    for c in _cases(golden_dir):
        got = pu.relevance_and_ndcg(c["ranked"], c["relevant"])
        assert got == pytest.approx(c["ndcg"], rel=RTOL, abs=0)
        if c["ndcg"] == 0:
            assert got == 0 and isinstance(got, int)               # the reference returns the int 0
        rel = [1 if x in set(c["relevant"]) else 0 for x in c["ranked"]]
        assert pu.discounted_cumulative_gain(rel) == pytest.approx(c["dcg"], rel=RTOL, abs=0)
        assert pu.normalized_discounted_cumulative_gain(rel, binary_relevance=True) == pytest.approx(c["ndcg_from_vector"], rel=RTOL, abs=0)
        assert pu.parallel_empirical_p_ndcg(c["ndcg"], c["permutations"], c["relevant"]) == c["p_value"]
        med = float(np.median(c["perm_ndcg"]))
        assert pu.parallel_empirical_p_ndcg(med, c["permutations"], c["relevant"]) == c["p_value_vs_median"]


def test_ndcg_batch_and_edge_cases(gpu):
    from node2vec2rank_b200 import post_utils as pu
    rng = np.random.default_rng(0)
    n, R = 50000, 64                                                # BASELINE full node count
    mask = (rng.random(n) < 0.03).astype(np.uint8)
    perms = np.stack([rng.permutation(n) for _ in range(R)]).astype(np.int32)
    got = pu.ndcg_of_rankings(perms, mask)
    disc = 1.0 / np.log2(np.arange(n) + 2)
    ref = (mask[perms] * disc).sum(axis=1) / disc[: int(mask.sum())].sum()
    assert np.allclose(got, ref, rtol=RTOL, atol=0)
    ideal = np.argsort(-mask.astype(np.int64), kind="stable")
    assert pu.ndcg_of_rankings(ideal, mask)[0] == pytest.approx(1.0, rel=1e-15)
    assert pu.ndcg_of_rankings(perms[:2], np.zeros(n, dtype=np.uint8)).tolist() == [0.0, 0.0]
    # ids outside the item universe count as not relevant; duplicates are scored by position like the reference
    assert pu.ndcg_of_rankings(np.array([[5, -1, 0, 99]]), np.array([1, 0, 0, 0, 0, 1]))[0] == pytest.approx(
        (1 / np.log2(2) + 1 / np.log2(4)) / (1 / np.log2(2) + 1 / np.log2(3)), rel=RTOL)
    assert pu.relevance_and_ndcg(["a", "b", "a"], ["a"]) == pytest.approx(port.relevance_and_ndcg(["a", "b", "a"], ["a"]), rel=RTOL)
    assert pu.relevance_and_ndcg(["a", "b"], ["zzz"]) == 0
    with pytest.raises(NotImplementedError):
        pu.relevance_and_ndcg(["a", "b"], ["a"], binary_relevance=False)
    with pytest.raises(NotImplementedError):
        pu.discounted_cumulative_gain([3, 2, 0])


def test_borda_ndcg_quality_gate(gpu):
    """Planted-signal SBM: the Borda table's nDCG against the changed communities beats every random permutation."""
    import torch
    from node2vec2rank_b200 import post_utils as pu
    from node2vec2rank_b200.model import N2V2R
    from node2vec2rank_b200.synthetic import block_matrices, sbm_layer_numpy
    n, K = 1500, 6
    Bs, changed = block_matrices(K, 2, seed=2)
    graphs = [sbm_layer_numpy(n, K, Bs[t], 5, t) for t in range(2)]
    names = np.array([f"n{i}" for i in range(n)])
    m = N2V2R(graphs, names, dict(embed_dimensions=[2, 4, 6], distance_metrics=["euclidean", "cosine"], comp_strategy="sequential",
                                  seed=1, save_dir=None, verbose=-1))
    m.fit_transform_rank()
    agg = m.aggregate_transform()["1"]
    relevant = names[np.isin(np.arange(n) % K, changed[1])]
    value, p = pu.borda_ndcg(agg, relevant, n_permutations=200, seed=3)
    order = names[np.argsort(port.descending_positions(agg.iloc[:, 0].to_numpy()))].tolist()      # ties by node position
    assert value == pytest.approx(port.relevance_and_ndcg(order, relevant), rel=RTOL)
    assert value > 0.95 and p == 0.0
