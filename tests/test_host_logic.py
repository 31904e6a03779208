"""CPU: host-side logic of the mirror (no device work)."""
import io
import os

import numpy as np
import pandas as pd
import pytest


def test_column_specs_order_and_skip():
    from node2vec2rank_b200 import device as D
    names, dims, mets = D.column_specs([1, 2, 4], ["euclidean", "cosine"])
    assert names == ["dim-1_distance-euclidean", "dim-2_distance-euclidean", "dim-2_distance-cosine",
                     "dim-4_distance-euclidean", "dim-4_distance-cosine"]          # dims outer, metrics inner, cosine skipped at d=1
    assert list(dims) == [1, 2, 2, 4, 4] and list(mets) == [0, 0, 1, 0, 1]
    with pytest.raises(NotImplementedError):
        D.column_specs([2], ["manhattan"])


def test_rows_panel_cols():
    """Panel width of a row-list layer by the solver block it is built for: the widest multiple of 64 (<= 512) whose stage
    (X panel + about 96 bytes of list per column) fits twice in the shared memory the product kernel can use."""
    from node2vec2rank_b200 import device as D
    assert [D.rows_panel_cols(l) for l in (None, 0, 16, 32, 48, 64, 80)] == [256, 256, 512, 512, 384, 320, 256]
    for l in (16, 32, 48, 64, 80):
        W = D.rows_panel_cols(l)
        assert W % 64 == 0 and 2 * W * (4 * l + 96) <= 2 * D.ROWS_STAGE_BYTES


def test_default_block_width():
    from node2vec2rank_b200 import device as D
    assert D.default_block_width(24, 50000) == 48
    assert D.default_block_width(24, 50000, "rows") == 32        # the row-list product is linear in l: 8 guard vectors
    assert D.default_block_width(70, 10**6, "rows") == 80 and D.default_block_width(24, 50000, "tiles") == 48
    assert D.default_block_width(8, 1000) == 32
    assert D.default_block_width(70, 10**6) == 80
    assert D.default_block_width(4, 20) == 16
    for k in (1, 4, 24, 56):
        l = D.default_block_width(k, 10**5)
        assert l % 16 == 0 and k <= l <= 80


def test_layer_partition():
    from node2vec2rank_b200.dist import layer_partition
    for T in (1, 2, 3, 8, 13):
        for world in (1, 2, 3, 4, 8):
            got = [list(layer_partition(T, world, r)) for r in range(world)]
            assert sum(got, []) == list(range(T))
            sizes = [len(g) for g in got]
            assert max(sizes) - min(sizes) <= 1 and sizes == sorted(sizes, reverse=True)
    with pytest.raises(ValueError):
        layer_partition(4, 2, 2)


def test_synthetic_twin_properties():
    from node2vec2rank_b200.synthetic import block_matrices, sbm_layer_numpy
    Bs, changed = block_matrices(6, 3, seed=4)
    assert len(Bs) == 3 and changed[0] == [] and len(changed[1]) == 3
    for B in Bs:
        assert B.dtype == np.float32 and np.array_equal(B, B.T) and B.min() >= 0.02 and B.max() <= 0.5
    diff = np.argwhere(Bs[0] != Bs[1])
    assert set(diff[:, 0]) | set(diff[:, 1]) >= set(changed[1])
    a = sbm_layer_numpy(211, 6, Bs[0], 9, 0)
    assert np.array_equal(a, a.T) and not a.diagonal().any() and set(np.unique(a)) <= {0.0, 1.0}
    assert np.array_equal(a, sbm_layer_numpy(211, 6, Bs[0], 9, 0))           # pure function of (seed, layer, i, j)
    assert not np.array_equal(a, sbm_layer_numpy(211, 6, Bs[0], 9, 1))
    assert np.array_equal(a[[5, 77]], sbm_layer_numpy(211, 6, Bs[0], 9, 0, rows=[5, 77]))
    w = sbm_layer_numpy(211, 6, Bs[0], 9, 0, weighted=True)
    assert np.array_equal(w != 0, a != 0) and w.max() <= 1.0 and w[w != 0].min() >= 2.0 ** -16
    assert np.all((w * 65536) % 1 == 0)                                       # 16-bit weights: exact in two bf16 planes
    dens = a.mean()
    assert abs(dens - np.mean(Bs[0])) < 0.03


def test_match_networks_deterministic_order():
    from node2vec2rank_b200.preprocessing_utils import match_networks
    a = pd.DataFrame(np.arange(16.0).reshape(4, 4), index=list("abcd"), columns=list("abcd"))
    b = pd.DataFrame(np.arange(16.0).reshape(4, 4), index=list("dcbx"), columns=list("xbcd"))
    ga, gb = match_networks([a, b])
    assert list(ga.index) == ["b", "c", "d"] and list(ga.columns) == ["b", "c", "d"]     # first graph's order
    assert list(gb.index) == list(ga.index) and list(gb.columns) == list(ga.columns)
    assert gb.loc["d", "b"] == b.loc["d", "b"]


def test_percentile_lerp_matches_numpy_host_side(monkeypatch):
    """device.percentile_cut does NumPy's 'linear' interpolation on the host from the two order statistics the
    device returns; feed it order statistics computed on the CPU and compare with np.percentile."""
    from node2vec2rank_b200 import device as D
    rng = np.random.default_rng(3)
    for dtype in (np.float32, np.float64):
        vals = np.sort(rng.normal(size=1001).astype(dtype))

        class FakeT:
            def __init__(self):
                import torch
                self.dtype = torch.float32 if dtype == np.float32 else torch.float64

        for q in (0, 10, 33.3, 50, 90, 99.5, 100):
            def fake_select(x, frac, absolute=False, threshold=None):
                pos = (np.float32(len(vals) - 1) * np.float32(frac)) if dtype == np.float32 else (len(vals) - 1) * frac
                lo = int(np.floor(pos))
                hi = min(lo + 1, len(vals) - 1)
                return len(vals), float(vals[lo]), float(vals[hi]), float(pos)
            monkeypatch.setattr(D, "select_nonzero", fake_select)
            got = D.percentile_cut(FakeT(), q)
            ref = np.percentile(vals, q)
            assert got == ref and got.dtype == ref.dtype, (dtype, q, got, ref)


def test_signed_transform_single_mirror():
    from node2vec2rank_b200.model_utils import signed_transform_single
    ranks = pd.Series([3.0, 1.0, 2.0, 5.0], index=["a", "b", "c", "z"])
    prior = pd.Series([0.5, -1.0, 0.0], index=["a", "b", "c"])
    out = signed_transform_single(ranks, prior)
    assert list(out.index) == ["a", "b", "c"] and list(out.values) == [3.0, -1.0, -2.0]      # <= 0 negates; 'z' dropped


def test_n2v2r_constructor_contract(tmp_path, capsys):
    from node2vec2rank_b200.model import N2V2R
    cfg = dict(embed_dimensions=[2, 4], distance_metrics=["euclidean"], comp_strategy="sequential", seed=42,
               save_dir=str(tmp_path), verbose=0)
    m = N2V2R([np.eye(5, dtype=np.float32)] * 2, list("abcde"), cfg)          # positional order (graphs, nodes, config)
    assert m.max_embed_dim == 4 and m.num_graphs == 2 and m.node_embeddings is None
    assert os.path.isdir(m.save_dir) and os.path.exists(os.path.join(m.save_dir, "config.json"))
    assert os.path.basename(m.save_dir).count("_") == 5                        # mm_dd_YYYY_HH_MM_SS
    assert m.save_dir in capsys.readouterr().out
    with pytest.raises(ValueError, match="No n2v2r embeddings found"):
        m.aggregate_transform()
    with pytest.raises(ValueError, match="Prior signed ranks needed"):
        m.signed_ranks_transform()
    with pytest.raises(KeyError):
        N2V2R([np.eye(3)], ["a"], {"embed_dimensions": [2]})


def test_cli_config_flattening(tmp_path):
    from node2vec2rank_b200.cli import load_config
    p = tmp_path / "c.json"
    p.write_text('{"data_io": {"save_dir": "x", "seed": 1}, "fitting": {"seed": 7, "embed_dimensions": [2]}}')
    cfg = load_config(str(p))
    assert cfg == {"save_dir": "x", "seed": 7, "embed_dimensions": [2]}       # section names irrelevant, later keys win
