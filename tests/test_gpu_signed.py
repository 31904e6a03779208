"""GPU: N2V2R.signed_ranks_transform (model.py:203-280, model_utils.py:7-19) against golden vectors generated from the
unmodified reference (oracle/make_golden_signed.py -> tests/golden/signed.npz): `<k>_signed.tsv`, `<k>_agg_signed.tsv`,
pairwise_signed_ranks, pairwise_signed_aggregate_ranks; the sign transfer runs on the device (K11b)."""
import contextlib
import io
import json
import os

import numpy as np
import pandas as pd
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def signed(golden_dir):
    z = np.load(os.path.join(golden_dir, "signed.npz"), allow_pickle=False)
    return {k: z[k] for k in z.files}


def test_signed_scores_kernel(gpu):
    from node2vec2rank_b200 import device as D
    rng = np.random.default_rng(3)
    n, cols = 1234, 7
    s = rng.normal(size=(cols, n))
    s[0, 5] = np.nan
    prior = rng.integers(-2, 3, size=n).astype(np.float64)
    prior[7] = np.nan
    got = D.signed_scores(torch.from_numpy(s).to(gpu), torch.from_numpy(prior).to(gpu)).cpu().numpy()
    want = np.where(prior > 0, s, -s)
    assert np.array_equal(got, want, equal_nan=True)
    b = rng.integers(1, 10 ** 9, size=(1, n))
    got = D.signed_scores(torch.from_numpy(b).to(gpu), torch.from_numpy(prior).to(gpu)).cpu().numpy()
    assert got.dtype == np.int64 and np.array_equal(got, np.where(prior > 0, b, -b))


def test_signed_ranks_transform_vs_reference_golden(gpu, demo, demo_graphs, signed, golden_dir, tmp_path):
    """The whole signed path of the reference's demo run: our model is handed the reference's own score table (so that the
    comparison is bit-exact and independent of the SVD's fp32 noise), then aggregate -> DeDi -> signed_ranks_transform."""
    from node2vec2rank_b200.model import N2V2R
    names = demo["file_cols"]
    cfg = dict(embed_dimensions=list(range(4, 25, 2)), distance_metrics=["euclidean", "cosine"], comp_strategy="sequential",
               seed=42, save_dir=str(tmp_path), verbose=1)
    with contextlib.redirect_stdout(io.StringIO()) as out:
        m = N2V2R(graphs=demo_graphs, nodes=names, config=cfg)
        with pytest.raises(ValueError, match="Prior signed ranks needed"):
            m.signed_ranks_transform()
        m.fit_transform_rank()
        # the reference's score table in the reference's node order
        m.pairwise_ranks = {"1": pd.DataFrame(demo["ref_scores"], index=demo["ref_nodes"], columns=demo["ref_columns"])}
        m.node_names = demo["ref_nodes"]
        agg = m.aggregate_transform()
        # DeDi in the reference's node order: the column sums are keyed by the model's own order
        where = {nm: i for i, nm in enumerate(names)}
        perm = np.asarray([where[nm] for nm in demo["ref_nodes"]])
        dedi = m.degree_difference_ranking()
        assert np.array_equal(dedi["1"]["DeDi"].to_numpy()[perm], demo["ref_dedi"])
        prior = pd.Series(dedi["1"]["DeDi"].to_numpy()[perm], index=demo["ref_nodes"])
        got = m.signed_ranks_transform(prior_signed_ranks=[prior])
    assert "Signed ranks transformation" in out.getvalue() and "Finished signed transformation" in out.getvalue()
    tie_free = all(len(np.unique(demo["ref_scores"][:, c])) == 1000 for c in range(22))
    assert list(got) == ["1"] and list(got["1"].columns) == list(signed["signed_columns"])
    assert np.array_equal(np.asarray([str(x) for x in got["1"].index]), signed["signed_index"])
    assert got["1"].to_numpy().dtype == np.float64 and np.array_equal(got["1"].to_numpy(), signed["signed_scores"])
    sa = m.pairwise_signed_aggregate_ranks["1"]
    assert list(sa.columns) == ["signed_agg_ranks"] and sa["signed_agg_ranks"].dtype == np.int64
    assert np.array_equal(np.asarray([str(x) for x in sa.index]), signed["signed_agg_index"])
    if tie_free:
        assert np.array_equal(sa["signed_agg_ranks"].to_numpy(), signed["signed_agg"])
    assert np.array_equal(np.abs(sa["signed_agg_ranks"].to_numpy()), agg["1"]["borda_ranks"].to_numpy())
    # files: names and the first bytes of the reference's own files
    ref_files = json.load(open(os.path.join(golden_dir, "signed_files.json")))
    files = sorted(os.listdir(m.save_dir))
    assert [f for f in files if "signed" in f] == [f for f in ref_files["files"] if "signed" in f] == ["1_agg_signed.tsv", "1_signed.tsv"]
    for fn in ("1_signed.tsv",) + (("1_agg_signed.tsv",) if tie_free else ()):
        with open(os.path.join(m.save_dir, fn)) as f:
            assert f.read(400) == ref_files["heads"][fn], fn
    # a user prior that lacks nodes and holds zeros: missing nodes are dropped, zero counts as negative
    prior2 = pd.Series(signed["prior_values"], index=signed["prior_index"])
    with contextlib.redirect_stdout(io.StringIO()):
        got2 = m.signed_ranks_transform(prior_signed_ranks=[prior2])
    assert np.array_equal(np.asarray([str(x) for x in got2["1"].index]), signed["signed2_index"])
    assert np.array_equal(got2["1"].to_numpy(), signed["signed2_scores"])
    if tie_free:
        assert np.array_equal(m.pairwise_signed_aggregate_ranks["1"]["signed_agg_ranks"].to_numpy(), signed["signed2_agg"])
