"""Golden vectors of N2V2R.signed_ranks_transform (model.py:203-280, model_utils.py:7-19) from the UNMODIFIED reference.

TEST INFRASTRUCTURE ONLY.  Needs /root/reference (not present on the GPU box); writes tests/golden/signed.npz and
tests/golden/signed_files.json, which are committed.

    PYTHONHASHSEED=0 python oracle/make_golden_signed.py [--ref /root/reference]

Runs the reference's demo (configs/config_demo_adj.json) exactly like oracle/make_golden.py does -- DataLoader ->
N2V2R.fit_transform_rank -> aggregate_transform -> degree_difference_ranking -- and then signed_ranks_transform(), first with the
prior the model stored (DeDi) and then with a user-supplied prior that lacks some nodes (those are dropped, model_utils.py:11).
"""
from __future__ import annotations

import argparse
import contextlib
import io
import json
import os
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)

from oracle import reference_port as port  # noqa: E402
from oracle.make_golden import flatten, import_reference  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ref", default="/root/reference")
    ap.add_argument("--out", default=os.path.join(ROOT, "tests", "golden"))
    a = ap.parse_args()
    DataLoader, N2V2R, mu, pu = import_reference(a.ref)
    import pandas as pd
    with open(os.path.join(a.ref, "configs", "config_demo_adj.json")) as f:
        cfg = flatten(json.load(f))
    cfg["data_dir"] = os.path.join(a.ref, "data", "networks", "demo")
    cfg["save_dir"] = tempfile.mkdtemp(prefix="n2v2r_golden_signed_")
    with contextlib.redirect_stdout(io.StringIO()):
        dl = DataLoader(config=cfg)
        model = N2V2R(graphs=dl.get_graphs(), config=cfg, nodes=dl.get_nodes())
        ranks = model.fit_transform_rank()
        agg = model.aggregate_transform()
        dedi = model.degree_difference_ranking()
        signed = model.signed_ranks_transform()
    demo = np.load(os.path.join(a.out, "demo.npz"))
    nodes = np.asarray([str(x) for x in dl.get_nodes()])
    assert np.array_equal(nodes, demo["ref_nodes"]) and np.array_equal(ranks["1"].to_numpy(), demo["ref_scores"]), \
        "the demo run differs from tests/golden/demo.npz (run with PYTHONHASHSEED=0 like make_golden.py)"
    out = dict(
        signed_index=np.asarray([str(x) for x in signed["1"].index]),
        signed_columns=np.asarray(list(signed["1"].columns)),
        signed_scores=signed["1"].to_numpy(),
        signed_agg_index=np.asarray([str(x) for x in model.pairwise_signed_aggregate_ranks["1"].index]),
        signed_agg=model.pairwise_signed_aggregate_ranks["1"]["signed_agg_ranks"].to_numpy(),
    )
    # the port's one-line restatement agrees
    assert np.array_equal(port.signed_transform(ranks["1"].to_numpy()[:, 3], dedi["1"]["DeDi"].to_numpy()), out["signed_scores"][:, 3])
    files = sorted(os.listdir(model.save_dir))
    heads = {}
    for fn in files:
        if "signed" in fn:
            with open(os.path.join(model.save_dir, fn)) as g:
                heads[fn] = g.read(400)
    # a user prior that lacks every 7th node and has zeros (zero -> negative, model_utils.py:14-17)
    rng = np.random.default_rng(7)
    keep = np.arange(len(nodes)) % 7 != 3
    vals = rng.integers(-2, 3, size=int(keep.sum())).astype(np.float64)
    prior = pd.Series(vals, index=dl.get_nodes()[keep])
    with contextlib.redirect_stdout(io.StringIO()):
        signed2 = model.signed_ranks_transform(prior_signed_ranks=[prior])
    out.update(prior_index=np.asarray([str(x) for x in prior.index]), prior_values=vals,
               signed2_index=np.asarray([str(x) for x in signed2["1"].index]),
               signed2_scores=signed2["1"].to_numpy(),
               signed2_agg=model.pairwise_signed_aggregate_ranks["1"]["signed_agg_ranks"].to_numpy())
    np.savez_compressed(os.path.join(a.out, "signed.npz"), **out)
    with open(os.path.join(a.out, "signed_files.json"), "w") as f:
        json.dump(dict(files=files, heads=heads), f, indent=1)
    print("signed golden written:", out["signed_scores"].shape, out["signed2_scores"].shape, files)


if __name__ == "__main__":
    main()
