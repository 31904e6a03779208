"""Generate tests/golden/* by running the UNMODIFIED reference in the build container.

TEST INFRASTRUCTURE ONLY.  Needs /root/reference (not present on the GPU box); the
fixtures it writes are committed so nothing at test time reads the reference.

    PYTHONHASHSEED=0 python oracle/make_golden.py [--ref /root/reference]

The reference modules are imported as they lie (nothing is copied) behind two
``sys.modules`` stubs:
  * ``plotly`` / ``plotly.graph_objects`` -- empty modules (unused import at
    preprocessing_utils.py:5);
  * ``spectral_embedding`` -- a module exposing ``UASE`` = oracle.reference_port.UASE
    (the real package is an un-vendored, un-pinned git dependency, environment.yaml:19).
"""
from __future__ import annotations

import argparse
import io
import json
import os
import sys
import tempfile
import types
import contextlib

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)

from oracle import reference_port as port  # noqa: E402


def import_reference(ref):
    for name in ("plotly", "plotly.graph_objects"):
        sys.modules.setdefault(name, types.ModuleType(name))
    se = types.ModuleType("spectral_embedding")
    se.UASE = port.UASE
    sys.modules["spectral_embedding"] = se
    sys.path.insert(0, ref)
    from node2vec2rank.dataloader import DataLoader
    from node2vec2rank.model import N2V2R
    from node2vec2rank import model_utils, preprocessing_utils
    return DataLoader, N2V2R, model_utils, preprocessing_utils


def flatten(cfg):
    return {k: v for _, sec in cfg.items() for k, v in sec.items()}     # node2vec2rank.py:59-60


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ref", default="/root/reference")
    ap.add_argument("--out", default=os.path.join(ROOT, "tests", "golden"))
    a = ap.parse_args()
    os.makedirs(a.out, exist_ok=True)
    DataLoader, N2V2R, mu, pu = import_reference(a.ref)
    import pandas as pd

    # ---- 1. network_transform known answers + random cases -------
    rng = np.random.default_rng(20261104)
    cases = []
    A4 = np.array([[0, -3, 2, .5], [-3, 0, 1, 4], [2, 1, 0, -.2], [.5, 4, -.2, 0]], dtype=np.float64)
    flagsets = [
        dict(),
        dict(absolute=True, threshold=1),
        dict(threshold=1),
        dict(top_percent_keep=50),
        dict(binarize=True),
        dict(absolute=True, top_percent_keep=30, binarize=True),
        dict(absolute=True, top_percent_keep=75),
    ]
    for fl in flagsets:
        cases.append(("A4", A4, fl))
    R = rng.normal(size=(37, 37))
    R[rng.random(R.shape) < 0.3] = 0.0
    for fl in flagsets + [dict(top_percent_keep=10), dict(top_percent_keep=1, absolute=True),
                          dict(threshold=0.25, top_percent_keep=60, binarize=True)]:
        cases.append(("R37", R, fl))
    Rf = rng.random((29, 29)).astype(np.float32)          # fp32 input: lerp happens in fp32
    for fl in [dict(top_percent_keep=10), dict(top_percent_keep=33), dict(top_percent_keep=90, binarize=True)]:
        cases.append(("R29f32", Rf, fl))
    W = rng.normal(size=(7, 19))
    for on in ("columns", "rows"):
        for fl in [dict(), dict(absolute=True, top_percent_keep=40), dict(threshold=0.1, binarize=True)]:
            cases.append(("W7x19", W, dict(fl, project_unipartite_on=on)))
    nt = {}
    meta = []
    for idx, (name, M, fl) in enumerate(cases):
        before = M.copy()
        out = pu.network_transform(M, **fl)
        assert np.array_equal(before, M), "reference mutated its input"
        assert out.dtype == np.float32
        nt[f"in_{idx}"] = M
        nt[f"out_{idx}"] = out
        meta.append(dict(name=name, flags=fl))
        # the port must agree bit for bit
        assert np.array_equal(port.network_transform(M, **fl), out), (name, fl)
    np.savez_compressed(os.path.join(a.out, "network_transform.npz"), **nt)
    with open(os.path.join(a.out, "network_transform.json"), "w") as f:
        json.dump(meta, f, indent=1)

    # ---- 2. compute_pairwise_distances ------------------------------------------------
    E1 = rng.normal(size=(64, 24))
    E2 = E1 + 0.1 * rng.normal(size=(64, 24))
    E1[5] = 0.0                      # zero row -> cosine NaN (scipy 1.18)
    E2[9] = E1[9]                    # identical rows -> 0 distance
    E1[11] = 0.0
    E2[11] = 0.0
    dd = {"E1": E1, "E2": E2}
    import warnings
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        for metric in ("euclidean", "cosine", "correlation"):
            for d in (2, 4, 9, 24):
                ref = np.asarray(mu.compute_pairwise_distances(E1[:, :d], E2[:, :d], metric), dtype=np.float64)
                dd[f"{metric}_{d}"] = ref
                mine = port.compute_pairwise_distances_vec(E1[:, :d], E2[:, :d], metric)
                assert np.allclose(mine, ref, rtol=1e-12, atol=1e-15, equal_nan=True), (metric, d)
    np.savez_compressed(os.path.join(a.out, "distances.npz"), **dd)

    # ---- 3. Borda ----------------------------------------------------------------------
    n = 50
    names = [f"g{i}" for i in range(n)]
    S = rng.normal(size=(n, 7))              # tie-free scores
    lists = []
    for c in range(S.shape[1]):
        s = pd.Series(S[:, c], index=names)
        s.sort_values(ascending=False, inplace=True)
        lists.append(s.index.to_list())
    bd = mu.borda_aggregate_parallel(lists)
    bd = bd.loc[names, :]
    ref_borda = bd["borda_ranks"].to_numpy()
    assert np.array_equal(ref_borda, port.borda_from_scores(S))
    np.savez_compressed(os.path.join(a.out, "borda.npz"), scores=S, borda=ref_borda)

    # ---- 4. the demo run (configs/config_demo_adj.json on data/networks/demo) -----------
    with open(os.path.join(a.ref, "configs", "config_demo_adj.json")) as f:
        cfg = flatten(json.load(f))
    cfg["data_dir"] = os.path.join(a.ref, "data", "networks", "demo")
    tmp = tempfile.mkdtemp(prefix="n2v2r_golden_")
    cfg["save_dir"] = tmp
    with contextlib.redirect_stdout(io.StringIO()):
        dl = DataLoader(config=cfg)
        graphs, nodes = dl.get_graphs(), dl.get_nodes()
        row_nodes = dl.graphs  # noqa
        model = N2V2R(graphs=graphs, config=cfg, nodes=nodes)
        ranks = model.fit_transform_rank()
        agg = model.aggregate_transform()
        dedi = model.degree_difference_ranking()
    # raw files in *file* order (deterministic), packed to bits: 2 x 125 KB
    raw = [pd.read_csv(os.path.join(cfg["data_dir"], fn), index_col=0, header=0) for fn in cfg["graph_filenames"]]
    for r in raw:
        v = r.values
        assert set(np.unique(v)) <= {0, 1}
    comm = pd.read_csv(os.path.join(cfg["data_dir"], "comm_asiggnments.csv"), index_col=0)
    node_names = np.asarray([str(x) for x in nodes])
    out = dict(
        adj_bits_0=np.packbits(raw[0].values.astype(np.uint8)),
        adj_bits_1=np.packbits(raw[1].values.astype(np.uint8)),
        file_cols=np.asarray([str(c) for c in raw[0].columns]),
        file_rows=np.asarray([str(c) for c in raw[0].index]),
        communities=comm.iloc[:, 0].to_numpy(),
        community_index=np.asarray([str(x) for x in comm.index]),
        ref_nodes=node_names,                                  # column-node order of the reference run
        ref_embeddings=model.node_embeddings.astype(np.float32),   # fp32-precision values (model.py:53)
        ref_columns=np.asarray(list(ranks["1"].columns)),
        ref_scores=ranks["1"].to_numpy(),
        ref_borda=agg["1"]["borda_ranks"].to_numpy(),
        ref_dedi=dedi["1"]["DeDi"].to_numpy(),
        ref_absdedi=dedi["1"]["absDeDi"].to_numpy(),
    )
    # fp64 ground truth on the same graphs (column-node order of the reference run)
    S64, V64, YA64 = port.uase_dense_fp64(graphs, 24)
    out["fp64_sigma"] = S64
    out["fp64_embeddings"] = YA64
    r64 = port.rank_columns(YA64, cfg["embed_dimensions"], cfg["distance_metrics"], cfg["comp_strategy"])
    out["fp64_scores"] = r64["1"][1]
    # the reference's own fp32-vs-fp64 noise floor per swept d
    out["ref_sin_angle"] = np.asarray([
        port.sin_principal_angle(port.unfolded_right_basis(YA64, d),
                                 port.unfolded_right_basis(model.node_embeddings, d))
        for d in cfg["embed_dimensions"]])
    # recall@|community 0| as in the demo notebook (cell 12)
    np.savez_compressed(os.path.join(a.out, "demo.npz"), **out)
    files = sorted(os.listdir(model.save_dir))
    with open(os.path.join(a.out, "demo_files.json"), "w") as f:
        heads = {}
        for fn in files:
            with open(os.path.join(model.save_dir, fn)) as g:
                heads[fn] = g.read(400)
        json.dump(dict(files=files, heads=heads, config=cfg["embed_dimensions"]), f, indent=1)

    # ---- 5. loader smoke inputs (input/*.csv, *.edgelist): tiny, stored verbatim --------
    small = {}
    for fn in ("test1.csv", "test2.csv", "test1_named.edgelist", "test2_named.edgelist", "sbm1.txt", "sbm2.txt"):
        with open(os.path.join(a.ref, "input", fn)) as f:
            small[fn] = f.read()
    loader_cases = []
    for fns, sep, is_edge in ((("test1.csv", "test2.csv"), ",", False),
                              (("sbm1.txt", "sbm2.txt"), ",", False),
                              (("test1_named.edgelist", "test2_named.edgelist"), " ", True)):
        c = dict(graph_filenames=list(fns), data_dir=os.path.join(a.ref, "input"), separator=sep,
                 is_edge_list=is_edge, transpose=False, project_unipartite_on=None, threshold=None,
                 top_percent_keep=100, binarize=False, absolute=False)
        with contextlib.redirect_stdout(io.StringIO()):
            d = DataLoader(config=c)
        g = d.get_graphs()
        nn = [str(x) for x in d.get_nodes()]
        # key by node name: store as dict-of-dict edge weights over column nodes
        # rows were permuted independently by the reference; recover row labels via match on square graphs
        loader_cases.append(dict(files=list(fns), sep=sep, is_edge_list=is_edge, nodes=sorted(nn),
                                 colsum={nm: [float(np.sum(gr[:, j])) for gr in g] for j, nm in enumerate(nn)},
                                 shape=list(g[0].shape), dtype=str(g[0].dtype)))
    with open(os.path.join(a.out, "loader.json"), "w") as f:
        json.dump(dict(inputs=small, cases=loader_cases), f, indent=1)

    # recall numbers for the record
    k = int(np.sum(out["communities"] == 0))
    comm_of = dict(zip(out["community_index"], out["communities"]))
    top = node_names[np.argsort(-out["ref_borda"], kind="stable")[:k]]
    rec = np.mean([comm_of[t] == 0 for t in top])
    print(f"demo: n={len(node_names)} borda sum={out['ref_borda'].sum()} recall@{k}={rec:.4f}")
    print("ref sin angles fp32 vs fp64:", np.array2string(out["ref_sin_angle"], precision=2))
    print("golden written to", a.out)


if __name__ == "__main__":
    main()
