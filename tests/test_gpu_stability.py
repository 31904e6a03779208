"""GPU parity of the stability driver (multi-seed runs + Borda across runs, BASELINE configs[4]) against
the oracle's restatement of the reference procedure (README.md:54,173; model.py:98-201 per run)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import reference_port as port

DIMS = [2, 4, 6]
METRICS = ["euclidean", "cosine"]
SCORE_RTOL = 1e-4          # north_star: per-node distance scores within 1e-4 relative (gapped cuts only)


def _graphs(n=900, K=6, T=3, weighted=False):
    from node2vec2rank_b200.synthetic import block_matrices, sbm_layer_numpy
    Bs, changed = block_matrices(K, T, seed=3)
    return [sbm_layer_numpy(n, K, Bs[t], 7, t, weighted) for t in range(T)], changed


def _cfg(**kw):
    c = dict(embed_dimensions=DIMS, distance_metrics=METRICS, comp_strategy="one_vs_rest", seed=None, save_dir=None,
             verbose=-1, block_width=32, solver_tol=1e-6)
    c.update(kw)
    return c


def test_stability_run_vs_oracle(gpu):
    from node2vec2rank_b200.stability import StabilityRun
    graphs, _ = _graphs()
    n = graphs[0].shape[0]
    seeds = [3, 11, 42]
    run = StabilityRun(graphs=graphs, nodes=np.arange(n), config=_cfg(), seeds=seeds)
    agg = run.run()
    tables, borda = port.stability_borda(graphs, DIMS, METRICS, "one_vs_rest", seeds, dtype=np.float64)
    assert sorted(agg) == sorted(borda) == ["1", "2", "3"]
    for key in agg:
        names, ref = tables[key]
        ours = run.pairwise_ranks[key]
        assert list(ours.columns) == names                         # seeds outer, then dims, then metrics
        o = ours.to_numpy()
        atol = 1e-5 * np.median(np.abs(ref))
        assert np.all(np.abs(o - ref) <= SCORE_RTOL * np.abs(ref) + atol), np.abs(o - ref).max()
        # Borda: bit-exact on identical score vectors (ours fed through the oracle's rank rule) ...
        assert np.array_equal(agg[key]["borda_ranks"].to_numpy(), port.borda_from_scores(o))
        # ... and statistically identical to the oracle's own end-to-end table
        c = np.corrcoef(agg[key]["borda_ranks"].to_numpy(), borda[key])[0, 1]
        assert c > 0.9999, c
        assert int(agg[key]["borda_ranks"].sum()) == len(names) * n * (n + 1) // 2
        per = run.per_seed_aggregate_ranks[key].to_numpy()
        for j in range(len(seeds)):
            assert np.array_equal(per[:, j], port.borda_from_scores(o[:, j * 6:(j + 1) * 6]))
    ag = run.agreement()
    assert all(v.shape == (3,) and v.min() > 0.999 for v in ag.values())
    assert [i["seed"] for i in run.solver_infos] == seeds and all(i["converged"] for i in run.solver_infos)


def test_stability_seeds_change_start_block_not_result(gpu):
    """Different seeds start from different random blocks (embeddings differ in the last bits) but converge to the
    same subspace; the same seed reproduces bit for bit."""
    from node2vec2rank_b200 import device as D
    graphs, _ = _graphs(n=700, T=2)
    layers = [D.make_layer(torch.from_numpy(g).to(gpu)) for g in graphs]
    a = D.uase(layers, 6, l=32, seed=1, check_dims=DIMS, tol=1e-6)[0]
    b = D.uase(layers, 6, l=32, seed=1, check_dims=DIMS, tol=1e-6)[0]
    c = D.uase(layers, 6, l=32, seed=2, check_dims=DIMS, tol=1e-6)[0]
    assert torch.equal(a, b)
    assert not torch.equal(a, c)
    ea, ec = a.cpu().numpy().astype(np.float64), c.cpu().numpy().astype(np.float64)
    ang = port.sin_principal_angle(port.unfolded_right_basis(ea, 6), port.unfolded_right_basis(ec, 6))
    assert ang <= 1e-4, ang


def test_stability_errors(gpu):
    from node2vec2rank_b200.stability import StabilityRun
    graphs, _ = _graphs(n=300, T=2)
    with pytest.raises(ValueError):
        StabilityRun(graphs, np.arange(300), _cfg(), seeds=[])
    run = StabilityRun(graphs, np.arange(300), _cfg(), seeds=[1])
    with pytest.raises(NotImplementedError):
        run.run(method="kemeny")
    with pytest.raises(ValueError):
        run.agreement()


@pytest.mark.parametrize("n_seeds", [2, 4, 7])
def test_seed_batched_solves_are_bit_identical_to_unbatched(gpu, n_seeds):
    """device.uase_batched: concurrent solves that share every pass over a layer (three operands per pass, csrc/batch.cu)
    return, seed by seed, exactly the bits of device.uase run alone -- including a non-symmetric and a two-plane layer."""
    import torch
    from node2vec2rank_b200 import device as D
    from node2vec2rank_b200.synthetic import block_matrices, sbm_layer_numpy
    n, K, T, k = 1800, 6, 3, 12
    Bs, _ = block_matrices(K, T, seed=2)
    graphs = [sbm_layer_numpy(n, K, Bs[t], 5, t, weighted=(t == 1)) for t in range(T)]
    graphs[2] = np.triu(graphs[2]) + 0.5 * np.tril(graphs[2])                 # not symmetric
    layers = [D.make_layer(torch.from_numpy(np.ascontiguousarray(g, dtype=np.float32)).to(gpu)) for g in graphs]
    assert [L.n_planes for L in layers] == [1, 2, 1] and not layers[2].symmetric
    seeds = list(range(11, 11 + n_seeds))
    solo = [D.uase(layers, k, l=48, tol=1e-6, max_iters=300, seed=s, check_dims=[2, 4, K, k]) for s in seeds]
    batched, stats = D.uase_batched(layers, k, seeds, l=48, tol=1e-6, max_iters=300, check_dims=[2, 4, K, k])
    assert len(batched) == n_seeds
    for (e1, s1, _, i1), (e2, s2, i2) in zip(solo, batched):
        assert torch.equal(e1.view(torch.int32), e2.view(torch.int32))
        assert np.array_equal(s1, s2) and i1["iterations"] == i2["iterations"] and i1["converged"] == i2["converged"]
    # fewer passes over the layers than the solves asked for: up to three operands per pass
    assert stats["unbatched_layer_streams"] == sum(i["layer_passes"] for _, _, _, i in solo)
    assert stats["layer_streams"] < stats["unbatched_layer_streams"]
    assert stats["layer_streams"] >= stats["unbatched_layer_streams"] / 3


def test_stability_run_with_and_without_seed_batching(gpu):
    import contextlib
    import io
    import torch
    from node2vec2rank_b200.stability import StabilityRun
    from node2vec2rank_b200.synthetic import block_matrices, sbm_layer_numpy
    n, K, T = 1200, 5, 3
    Bs, _ = block_matrices(K, T, seed=3)
    graphs = [sbm_layer_numpy(n, K, Bs[t], 9, t) for t in range(T)]
    out = {}
    for batch in (1, 16):
        cfg = dict(embed_dimensions=[2, 4, 6], distance_metrics=["euclidean", "cosine"], comp_strategy="sequential", seed=None,
                   save_dir=None, verbose=-1, seed_batch=batch)
        with contextlib.redirect_stdout(io.StringIO()):
            run = StabilityRun(graphs=graphs, nodes=np.arange(n), config=cfg, seeds=[1, 2, 3, 4, 5])
            agg = run.run()
        out[batch] = (agg, run.pairwise_ranks, run.batch_stats)
    assert out[16][2] and not out[1][2]
    for key in out[1][0]:
        assert np.array_equal(out[1][0][key].to_numpy(), out[16][0][key].to_numpy())
        assert np.array_equal(out[1][1][key].to_numpy(), out[16][1][key].to_numpy())
