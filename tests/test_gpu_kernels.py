"""GPU parity tests: every call goes through the C ABI (libn2v2r_b200.so) and is checked against
the CPU oracle (oracle/reference_port.py) or the golden vectors generated from the reference."""
import json
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import reference_port as port


def _t(a, gpu):
    return torch.from_numpy(np.ascontiguousarray(a)).to(gpu)


# ---- K0 ---------------------------------------------------------------------------------------
@pytest.mark.parametrize("weighted", [False, True, 2])
def test_sbm_generator_matches_numpy_twin(gpu, weighted):
    from node2vec2rank_b200 import device as D
    from node2vec2rank_b200.synthetic import block_matrices, sbm_layer_numpy
    n, K = 517, 7
    Bs, _ = block_matrices(K, 2, seed=3)
    for t in range(2):
        got = D.sbm_generate(n, K, Bs[t], seed=11, layer=t, weighted=weighted).cpu().numpy()
        ref = sbm_layer_numpy(n, K, Bs[t], 11, t, weighted)
        assert np.array_equal(got, ref)
        assert np.array_equal(got, got.T) and not got.diagonal().any()


# ---- K1 -----------------------------------------------------------------------------------------
def test_network_transform_golden_bit_exact(gpu, golden_dir):
    from node2vec2rank_b200.preprocessing_utils import network_transform
    z = np.load(os.path.join(golden_dir, "network_transform.npz"))
    meta = json.load(open(os.path.join(golden_dir, "network_transform.json")))
    for i, m in enumerate(meta):
        x = z[f"in_{i}"]
        before = x.copy()
        fl = dict(m["flags"])
        projected = x.shape[0] != x.shape[1]
        out = network_transform(x, **fl)
        assert np.array_equal(before, x)
        assert out.dtype == np.float32 and out.shape == z[f"out_{i}"].shape
        if projected:       # fp64 dgemm vs split-bf16 tensor-core GEMM: fp32-level agreement, same support
            ref = z[f"out_{i}"]
            assert np.allclose(out, ref, rtol=2e-6, atol=1e-6), (i, m)
        else:
            assert np.array_equal(out, z[f"out_{i}"]), (i, m)


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("keep", [100, 50, 10, 1, 0.5])
def test_percentile_cut_matches_numpy(gpu, dtype, keep):
    from node2vec2rank_b200 import device as D
    rng = np.random.default_rng(5)
    x = rng.normal(size=(301, 301)).astype(dtype)
    x[rng.random(x.shape) < 0.4] = 0
    x[:20] = np.round(x[:20], 1)                     # ties
    for absolute, thr in [(False, None), (True, None), (True, 0.3), (False, 0.1)]:
        y = np.abs(x) if absolute else x.copy()
        if thr is not None:
            y[y < thr] = 0
        ref = np.percentile(y[y.nonzero()], 100 - keep)
        got = D.percentile_cut(_t(x, gpu), 100 - keep, absolute, thr)
        assert got == ref and type(got) == type(ref), (dtype, keep, absolute, thr, got, ref)


@pytest.mark.parametrize("shape", [(300, 1200), (1111, 257), (64, 4100)])
def test_projection_vs_fp64(gpu, shape):
    """K2 (preprocessing_utils.py:16-32): W^T W / W W^T.  The reference multiplies in float64 and casts to float32;
    the tensor-core path carries 24-bit operands and accumulates in fp32 (truncating adder, 512-long chains):
    tolerance 4e-6 of the largest entry, output exactly symmetric."""
    from node2vec2rank_b200.preprocessing_utils import bipartite_to_unipartite_projection
    rng = np.random.default_rng(shape[0])
    W = rng.normal(size=shape) + 0.5
    for on in ("columns", "rows", "Rows"):
        out = bipartite_to_unipartite_projection(W, on)
        ref = port.bipartite_to_unipartite_projection(W, on)
        assert out.shape == ref.shape and out.dtype == np.float32
        assert np.abs(out - ref).max() <= 4e-6 * np.abs(ref).max()
        # per entry: |out - ref| <= c eps32 sum_k |w||w| with c = 64 (measured 17-21: the tensor core's fp32 adder truncates, over
        # chains of 512 products; the worst case of a plain fp32 dot product is K eps32, the float32 cast of the reference's
        # float64 product alone costs eps32/2 |ref|)
        Wa = np.abs(W)
        bound = (Wa.T @ Wa) if on.casefold() == "columns" else (Wa @ Wa.T)
        ratio = float((np.abs(out.astype(np.float64) - ref) / (np.finfo(np.float32).eps * bound)).max())
        assert ratio <= 64.0, (shape, on, ratio)
        assert np.array_equal(out, out.T)
    with pytest.raises(ValueError):
        bipartite_to_unipartite_projection(W, "diagonal")
    with pytest.raises(AssertionError):
        bipartite_to_unipartite_projection(np.eye(5), "rows")


def test_transform_planes_colsum_and_symmetry(gpu):
    from node2vec2rank_b200 import device as D
    rng = np.random.default_rng(9)
    n = 333
    x = rng.random((n, n)).astype(np.float32)
    x = (x + x.T) / 2
    x[x < 0.5] = 0
    L = D.make_layer(_t(x, gpu))
    assert L.n_planes == 3 and L.symmetric and L.planes_t is None
    assert np.array_equal(L.numpy(), x)
    assert np.allclose(L.colsum.cpu().numpy(), x.astype(np.float64).sum(0), rtol=1e-12)
    b = (x != 0).astype(np.float32)
    Lb = D.make_layer(_t(b, gpu))
    assert Lb.n_planes == 1 and np.array_equal(Lb.numpy(), b)
    a = np.triu(x)
    La = D.make_layer(_t(a, gpu))
    assert not La.symmetric and La.planes_t is not None
    assert np.array_equal(La.planes_t[:, :, :n].float().sum(0).cpu().numpy(), a.T)


@pytest.mark.parametrize("n", [64, 127, 200, 1000, 1284, 2049])
def test_symmetry_and_transpose_tiles(gpu, n):
    """64 x 64-tile kernels at sizes around the tile edges: a single asymmetric entry anywhere (corners, last partial
    tile) must be seen, garbage in the column padding must not matter, and the transpose must match torch bit for bit."""
    import ctypes as C
    from node2vec2rank_b200 import _native as N
    from node2vec2rank_b200 import device as D
    lib = N.load()
    rng = np.random.default_rng(n)
    x = rng.random((n, n)).astype(np.float32)
    x = np.where(rng.random((n, n)) < 0.3, x, 0).astype(np.float32)
    x = np.maximum(x, x.T)
    _, planes, ld, _ = D.transform(_t(x, gpu), n_planes=2)
    sym = C.c_int(0)

    def is_sym(p):
        N.check(lib.n2v2r_is_symmetric(N.stream_ptr(), N.ptr(p), 2, n, ld, C.byref(sym)))
        return bool(sym.value)

    assert is_sym(planes)
    if ld > n:                                     # padding columns are not part of the matrix
        dirty = planes.clone()
        dirty[:, :, n:] = 1.0
        assert is_sym(dirty)
    for (i, j) in [(0, n - 1), (n - 1, 0), (n // 2, n // 2 - 1), (n - 1, n - 2), (63 % n, 64 % n), (1, 0)]:
        if i == j:
            continue
        bad = planes.clone()
        bad[1, i, j] = bad[1, i, j] + 1.0           # second plane only
        assert not is_sym(bad), (i, j)
    a = np.triu(x) + np.tril(rng.random((n, n)).astype(np.float32), -1)
    _, pa, ld, _ = D.transform(_t(a, gpu), n_planes=2)
    pt = torch.full_like(pa, 7.0)
    N.check(lib.n2v2r_transpose_planes(N.stream_ptr(), N.ptr(pa), 2, n, ld, N.ptr(pt)))
    assert torch.equal(pt[:, :, :n], pa[:, :, :n].transpose(1, 2))
    assert float(pt[:, :, n:].abs().sum()) == 0.0  # padding written as zeros


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_transform_fast_and_general_paths_agree(gpu, dtype):
    """The 4-column cp.async kernel (cols % 4 == 0) against the 2-column kernel (odd leading dimension) and NumPy, for every
    option combination, ragged row counts included."""
    from node2vec2rank_b200 import device as D
    rng = np.random.default_rng(3)
    rows, cols = 1237, 2052
    x = (rng.standard_normal((rows, cols)) * (rng.random((rows, cols)) < 0.4)).astype(dtype)
    wide = torch.zeros((rows, cols + 1), dtype=_t(x, gpu).dtype, device=gpu)       # odd ld: general path
    wide[:, :cols] = _t(x, gpu)
    for absolute in (False, True):
        for thr in (None, 0.3):
            for cut in (None, 0.8):
                for binarize in (False, True):
                    y = np.abs(x) if absolute else x.copy()
                    if thr is not None:
                        y[y < dtype(thr)] = 0
                    if cut is not None:
                        y[y < dtype(cut)] = 0
                    if binarize:
                        y[y != 0] = 1
                    want = y.astype(np.float32)
                    kw = dict(absolute=absolute, threshold=thr, cut=cut, binarize=binarize)
                    f, p, ld, cs = D.transform(_t(x, gpu), want_f32=True, n_planes=3, want_colsum=True, **kw)
                    g, q, _, cg = D.transform(wide[:, :cols], want_f32=True, n_planes=3, want_colsum=True, **kw)
                    assert np.array_equal(f.cpu().numpy(), want) and torch.equal(f, g)
                    assert torch.equal(p, q) and np.array_equal(p[:, :, :cols].float().sum(0).cpu().numpy(), want)
                    assert float(p[:, :, cols:].abs().sum()) == 0.0
                    assert np.allclose(cs.cpu().numpy(), want.astype(np.float64).sum(0), rtol=1e-12, atol=1e-9)
                    assert np.allclose(cg.cpu().numpy(), cs.cpu().numpy(), rtol=1e-12, atol=1e-9)


# ---- K3/K4 ----------------------------------------------------------------------------------------
@pytest.mark.parametrize("n,l,planes", [(1000, 64, 1), (777, 32, 3), (2048, 48, 2), (4500, 64, 1), (300, 16, 3), (1030, 80, 1)])
def test_layer_apply_tcgen05_vs_fp64(gpu, n, l, planes):
    from node2vec2rank_b200 import _native as N
    from node2vec2rank_b200 import device as D
    rng = np.random.default_rng(n + l)
    if planes == 1:
        a = (rng.random((n, n)) < 0.2).astype(np.float32)
    elif planes == 2:
        a = (rng.integers(0, 65536, size=(n, n)).astype(np.float32) / 65536.0) * (rng.random((n, n)) < 0.3)
        a = a.astype(np.float32)
    else:
        a = (rng.random((n, n)).astype(np.float32)) * (rng.random((n, n)) < 0.3).astype(np.float32)
    L = D.make_layer(_t(a, gpu), storage="planes")          # the dense-plane kernels (tiles / e4m3 have their own tests)
    assert L.n_planes == planes
    x = rng.normal(size=(n, l)).astype(np.float32)
    xd = _t(x, gpu)
    for transpose in (False, True):
        at = (a.T if transpose else a).astype(np.float64)
        ref = at @ x.astype(np.float64)
        scale = (np.abs(at) @ np.abs(x).astype(np.float64)).max()      # fp32-level: error <= c eps32 sum |a||x|
        simt = D.layer_apply(L, xd, transpose, algo=N.ALGO_SIMT).cpu().numpy()
        assert np.abs(simt - ref).max() <= 1e-6 * scale
        for k_chunk in (0, 512):
            got = D.layer_apply(L, xd, transpose, algo=N.ALGO_TCGEN05, k_chunk=k_chunk).cpu().numpy()
            err = np.abs(got - ref).max() / scale
            assert err <= 1e-6, (n, l, planes, transpose, k_chunk, err)   # operands carry 24 bits, fp32 accumulate


# ---- K9-K11 ----------------------------------------------------------------------------------------
def test_distances_golden(gpu, golden_dir):
    from node2vec2rank_b200.model_utils import compute_pairwise_distances
    z = np.load(os.path.join(golden_dir, "distances.npz"))
    for metric in ("euclidean", "cosine", "correlation"):
        for d in (2, 4, 9, 24):
            got = np.asarray(compute_pairwise_distances(z["E1"][:, :d], z["E2"][:, :d], metric))
            ref = z[f"{metric}_{d}"]
            assert np.array_equal(np.isnan(got), np.isnan(ref))
            assert np.allclose(got, ref, rtol=1e-12, atol=1e-15, equal_nan=True), (metric, d)
    with pytest.raises(NotImplementedError):
        compute_pairwise_distances(z["E1"], z["E2"], "manhattan")


@pytest.mark.parametrize("strategy", ["sequential", "one_vs_before", "one_vs_rest"])
def test_distance_strategies_vs_oracle(gpu, strategy):
    from node2vec2rank_b200 import _native as N
    from node2vec2rank_b200 import device as D
    rng = np.random.default_rng(2)
    T, n, k = 4, 257, 24
    emb = rng.normal(size=(T, n, k)).astype(np.float32)
    emb[2, 7] = 0
    dims = [1, 2, 4, 13, 24]
    mets = ["euclidean", "cosine", "correlation"]
    names, cd, cm = D.column_specs(dims, mets)
    got = D.distances(_t(emb, gpu), cd, cm, N.STRATEGIES[strategy]).cpu().numpy()
    ref = port.rank_columns(emb.astype(np.float64), dims, mets, strategy)
    assert got.shape[0] == len(ref)
    for c, key in enumerate(sorted(ref, key=int)):
        rn, rs = ref[key]
        assert rn == names
        assert np.allclose(got[c].T, rs, rtol=1e-11, atol=1e-14, equal_nan=True), (strategy, key)


def test_rank_borda_golden_and_ties(gpu, golden_dir):
    from node2vec2rank_b200 import device as D
    z = np.load(os.path.join(golden_dir, "borda.npz"))
    b, _ = D.rank_borda(_t(z["scores"].T, gpu))
    assert np.array_equal(b.cpu().numpy(), z["borda"])
    rng = np.random.default_rng(4)
    for n, cols in [(1, 3), (2, 1), (1000, 22), (5003, 7), (70000, 3)]:
        s = rng.normal(size=(n, cols))
        s[rng.random(s.shape) < 0.2] = 0.5          # heavy ties
        s[rng.random(s.shape) < 0.05] = np.nan
        s[rng.random(s.shape) < 0.02] = -0.0
        s[rng.random(s.shape) < 0.02] = 0.0
        s[rng.random(s.shape) < 0.01] = np.inf
        s[rng.random(s.shape) < 0.01] = -np.inf
        borda, pos = D.rank_borda(_t(s.T, gpu), want_positions=True)
        ref_pos = np.stack([port.descending_positions(s[:, c]) for c in range(cols)])
        assert np.array_equal(pos.cpu().numpy(), ref_pos), (n, cols)
        assert np.array_equal(borda.cpu().numpy(), port.borda_from_scores(s))
        assert borda.sum().item() == cols * n * (n + 1) // 2


def test_rank_borda_grouped(gpu):
    from node2vec2rank_b200 import device as D
    rng = np.random.default_rng(6)
    G, cols, n = 5, 22, 3001
    s = rng.normal(size=(G, cols, n))
    s[rng.random(s.shape) < 0.1] = 1.25
    borda, pos = D.rank_borda(_t(s, gpu), want_positions=True)
    assert borda.shape == (G, n) and pos.shape == (G, cols, n)
    for g in range(G):
        assert np.array_equal(borda[g].cpu().numpy(), port.borda_from_scores(s[g].T))
    assert np.array_equal(pos[3, 7].cpu().numpy(), port.descending_positions(s[3, 7]))


def test_borda_aggregate_parallel_mirror(gpu):
    from node2vec2rank_b200.model_utils import borda_aggregate_parallel
    rng = np.random.default_rng(8)
    names = [f"g{i}" for i in range(300)]
    lists = [list(rng.permutation(names)) for _ in range(5)]
    got = borda_aggregate_parallel(lists)
    idx, ref = port.borda_aggregate(lists, faithful=False)
    assert list(got.index) == idx and list(got.columns) == ["borda_ranks"]
    assert got["borda_ranks"].dtype == np.int64 and np.array_equal(got["borda_ranks"].to_numpy(), ref)
