"""GPU: one-plane e4m3 storage of binarised layers (N2V2R_STORE_E4M3, 1 byte per entry) and the kind::f8f6f4 product kernel
(k3_fp8.cu) whose fp32 operand is split into six scaled e4m3 terms."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _binary(n, density, seed, symmetric=True):
    rng = np.random.default_rng(seed)
    a = (rng.random((n, n)) < density).astype(np.float32)
    if symmetric:
        a = np.triu(a, 1)
        a = a + a.T
    return a


@pytest.mark.parametrize("n,symmetric", [(1000, True), (777, False), (130, True)])
def test_e4m3_layer_round_trip(gpu, n, symmetric):
    from node2vec2rank_b200 import device as D
    a = _binary(n, 0.3, n, symmetric)
    L = D.make_layer(torch.from_numpy(a).to(gpu), storage="e4m3")
    assert L.storage == "e4m3" and L.n_planes == 1 and L.ld % 128 == 0 and L.planes.dtype == torch.uint8
    assert L.symmetric == symmetric and (L.planes_t is None) == symmetric
    assert np.array_equal(L.numpy(), a)
    assert L.nbytes == n * L.ld * (1 if symmetric else 2)
    if not symmetric:
        assert np.array_equal(L.planes_t[0, :, :n].view(torch.float8_e4m3fn).float().cpu().numpy(), a.T)
    assert not bool(L.planes[0, :, n:].any())                      # padding columns are zero
    # e4m3 holds 0.5, 1.5, 448 ... but not 0.3
    ok = D.make_layer(torch.from_numpy(a * np.float32(1.5)).to(gpu), storage="e4m3")
    assert np.array_equal(ok.numpy(), a * np.float32(1.5))
    with pytest.raises(ValueError, match="e4m3"):
        D.make_layer(torch.from_numpy(a * np.float32(0.3)).to(gpu), storage="e4m3")


@pytest.mark.parametrize("n,l", [(1000, 48), (2100, 32), (640, 80), (3000, 64), (700, 16)])
@pytest.mark.parametrize("symmetric", [True, False])
def test_e4m3_product_vs_fp64(gpu, n, l, symmetric):
    """A x and A^T x on the e4m3 layer against the float64 product; the bf16-plane kernel on the same layer for scale."""
    from node2vec2rank_b200 import device as D
    a = _binary(n, 0.25, 3 + n + l, symmetric)
    x32 = torch.from_numpy(a).to(gpu)
    L8 = D.make_layer(x32, storage="e4m3")
    Lp = D.make_layer(x32, storage="planes")
    g = torch.Generator(device="cpu").manual_seed(n + l)
    x = torch.randn((n, l), generator=g, dtype=torch.float32)
    # columns of very different scale, a zero column, and entries far below their column's maximum
    x[:, 0] *= 1e-12
    x[:, 1] *= 1e9
    x[:, 2] = 0.0
    x[::3, 3] *= 1e-7
    x = x.to(gpu)
    A64 = torch.from_numpy(a.astype(np.float64)).to(gpu)
    for tr in (False, True):
        M = A64.T if tr else A64
        ref = M @ x.double()
        scale = (M @ x.double().abs()).clamp_min(1e-300)
        y8 = D.layer_apply(L8, x, transpose=tr)
        yp = D.layer_apply(Lp, x, transpose=tr)
        e8 = float(((y8.double() - ref).abs() / scale).max())
        ep = float(((yp.double() - ref).abs() / scale).max())
        # error relative to sum |a||x| per entry: fp32 accumulation + the 24-bit operand split
        colmax = x.abs().max(dim=0).values.double().clamp_min(1e-300)
        rel_to_colmax = float(((y8.double() - ref).abs() / ((M @ torch.ones_like(x, dtype=torch.float64)) * colmax + 1e-300)).max())
        assert e8 <= max(2e-6, 4 * ep), (tr, e8, ep)
        assert rel_to_colmax <= 2e-7, (tr, rel_to_colmax)
        assert torch.equal(y8[:, 2], torch.zeros_like(y8[:, 2]))
    y8b = D.layer_apply(L8, x)
    assert torch.equal(y8b.view(torch.int32), D.layer_apply(L8, x).view(torch.int32))     # deterministic


def test_e4m3_solver_vs_fp64_oracle(gpu):
    from oracle import reference_port as port
    from node2vec2rank_b200 import device as D
    from node2vec2rank_b200.synthetic import block_matrices, sbm_layer_numpy
    n, K, T, k = 2000, 6, 3, 12
    Bs, _ = block_matrices(K, T, seed=2)
    graphs = [sbm_layer_numpy(n, K, Bs[t], 5, t) for t in range(T)]
    S, V, YA = port.uase_dense_fp64(graphs, k)
    res = {}
    for storage in ("planes", "e4m3"):
        layers = [D.make_layer(torch.from_numpy(g).to(gpu), storage=storage) for g in graphs]
        emb, sigma, _, info = D.uase(layers, k, l=48, tol=1e-6, max_iters=500, seed=1, check_dims=[2, 4, K])
        assert np.allclose(sigma, S, rtol=5e-6), (storage, sigma, S)
        e = emb.cpu().numpy().astype(np.float64)
        for d in (2, 4, K):
            ang = port.sin_principal_angle(port.unfolded_right_basis(YA, d), port.unfolded_right_basis(e, d))
            assert ang <= 1e-4, (storage, d, ang, info)
        res[storage] = info
    assert abs(res["e4m3"]["iterations"] - res["planes"]["iterations"]) <= 2


def test_preprocess_binarize_to_e4m3(gpu):
    from node2vec2rank_b200 import device as D
    a = np.random.default_rng(1).normal(size=(600, 600)).astype(np.float32)
    a = a + a.T
    L, f32 = D.preprocess_to_layer(torch.from_numpy(a).to(gpu), absolute=True, top_percent_keep=30, binarize=True, want_f32=True,
                                   storage="e4m3")
    assert L.storage == "e4m3" and np.array_equal(L.numpy(), f32.cpu().numpy()) and set(np.unique(L.numpy())) <= {0.0, 1.0}
    assert torch.allclose(L.colsum, f32.sum(dim=0, dtype=torch.float64))
    with pytest.raises(ValueError, match="binarised"):
        D.preprocess_to_layer(torch.from_numpy(a).to(gpu), absolute=True, top_percent_keep=30, storage="e4m3")
