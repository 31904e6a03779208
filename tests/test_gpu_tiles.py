"""GPU: compressed-tile storage of sparse layers (K1c) and the product kernel that expands the tiles in shared
memory (k3_sparse.cu).  The tiles must hold exactly the matrix they were built from, and the product must be
bit-identical to the dense-plane tcgen05 kernel for the same K split (zero entries add exact zeros)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _sparse(n, density, seed, planes, symmetric=True):
    """Random sparse fp32 matrix whose non-zeros need exactly `planes` bf16 terms."""
    rng = np.random.default_rng(seed)
    mask = rng.random((n, n)) < density
    if planes == 1:
        vals = np.ones((n, n), dtype=np.float32)
    elif planes == 2:
        vals = ((rng.integers(0, 65536, size=(n, n)) + 1) / 65536.0).astype(np.float32)        # the C3 benchmark's 16-bit weights
    else:
        vals = (rng.random((n, n)).astype(np.float32) + np.float32(2.0 ** -24))
    a = np.where(mask, vals, np.float32(0)).astype(np.float32)
    if symmetric:
        a = np.triu(a, 1)
        a = a + a.T
    return a


@pytest.mark.parametrize("n,density,planes,symmetric", [(1000, 0.03, 2, True), (777, 0.05, 3, False), (1300, 0.02, 1, True),
                                                        (520, 0.30, 2, False), (64, 0.1, 3, True), (257, 0.0, 1, True)])
def test_tiles_round_trip(gpu, n, density, planes, symmetric):
    from node2vec2rank_b200 import device as D
    a = _sparse(n, density, 11 + n, planes, symmetric)
    x = torch.from_numpy(a).to(gpu)
    L, P, sym = D.make_tiles_layer(x)
    nnz = int(np.count_nonzero(a))
    assert L.storage == "tiles" and L.planes is None
    assert P == (planes if nnz else 1) and L.n_planes == P
    assert sym == bool(np.array_equal(a, a.T)) and L.symmetric == sym
    assert L.tile_rows == (256 if P <= 2 else 128)
    # every list is padded to a multiple of 4 entries: between nnz and nnz + 3 x number of tiles
    n_tiles = L.tile_ptr.numel() - 1
    assert nnz <= L.n_entries <= nnz + 3 * n_tiles and L.n_entries % 4 == 0 and L.entries.shape[0] >= L.n_entries
    tp = L.tile_ptr.cpu().numpy().astype(np.int64)
    assert tp[0] == 0 and np.all(np.diff(tp) >= 0) and np.all(tp % 4 == 0) and tp[-1] == L.n_entries
    assert np.array_equal(L.numpy().view(np.uint32), a.view(np.uint32))
    if not sym:
        assert L.tile_ptr_t is not None
        Lt = D.DeviceLayer(planes=None, n=n, ld=L.ld, n_planes=P, storage="tiles", tile_rows=L.tile_rows,
                           tile_ptr=L.tile_ptr_t, entries=L.entries_t, n_entries=L.n_entries_t)
        assert np.array_equal(Lt.numpy(), a.T)
    else:
        assert L.tile_ptr_t is None


@pytest.mark.parametrize("n,density,planes,l", [(1000, 0.03, 2, 48), (1500, 0.026, 2, 64), (900, 0.05, 3, 48), (2100, 0.02, 1, 32),
                                                (1024, 0.20, 2, 48), (640, 0.08, 3, 80), (700, 0.04, 1, 16), (3000, 0.026, 2, 48)])
@pytest.mark.parametrize("symmetric", [True, False])
def test_tiles_product_is_bit_identical_to_the_dense_planes(gpu, n, density, planes, l, symmetric):
    """layer_apply A/B: same layer stored as planes and as tiles, same K chunk -> identical bits, A x and A^T x.
    density 0.20 exceeds the undo list (whole-stage clear), 0.04-0.08 needs several landing chunks per tile."""
    from node2vec2rank_b200 import device as D
    a = _sparse(n, density, 5 + n + l, planes, symmetric)
    x32 = torch.from_numpy(a).to(gpu)
    Lp = D.make_layer(x32, storage="planes")
    Lt = D.make_layer(x32, storage="tiles")
    assert Lp.storage == "planes" and Lt.storage == "tiles" and Lp.n_planes == Lt.n_planes == planes
    g = torch.Generator(device="cpu").manual_seed(n + l)
    x = torch.randn((n, l), generator=g, dtype=torch.float32).to(gpu)
    for k_chunk in (512, 4096):
        for tr in (False, True):
            yp = D.layer_apply(Lp, x, transpose=tr, k_chunk=k_chunk)
            yt = D.layer_apply(Lt, x, transpose=tr, k_chunk=k_chunk)
            assert torch.equal(yp.view(torch.int32), yt.view(torch.int32)), (k_chunk, tr, float((yp - yt).abs().max()))
    # the kernel's own K split (balanced over the SMs): same product up to fp32 summation order
    ref = torch.from_numpy(a.astype(np.float64)).to(gpu) @ x.double()
    yt = D.layer_apply(Lt, x, transpose=False)
    scale = (torch.from_numpy(np.abs(a).astype(np.float64)).to(gpu) @ x.double().abs()).clamp_min(1e-30)
    assert float(((yt.double() - ref).abs() / scale).max()) <= 4e-6       # fp32 accumulation in TMEM over the non-zeros of a row
    # repeated calls reuse the zeroed stages correctly (persistent CTAs walk several units)
    yt2 = D.layer_apply(Lt, x, transpose=False)
    assert torch.equal(yt.view(torch.int32), yt2.view(torch.int32))


def test_tiles_solver_matches_planes_solver(gpu):
    """The UASE solve on tiles layers gives the embedding of the planes layers (same operator, other summation order)."""
    from oracle import reference_port as port
    from node2vec2rank_b200 import device as D
    from node2vec2rank_b200.synthetic import block_matrices, sbm_layer_numpy
    n, K, T, k = 2200, 6, 3, 12
    Bs, _ = block_matrices(K, T, seed=2)
    graphs = []
    for t in range(T):
        g = sbm_layer_numpy(n, K, Bs[t], 5, t, True)
        graphs.append(port.network_transform(g.astype(np.float64), top_percent_keep=20))
    out = {}
    for storage in ("planes", "tiles"):
        layers = [D.make_layer(torch.from_numpy(g).to(gpu), storage=storage) for g in graphs]
        assert all(L.storage == storage for L in layers)
        emb, sigma, _, info = D.uase(layers, k, l=48, tol=1e-6, max_iters=500, seed=1, check_dims=[2, 4, K])
        out[storage] = (emb.cpu().numpy().astype(np.float64), sigma, info)
    S, V, YA = port.uase_dense_fp64(graphs, k)
    for storage in out:
        e, sigma, info = out[storage]
        assert np.allclose(sigma, S, rtol=5e-6), (storage, sigma, S)
        for d in (2, 4, K):
            ang = port.sin_principal_angle(port.unfolded_right_basis(YA, d), port.unfolded_right_basis(e, d))
            assert ang <= 1e-4, (storage, d, ang, info)
    assert out["tiles"][2]["iterations"] == out["planes"][2]["iterations"]


def test_auto_storage_picks_tiles_for_large_sparse_layers(gpu, monkeypatch):
    from node2vec2rank_b200 import device as D
    monkeypatch.setattr(D, "TILES_MIN_N", 512)
    monkeypatch.setattr(D, "ROWS_MAX_DENSITY", 0.0)          # the row lists would take this layer first (tests/test_gpu_rows.py)
    sparse = torch.from_numpy(_sparse(1000, 0.03, 1, 2)).to(gpu)
    dense = torch.from_numpy(_sparse(1000, 0.30, 2, 2, symmetric=False)).to(gpu)
    Ls, Ld = D.make_layer(sparse), D.make_layer(dense)
    assert Ls.storage == "tiles" and Ld.storage == "planes"
    assert Ld.n_planes == 2 and not Ld.symmetric and Ld.planes_t is not None       # planes / symmetry taken from the count pass
    assert np.array_equal(Ld.numpy(), dense.cpu().numpy())
    assert torch.allclose(Ls.colsum, sparse.sum(dim=0, dtype=torch.float64))
    small = D.make_layer(sparse[:300, :300].contiguous())
    assert small.storage == "planes"
