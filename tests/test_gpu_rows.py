"""GPU: row-list storage of sparse layers (K1d) and its product on the FMA pipes (k3_rows.cu).  The lists must hold exactly
the matrix they were built from (fp32 values, no plane split), the product must agree with the float64 product within the
fp32 accumulation bound, be deterministic, and drive the solver to the embedding of the dense-plane layers."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _sparse(n, density, seed, symmetric=True, bits=24):
    rng = np.random.default_rng(seed)
    mask = rng.random((n, n)) < density
    if bits == 1:
        vals = np.ones((n, n), dtype=np.float32)
    elif bits == 16:
        vals = ((rng.integers(0, 65536, size=(n, n)) + 1) / 65536.0).astype(np.float32)
    else:
        vals = (rng.random((n, n)).astype(np.float32) + np.float32(2.0 ** -24))          # full fp32 mantissas
    a = np.where(mask, vals, np.float32(0)).astype(np.float32)
    if symmetric:
        a = np.triu(a, 1)
        a = a + a.T
    return a


@pytest.mark.parametrize("n,density,symmetric,rb,W", [(1000, 0.03, True, None, None), (777, 0.05, False, 100, 64), (1300, 0.02, True, 528, 256),
                                                      (520, 0.30, False, 7, 32), (64, 0.1, True, 64, 96), (257, 0.0, True, None, None),
                                                      (2100, 0.01, False, 333, 128)])
def test_rows_round_trip(gpu, n, density, symmetric, rb, W):
    from node2vec2rank_b200 import device as D
    a = _sparse(n, density, 11 + n, symmetric)
    x = torch.from_numpy(a).to(gpu)
    L = D.make_rows_layer(x, rows_block=rb, panel_cols=W, max_list_bytes=None)
    nnz = int(np.count_nonzero(a))
    assert L.storage == "rows" and L.planes is None
    assert L.n_entries == nnz and L.symmetric == bool(np.array_equal(a, a.T))
    rb_, W_ = L.tile_rows, L.panel_cols
    n_lists = -(-n // rb_) * -(-n // W_)
    lp = L.tile_ptr.cpu().numpy().astype(np.int64)
    hdr_units = ((2 * (rb_ + 1) + 15) // 16)
    assert lp.shape[0] == n_lists + 1 and lp[0] == 0 and lp[-1] == L.n_units and np.all(np.diff(lp) >= hdr_units)
    assert 16 * int(np.diff(lp).max()) == L.list_cap
    # every list: header offsets (in pairs of entries) ascending from 0, every row padded to whole pairs with {0xFFFFFFFF, 0},
    # inside a row the columns alternate in parity (even first for the rows of even lane groups), ascending in each class,
    # the surplus of the larger class last
    raw = L.entries.cpu().numpy().reshape(-1).view(np.uint8)
    rpg = -(-rb_ // 176)
    for li in (0, n_lists // 2, n_lists - 1):
        blob = raw[16 * lp[li]:16 * lp[li + 1]]
        hdr = blob[:2 * (rb_ + 1)].view(np.uint16).astype(np.int64)
        assert hdr[0] == 0 and np.all(np.diff(hdr) >= 0) and 16 * (hdr_units + hdr[-1]) == blob.shape[0]
        ent = blob[16 * hdr_units:].view(np.uint32).reshape(-1, 2)
        b, p = divmod(li, -(-n // W_))
        block = a[b * rb_:(b + 1) * rb_, p * W_:(p + 1) * W_]
        for r in range(rb_):
            cols = np.nonzero(block[r])[0] if r < block.shape[0] else np.zeros(0, dtype=np.int64)
            first = (r // rpg) & 1
            c0, c1 = cols[(cols & 1) == first], cols[(cols & 1) != first]
            m = min(len(c0), len(c1))
            seq = np.empty(2 * m, dtype=np.int64)
            seq[0::2], seq[1::2] = c0[:m], c1[:m]
            want = np.concatenate([seq, c0[m:], c1[m:]])
            seg = ent[2 * hdr[r]:2 * hdr[r + 1]]
            assert seg.shape[0] == 2 * ((len(cols) + 1) // 2)
            assert np.array_equal(seg[:len(cols), 0], want.astype(np.uint32))
            assert np.array_equal(seg[:len(cols), 1], block[r, want].view(np.uint32) if len(cols) else np.zeros(0, dtype=np.uint32))
            if len(cols) & 1:
                assert seg[-1, 0] == 0xFFFFFFFF and seg[-1, 1] == 0
    assert np.array_equal(L.numpy().view(np.uint32), a.view(np.uint32))
    if not L.symmetric:
        Lt = D.DeviceLayer(planes=None, n=n, ld=L.ld, n_planes=0, storage="rows", tile_rows=rb_, panel_cols=W_, tile_ptr=L.tile_ptr_t,
                           entries=L.entries_t, n_entries=L.n_entries_t, n_units=L.n_units_t, list_cap=L.list_cap_t)
        assert np.array_equal(Lt.numpy(), a.T)
    else:
        assert L.tile_ptr_t is None


@pytest.mark.parametrize("n,density,l,rb,bits,W", [(1000, 0.03, 48, None, 24, None), (1500, 0.026, 64, 120, 16, None), (900, 0.05, 48, 300, 24, None),
                                                   (2100, 0.02, 32, 500, 1, None), (1024, 0.20, 48, 64, 24, None), (640, 0.08, 80, None, 24, None),
                                                   (700, 0.04, 16, 9, 16, 64), (3000, 0.026, 48, 352, 24, 320), (1800, 0.03, 80, 528, 24, 128),
                                                   (3000, 0.026, 32, 16, 24, None), (2500, 0.03, 48, 12, 16, 128),
                                                   (3000, 0.026, 32, 338, 24, None), (2000, 0.03, 64, 250, 24, 128), (1200, 0.05, 32, 40, 16, 64)])
@pytest.mark.parametrize("symmetric", [True, False])
def test_rows_product_against_float64(gpu, n, density, l, rb, bits, W, symmetric):
    """A x and A^T x of a row-list layer: within the fp32 accumulation bound of the float64 product, per entry; one, two and
    three rows per lane group (rb up to 176 / 352 / 528), partial last block, panels that end inside the matrix; the last two
    cases before the last three have more row blocks than SMs and one row per group: two CTAs share an SM; l = 64 runs with
    groups of 8 lanes (1 to 4 rows per group) when a block has at most 352 rows."""
    from node2vec2rank_b200 import device as D
    a = _sparse(n, density, 5 + n + l, symmetric, bits)
    x32 = torch.from_numpy(a).to(gpu)
    L = D.make_rows_layer(x32, rows_block=rb, panel_cols=W, max_list_bytes=None)
    g = torch.Generator(device="cpu").manual_seed(n + l)
    x = torch.randn((n, l), generator=g, dtype=torch.float32).to(gpu)
    a64 = torch.from_numpy(a.astype(np.float64)).to(gpu)
    for tr in (False, True):
        y = D.layer_apply(L, x, transpose=tr)
        m = a64.t() if tr else a64
        ref = m @ x.double()
        scale = (m.abs() @ x.double().abs()).clamp_min(1e-30)
        err = float(((y.double() - ref).abs() / scale).max())
        assert err <= 4e-6, (tr, err)        # sequential fp32 FMA over the non-zeros of a row: ~ sqrt(nnz_row) eps32 typical
        y2 = D.layer_apply(L, x, transpose=tr)
        assert torch.equal(y.view(torch.int32), y2.view(torch.int32))        # fixed order of accumulation: same bits every call
    # and against the dense-plane tensor-core product of the same layer: two fp32-exact evaluations of the same sums
    yp = D.layer_apply(D.make_layer(x32, storage="planes"), x)
    yr = D.layer_apply(L, x)
    assert float((yp - yr).abs().max() / yp.abs().max()) <= 5e-6


def test_rows_groups_of_eight_lanes_at_l32(gpu):
    """l = 32 with groups of 8 lanes (off by default: slower there; kept behind the debug switch): same product."""
    import ctypes
    from node2vec2rank_b200 import _native as N, device as D
    a = _sparse(2600, 0.03, 17)
    L = D.make_rows_layer(torch.from_numpy(a).to(gpu), rows_block=300, max_list_bytes=None)
    x = torch.randn((2600, 32), device=gpu)
    y4 = D.layer_apply(L, x)
    N.load().n2v2r__debug_rows_lanes8(ctypes.c_int(2))
    try:
        y8 = D.layer_apply(L, x)
    finally:
        N.load().n2v2r__debug_rows_lanes8(ctypes.c_int(1))
    ref = torch.from_numpy(a.astype(np.float64)).to(gpu) @ x.double()
    scale = (torch.from_numpy(np.abs(a).astype(np.float64)).to(gpu) @ x.double().abs()).clamp_min(1e-30)
    assert float(((y8.double() - ref).abs() / scale).max()) <= 4e-6
    assert torch.equal(y4.view(torch.int32), y8.view(torch.int32))      # same entries in the same order per row: same bits


def test_rows_solver_matches_planes_solver(gpu):
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
    for storage in ("planes", "rows"):
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
    # same operator, other summation order: the residual-driven stopping rule may differ by a filter step
    assert abs(out["rows"][2]["iterations"] - out["planes"][2]["iterations"]) <= 6


def test_rows_stage_that_does_not_fit_fails_loudly(gpu):
    """528 rows per block x 512 columns at 3 % and l = 80: not even one stage (X panel + longest list) fits in shared memory."""
    from node2vec2rank_b200 import device as D
    from node2vec2rank_b200._native import NativeError
    a = torch.from_numpy(_sparse(1800, 0.03, 3)).to(gpu)
    L = D.make_rows_layer(a, rows_block=528, panel_cols=512, max_list_bytes=None)
    with pytest.raises(NativeError, match="does not fit in shared memory"):
        D.layer_apply(L, torch.zeros((1800, 80), device=gpu))
    assert D.make_rows_layer(a, rows_block=528, panel_cols=256, max_list_bytes=16 * 1024) is None      # what storage="auto" does


def test_rows_layer_built_for_a_block_width(gpu):
    """A layer built for the solver's block width gets the widest panels whose stage fits twice (512 columns for l = 32); it still
    multiplies wider blocks, with one stage (no overlap of copy and compute) when two no longer fit."""
    from node2vec2rank_b200 import device as D
    assert [D.rows_panel_cols(l) for l in (None, 16, 32, 48, 64, 80)] == [256, 512, 512, 384, 320, 256]
    n = 3000
    a = _sparse(n, 0.026, 23)
    x32 = torch.from_numpy(a).to(gpu)
    L = D.make_layer(x32, storage="rows", block_width=32)
    assert L.storage == "rows" and L.panel_cols == 512
    a64 = torch.from_numpy(a.astype(np.float64)).to(gpu)
    for l in (32, 48, 80):
        x = torch.randn((n, l), device=gpu)
        y = D.layer_apply(L, x)
        ref = a64 @ x.double()
        scale = (a64.abs() @ x.double().abs()).clamp_min(1e-30)
        assert float(((y.double() - ref).abs() / scale).max()) <= 4e-6, l
    # the solve on such layers
    emb, sigma, _, info = D.uase([L], 8, l=32, seed=3, check_dims=[8])
    emb2, sigma2, _, _ = D.uase([D.make_layer(x32, storage="planes")], 8, l=48, seed=4, check_dims=[8])
    assert np.allclose(sigma, sigma2, rtol=2e-5)


def test_auto_storage_picks_rows_then_tiles_then_planes(gpu, monkeypatch):
    from node2vec2rank_b200 import device as D
    monkeypatch.setattr(D, "TILES_MIN_N", 512)
    sparse = torch.from_numpy(_sparse(1000, 0.03, 1, bits=16)).to(gpu)
    mid = torch.from_numpy(_sparse(1000, 0.06, 3, bits=16)).to(gpu)
    dense = torch.from_numpy(_sparse(1000, 0.30, 2, symmetric=False, bits=16)).to(gpu)
    Ls, Lm, Ld = D.make_layer(sparse), D.make_layer(mid), D.make_layer(dense)
    assert (Ls.storage, Lm.storage, Ld.storage) == ("rows", "tiles", "planes")
    assert torch.allclose(Ls.colsum, sparse.sum(dim=0, dtype=torch.float64))
    assert np.array_equal(Ls.numpy(), sparse.cpu().numpy())
    # a layer whose longest list does not fit a stage falls through to the tiles
    monkeypatch.setattr(D, "ROWS_MAX_LIST_BYTES", 64)
    assert D.make_layer(sparse).storage == "tiles"
    small = D.make_layer(sparse[:300, :300].contiguous())
    assert small.storage == "planes"
    with pytest.raises(Exception):
        D.layer_apply(D.make_rows_layer(sparse, max_list_bytes=None), torch.zeros((1000, 96), device=gpu))     # l > 80
