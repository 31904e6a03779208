"""K0c on the GPU: host pack -> H2D -> device expansion is bit-exact, and the packed host->device path of
layers_from_host builds exactly the layers the dense copies build."""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _matrix(rows, cols, density, seed):
    rng = np.random.default_rng(seed)
    a = rng.standard_normal((rows, cols)).astype(np.float32)
    a[rng.random((rows, cols)) >= density] = 0
    return a


@pytest.mark.parametrize("rows,cols,pad,density", [(70, 1000, 0, 0.1), (33, 2100, 4, 0.3), (5, 31, 1, 1.0), (129, 64, 0, 0.0),
                                                    (17, 40000, 0, 0.05)])
def test_unpack_round_trip_is_bit_exact(gpu, rows, cols, pad, density):
    from node2vec2rank_b200 import _native as N
    lib = N.load()
    a = _matrix(rows, cols, density, rows + cols)
    if density > 0 and cols > 40:
        a[1, 2] = -0.0
        a[2, 33] = np.nan
    words = (cols + 31) // 32
    bitmap = np.zeros((rows, words), np.uint32)
    values = np.zeros(rows * cols + 8, np.float32)
    row_start = np.zeros(rows, np.int64)
    nnz = C.c_int64(0)
    p = lambda x: C.c_void_p(x.ctypes.data)  # noqa: E731
    N.check(lib.n2v2r_host_pack_nonzeros(p(a), rows, cols, cols, p(bitmap), p(values), rows * cols, p(row_start), 4, C.byref(nnz)))
    assert nnz.value == np.count_nonzero(a.view(np.uint32))
    d_bitmap = torch.from_numpy(bitmap.view(np.int32)).to(gpu)
    d_values = torch.from_numpy(values).to(gpu)
    d_rows = torch.from_numpy(row_start).to(gpu)
    out = torch.full((rows, cols + pad), 7.0, dtype=torch.float32, device=gpu)
    N.check(lib.n2v2r_unpack_nonzeros(N.stream_ptr(), N.ptr(d_bitmap), N.ptr(d_values), N.ptr(d_rows), rows, cols, N.ptr(out),
                                      cols + pad))
    got = out.cpu().numpy()
    assert np.array_equal(got[:, :cols].view(np.uint32), a.view(np.uint32))
    assert np.all(got[:, cols:] == 7.0)                     # the padding columns are left alone


@pytest.mark.parametrize("chunk_entries,max_density,layers", [(300_000_000, 0.125, 3), (50_000, 0.125, 4), (50_000, 0.01, 3), (50_000, 0.125, 1)])
def test_packed_and_dense_transfers_build_the_same_layers(gpu, monkeypatch, chunk_entries, max_density, layers):
    from node2vec2rank_b200 import device as D
    monkeypatch.setattr(D, "PACK_CHUNK_ENTRIES", chunk_entries)
    monkeypatch.setattr(D, "PACK_MAX_DENSITY", max_density)
    n = 700
    graphs = []
    for t in range(layers):
        g = np.abs(_matrix(n, n, 0.06 if t != 1 else 0.10, 50 + t))
        g = np.triu(g, 1)
        graphs.append(g + g.T if t != 2 else g + 0.5 * g.T)         # layer 2 is not symmetric
    dense = D.layers_from_host(graphs, transfer="dense")
    assert D.layers_from_host.last_stats["packed_chunks"] == 0
    packed = D.layers_from_host(graphs, transfer="packed")
    stats = D.layers_from_host.last_stats
    if max_density >= 0.1:
        assert stats["dense_chunks"] == 0 and stats["packed_chunks"] == layers * -(-n // max(1, min(n, chunk_entries // n)))
        assert stats["h2d_bytes"] < 0.25 * layers * n * n * 4
    else:
        assert stats["dense_chunks"] > 0                          # too dense for the staging capacity: copied dense instead
    for a, b, g in zip(dense, packed, graphs):
        assert a.n_planes == b.n_planes and a.symmetric == b.symmetric
        assert torch.equal(a.planes.view(torch.int16), b.planes.view(torch.int16))
        assert (a.planes_t is None) == (b.planes_t is None)
        assert np.array_equal(b.numpy(), g) or a.n_planes < 3
    # "auto" keeps small layers on the dense path
    D.layers_from_host(graphs[:1])
    assert D.layers_from_host.last_stats["packed_chunks"] == 0


def test_packed_transfer_through_the_model(gpu, demo, demo_graphs, monkeypatch):
    """N2V2R with host_transfer = "packed" returns what the dense transfer returns, bit for bit."""
    import contextlib
    import io
    from node2vec2rank_b200 import device as D
    from node2vec2rank_b200.model import N2V2R
    monkeypatch.setattr(D, "PACK_MAX_DENSITY", 0.5)            # the demo graphs are 24 % dense
    out = {}
    for mode in ("dense", "packed"):
        cfg = dict(embed_dimensions=[4, 8], distance_metrics=["euclidean", "cosine"], comp_strategy="sequential", seed=42,
                   save_dir=None, verbose=-1, host_transfer=mode)
        with contextlib.redirect_stdout(io.StringIO()):
            m = N2V2R(graphs=demo_graphs, nodes=demo["file_cols"], config=cfg)
            ranks = m.fit_transform_rank()
            agg = m.aggregate_transform()
        out[mode] = (m.node_embeddings, ranks["1"].to_numpy(), agg["1"].to_numpy(), m.transfer_stats)
    assert out["packed"][3]["packed_chunks"] > 0 and out["dense"][3]["packed_chunks"] == 0
    for i in range(3):
        assert np.array_equal(out["dense"][i], out["packed"][i])
