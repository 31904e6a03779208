"""GPU: n2v2r_rank_borda_noalloc -- rank + Borda with the sort buffers in a static device array inside the library (for
callers that cannot allocate device memory).  Called through the C-ABI and compared with n2v2r_rank_borda (caller workspace):
positions and Borda points are integers, so the comparison is exact, on two calls in a row (the second finds the first
call's contents in the shared scratch).  Collected last on purpose (file name)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _scores(n, n_cols, seed, dev):
    """float64 score columns with ties, zeros of both signs and a few NaN (the cases the rank kernel orders specially)."""
    rng = np.random.default_rng(seed)
    s = rng.standard_normal((n_cols, n))
    s[:, ::7] = np.round(s[:, ::7], 1)              # ties
    s[0, :5] = [0.0, -0.0, np.nan, 1.0, 1.0]
    s[-1, -3:] = np.nan
    return torch.from_numpy(s).to(dev)


def _both(scores, groups):
    """(borda, positions) from the workspace entry point and from the no-allocation entry point, same inputs."""
    from node2vec2rank_b200 import _native as N
    lib = N.load()
    n_cols, n = scores.shape
    dev = scores.device
    ws = torch.empty(lib.n2v2r_rank_workspace_bytes(n, n_cols), dtype=torch.uint8, device=dev)
    out = []
    bufs = [(torch.full((groups, n), -1, dtype=torch.int64, device=dev), torch.full((n_cols, n), -1, dtype=torch.int32, device=dev))
            for _ in range(2)]
    torch.cuda.synchronize()
    N.check(lib.n2v2r_rank_borda(N.stream_ptr(), N.ptr(scores), n, n_cols, groups, N.ptr(bufs[0][1]), N.ptr(bufs[0][0]),
                                 N.ptr(ws), ws.numel()))
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    N.check(lib.n2v2r_rank_borda_noalloc(N.stream_ptr(), N.ptr(scores), n, n_cols, groups, N.ptr(bufs[1][1]), N.ptr(bufs[1][0])))
    torch.cuda.synchronize()
    free1 = torch.cuda.mem_get_info()[0]
    return bufs[0], bufs[1], free0, free1


def test_rank_borda_noalloc_matches_rank_borda_twice_in_a_row(gpu):
    # first call: 12 columns in 3 groups; second call: another shape and other data over the same shared scratch
    for call, (n, n_cols, groups, seed) in enumerate([(3001, 12, 3, 1), (1777, 20, 4, 2)]):
        s = _scores(n, n_cols, seed, gpu)
        (b_ref, p_ref), (b_new, p_new), free0, free1 = _both(s, groups)
        assert torch.equal(p_ref, p_new), (call, int((p_ref != p_new).sum()))
        assert torch.equal(b_ref, b_new), (call, int((b_ref != b_new).sum()))
        # sanity of the result itself: every column's positions are a permutation, Borda = sum over the group of (n - position)
        assert torch.equal(torch.sort(p_new, dim=1).values, torch.arange(n, device=gpu, dtype=torch.int32).expand(n_cols, n))
        want = (n - p_new.to(torch.int64)).reshape(groups, n_cols // groups, n).sum(dim=1)
        assert torch.equal(b_new, want)
        if call == 1:
            assert free1 == free0, "the no-allocation entry point changed the free device memory"


def test_rank_borda_noalloc_without_positions_and_on_a_side_stream(gpu):
    from node2vec2rank_b200 import _native as N
    lib = N.load()
    n, n_cols = 2500, 6
    s = _scores(n, n_cols, 5, gpu)
    (b_ref, _), _, _, _ = _both(s, 1)
    b = torch.empty((1, n), dtype=torch.int64, device=gpu)
    st = torch.cuda.Stream()
    st.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(st):
        N.check(lib.n2v2r_rank_borda_noalloc(N.stream_ptr(), N.ptr(s), n, n_cols, 1, None, N.ptr(b)))
    st.synchronize()
    assert torch.equal(b, b_ref)


def test_rank_borda_noalloc_rejects_shapes_beyond_the_built_in_scratch(gpu):
    from node2vec2rank_b200 import _native as N
    lib = N.load()
    limit = 1 << 20                                   # N2V2R_RANK_NOALLOC_MAX_ELEMS (include/n2v2r_b200.h)
    small = torch.zeros(16, dtype=torch.float64, device=gpu)
    out = torch.zeros(16, dtype=torch.int64, device=gpu)
    # the shape is checked before anything is read, so small buffers stand in for the oversized ones
    for n, n_cols in [(limit + 1, 1), (limit // 4 + 1, 4), (50000, 176)]:
        rc = lib.n2v2r_rank_borda_noalloc(N.stream_ptr(), N.ptr(small), n, n_cols, 1, None, N.ptr(out))
        assert rc == N.EINVAL
        msg = lib.n2v2r_last_error().decode()
        assert "exceeds the built-in scratch" in msg and "n2v2r_rank_borda" in msg, msg
    with pytest.raises(N.NativeError):
        N.check(lib.n2v2r_rank_borda_noalloc(N.stream_ptr(), None, 10, 1, 1, None, None))
    # exactly at the limit is accepted
    n, n_cols = limit // 8, 8
    s = _scores(n, n_cols, 9, gpu)
    (b_ref, p_ref), (b_new, p_new), _, _ = _both(s, 2)
    assert torch.equal(b_ref, b_new) and torch.equal(p_ref, p_new)
