"""Synthetic multi-layer stochastic block models (benchmark / test input).

`sbm_layer_numpy` is the bit-identical NumPy twin of the device generator K0
(csrc/api.cu: sbm_kernel, n2v2r_sbm_generate): entry (i, j), i < j, of layer t is a pure
function of (seed, t, i, j); A[j, i] = A[i, j]; zero diagonal; community(i) = i % K.
"""
from __future__ import annotations

import numpy as np

_M64 = np.uint64(0xFFFFFFFFFFFFFFFF)


def _mix64(z):
    z = (z + np.uint64(0x9E3779B97F4A7C15)) & _M64
    z = ((z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)) & _M64
    z = ((z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)) & _M64
    return z ^ (z >> np.uint64(31))


def _hash4(seed, a, b, c):
    with np.errstate(over="ignore"):
        h = _mix64(np.uint64(seed) ^ (np.uint64(a) * np.uint64(0xD6E8FEB86659FD93)))
        h = _mix64(h ^ b.astype(np.uint64))
        h = _mix64(h ^ c.astype(np.uint64))
    return h


def sbm_layer_numpy(n, K, B, seed, layer, weighted=False, rows=None):
    """fp32 (n, n) layer (or the given rows of it), identical to device.sbm_generate."""
    B = np.asarray(B, dtype=np.float32).reshape(K, K)
    ii = np.arange(n, dtype=np.int64) if rows is None else np.asarray(rows, dtype=np.int64)
    jj = np.arange(n, dtype=np.int64)
    I, J = np.meshgrid(ii, jj, indexing="ij")
    a = np.minimum(I, J).astype(np.uint64)
    b = np.maximum(I, J).astype(np.uint64)
    with np.errstate(over="ignore"):
        h = _hash4(seed, layer, a, b)
        u = (h >> np.uint64(40)).astype(np.uint32).astype(np.float32) * np.float32(1.0 / 16777216.0)
        p = B[(I % K), (J % K)]
        edge = (u < p) & (I != J)
        if weighted == 2:                 # 24-bit weights fp32(u') + 2^-24
            h2 = _mix64(h ^ np.uint64(0xA5A5A5A5A5A5A5A5))
            w = (h2 >> np.uint64(40)).astype(np.uint32).astype(np.float32) * np.float32(1.0 / 16777216.0) + np.float32(1.0 / 16777216.0)
            out = np.where(edge, w, np.float32(0))
        elif weighted:
            h2 = _mix64(h ^ np.uint64(0xA5A5A5A5A5A5A5A5))
            w = ((h2 >> np.uint64(48)).astype(np.uint32) + np.uint32(1)).astype(np.float32) * np.float32(1.0 / 65536.0)
            out = np.where(edge, w, np.float32(0))
        else:
            out = edge.astype(np.float32)
    return out.astype(np.float32)


def block_matrices(K, T, seed=1, lo=0.02, hi=0.5, n_changed=3):
    """T symmetric K x K block-probability matrices, entries uniform in [lo, hi].

    A random symmetric matrix has K distinct eigenvalues (top ~ K*mean, the rest spread over a
    semicircle of radius ~ (hi-lo)*sqrt(K/3)), so the expected unfolded matrix has K separated
    singular values.  Layer t > 0 is the base matrix with the rows/columns of `n_changed`
    communities re-drawn: nodes of those communities are the known ranking signal.
    Returns (list of float32 (K, K), list of changed communities per layer).
    """
    rng = np.random.default_rng(seed)

    def draw():
        M = rng.uniform(lo, hi, size=(K, K))
        return np.triu(M) + np.triu(M, 1).T

    base = draw()
    out, changed = [base.astype(np.float32)], [[]]
    for t in range(1, T):
        comm = sorted(int(c) for c in rng.choice(K, size=min(n_changed, K), replace=False))
        alt = draw()
        cur = base.copy()
        for c in comm:
            cur[c, :] = alt[c, :]
            cur[:, c] = alt[c, :]
        out.append(cur.astype(np.float32))
        changed.append(comm)
    return out, changed


def bipartite_layer_numpy(rows, cols, K, seed, layer, signal=1.0):
    """PANDA-shaped bipartite layer (rows x cols, fp32): block signal + N(0,1) noise (host-side,
    NumPy Generator; used at sizes where host generation is cheap: 2k x 30k = 6e7 entries)."""
    rng = np.random.default_rng([seed, layer])
    zr = np.arange(rows) % K
    zc = np.arange(cols) % K
    S = np.random.default_rng([seed, 1000 + layer // 1]).normal(size=(K, K)) * signal
    if layer > 0:
        S0 = np.random.default_rng([seed, 1000]).normal(size=(K, K)) * signal
        S = S0 + 0.3 * S
    W = S[zr][:, zc] + rng.standard_normal((rows, cols))
    return W.astype(np.float32)


__all__ = ["sbm_layer_numpy", "block_matrices", "bipartite_layer_numpy"]
