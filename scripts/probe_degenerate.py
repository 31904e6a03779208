"""Degenerate inputs through the solver: rank-deficient, all-zero, tiny.  Prints what comes back."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from node2vec2rank_b200 import _native as N
from node2vec2rank_b200 import device as D

dev = N.require_gpu()


def run(name, graphs, k, l=None):
    try:
        layers = [D.make_layer(torch.from_numpy(g).to(dev)) for g in graphs]
        n = graphs[0].shape[0]
        l = l or D.default_block_width(k, n)
        emb, sigma, _, info = D.uase(layers, k, l=l, seed=1)
        e = emb.cpu().numpy()
        A = np.hstack([g.astype(np.float64) for g in graphs])
        S = np.linalg.svd(A, compute_uv=False)[:k]
        print(name, "l", l, "conv", info["converged"], "iters", info["iterations"], "finite", bool(np.isfinite(e).all()),
              "sigma", np.round(sigma, 4).tolist(), "truth", np.round(S, 4).tolist(), flush=True)
    except Exception as ex:  # noqa: BLE001
        print(name, "EXC", type(ex).__name__, str(ex)[:200], flush=True)


a = np.zeros((40, 40), np.float32); a[0, 1] = a[1, 0] = 1
run("single-edge", [a, a], 8)
run("all-zero", [np.zeros((64, 64), np.float32)] * 2, 4)
z = np.arange(512) % 3
B = np.array([[.9, .1, .2], [.1, .8, .3], [.2, .3, .7]], np.float32)
run("rank3", [B[z][:, z].copy(), (B * 0.5)[z][:, z].copy()], 8)
rng = np.random.default_rng(0)
g = (rng.random((17, 17)) < 0.4).astype(np.float32); g = np.triu(g, 1); g = g + g.T
run("n17", [g, g.T.copy()], 4)
g = (rng.random((10, 10)) < 0.5).astype(np.float32); g = np.triu(g, 1); g = g + g.T
run("n10", [g, g], 4)
run("n10-k9", [g, g], 9)
