"""Developer probe: per-solve timing split on the C5 layers (n=30k, T=4, binarised) for several block widths. Needs a GPU."""
import os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from node2vec2rank_b200 import device as D, _native as N
from node2vec2rank_b200.synthetic import block_matrices
dev = N.require_gpu()
n, T, K = 30000, 4, 32
Bs = block_matrices(K, T, seed=1)[0]
layers = []
for t in range(T):
    raw = D.sbm_generate(n, K, Bs[t], seed=1, layer=t)
    layers.append(D.preprocess_to_layer(raw, binarize=True)[0])
    del raw
dims = list(range(4, 25, 2))
for l in (64, 48, 64, 48, 32):
    for seed in range(4):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        emb, sigma, _, info = D.uase(layers, 24, l=l, seed=seed, check_dims=dims)
        torch.cuda.synchronize()
        wall = (time.perf_counter() - t0) * 1e3
        print(f"l={l} seed={seed}: wall {wall:.1f} ms, solver {info['total_ms']:.1f}, stream {info['stream_ms']:.1f}, mma {info['mma_ms']:.1f} "
              f"({info['mma_launches']} launches), apps {info['iterations']}, outer {info['outer_iterations']}, launches {info['kernel_launches']}", flush=True)
