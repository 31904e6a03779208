"""Developer probe: fp32 residual floor of the solver vs the K-split length (needs a GPU)."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from node2vec2rank_b200 import device as D, _native as N
from node2vec2rank_b200.synthetic import block_matrices
n, T, K = int(sys.argv[1]), int(sys.argv[2]), 32
weighted = int(sys.argv[3])
dev = N.require_gpu()
Bs, _ = block_matrices(K, T, seed=1)
layers = []
for t in range(T):
    raw = D.sbm_generate(n, K, Bs[t], seed=1, layer=t, weighted=bool(weighted))
    layers.append(D.preprocess_to_layer(raw, top_percent_keep=10 if weighted else 100, binarize=not weighted)[0])
    del raw
dims = list(range(4, 25, 2))
for kc in (256, 1024, 4096, 16384):
    for algo in (N.ALGO_TCGEN05,):
        emb, sigma, _, info = D.uase(layers, 24, l=64, tol=0.0, max_iters=60, seed=42, check_dims=dims, k_chunk=kc, algo=algo)
        rel = info["resid"][:24] * info["ritz"][0] / info["ritz"][:24]
        print(f"k_chunk={kc} apps={info['iterations']} conv={info['converged']} max res/lam_j={rel.max():.2e} median={np.median(rel):.2e} "
              f"res/lam1 max={info['resid'][:24].max():.2e} stream_ms={info['stream_ms']:.1f}", flush=True)
