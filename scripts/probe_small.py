"""Developer probe: one small solve (the l x l kernels do not depend on n) -- run under ncu for per-launch times."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from node2vec2rank_b200 import device as D, _native as N
from node2vec2rank_b200.synthetic import block_matrices
dev = N.require_gpu()
n, T, K = 6000, 2, 32
l = int(sys.argv[1]) if len(sys.argv) > 1 else 64
Bs = block_matrices(K, T, seed=1)[0]
layers = [D.preprocess_to_layer(D.sbm_generate(n, K, Bs[t], seed=1, layer=t), binarize=True)[0] for t in range(T)]
emb, sigma, _, info = D.uase(layers, 24, l=l, seed=1, check_dims=list(range(4, 25, 2)))
torch.cuda.synchronize()
print(info["iterations"], info["outer_iterations"], info["total_ms"])
