"""Developer probe: solver trace on the C4 workload (bipartite 2k x 30k -> W^T W -> UASE).  Needs a GPU.
    N2V2R_TRACE=1 python scripts/probe_c4.py [cols] [max_iters] [block_width]"""
import os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from node2vec2rank_b200 import device as D, _native as N
from node2vec2rank_b200.synthetic import bipartite_layer_numpy
dev = N.require_gpu()
c = int(sys.argv[1]) if len(sys.argv) > 1 else 30000
mi = int(sys.argv[2]) if len(sys.argv) > 2 else 60
l = int(sys.argv[3]) if len(sys.argv) > 3 else 64
layers = []
for t in range(2):
    W = torch.from_numpy(bipartite_layer_numpy(2000, c, 32, 1, t)).to(dev)
    layers.append(D.preprocess_to_layer(D.project(W, "columns"))[0])
    # spectrum of W itself (2000 x c): singular values^2 are the eigenvalues of the projected layer
    sv = torch.linalg.svdvals(W.double())
    print(f"layer {t}: planes {layers[-1].n_planes}, sv(W)^2 [0,1,2,23,24,31,32,33,63,64]:",
          [f"{float(sv[i])**2:.4e}" for i in (0, 1, 2, 23, 24, 31, 32, 33, 63, 64)], flush=True)
dims = list(range(4, 25, 2))
t0 = time.perf_counter()
emb, sigma, _, info = D.uase(layers, 24, l=l, tol=1e-5, max_iters=mi, seed=42, check_dims=dims)
torch.cuda.synchronize()
print(f"solve {time.perf_counter() - t0:.3f} s, apps {info['iterations']}, converged {info['converged']}, status {info['status']}, bound {info['angle_bound']:.3e}")
print("sigma", np.array2string(sigma, precision=4))
print("resid/lam1", np.array2string(info["resid"][:28], precision=2))
print("ritz", np.array2string(info["ritz"][:66:4], precision=3))
