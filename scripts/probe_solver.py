"""Developer probe: per-column residual floors of the solver on the demo graphs (needs a GPU)."""
import sys, os
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from node2vec2rank_b200 import device as D, _native as N
from oracle import reference_port as port
z = np.load("tests/golden/demo.npz"); n = 1000
graphs = [np.unpackbits(z[f"adj_bits_{i}"])[:n*n].reshape(n, n).astype(np.float32) for i in range(2)]
dev = N.require_gpu()
layers = [D.make_layer(torch.from_numpy(g).to(dev)) for g in graphs]
S, V, YA = port.uase_dense_fp64(graphs, 24)
dims = list(range(4, 25, 2))
for l in (48, 64):
    for mi in (int(a) for a in sys.argv[1:] or ["400"]):
        emb, sigma, _, info = D.uase(layers, 24, l=l, tol=0.0, max_iters=mi, seed=42, check_dims=dims)
        e = emb.cpu().numpy().astype(np.float64)
        ang = [port.sin_principal_angle(port.unfolded_right_basis(YA, d), port.unfolded_right_basis(e, d)) for d in dims]
        print(f"l={l} max_iters={mi} apps={info['iterations']} outer={info['outer_iterations']} locked={info['locked']} conv={info['converged']} ms={info['total_ms']:.1f} stream_ms={info['stream_ms']:.1f}")
        print(" resid/lam1:", np.array2string(info["resid"][:26], precision=1))
        print(" resid/lamj:", np.array2string(info["resid"][:26] * info["ritz"][0] / info["ritz"][:26], precision=1))
        print(" angles:", np.array2string(np.asarray(ang), precision=1))
