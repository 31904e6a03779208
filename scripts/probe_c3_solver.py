"""Developer probe: the C3 solve at full size on row-list layers under solver variants (block width, tolerance on the angle
bound, degree of the first filter): operator applications, time, and the angle against a tight reference solve (l = 80)."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench
from node2vec2rank_b200 import device as D, _native as N

n = int(sys.argv[1]) if len(sys.argv) > 1 else 50000
dev = N.require_gpu()
w = dict(bench.WORKLOADS["c3"])
layers, _, _ = bench.build_workload_layers(w, n, range(w["T"]), storage="auto")
dims = bench.DIMS
k = max(dims)
ref, _, _, iref = D.uase(layers, k, l=80, tol=1e-7, max_iters=400, seed=1000045, check_dims=dims)
print(json.dumps({"reference": {"l": 80, "apps": iref["iterations"], "status": iref["converged_status"], "bound": iref["angle_bound"]}}), flush=True)
variants = [dict(l=48, tol=1e-5), dict(l=48, tol=1e-4), dict(l=48, tol=1e-5, first_degree=3), dict(l=48, tol=1e-5, first_degree=4),
            dict(l=32, tol=1e-5), dict(l=64, tol=1e-5), dict(l=48, tol=1e-5, trace=True)]
for v in variants:
    kw = dict(v)
    l = kw.pop("l")
    for rep in range(2):
        torch.cuda.synchronize()
        emb, sigma, _, info = D.uase(layers, k, l=l, max_iters=400, seed=42, check_dims=dims, **kw)
    ang = D.subspace_sin_angles(emb, ref, dims)
    print(json.dumps({"variant": v, "apps": info["iterations"], "outer": info["outer_iterations"], "passes": info["layer_passes"],
                      "status": info["converged_status"], "bound_met": info.get("bound_met"), "bound": info["angle_bound"],
                      "total_ms": round(info["total_ms"], 2), "mma_ms": round(info["mma_ms"], 2), "max_angle_vs_reference": max(ang.values())}), flush=True)
