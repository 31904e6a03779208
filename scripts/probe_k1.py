"""Developer probe: HBM throughput of the load-pass kernels (K1a select, K1b transform, plane utilities) at a C3-sized
layer (n x n fp32).  Needs a GPU.   python scripts/probe_k1.py [n]"""
import os, sys, json
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from node2vec2rank_b200 import device as D, _native as N
from node2vec2rank_b200.synthetic import block_matrices
dev = N.require_gpu()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 50000
Bs = block_matrices(32, 2, seed=1)[0]
raw = D.sbm_generate(n, 32, Bs[0], seed=1, layer=0, weighted=True)
peak = json.load(open("MEASURED_PEAKS.json"))["hbm_gbs"] if os.path.exists("MEASURED_PEAKS.json") else 6650.0
out = {}


def timed(name, fn, nbytes, reps=3):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        r = fn()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    gbs = nbytes / ms / 1e6
    out[name] = dict(ms=round(ms, 3), algorithmic_gb=round(nbytes / 1e9, 3), gbs=round(gbs, 1), frac=round(gbs / peak, 3))
    print(f"{name:42s} {ms:8.3f} ms  {nbytes / 1e9:7.2f} GB  {gbs:7.1f} GB/s  {gbs / peak:5.2f} of peak", flush=True)
    return r


e = float(n) * n
# K1a: 3 histogram passes (11+11+10 bits of the fp32 key) + 1 adjacent-order-statistic pass, each one read of the matrix
timed("select_nonzero fp32 (4 read passes)", lambda: D.select_nonzero(raw, 0.9), 4 * 4 * e)
cut = D.percentile_cut(raw, 90)
timed("transform fp32 -> fp32 + colsum", lambda: D.transform(raw, cut=float(cut), want_f32=True, want_colsum=True), 8 * e)
timed("transform fp32 -> 2 bf16 planes + colsum", lambda: D.transform(raw, cut=float(cut), n_planes=2, want_colsum=True), (4 + 4) * e)
timed("transform fp32 -> 1 bf16 plane (binarize)", lambda: D.transform(raw, cut=float(cut), binarize=True, n_planes=1), (4 + 2) * e)
x32 = D.transform(raw, cut=float(cut), want_f32=True)[0]
timed("planes_needed (one read)", lambda: D.planes_needed(x32), 4 * e)
_, planes, ld, _ = D.transform(raw, cut=float(cut), n_planes=2)
lib = N.load()
import ctypes as C
sym = C.c_int(0)
timed("is_symmetric, 2 planes (one read)", lambda: N.check(lib.n2v2r_is_symmetric(N.stream_ptr(), N.ptr(planes), 2, n, ld, C.byref(sym))), 2 * 2 * n * ld)
pt = torch.empty_like(planes)
timed("transpose_planes, 2 planes (read + write)", lambda: N.check(lib.n2v2r_transpose_planes(N.stream_ptr(), N.ptr(planes), 2, n, ld, N.ptr(pt))), 2 * 2 * 2 * n * ld)
assert sym.value == 1 and torch.equal(pt, planes)
r64 = raw[: n // 2].double().contiguous()
timed("transform fp64 -> 2 bf16 planes (half layer)", lambda: D.transform(r64, cut=float(cut), n_planes=2), (8 + 4) * e / 2)
os.makedirs("gpurun_out", exist_ok=True)
json.dump({"n": n, "hbm_peak_gbs": peak, "kernels": out}, open("gpurun_out/k1_roofline.json", "w"), indent=1)
