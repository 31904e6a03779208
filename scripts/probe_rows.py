"""Developer probe: one C3-sized layer stored as row lists; build time, product time (CUDA events) against the compressed tiles.
    python scripts/probe_rows.py [n] [l] [keep_percent] [reps] [--w24] [--rpg N] [--W cols] [--rb rows]"""
import ctypes
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from node2vec2rank_b200 import _native as N, device as D
from node2vec2rank_b200.synthetic import block_matrices

args = [a for a in sys.argv[1:] if not a.startswith("--")]
n = int(args[0]) if len(args) > 0 else 50000
l = int(args[1]) if len(args) > 1 else 48
keep = float(args[2]) if len(args) > 2 else 10.0
reps = int(args[3]) if len(args) > 3 else 20


def opt(name, default):
    return int(sys.argv[sys.argv.index(name) + 1]) if name in sys.argv else default


dev = N.require_gpu()
Bs, _ = block_matrices(32, 2, seed=1)
g = D.sbm_generate(n, 32, Bs[0], seed=1, layer=0, weighted=2 if "--w24" in sys.argv else True)
layer_p, x32 = D.preprocess_to_layer(g, top_percent_keep=keep, want_f32=True, storage="planes")
del g
out = {"n": n, "l": l, "keep": keep, "planes": layer_p.n_planes}
del layer_p
torch.cuda.synchronize()
if "--lanes4" in sys.argv:
    N.load().n2v2r__debug_rows_lanes8(ctypes.c_int(0))
if "--one-per-sm" in sys.argv:
    N.load().n2v2r__debug_rows_two_per_sm(ctypes.c_int(0))
if "--no-parity" in sys.argv:
    N.load().n2v2r__debug_rows_parity(ctypes.c_int(0))
t0 = time.perf_counter()
Lr = D.make_rows_layer(x32, rows_block=opt("--rb", None), panel_cols=opt("--W", None), max_list_bytes=None)
torch.cuda.synchronize()
out.update(build_rows_s=time.perf_counter() - t0, rows_per_block=Lr.tile_rows, panel_cols=Lr.panel_cols, list_cap=Lr.list_cap,
           nonzeros=Lr.n_entries, density=Lr.n_entries / n / n, rows_bytes=Lr.nbytes)
x = torch.randn((n, l), device=dev)
ws = torch.empty(N.load().n2v2r_apply_workspace_bytes(n, l, 0), dtype=torch.uint8, device=dev)


def timed(L):
    for _ in range(3):
        y = D.layer_apply(L, x, ws=ws)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        y = D.layer_apply(L, x, ws=ws)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps, y


for rpg in ([opt("--rpg", 0)] if "--rpg" in sys.argv else [0, 1, 2, 3]):
    N.load().n2v2r__debug_rows_rpg(ctypes.c_int(rpg))
    try:
        ms, y_r = timed(Lr)
        out[f"rows_ms_rpg{rpg}"] = ms
    except Exception as e:          # rows per block too large for this many rows per group
        out[f"rows_ms_rpg{rpg}"] = str(e)[:80]
N.load().n2v2r__debug_rows_rpg(ctypes.c_int(0))
ms, y_r = timed(Lr)
out["rows_ms"] = ms
if "--no-tiles" not in sys.argv:
    Lt, P, _ = D.make_tiles_layer(x32)
    ms_t, y_t = timed(Lt)
    out.update(tiles_ms=ms_t, max_rel_diff_rows_vs_tiles=float(((y_r - y_t).abs().max() / y_t.abs().max()).item()))
print(json.dumps(out))
