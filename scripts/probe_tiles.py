"""Developer probe: one C3-sized layer stored as dense bf16 planes and as compressed tiles, both products timed with CUDA events
(layer_apply = split + MMA kernel + reduce), bit-compared at a common K chunk.  Needs a GPU.
    python scripts/probe_tiles.py [n] [l] [keep_percent] [reps] [--only-tiles]"""
import json, os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from node2vec2rank_b200 import device as D, _native as N
from node2vec2rank_b200.synthetic import block_matrices
dev = N.require_gpu()
args = [a for a in sys.argv[1:] if not a.startswith("--")]
n = int(args[0]) if len(args) > 0 else 50000
l = int(args[1]) if len(args) > 1 else 48
keep = float(args[2]) if len(args) > 2 else 10.0
reps = int(args[3]) if len(args) > 3 else 20
only_tiles = "--only-tiles" in sys.argv
Bs, _ = block_matrices(32, 2, seed=1)
raw = D.sbm_generate(n, 32, Bs[0], seed=1, layer=0, weighted=True)
t0 = time.perf_counter()
Lt, f32 = D.preprocess_to_layer(raw, top_percent_keep=keep, want_f32=True, storage="tiles")
torch.cuda.synchronize()
t_tiles = time.perf_counter() - t0
del raw
out = {"n": n, "l": l, "keep": keep, "planes": Lt.n_planes, "tile_rows": Lt.tile_rows, "entries": int(Lt.n_entries),
       "density": float(Lt.n_entries) / n / n, "tiles_bytes": Lt.nbytes, "build_tiles_s": t_tiles}
x = torch.randn((n, l), device=dev)
ws = torch.empty(N.load().n2v2r_apply_workspace_bytes(n, l, 0), dtype=torch.uint8, device=dev)


def timed(layer, k_chunk=0):
    for _ in range(3):
        y = D.layer_apply(layer, x, ws=ws if k_chunk == 0 else None, k_chunk=k_chunk)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        y = D.layer_apply(layer, x, ws=ws if k_chunk == 0 else None, k_chunk=k_chunk)
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps, y


ms_t, y_t = timed(Lt)
out["tiles_ms"] = ms_t
if "--ts" in sys.argv:                          # A tile in tensor memory (k3_sparse_ts.cu) against the shared-memory kernel
    import ctypes
    variants = [("tr256", Lt)]
    if Lt.n_planes <= 2:
        L128, _, _ = D.make_tiles_layer(f32, tile_rows=128)
        variants.append(("tr128", L128))
    for nm, Lx in variants:
        y_ss = timed(Lx)[1] if nm != "tr256" else y_t
        if nm != "tr256":
            out[f"tiles_{nm}_ms"] = timed(Lx)[0]
        N.load().n2v2r__debug_sparse_ts(ctypes.c_int(1))
        try:
            ms_ts, y_ts = timed(Lx)
            out[f"ts_{nm}_ms"] = ms_ts
            out[f"ts_{nm}_bit_identical"] = bool(torch.equal(y_ts.view(torch.int32), y_ss.view(torch.int32)))
            out[f"ts_{nm}_max_abs_diff"] = float((y_ts - y_ss).abs().max())
            out[f"ts_{nm}_ref_max"] = float(y_ss.abs().max())
        finally:
            N.load().n2v2r__debug_sparse_ts(ctypes.c_int(0))
if "--cta2" in sys.argv:                        # the CTA-pair kernel (cta_group::2) against the one-CTA kernel
    import ctypes
    N.load().n2v2r__debug_sparse_cta2(ctypes.c_int(1))
    try:
        ms_2, y_2 = timed(Lt)
        out["tiles_cta2_ms"] = ms_2
        out["cta2_bit_identical"] = bool(torch.equal(y_2.view(torch.int32), y_t.view(torch.int32)))
        out["cta2_max_abs_diff"] = float((y_2 - y_t).abs().max())
        names2 = ["mma:wait_a_full", "mma:wait_x_full+peer", "mma:wait_tmem_empty", "-", "conv:wait_a_empty", "conv:restore", "conv:bar",
                  "conv:wait_landing", "conv:scatter", "conv:fence+arrive"]
        for mode, name in ((0, "full"), (1, "no_scatter"), (2, "no_mma"), (3, "skeleton")):
            N.load().n2v2r__debug_sparse_mode(ctypes.c_int(mode))
            out[f"cta2_ms_{name}"] = timed(Lt)[0]
            clk = torch.zeros(148 * 16, dtype=torch.int64, device=dev)
            N.load().n2v2r__debug_sparse_clk(ctypes.c_void_p(clk.data_ptr()))
            D.layer_apply(Lt, x, ws=ws)
            torch.cuda.synchronize()
            N.load().n2v2r__debug_sparse_clk(ctypes.c_void_p(0))
            c = clk.view(74, 2, 16).double().mean(dim=0).cpu().numpy()
            out[f"cta2_cycles_{name}"] = {"leader": {nm: round(float(v)) for nm, v in zip(names2, c[0]) if nm != "-"},
                                          "peer": {nm: round(float(v)) for nm, v in zip(names2, c[1]) if nm != "-" and not nm.startswith("mma")}}
        N.load().n2v2r__debug_sparse_mode(ctypes.c_int(0))
    finally:
        N.load().n2v2r__debug_sparse_cta2(ctypes.c_int(0))
if "--debug-modes" in sys.argv:                 # profiling: the kernel without its converters / without its MMAs (wrong results)
    import ctypes
    names = ["mma:wait_a_full", "mma:wait_x_full", "mma:wait_tmem_empty", "-", "conv:wait_a_empty", "conv:restore", "conv:bar", "conv:wait_landing",
             "conv:scatter", "conv:fence+arrive"]
    for mode, name in ((0, "full"), (1, "no_scatter"), (2, "no_mma"), (3, "skeleton")):
        N.load().n2v2r__debug_sparse_mode(ctypes.c_int(mode))
        out[f"tiles_ms_{name}"] = timed(Lt)[0]
        clk = torch.zeros(148 * 16, dtype=torch.int64, device=dev)
        N.load().n2v2r__debug_sparse_clk(ctypes.c_void_p(clk.data_ptr()))
        D.layer_apply(Lt, x, ws=ws)
        torch.cuda.synchronize()
        N.load().n2v2r__debug_sparse_clk(ctypes.c_void_p(0))
        c = clk.view(148, 16).double().mean(dim=0).cpu().numpy()
        out[f"cycles_{name}"] = {nm: round(float(v)) for nm, v in zip(names, c) if nm != "-"}
    N.load().n2v2r__debug_sparse_mode(ctypes.c_int(0))
if not only_tiles:
    Lp = D.make_layer(f32, storage="planes")
    del f32
    ms_p, y_p = timed(Lp)
    out["planes_ms"] = ms_p
    out["speedup"] = ms_p / ms_t
    out["max_rel_diff_default_splits"] = float(((y_t - y_p).abs().max() / y_p.abs().max()).item())
    _, a = timed(Lt, 4096)
    _, b = timed(Lp, 4096)
    out["bit_identical_k4096"] = bool(torch.equal(a.view(torch.int32), b.view(torch.int32)))
print(json.dumps(out), flush=True)
