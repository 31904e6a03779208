"""Developer probe: K2 projection (W^T W on tcgen05, split-bf16) at the C4 shape: time, TFLOP/s, accuracy (needs a GPU)."""
import os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from node2vec2rank_b200 import device as D, _native as N
from node2vec2rank_b200.synthetic import bipartite_layer_numpy
dev = N.require_gpu()
r, c = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (2000, 30000)
W = torch.from_numpy(bipartite_layer_numpy(r, c, 16, 1, 0)).to(dev)
import ctypes
variants = [(512, 1), (512, 2), (1024, 2), (2048, 2)] if "--variants" in sys.argv else [(512, 2)]
for chunk, nb in variants:
  N.load().n2v2r__debug_project(ctypes.c_int(chunk), ctypes.c_int(nb))
  print(f"--- K chunk {chunk}, column blocks per A tile {nb}", flush=True)
  for on in (("columns",) if "--variants" in sys.argv else ("columns", "rows")):
      out = D.project(W, on)
      torch.cuda.synchronize()
      e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
      e0.record()
      for _ in range(3):
          out = D.project(W, on)
      e1.record(); torch.cuda.synchronize()
      Wa = W.double().abs()
      scale = ((Wa.T @ Wa) if on == "columns" else (Wa @ Wa.T)).clamp_min(1e-300)
      ms = e0.elapsed_time(e1) / 3
      m = c if on == "columns" else r
      flop = 2.0 * r * c * m
      W64 = W.double()
      ref = (W64.T @ W64) if on == "columns" else (W64 @ W64.T)
      err = (out.double() - ref).abs().max().item() / ref.abs().max().item()
      sym = (out - out.T).abs().max().item()
      ratio = ((out.double() - ref).abs() / scale).max().item() / 2.0 ** -24
      del scale, Wa
      print(f"project on={on} ({r}x{c}) -> {m}x{m}: {ms:.2f} ms, {flop / ms / 1e9:.1f} algorithmic TFLOP/s, max rel err {err:.2e}, per-entry error {ratio:.1f} eps32 sum|w||w|, asym {sym:.2e}", flush=True)
