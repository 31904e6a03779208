"""Development probe: where the end-to-end (host-buffer) step spends its time at C3 scale.

    python scripts/probe_e2e.py [n] [T]
Prints the pinned H2D bandwidth of one fp32 layer and a per-phase breakdown of
N2V2R(graphs=<pinned host float32>).fit_transform_rank() + aggregate_transform().
"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from node2vec2rank_b200 import device as D
from node2vec2rank_b200.model import N2V2R

n = int(sys.argv[1]) if len(sys.argv) > 1 else 50000
T = int(sys.argv[2]) if len(sys.argv) > 2 else 8
dev = torch.device("cuda:0")
if os.environ.get("PROBE_H2D"):
    h = torch.empty((n, n), dtype=torch.float32, pin_memory=True)
    h.zero_()
    d = torch.empty((n, n), dtype=torch.float32, device=dev)
    for chunk_rows in (n, 4096, 1024):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for r0 in range(0, n, chunk_rows):
            d[r0:r0 + chunk_rows].copy_(h[r0:r0 + chunk_rows], non_blocking=True)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        print(f"H2D {n}x{n} fp32 in {chunk_rows}-row chunks: {dt * 1e3:.1f} ms = {n * n * 4 / dt / 1e9:.1f} GB/s", flush=True)
    t0 = time.perf_counter()
    h.copy_(d)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    print(f"D2H: {n * n * 4 / dt / 1e9:.1f} GB/s", flush=True)
    del d, h


def clocks():
    try:
        import pynvml
        pynvml.nvmlInit()
        hd = pynvml.nvmlDeviceGetHandleByIndex(0)
        return (pynvml.nvmlDeviceGetClockInfo(hd, pynvml.NVML_CLOCK_SM), pynvml.nvmlDeviceGetClockInfo(hd, pynvml.NVML_CLOCK_MEM),
                pynvml.nvmlDeviceGetPowerUsage(hd) // 1000)
    except Exception as ex:  # noqa: BLE001
        return str(ex)


from node2vec2rank_b200.synthetic import block_matrices
Bs = block_matrices(32, T, seed=1)[0]
hosts = []
for t in range(T):
    raw = D.sbm_generate(n, 32, Bs[t], seed=1, layer=t, weighted=True)
    layer, f32 = D.preprocess_to_layer(raw, top_percent_keep=10, want_f32=True)
    hh = torch.empty((n, n), dtype=torch.float32, pin_memory=True)
    hh.copy_(f32)
    hosts.append(hh.numpy())
    del raw, layer, f32
torch.cuda.synchronize()
torch.cuda.empty_cache()
for mode in ("dense", "packed", "packed"):
    cfg = dict(embed_dimensions=list(range(4, 25, 2)), distance_metrics=["euclidean", "cosine"], comp_strategy="one_vs_rest",
               seed=42, save_dir=None, verbose=-1, block_width=48, host_transfer=mode)
    for rep in range(2):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        m = N2V2R(graphs=hosts, nodes=np.arange(n), config=cfg)
        m._device_layers()
        torch.cuda.synchronize()
        t1 = time.perf_counter()
        ck = clocks()
        m.fit_transform_rank()
        torch.cuda.synchronize()
        t2 = time.perf_counter()
        m.aggregate_transform()
        torch.cuda.synchronize()
        t3 = time.perf_counter()
        si = m.solver_info
        print(f"{mode} rep {rep}: layers to device {t1 - t0:.3f} s | fit+rank {t2 - t1:.3f} s (solver {si['total_ms']:.0f} ms, "
              f"mma {si.get('mma_ms', 0):.0f} ms, apps {si['iterations']}) | borda {t3 - t2:.3f} s | total {t3 - t0:.3f} s | "
              f"clocks before fit (sm, mem, W) {ck} | {m.transfer_stats}", flush=True)
        del m
