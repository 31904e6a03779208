"""Host pack throughput (n2v2r_host_pack_nonzeros) by thread count, from pageable and from pinned memory."""
import ctypes as C
import os
import sys
import time

import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

from node2vec2rank_b200 import _native as N

lib = N.load()
rng = np.random.default_rng(0)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 24000
density = float(sys.argv[2]) if len(sys.argv) > 2 else 0.026
cols = n
a = rng.random((n, cols), dtype=np.float32)
a[a < 1 - density] = 0
words = (cols + 31) // 32
bm = np.zeros((n, words), np.uint32)
cap = n * cols // 4
vals = np.zeros(cap, np.float32)
rs = np.zeros(n, np.int64)
nnz = C.c_int64(0)


def p(x):
    return C.c_void_p(x.ctypes.data)


def run(src, label, threads):
    best = 1e9
    for _ in range(3):
        t = time.perf_counter()
        lib.n2v2r_host_pack_nonzeros(p(src), n, cols, cols, p(bm), p(vals), cap, p(rs), threads, C.byref(nnz))
        best = min(best, time.perf_counter() - t)
    print(f"{label:9s} threads={threads:3d} nnz={nnz.value} {best * 1e3:7.1f} ms {src.nbytes / best / 1e9:6.1f} GB/s", flush=True)


cores = len(os.sched_getaffinity(0))
for th in (1, 4, 8, cores):
    run(a, "pageable", th)
try:
    import torch
    if torch.cuda.is_available():
        h = torch.empty((n, cols), dtype=torch.float32, pin_memory=True)
        h.copy_(torch.from_numpy(a))
        for th in (8, cores):
            run(h.numpy(), "pinned", th)
except Exception as ex:  # noqa: BLE001
    print("pinned leg skipped:", ex)
