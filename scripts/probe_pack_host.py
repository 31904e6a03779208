import numpy as np, time, ctypes as C
from node2vec2rank_b200 import _native as N
lib = N.load()
rng = np.random.default_rng(0)
n = 24000; cols = n + 13
a = rng.random((n, cols), dtype=np.float32)
a[a < 0.9] = 0
a[3, 5] = -0.0
words = (cols + 31)//32
bm = np.zeros((n, words), np.uint32); cap = n*cols//4
vals = np.zeros(cap, np.float32); rs = np.zeros(n, np.int64); nnz = C.c_int64(0)
def p(x): return C.c_void_p(x.ctypes.data)
import os
for th in (1, 4, 8, 16, 32, 64, len(os.sched_getaffinity(0))):
    t = time.perf_counter()
    rc = lib.n2v2r_host_pack_nonzeros(p(a), n, cols, cols, p(bm), p(vals), cap, p(rs), th, C.byref(nnz))
    dt = time.perf_counter() - t
    print(th, rc, nnz.value, f"{dt*1e3:.1f} ms  {a.nbytes/dt/1e9:.1f} GB/s")
bits = np.unpackbits(bm.view(np.uint8).reshape(n, -1), axis=1, bitorder='little')[:, :cols].astype(bool)
out = np.zeros_like(a)
for r in range(n):
    k = int(bits[r].sum())
    out[r, bits[r]] = vals[rs[r]:rs[r]+k]
print("roundtrip", np.array_equal(out.view(np.uint32), a.view(np.uint32)), "nnz", int(np.count_nonzero(a.view(np.uint32))))
rc = lib.n2v2r_host_pack_nonzeros(p(a), n, cols, cols, p(bm), p(vals), 1000, p(rs), 8, C.byref(nnz)); print("overflow ->", rc, nnz.value)
