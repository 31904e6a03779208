"""Build libn2v2r_b200.so (the C-ABI library, include/n2v2r_b200.h) in-tree with nvcc for sm_100a.

    python -m node2vec2rank_b200.csrc.build [--force]

Every .cu is compiled to an object under build/ (in parallel) and linked into
node2vec2rank_b200/libn2v2r_b200.so.  nvcc cross-compiles without a GPU.
"""
from __future__ import annotations

import concurrent.futures
import hashlib
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.dirname(HERE)
ROOT = os.path.dirname(PKG)
OUT = os.path.join(PKG, "libn2v2r_b200.so")
OBJ = os.path.join(ROOT, "build", "obj")
SOURCES = ["api.cu", "batch.cu", "comm.cu", "k0_pack.cu", "k1_preprocess.cu", "k3_stream.cu", "k3_tcgen05.cu", "k3_sparse.cu", "k3_sparse_ts.cu", "k3_rows.cu", "k3_fp8.cu", "k5_orth.cu", "k9_rank.cu", "uase_solver.cu"]
HOST_SOURCES = ["host_pack.cpp"]          # plain C++ (g++): host-side packing for the sparse PCIe transfer
HOST_FLAGS = ["-O3", "-std=c++17", "-fPIC", "-pthread", "-fvisibility=default"]
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
         "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=default"]


def nvcc():
    exe = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(exe):
        raise RuntimeError("nvcc not found; the CUDA extension cannot be built")
    return exe


def _digest():
    h = hashlib.sha256()
    names = sorted(f for f in os.listdir(HERE) if f.endswith((".cu", ".cuh", ".cpp")))
    for f in names + [os.path.join(ROOT, "include", "n2v2r_b200.h")]:
        with open(os.path.join(HERE, f) if not os.path.isabs(f) else f, "rb") as fh:
            h.update(fh.read())
    h.update(" ".join(FLAGS + HOST_FLAGS).encode())
    return h.hexdigest()


def build(force: bool = False, verbose: bool = True) -> str:
    os.makedirs(OBJ, exist_ok=True)
    stamp = os.path.join(OBJ, "digest.txt")
    dig = _digest()
    if not force and os.path.exists(OUT) and os.path.exists(stamp) and open(stamp).read() == dig:
        return OUT
    exe = nvcc()

    def one(src):
        obj = os.path.join(OBJ, src.replace(".cu", ".o"))
        cmd = [exe, *FLAGS, "-c", os.path.join(HERE, src), "-o", obj]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed on {src}:\n{r.stdout}\n{r.stderr}")
        return obj

    def host_one(src):
        obj = os.path.join(OBJ, src.replace(".cpp", ".o"))
        cxx = shutil.which("g++") or "g++"
        r = subprocess.run([cxx, *HOST_FLAGS, "-c", os.path.join(HERE, src), "-o", obj], capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"g++ failed on {src}:\n{r.stdout}\n{r.stderr}")
        return obj

    with concurrent.futures.ThreadPoolExecutor(max_workers=min(8, len(SOURCES))) as ex:
        objs = list(ex.map(one, SOURCES)) + [host_one(s) for s in HOST_SOURCES]
    cmd = [exe, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-Xcompiler", "-pthread", "-o", OUT, *objs, "-ldl"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    with open(stamp, "w") as f:
        f.write(dig)
    if verbose:
        print(f"built {OUT}")
    return OUT


if __name__ == "__main__":
    build(force="--force" in sys.argv)
