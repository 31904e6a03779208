"""CPU: the C-ABI library builds for sm_100a, loads, and exports every symbol include/n2v2r_b200.h declares."""
import ctypes
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "n2v2r_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(n2v2r_[a-z0-9_]+)\s*\(", src)) - {"n2v2r_allreduce_fn"})


def test_header_symbols_exported(lib):
    from node2vec2rank_b200 import _native
    names = _declared()
    assert len(names) >= 19
    for name in names:
        assert hasattr(lib, name), f"{name} declared in include/n2v2r_b200.h but not exported"
        assert name in _native.SIGNATURES, f"{name} has no ctypes signature"
    assert sorted(_native.SIGNATURES) == names
    assert lib.n2v2r_abi_version() == 1


def test_struct_layout_matches_header(lib):
    from node2vec2rank_b200 import _native as N
    assert ctypes.sizeof(N.Layer) == 32
    assert N.UaseParams.tol.offset % 8 == 0 and N.UaseStats.stream_ms.offset == 24
    assert ctypes.sizeof(N.UaseStats) == 24 + 8 * 2 + 8 + 8 * 3 + 8 * 96 * 2


def test_sass_has_tcgen05_and_tma(lib):
    from node2vec2rank_b200 import _native as N
    cuobjdump = "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(cuobjdump):
        pytest.skip("cuobjdump not available")
    sass = subprocess.run([cuobjdump, "-sass", N.LIB_PATH], capture_output=True, text=True).stdout
    assert "sm_100a" in sass
    assert "UTCHMMA" in sass, "tcgen05.mma missing from SASS"
    assert "UTMALDG" in sass, "TMA loads missing from SASS"
    assert "LDTM" in sass, "tcgen05.ld missing from SASS"


def test_no_gpu_fails_loudly(lib):
    import numpy as np
    import torch
    from node2vec2rank_b200 import _native as N
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        N.require_gpu()
    from node2vec2rank_b200.preprocessing_utils import network_transform
    with pytest.raises(RuntimeError):
        network_transform(np.eye(4))


def test_argument_validation_without_gpu(lib):
    """Entry points reject bad arguments before touching the device."""
    from node2vec2rank_b200 import _native as N
    rc = lib.n2v2r_rank_borda(None, None, 10, 1, 1, None, None, None, 0)
    assert rc == N.EINVAL and b"null pointer" in lib.n2v2r_last_error()
    p = N.UaseParams(n=100, T_local=1, T_global=1, layer_offset=0, k=8, l=40, max_iters=10)
    assert lib.n2v2r_uase_workspace_bytes(ctypes.byref(p)) == 0
    assert b"l must be" in lib.n2v2r_last_error()
    p.l = 32
    assert lib.n2v2r_uase_workspace_bytes(ctypes.byref(p)) > 0
    p.k = 100
    assert lib.n2v2r_uase_workspace_bytes(ctypes.byref(p)) == 0


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, "node2vec2rank_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh")):
                text = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in text and "from oracle" not in text, f
                assert "/root/reference" not in text, f
