"""CPU: the C-ABI library builds for sm_100a, loads, and exports every symbol include/n2v2r_b200.h declares."""
import ctypes
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "n2v2r_b200.h")


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
    assert lib.n2v2r_abi_version() == 3


def test_struct_layout_matches_header(lib, tmp_path):
    """sizeof / offsetof of every struct of the header, printed by a C program compiled from the header itself, against
    the ctypes mirrors in _native.py."""
    from node2vec2rank_b200 import _native as N
    structs = {"n2v2r_layer_t": N.Layer, "n2v2r_uase_params_t": N.UaseParams, "n2v2r_uase_stats_t": N.UaseStats}
    lines = ['#include <stdio.h>', '#include <stddef.h>', f'#include "{HEADER}"', 'int main(void) {']
    for cname, cls in structs.items():
        lines.append(f'  printf("{cname} %zu\\n", sizeof({cname}));')
        for fname, _ in cls._fields_:
            lines.append(f'  printf("{cname}.{fname} %zu\\n", offsetof({cname}, {fname}));')
    lines += ['  return 0;', '}']
    src = tmp_path / "layout.c"
    src.write_text("\n".join(lines))
    exe = tmp_path / "layout"
    subprocess.run(["gcc", "-o", str(exe), str(src)], check=True)
    got = dict(line.split() for line in subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.splitlines())
    for cname, cls in structs.items():
        assert int(got[cname]) == ctypes.sizeof(cls), cname
        for fname, _ in cls._fields_:
            assert int(got[f"{cname}.{fname}"]) == getattr(cls, fname).offset, (cname, fname)


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
