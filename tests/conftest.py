import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with `-m gpu` on the GPU box)")


@pytest.fixture(scope="session")
def golden_dir():
    return GOLDEN


@pytest.fixture(scope="session")
def demo():
    """Golden vectors of the reference's demo run (oracle/make_golden.py)."""
    z = np.load(os.path.join(GOLDEN, "demo.npz"), allow_pickle=False)
    return {k: z[k] for k in z.files}


@pytest.fixture(scope="session")
def demo_graphs(demo):
    """The two demo adjacency matrices in FILE order (rows/cols = file order), float32."""
    n = len(demo["file_cols"])
    return [np.unpackbits(demo[f"adj_bits_{i}"])[: n * n].reshape(n, n).astype(np.float32) for i in range(2)]


@pytest.fixture(scope="session")
def lib():
    """The C-ABI library, built on demand (nvcc cross-compiles without a GPU)."""
    from node2vec2rank_b200.csrc.build import build
    build(verbose=False)
    from node2vec2rank_b200 import _native
    return _native.load()


@pytest.fixture(scope="session")
def gpu(lib):
    import torch
    from node2vec2rank_b200 import _native
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return _native.require_gpu()
