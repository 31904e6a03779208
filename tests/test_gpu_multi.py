"""GPU, >= 2 devices: the layer-sharded run (NCCL all-reduce of Y issued by the library, sweep columns sharded, Borda
all-reduced) gives the rankings of the one-GPU run: Spearman > 0.9999 per comparison, identical top-100 sets, subspace angle
<= 1e-4 at every swept d.  Skipped on one-GPU boxes; profiles/ holds the log of the 2-GPU run (gpurun --gpus 2)."""
import json
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_two_gpu_rankings_match_one_gpu(gpu, tmp_path):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    out = tmp_path / "inv.json"
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29517", os.path.join(ROOT, "scripts", "multi_gpu_invariance.py"), "c3", "12000", str(out)]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    res = json.load(open(out))
    assert res["allreduce_path"] == "nccl-native"
    assert res["borda_checksum_ok"] and res["borda_spearman_min"] > 0.9999, res
    assert min(res["top100_overlap_per_comparison"]) >= 98, res        # identical up to exact near-ties at the 100th place
    assert res["max_sin_angle"] <= 1e-4 and res["sigma_rel_diff_max"] <= 2e-6, res


def test_row_slabs_of_one_layer_match_the_whole_layer(gpu, tmp_path):
    """Fewer layers than GPUs: ONE binarised layer cut into two row slabs on two GPUs (Z_t partial sums all-reduced
    inside the layer's group, Y assembled by the global all-reduce) against the whole layer on one GPU."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    out = tmp_path / "slab.json"
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29519", os.path.join(ROOT, "scripts", "multi_gpu_invariance.py"), "c2", "6000", str(out), "--one-layer"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    res = json.load(open(out))
    assert res["row_slabs_per_layer"] == 2 and res["allreduce_path"] == "nccl-native"
    assert res["borda_checksum_ok"] and res["borda_spearman_min"] > 0.9999, res
    assert res["max_sin_angle"] <= 1e-4 and res["sigma_rel_diff_max"] <= 2e-6, res
