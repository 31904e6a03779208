"""N-GPU result invariance: the layer-sharded solve + sharded sweep on WORLD_SIZE ranks against the
unsharded solve on one GPU, same workload, same seeds.  Run under torchrun:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
        scripts/multi_gpu_invariance.py [workload] [n] [out.json]

Rank 0 additionally builds ALL layers and solves alone; it then compares the Borda tables (Spearman per comparison, top-100
sets), the singular values and the subspace angle at every swept d, and writes one JSON object."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np
import torch
import torch.distributed as tdist

import bench
from node2vec2rank_b200 import _native as N
from node2vec2rank_b200 import device as D
from node2vec2rank_b200 import dist as ND


def spearman(a, b):
    ra = np.argsort(np.argsort(a, kind="stable"), kind="stable").astype(np.float64)
    rb = np.argsort(np.argsort(b, kind="stable"), kind="stable").astype(np.float64)
    return float(np.corrcoef(ra, rb)[0, 1])


def main():
    args = [a for a in sys.argv[1:] if not a.startswith("--")]
    workload = args[0] if len(args) > 0 else "c3"
    n = int(args[1]) if len(args) > 1 else 20000
    out_path = args[2] if len(args) > 2 else os.path.join(ROOT, "gpurun_out", "multi_gpu_invariance.json")
    one_layer = "--one-layer" in sys.argv           # row slabs on two GPUs: a single layer cut into WORLD_SIZE slabs
    world, rank, local = int(os.environ["WORLD_SIZE"]), int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = N.require_gpu()
    tdist.init_process_group("nccl", device_id=dev)
    w = dict(bench.WORKLOADS[workload])
    if one_layer:
        w["T"], w["strategy"] = 1, "one_vs_rest"
    T, l = w["T"], 48
    shard = (rank, world, lambda t: tdist.all_reduce(t))
    if world > T:                                   # row slabs: rank r holds slab r % S of layer r // S
        S = world // T
        mine = range(rank // S, rank // S + 1)
        rows = ND.layer_partition(n, S, rank % S)
        group = None
        for t in range(T):
            g = tdist.new_group(list(range(t * S, (t + 1) * S)))
            if t == rank // S:
                group = g
        full, f32s, _ = bench.build_workload_layers(w, n, mine, want_f32=True, storage="planes")
        layers = [D.make_slab_layer(f, rows.start, len(rows), n_planes=L.n_planes) for f, L in zip(f32s, full)]
        del full, f32s

        def gather(e):
            out = torch.empty((world,) + tuple(e.shape[1:]), dtype=e.dtype, device=e.device)
            tdist.all_gather_into_tensor(out, e.contiguous())
            return out[::S].contiguous()
        info, emb, scores, bordas = bench.hot_path(layers, w, l, T_global=T, layer_offset=mine.start, gather=gather, shard=shard,
                                                   slab_group=group)
    else:
        mine = ND.layer_partition(T, world, rank)
        layers, _, _ = bench.build_workload_layers(w, n, mine)
        gather = lambda e: ND.gather_layers(e, T)                                # noqa: E731
        info, emb, scores, bordas = bench.hot_path(layers, w, l, T_global=T, layer_offset=mine.start, gather=gather, shard=shard)
    torch.cuda.synchronize()
    result = None
    if rank == 0:
        del layers
        all_layers, _, _ = bench.build_workload_layers(w, n, range(T), storage="planes" if world > T else "auto")
        info1, emb1, scores1, bordas1 = bench.hot_path(all_layers, w, l)
        b_n, b_1 = bordas.cpu().numpy(), bordas1.cpu().numpy()
        ang = D.subspace_sin_angles(emb, emb1, bench.DIMS)
        sp = [spearman(b_n[c], b_1[c]) for c in range(b_1.shape[0])]
        tops = [(set(np.argsort(-b_n[c], kind="stable")[:100]), set(np.argsort(-b_1[c], kind="stable")[:100])) for c in range(b_1.shape[0])]
        top_same = [bool(a == b) for a, b in tops]
        top_overlap = [len(a & b) for a, b in tops]
        result = {"workload": w["name"], "n": n, "world": world, "storage": all_layers[0].storage, "layers": T,
                  "row_slabs_per_layer": max(1, world // T),
                  "allreduce_path": info["allreduce_path"],
                  "operator_applications": [info1["iterations"], info["iterations"]],
                  "sigma_rel_diff_max": float(np.max(np.abs(info["sigma"] - info1["sigma"]) / info1["sigma"])),
                  "sin_angle_per_d": {str(d): float(f"{v:.3e}") for d, v in ang.items()}, "max_sin_angle": max(ang.values()),
                  "borda_spearman_per_comparison": sp, "borda_spearman_min": min(sp),
                  "borda_identical": bool(np.array_equal(b_n, b_1)),
                  "top100_identical_per_comparison": top_same, "top100_overlap_per_comparison": top_overlap,
                  "borda_max_abs_diff": int(np.abs(b_n - b_1).max()),
                  "borda_checksum_ok": bool((b_n.sum(axis=1) == 22 * n * (n + 1) // 2).all())}
        os.makedirs(os.path.dirname(out_path), exist_ok=True)
        with open(out_path, "w") as f:
            json.dump(result, f, indent=1)
        print(json.dumps(result), flush=True)
    tdist.barrier()
    ND.destroy_native_comms()
    tdist.destroy_process_group()


if __name__ == "__main__":
    main()
