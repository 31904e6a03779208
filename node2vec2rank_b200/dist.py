"""Layer sharding across GPUs: one process per GPU, torch.distributed for the plumbing.

The unfolded operator splits by layer: [A_1|...|A_T] X = sum_t A_t X_t and (A^T Q)_t = A_t^T Q.
Rank r owns a contiguous block of layers; Z_t = A_t^T Q and the per-layer embeddings stay local,
and the only exchange per operator application is an all-reduce of Y = sum_t A_t Z_t (n x l fp32)
-- the solver calls back into `make_allreduce_callback` for it.  Every rank then runs the same
(deterministic) small dense steps on the same reduced data, so Q stays replicated bit for bit.
The per-layer embeddings are all-gathered once at the end for the ranking kernels.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _native as N


def is_distributed():
    return torch.distributed.is_available() and torch.distributed.is_initialized() and torch.distributed.get_world_size() > 1


def layer_partition(T, world_size, rank):
    """Contiguous block of layers owned by `rank`: the first T % world ranks get one extra layer."""
    if world_size < 1 or not (0 <= rank < world_size):
        raise ValueError("bad world_size / rank")
    base, extra = divmod(T, world_size)
    start = rank * base + min(rank, extra)
    return range(start, start + base + (1 if rank < extra else 0))


_TORCH_DTYPE = {N.F32: torch.float32, N.F64: torch.float64}


def make_allreduce_callback(workspace: torch.Tensor, group=None):
    """ctypes callback (n2v2r_allreduce_fn) summing a sub-range of `workspace` over all ranks, in place.

    The solver hands back raw pointers into the workspace it was given; the callback turns
    (pointer, count, dtype) into a view of that tensor and calls torch.distributed.all_reduce on it
    (NCCL on GPU tensors, ordered after the kernels already queued on the current stream; gloo on
    CPU tensors in the tests).  Returns (callback, state) -- keep both alive during the solve."""
    base = workspace.data_ptr()
    nbytes = workspace.numel() * workspace.element_size()
    flat = workspace.view(torch.uint8).reshape(-1)
    state = {"calls": 0, "bytes": 0, "error": None}

    def _cb(_user, ptr, count, dtype, _stream):
        try:
            tdt = _TORCH_DTYPE[dtype]
            size = torch.empty((), dtype=tdt).element_size()
            off = int(ptr) - base
            if off < 0 or off + count * size > nbytes or off % size:
                raise ValueError("all-reduce range outside the workspace")
            view = flat[off:off + count * size].view(tdt)
            torch.distributed.all_reduce(view, op=torch.distributed.ReduceOp.SUM, group=group)
            state["calls"] += 1
            state["bytes"] += count * size
            return 0
        except Exception as exc:            # never let an exception cross the C boundary
            state["error"] = exc
            return 1

    return N.ALLREDUCE_FN(_cb), state


_native_comms = {}


def native_comm(group=None):
    """ncclComm_t (as a ctypes void pointer) of a communicator the C library owns, spanning the ranks of `group`, or None
    when the process group is not NCCL (gloo in the CPU tests) / not initialised.  Created once per group: rank 0 draws the
    unique id, torch.distributed broadcasts its 128 bytes, every rank calls n2v2r_comm_init.  With it the solver issues
    ncclAllReduce itself (comm_mode = 1) and no Python runs inside the solver loop."""
    import os
    if not is_distributed() or os.environ.get("N2V2R_ALLREDUCE", "native") != "native":
        return None
    try:
        if torch.distributed.get_backend(group) != "nccl":
            return None
    except Exception:
        return None
    key = id(group) if group is not None else 0
    if key not in _native_comms:
        lib = N.load()
        world, rank = torch.distributed.get_world_size(group), torch.distributed.get_rank(group)
        buf = (C.c_ubyte * 128)()
        if rank == 0:
            N.check(lib.n2v2r_comm_unique_id(buf))
        t = torch.tensor(list(buf), dtype=torch.uint8, device=N.require_gpu())
        src = torch.distributed.get_global_rank(group, 0) if group is not None else 0
        torch.distributed.broadcast(t, src=src, group=group)
        raw = bytes(t.cpu().tolist())
        comm = C.c_void_p()
        N.check(lib.n2v2r_comm_init(raw, world, rank, C.byref(comm)))
        _native_comms[key] = comm
    return _native_comms[key]


def destroy_native_comms():
    lib = N.load()
    for comm in _native_comms.values():
        lib.n2v2r_comm_destroy(comm)
    _native_comms.clear()


def gather_layers(local: torch.Tensor, T, group=None):
    """All-gather the per-layer blocks (T_local, n, k) of every rank into (T, n, k), in layer order."""
    world = torch.distributed.get_world_size(group)
    rank = torch.distributed.get_rank(group)
    counts = [len(layer_partition(T, world, r)) for r in range(world)]
    if local.shape[0] != counts[rank]:
        raise ValueError(f"rank {rank} holds {local.shape[0]} layers, partition says {counts[rank]}")
    width = max(counts)
    n, k = local.shape[1], local.shape[2]
    if min(counts) == width:                        # even blocks: one collective straight into the result, no padding, no cat
        out = torch.empty((world * width, n, k), dtype=local.dtype, device=local.device)
        torch.distributed.all_gather_into_tensor(out, local.contiguous(), group=group)
        return out
    pad = torch.zeros((width, n, k), dtype=local.dtype, device=local.device)
    pad[:counts[rank]] = local
    out = torch.empty((world * width, n, k), dtype=local.dtype, device=local.device)
    torch.distributed.all_gather_into_tensor(out, pad, group=group)
    return torch.cat([out[r * width:r * width + c] for r, c in enumerate(counts)], dim=0)


def gather_blocks(local: torch.Tensor, total, group=None):
    """gather_layers for blocks with any number of trailing dimensions: (B_local, ...) -> (total, ...),
    the blocks split over the ranks by layer_partition (used for the seeds of a stability run)."""
    tail = tuple(local.shape[1:])
    flat = local.reshape(local.shape[0], 1, -1) if local.shape[0] else local.new_zeros((0, 1, int(torch.Size(tail).numel())))
    return gather_layers(flat, total, group).reshape((total,) + tail)


__all__ = ["is_distributed", "layer_partition", "make_allreduce_callback", "native_comm", "destroy_native_comms", "gather_layers", "gather_blocks"]
