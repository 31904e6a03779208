"""CPU, world_size 2, gloo: the host-side sharding logic (all-reduce callback over a workspace view,
embedding all-gather) that the NCCL path uses on the GPUs."""
import ctypes
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as tdist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    try:
        os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
        tdist.init_process_group("gloo", rank=rank, world_size=world)
        from node2vec2rank_b200 import _native as N
        from node2vec2rank_b200 import dist as ND
        # the solver hands back raw pointers into its workspace: the callback must map them to views
        ws = torch.zeros(4096, dtype=torch.uint8)
        f = ws[256:256 + 40].view(torch.float32)
        f.copy_(torch.arange(10, dtype=torch.float32) * (rank + 1))
        d = ws[1024:1024 + 64].view(torch.float64)
        d.copy_(torch.arange(8, dtype=torch.float64) + rank)
        cb, state = ND.make_allreduce_callback(ws)
        assert cb(None, ws.data_ptr() + 256, 10, N.F32, None) == 0
        assert cb(None, ws.data_ptr() + 1024, 8, N.F64, None) == 0
        assert torch.equal(f, torch.arange(10, dtype=torch.float32) * 3)
        assert torch.equal(d, torch.arange(8, dtype=torch.float64) * 2 + 1)
        assert state["calls"] == 2 and state["bytes"] == 104
        assert cb(None, ws.data_ptr() + 4090, 10, N.F32, None) == 1 and isinstance(state["error"], ValueError)   # out of range
        # uneven layer blocks: T = 3 over 2 ranks
        T, n, k = 3, 5, 4
        full = torch.arange(T * n * k, dtype=torch.float32).reshape(T, n, k)
        mine = ND.layer_partition(T, world, rank)
        got = ND.gather_layers(full[mine.start:mine.stop].clone(), T)
        assert torch.equal(got, full)
        with pytest.raises(ValueError):
            ND.gather_layers(full[:0].clone() if rank == 0 else full[:3].clone(), T)
        # even layer blocks: T = 4 over 2 ranks (single collective into the result)
        full4 = torch.arange(4 * n * k, dtype=torch.float32).reshape(4, n, k)
        m4 = ND.layer_partition(4, world, rank)
        assert torch.equal(ND.gather_layers(full4[m4.start:m4.stop].clone(), 4), full4)
        # seeds of a stability run: 3 seeds over 2 ranks, score blocks (S_local, C, cols, n); a rank may hold none
        S, Cc, cols = 3, 2, 4
        sc = torch.arange(S * Cc * cols * n, dtype=torch.float64).reshape(S, Cc, cols, n)
        sm = ND.layer_partition(S, world, rank)
        assert torch.equal(ND.gather_blocks(sc[sm.start:sm.stop].clone(), S), sc)
        one = ND.layer_partition(1, world, rank)
        assert torch.equal(ND.gather_blocks(sc[one.start:one.stop].clone(), 1), sc[:1])
        tdist.barrier()
        tdist.destroy_process_group()
        q.put((rank, "ok"))
    except Exception as exc:  # pragma: no cover
        q.put((rank, repr(exc)))


def test_sharding_plumbing_world2(lib):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = dict(q.get(timeout=180) for _ in procs)
    for p in procs:
        p.join(timeout=60)
    assert res == {0: "ok", 1: "ok"}, res
