"""Row-sharded search on real GPUs: two NCCL ranks, each a B200Index over its contiguous row range,
merged by the CUDA K10 kernel -- must equal one index holding every row.  Skipped on a 1-GPU box
(the CPU gloo test in test_sharded_cpu.py covers the host logic there)."""
import os
import socket

import numpy as np
import pytest

from conftest import pkg

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank: int, world: int, port: int, ret):
    import torch.distributed as dist

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        vs, sh = pkg("vectorstore"), pkg("sharded")
        g = torch.Generator(device="cpu").manual_seed(3)
        rows = torch.randn((20_001, 384), generator=g).bfloat16().float()
        rows[15_000] = rows[40]                                   # a tie across the shard boundary
        q_all = torch.randn((2 * 70, 384), generator=g).bfloat16().float()
        q_all[0] = rows[40]
        r0, r1 = sh.shard_range(rows.shape[0], world, rank)
        shard = vs.B200Index(384, device=rank, capacity=r1 - r0)
        shard.add(rows[r0:r1].to(dev).contiguous())
        ss = sh.ShardedSearch.over_index(shard, r0)
        full = vs.B200Index(384, device=rank, capacity=rows.shape[0])
        full.add(rows.to(dev).contiguous())
        want_i, want_s = full.search(q_all.to(dev).contiguous(), 10)
        ids, sc = ss.search_replicated(q_all.to(dev).contiguous(), 10)
        ok1 = torch.equal(ids, want_i) and torch.equal(sc.view(torch.int32), want_s.view(torch.int32))
        mine = slice(rank * 70, (rank + 1) * 70)
        ids, sc = ss.search_local_queries(q_all[mine].to(dev).contiguous(), 10)
        ok2 = torch.equal(ids, want_i[mine]) and torch.equal(sc.view(torch.int32), want_s[mine].view(torch.int32))
        torch.cuda.synchronize()
        ret[rank] = (bool(ok1), bool(ok2), int(want_i[0, 0]), int(want_i[0, 1]))
    finally:
        dist.destroy_process_group()


def test_two_gpu_sharded_search_equals_single_index():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    import torch.multiprocessing as mp

    ret = mp.Manager().dict()
    mp.spawn(_worker, args=(2, _free_port(), ret), nprocs=2, join=True)
    for rank in range(2):
        ok1, ok2, first, second = ret[rank]
        assert ok1 and ok2, f"rank {rank}: sharded result differs from the single index"
        assert (first, second) == (40, 15_000)
