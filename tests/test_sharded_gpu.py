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
        # the same two searches with the hit records stored into NVLink peer memory instead of all-gathered, several
        # times in a row (the receive buffers alternate, flags only grow)
        sp = sh.ShardedSearch.over_index(shard, r0, exchange="peer")
        ok3 = True
        for _rep in range(5):
            ids, sc = sp.search_replicated(q_all.to(dev).contiguous(), 10)
            ok3 = ok3 and torch.equal(ids, want_i) and torch.equal(sc.view(torch.int32), want_s.view(torch.int32))
            ids, sc = sp.search_local_queries(q_all[mine].to(dev).contiguous(), 10)
            ok3 = ok3 and torch.equal(ids, want_i[mine]) and torch.equal(sc.view(torch.int32), want_s[mine].view(torch.int32))
        torch.cuda.synchronize()
        ret[rank] = (bool(ok1), bool(ok2 and ok3), int(want_i[0, 0]), int(want_i[0, 1]))
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


@pytest.mark.parametrize("k,nq,n_shards", [(10, 300, 2), (24, 33, 3), (48, 5, 8), (50, 3, 2)])
def test_packed_hits_and_merge_kernel_on_one_gpu(k, nq, n_shards):
    """The multi-GPU data path minus the collective, on one GPU: every shard's b200_index_search_packed writes its
    16-byte hit records into its slot of the would-be all-gather buffer, b200_merge_hits merges a query sub-range --
    ids and fp32 score bits must equal one index over every row (including the exhaustive k > 48 path, an empty
    shard, and a tie across a shard boundary)."""
    vs, sh = pkg("vectorstore"), pkg("sharded")
    dev = torch.device("cuda", 0)
    g = torch.Generator(device="cpu").manual_seed(k)
    n = 30_011
    rows = torch.randn((n, 384), generator=g).bfloat16().float()
    rows[n - 7] = rows[40]
    qs = torch.randn((nq, 384), generator=g).bfloat16().float()
    qs[0] = rows[40]
    qd = qs.to(dev).contiguous()
    full = vs.B200Index(384, device=0, capacity=n)
    full.add(rows.to(dev).contiguous())
    want_i, want_s = full.search(qd, k)
    gathered = torch.empty((n_shards + 1, nq, k, 2), dtype=torch.int64, device=dev)
    shards = []
    for r in range(n_shards):
        r0, r1 = sh.shard_range(n, n_shards, r)
        ix = vs.B200Index(384, device=0, capacity=max(r1 - r0, 1))
        ix.add(rows[r0:r1].to(dev).contiguous())
        ix.search_packed(qd, k, r0, gathered[r])
        shards.append(ix)
    empty = vs.B200Index(384, device=0, capacity=16)               # a rank whose shard holds no rows
    empty.search_packed(qd, k, n, gathered[n_shards])
    ids, sc = sh.merge_hits_cuda(gathered, 0, nq, k)
    assert torch.equal(ids, want_i) and torch.equal(sc.view(torch.int32), want_s.view(torch.int32))
    lo, m = nq // 3, max(1, nq // 2)
    ids2, sc2 = sh.merge_hits_cuda(gathered, lo, m, k)              # the sub-range a rank merges for its own queries
    assert torch.equal(ids2, want_i[lo:lo + m]) and torch.equal(sc2.view(torch.int32), want_s[lo:lo + m].view(torch.int32))
    assert int(want_i[0, 0]) == 40 and int(want_i[0, 1]) == n - 7
    # single-rank ShardedSearch through the same two entry points (production wiring short-cuts to the plain search)
    ss = sh.ShardedSearch(lambda q, kk, base, out: full.search_packed(q, kk, base, out), 0)
    ids3, sc3 = ss.search_local_queries(qd, k)
    assert torch.equal(ids3, want_i) and torch.equal(sc3.view(torch.int32), want_s.view(torch.int32))
    for ix in shards + [empty, full]:
        ix.close()
