"""Host-side logic of the row-sharded search on CPU: two ``gloo`` ranks, the CUDA local search and
merge kernel replaced by numpy stand-ins through ShardedSearch's injection points, so that shard
ranges, id offsets, gather layouts and the per-rank query slicing are exercised without a GPU.
(The CUDA merge kernel itself is checked on the GPU in tests/test_search_gpu.py::test_merge_topk_abi.)"""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import pkg
from oracle import search_oracle as so


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _merge_numpy(ids_in: torch.Tensor, sc_in: torch.Tensor, k: int):
    ids_in, sc_in = ids_in.numpy(), sc_in.numpy()
    nl, nq, kin = ids_in.shape
    out_i = np.full((nq, k), -1, np.int64)
    out_s = np.zeros((nq, k), np.float32)
    for q in range(nq):
        cand = sorted((-float(sc_in[l, q, e]), int(ids_in[l, q, e])) for l in range(nl) for e in range(kin) if ids_in[l, q, e] >= 0)
        for j, (ns, i) in enumerate(cand[:k]):
            out_i[q, j], out_s[q, j] = i, -ns
    return torch.from_numpy(out_i), torch.from_numpy(out_s)


def _worker(rank: int, world: int, port: int, n_rows: int, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        sh = pkg("sharded")
        rng = np.random.default_rng(0)
        rows = so.round_bf16(rng.standard_normal((n_rows, 64), dtype=np.float32))
        rows[n_rows // 2 + 3] = rows[5]                    # an exact tie across the shard boundary
        q_all = so.round_bf16(rng.standard_normal((2 * 6, 64), dtype=np.float32))
        q_all[0] = rows[5]
        r0, r1 = sh.shard_range(n_rows, world, rank)
        shard = rows[r0:r1]

        def local_search(q, k, base):
            ids, sc = so.topk_canonical(q.numpy(), shard, k)
            ids = np.where(ids >= 0, ids + base, ids)
            return torch.from_numpy(ids), torch.from_numpy(sc)

        ss = sh.ShardedSearch(local_search, r0, merge=_merge_numpy)
        want_i, want_s = so.topk_canonical(q_all, rows, 10)
        ids, sc = ss.search_replicated(torch.from_numpy(q_all), 10)
        ok1 = np.array_equal(ids.numpy(), want_i) and np.array_equal(sc.numpy(), want_s)
        mine = slice(rank * 6, (rank + 1) * 6)
        ids, sc = ss.search_local_queries(torch.from_numpy(q_all[mine]), 10)
        ok2 = np.array_equal(ids.numpy(), want_i[mine]) and np.array_equal(sc.numpy(), want_s[mine])
        ret[rank] = (ok1, ok2, int(want_i[0, 0]), int(want_i[0, 1]))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n_rows", [1001, 64])
def test_two_rank_sharded_search_equals_single_store(n_rows):
    mgr = mp.Manager()
    ret = mgr.dict()
    port = _free_port()
    mp.spawn(_worker, args=(2, port, n_rows, ret), nprocs=2, join=True)
    assert len(ret) == 2
    for rank in range(2):
        ok1, ok2, first, second = ret[rank]
        assert ok1, f"rank {rank}: replicated-query search differs from the single-store oracle"
        assert ok2, f"rank {rank}: local-query search differs from the single-store oracle"
        assert (first, second) == (5, n_rows // 2 + 3)     # cross-shard tie resolved to the lower global row


def test_shard_range_is_contiguous_and_balanced():
    sh = pkg("sharded")
    for n, w in [(10, 3), (1_000_000, 8), (7, 8), (0, 2), (400_000_000, 8)]:
        spans = [sh.shard_range(n, w, r) for r in range(w)]
        assert spans[0][0] == 0 and spans[-1][1] == n
        assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
        sizes = [b - a for a, b in spans]
        assert max(sizes) - min(sizes) <= 1
