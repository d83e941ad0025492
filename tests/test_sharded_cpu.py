"""Host-side logic of the row-sharded search on CPU: two ``gloo`` ranks, the CUDA local search and
merge kernel replaced by numpy stand-ins through ShardedSearch's injection points, so that shard
ranges, id offsets, gather layouts and the per-rank query slicing are exercised without a GPU.
The stand-ins speak the same 16-byte hit-record format as b200_index_search_packed / b200_merge_hits; the CUDA
kernels behind those two are checked on the GPU (tests/test_sharded_gpu.py, also with a single rank)."""
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


def _orderable(sc: np.ndarray) -> np.ndarray:
    """fp32 -> uint32 whose unsigned order is the float order (common.cuh f32_orderable)."""
    u = np.ascontiguousarray(sc, np.float32).view(np.uint32)
    return np.where(u & 0x80000000, ~u, u | 0x80000000).astype(np.uint32)


def _unorderable(o: np.ndarray) -> np.ndarray:
    o = o.astype(np.uint32)
    return np.where(o & 0x80000000, o & 0x7FFFFFFF, ~o).astype(np.uint32).view(np.float32)


def _pack_numpy(ids: np.ndarray, sc: np.ndarray) -> np.ndarray:
    """The 16-byte hit record of b200_index_search_packed, as [nq, k, 2] int64."""
    x = (_orderable(sc).astype(np.uint64) << np.uint64(32))
    x = np.where(ids >= 0, x, np.uint64(0))
    return np.stack([x.view(np.int64), ids.astype(np.int64)], axis=-1)


def _merge_numpy(gathered: torch.Tensor, q_first: int, nq: int, k: int):
    g = gathered.numpy()
    nl, _per, kin, _ = g.shape
    out_i = np.full((nq, k), -1, np.int64)
    out_s = np.zeros((nq, k), np.float32)
    for q in range(nq):
        rec = g[:, q_first + q].reshape(nl * kin, 2)
        rec = rec[rec[:, 1] >= 0]
        order = np.lexsort((rec[:, 1], -(rec[:, 0].view(np.uint64) >> np.uint64(32)).astype(np.int64)))[:k]
        out_i[q, : len(order)] = rec[order, 1]
        out_s[q, : len(order)] = _unorderable(rec[order, 0].view(np.uint64) >> np.uint64(32))
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

        def local_search(q, k, base, hits_out):
            ids, sc = so.topk_canonical(q.numpy(), shard, k)
            ids = np.where(ids >= 0, ids + base, ids)
            hits_out.copy_(torch.from_numpy(_pack_numpy(ids, sc)))

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
