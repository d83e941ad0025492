"""Row-sharded top-k across the GPUs of one box.

One process per GPU.  The corpus is split into contiguous row ranges, rank r owning global rows
``[offset_r, offset_r + n_r)`` so that "lower row id" keeps meaning "earlier insertion" across
shards.  A search is: every rank scores ALL queries against its own shard (local top-k through
the C ABI) -> one all-gather of the ``[nq, k]`` (score, global id) records (NCCL over NVLink;
0.66 MB per rank at nq=4096, latency-bound) -> a per-query merge of ``world x k`` candidates under
the same (score desc, id asc) order (``b200_merge_hits``, kernel K10).  Embedding needs no
collective at all: each rank encodes its own slice of the batch.

The reference has no counterpart (it is single-process, inmemory.py); the merged result equals
what one store holding every row would return.

The local-search and merge callables are injectable so that the plumbing (layouts, offsets,
collectives) is exercised by CPU ``gloo`` tests with world_size 2; production wiring uses the CUDA
paths and nothing else."""
from __future__ import annotations

import ctypes
from typing import Any, Callable

from . import _native as nat


def shard_range(n_total: int, world: int, rank: int) -> tuple[int, int]:
    """Contiguous, balanced split: the first ``n_total % world`` ranks hold one extra row."""
    base, extra = divmod(int(n_total), int(world))
    start = rank * base + min(rank, extra)
    return start, start + base + (1 if rank < extra else 0)


def merge_topk_cuda(ids_in, scores_in, k_out: int):
    """``[nlists, nq, k_in]`` CUDA tensors -> ``([nq, k_out], [nq, k_out])`` through ``b200_merge_topk``."""
    import torch

    nl, nq, kin = ids_in.shape
    dev = ids_in.device
    ids = torch.empty((nq, k_out), dtype=torch.int64, device=dev)
    sc = torch.empty((nq, k_out), dtype=torch.float32, device=dev)
    nat.check(nat.lib().b200_merge_topk(
        ctypes.c_void_p(ids_in.data_ptr()), ctypes.c_void_p(scores_in.data_ptr()), nq, nl, kin, k_out,
        ctypes.c_void_p(ids.data_ptr()), ctypes.c_void_p(sc.data_ptr()), dev.index,
        ctypes.c_void_p(nat.current_stream_ptr(dev.index))))
    return ids, sc


def merge_hits_cuda(gathered, q_first: int, nq: int, k_out: int):
    """``b200_merge_hits``: gathered records ``[nlists, nq_per_list, k_in, 2]`` (int64 CUDA) -> the ``k_out`` best of
    queries ``[q_first, q_first + nq)`` as ``(ids[nq, k_out] int64, scores[nq, k_out] fp32)``."""
    import torch

    nl, nq_per_list, kin, _two = gathered.shape
    dev = gathered.device
    ids = torch.empty((nq, k_out), dtype=torch.int64, device=dev)
    sc = torch.empty((nq, k_out), dtype=torch.float32, device=dev)
    nat.check(nat.lib().b200_merge_hits(
        ctypes.c_void_p(gathered.data_ptr()), nl, nq_per_list, kin, q_first, nq, k_out, ctypes.c_void_p(ids.data_ptr()),
        ctypes.c_void_p(sc.data_ptr()), dev.index, ctypes.c_void_p(nat.current_stream_ptr(dev.index))))
    return ids, sc


class ShardedSearch:
    """Top-k over a corpus row-sharded across the ranks of a ``torch.distributed`` group.

    A hit travels as ONE 16-byte record ``(orderable(score) << 32, global id)`` (``b200_index_search_packed``): the
    shard's rescoring kernel writes the records straight into the all-gather send buffer and the merge kernel reads
    the gathered buffer, so a search is  [queries all-gather ->] K8 -> K9 -> all-gather -> K10  with no other kernel
    in between (the first version re-packed ids and scores with six to eight eager PyTorch kernels around the
    collective; at 128 queries over 8 GPUs those launches were half of the time).

    ``local_search(queries, k, id_base, hits_out)`` fills ``hits_out [nq, k, 2]`` (int64) for this rank's shard;
    ``merge(gathered [world, nq_per_list, k, 2], q_first, nq, k_out)`` returns ``(ids, scores)``."""

    def __init__(self, local_search: Callable[[Any, int, int, Any], None], row_offset: int,
                 merge: Callable[[Any, int, int, int], tuple[Any, Any]] = merge_hits_cuda, group: Any = None,
                 single: Callable[[Any, int, int], tuple[Any, Any]] | None = None):
        import torch.distributed as dist

        self._dist = dist
        self._search = local_search
        self._merge = merge
        self._group = group
        self.row_offset = int(row_offset)
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self._bufs: dict[tuple, Any] = {}
        self._single = single        # world == 1: the plain search, no records to exchange or merge

    @classmethod
    def over_index(cls, index: Any, row_offset: int, group: Any = None) -> "ShardedSearch":
        """Production wiring: a :class:`B200Index` shard + the CUDA merge kernel."""
        return cls(lambda q, k, base, out: index.search_packed(q, k, base, out), row_offset, merge_hits_cuda, group,
                   single=lambda q, k, base: index.search(q, k, id_base=base))

    def _buf(self, name: str, shape: tuple, dtype: Any, device: Any):
        """Exchange buffers are kept per shape: no allocation (and no fill kernel) on the query path."""
        import torch

        key = (name, tuple(shape), dtype, str(device))
        b = self._bufs.get(key)
        if b is None:
            b = self._bufs[key] = torch.empty(shape, dtype=dtype, device=device)
        return b

    def _search_and_gather(self, queries: Any, k: int):
        import torch

        n = queries.shape[0]
        send = self._buf("send", (n, k, 2), torch.int64, queries.device)
        self._search(queries, k, self.row_offset, send)
        if self.world == 1:
            return send.view(1, n, k, 2)
        recv = self._buf("recv", (self.world, n, k, 2), torch.int64, queries.device)
        self._dist.all_gather_into_tensor(recv.view(self.world * n, k, 2), send, group=self._group)
        return recv

    def search_replicated(self, queries: Any, k: int = 10):
        """Every rank passes the SAME ``[nq, dim]`` queries; every rank gets the merged global top-k."""
        if self.world == 1 and self._single is not None:
            return self._single(queries, k, self.row_offset)
        gathered = self._search_and_gather(queries, k)
        return self._merge(gathered, 0, queries.shape[0], k)

    def search_local_queries(self, queries: Any, k: int = 10):
        """Each rank passes its OWN ``[nq_local, dim]`` queries (same count on every rank, e.g. the
        embeddings it just computed); returns the merged global top-k for those queries only."""
        nq = queries.shape[0]
        if self.world == 1:
            if self._single is not None:
                return self._single(queries, k, self.row_offset)
            return self._merge(self._search_and_gather(queries, k), 0, nq, k)
        q_all = self._buf("q_all", (self.world * nq, queries.shape[1]), queries.dtype, queries.device)
        self._dist.all_gather_into_tensor(q_all, queries.contiguous(), group=self._group)
        gathered = self._search_and_gather(q_all, k)               # [world, world * nq, k, 2]
        return self._merge(gathered, self.rank * nq, nq, k)
