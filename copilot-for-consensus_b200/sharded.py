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


class PeerExchange:
    """Receive buffers and flags in NVLink peer memory (torch symmetric memory): every rank's search kernels store
    their hit records straight into every rank's buffer (``b200_index_search_exchange``) and the merge kernel waits
    on flags (``b200_merge_hits_wait``) -- no collective, one launch fewer, nothing for the host to do in between.
    Two buffers per shape, used alternately; epochs grow by one per search (see the header for why that suffices)."""

    def __init__(self, index: Any, row_offset: int, group: Any = None):
        import torch
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm

        self._torch, self._symm, self._dist = torch, symm, dist
        self._group = group if group is not None else dist.group.WORLD
        self.world, self.rank = dist.get_world_size(self._group), dist.get_rank(self._group)
        if self.world > 8:
            raise RuntimeError("the peer exchange covers at most 8 ranks")
        self._index, self._offset = index, int(row_offset)
        self._dev = torch.device("cuda", index.device)
        self._epoch = 0
        self._shapes: dict[tuple[int, int], tuple] = {}
        flags = symm.empty((64,), dtype=torch.int32, device=self._dev)      # uint32[world] per rank, zero = no search yet
        flags.zero_()
        self._flags = flags
        self._flags_h = symm.rendezvous(flags, self._group)
        self._flag_ptrs = (ctypes.c_void_p * self.world)(*[int(p) for p in self._flags_h.buffer_ptrs])
        torch.cuda.synchronize(self._dev)
        dist.barrier(group=self._group)

    def _buffers(self, nq: int, k: int):
        """Two symmetric ``[world, nq, k, 2]`` int64 buffers for this shape (allocation is collective: every rank
        asks for the same shapes in the same order, as they run the same searches)."""
        key = (nq, k)
        if key not in self._shapes:
            torch, symm = self._torch, self._symm
            entry = []
            for _ in range(2):
                buf = symm.empty((self.world, nq, k, 2), dtype=torch.int64, device=self._dev)
                h = symm.rendezvous(buf, self._group)
                stride = nq * k * 16                                        # bytes per shard slot
                ptrs = (ctypes.c_void_p * self.world)(*[int(p) + self.rank * stride for p in h.buffer_ptrs])
                entry.append((buf, h, ptrs))
            self._shapes[key] = tuple(entry)
        return self._shapes[key]

    def search_merge(self, queries: Any, k: int, q_first: int, nq_out: int):
        """Search this shard for ALL ``queries`` and return the merged top-k of queries ``[q_first, q_first + nq_out)``."""
        torch = self._torch
        nq = queries.shape[0]
        self._epoch += 1
        buf, _h, ptrs = self._buffers(nq, k)[self._epoch & 1]
        lib, st = nat.lib(), ctypes.c_void_p(nat.current_stream_ptr(self._dev.index))
        nat.check(lib.b200_index_search_exchange(
            self._index._h, ctypes.c_void_p(queries.data_ptr()), 0 if queries.dtype == torch.float32 else 1, nq, k,
            self._offset, ptrs, self._flag_ptrs, self.world, self.rank, self._epoch & 0xFFFFFFFF, st))
        ids = torch.empty((nq_out, k), dtype=torch.int64, device=self._dev)
        sc = torch.empty((nq_out, k), dtype=torch.float32, device=self._dev)
        nat.check(lib.b200_merge_hits_wait(
            ctypes.c_void_p(buf.data_ptr()), ctypes.c_void_p(self._flags.data_ptr()), self._epoch & 0xFFFFFFFF, self.world, nq, k,
            q_first, nq_out, k, ctypes.c_void_p(ids.data_ptr()), ctypes.c_void_p(sc.data_ptr()), self._dev.index, st))
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
                 single: Callable[[Any, int, int], tuple[Any, Any]] | None = None, peer: Any = None):
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
        self._peer = peer            # PeerExchange: records stored into peer memory instead of an all-gather
        self.exchange = "peer" if peer is not None else "nccl"

    @classmethod
    def over_index(cls, index: Any, row_offset: int, group: Any = None, exchange: str = "auto") -> "ShardedSearch":
        """Production wiring: a :class:`B200Index` shard + the CUDA merge kernel.  ``exchange``: ``"peer"`` = hit records
        stored into NVLink peer memory (:class:`PeerExchange`; measured 11 % faster than the collective at 128 queries
        on two GPUs, equal elsewhere), ``"nccl"`` = one all-gather, ``"auto"`` (default) = peer when torch's symmetric
        memory can be set up for the group, else nccl."""
        import torch.distributed as dist

        if exchange not in ("auto", "nccl", "peer"):
            raise ValueError("exchange must be 'auto', 'nccl' or 'peer'")
        peer = None
        if exchange != "nccl" and dist.is_initialized() and 1 < dist.get_world_size(group) <= 8:
            try:
                peer = PeerExchange(index, row_offset, group)
            except Exception:
                if exchange == "peer":
                    raise
                peer = None
        return cls(lambda q, k, base, out: index.search_packed(q, k, base, out), row_offset, merge_hits_cuda, group,
                   single=lambda q, k, base: index.search(q, k, id_base=base), peer=peer)

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
        if self._peer is not None:
            return self._peer.search_merge(queries, k, 0, queries.shape[0])
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
        if self._peer is not None:
            return self._peer.search_merge(q_all, k, self.rank * nq, nq)
        gathered = self._search_and_gather(q_all, k)               # [world, world * nq, k, 2]
        return self._merge(gathered, self.rank * nq, nq, k)
