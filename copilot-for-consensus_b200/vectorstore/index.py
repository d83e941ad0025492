"""B200Index -- thin array-level wrapper over the ``b200_index_*`` C ABI.

Rows are addressed by position (insertion order); no string ids, no metadata.  This is what the
benchmarks and the row-sharded multi-GPU search drive directly; :class:`B200VectorStore` adds the
reference's id/metadata/exception contract on top of the same calls."""
from __future__ import annotations

import ctypes
from typing import Any

import numpy as np

from .. import _native as nat


class B200Index:
    def __init__(self, dim: int, device: str | int = "cuda:0", master_dtype: str = "float32", capacity: int = 4096):
        if master_dtype not in ("float32", "bfloat16"):
            raise ValueError("master_dtype must be 'float32' or 'bfloat16'")
        self.dim = int(dim)
        self.device = nat.device_index(device)
        self.master_bf16 = master_dtype == "bfloat16"
        self._lib = nat.lib()
        self._h = ctypes.c_void_p()
        nat.check(self._lib.b200_index_create(self.dim, int(capacity), self.device,
                                              nat.MASTER_BF16 if self.master_bf16 else nat.MASTER_F32, ctypes.byref(self._h)))

    def close(self) -> None:
        if self._h:
            self._lib.b200_index_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self) -> None:  # pragma: no cover
        try:
            self.close()
        except Exception:
            pass

    def _stream(self) -> ctypes.c_void_p:
        return ctypes.c_void_p(nat.current_stream_ptr(self.device))

    def __len__(self) -> int:
        return int(self._lib.b200_index_count(self._h))

    def add(self, rows: Any) -> None:
        """Append ``[n, dim]`` rows: a CUDA tensor on this GPU (fp32, or bf16 for bf16-master indexes)
        or a host array (fp32; converted to bf16 first for bf16-master indexes)."""
        import torch

        if isinstance(rows, torch.Tensor) and rows.is_cuda:
            want = torch.bfloat16 if self.master_bf16 else torch.float32
            if rows.device.index != self.device or rows.dtype != want or not rows.is_contiguous() or rows.shape[1] != self.dim:
                raise ValueError(f"device rows must be contiguous {want} [n, {self.dim}] on cuda:{self.device}")
            nat.check(self._lib.b200_index_add(self._h, ctypes.c_void_p(rows.data_ptr()), rows.shape[0], 1, self._stream()))
            return
        a = np.ascontiguousarray(rows.numpy() if isinstance(rows, torch.Tensor) else rows, dtype=np.float32)
        if a.ndim != 2 or a.shape[1] != self.dim:
            raise ValueError(f"rows must be [n, {self.dim}]")
        if self.master_bf16:
            a = torch.from_numpy(a).to(torch.bfloat16).view(torch.int16).numpy()
        nat.check(self._lib.b200_index_add(self._h, ctypes.c_void_p(a.ctypes.data), a.shape[0], 0, self._stream()))

    def search(self, queries: Any, k: int = 10, id_base: int = 0):
        """``(ids[nq, k] int64, scores[nq, k] fp32)``; CUDA tensors in -> CUDA tensors out (asynchronous on
        torch's current stream), host arrays in -> numpy out."""
        import torch

        if isinstance(queries, torch.Tensor) and queries.is_cuda:
            if queries.device.index != self.device or not queries.is_contiguous() or queries.shape[1] != self.dim:
                raise ValueError(f"device queries must be contiguous [nq, {self.dim}] on cuda:{self.device}")
            if queries.dtype not in (torch.float32, torch.bfloat16):
                raise ValueError("device queries must be float32 or bfloat16")
            nq = queries.shape[0]
            ids = torch.empty((nq, k), dtype=torch.int64, device=queries.device)
            sc = torch.empty((nq, k), dtype=torch.float32, device=queries.device)
            nat.check(self._lib.b200_index_search(
                self._h, ctypes.c_void_p(queries.data_ptr()), 0 if queries.dtype == torch.float32 else 1, 1, nq, k,
                id_base, ctypes.c_void_p(ids.data_ptr()), ctypes.c_void_p(sc.data_ptr()), 1, self._stream()))
            return ids, sc
        q = np.ascontiguousarray(queries, dtype=np.float32)
        if q.ndim != 2 or q.shape[1] != self.dim:
            raise ValueError(f"queries must be [nq, {self.dim}]")
        ids = nat.host_out((q.shape[0], k), np.int64, self.device)
        sc = nat.host_out((q.shape[0], k), np.float32, self.device)
        nat.check(self._lib.b200_index_search(self._h, ctypes.c_void_p(q.ctypes.data), 0, 0, q.shape[0], k, id_base,
                                              ctypes.c_void_p(ids.ctypes.data), ctypes.c_void_p(sc.ctypes.data), 0,
                                              self._stream()))
        return ids, sc

    def search_packed(self, queries: Any, k: int, id_base: int, hits_out: Any) -> None:
        """``b200_index_search_packed``: hits as 16-byte records into ``hits_out`` -- a contiguous CUDA int64 tensor
        ``[nq, k, 2]`` (``[..., 0]`` = orderable score bits << 32, ``[..., 1]`` = global id), typically the send buffer of
        an all-gather (see ``sharded.py``)."""
        import torch

        if not (isinstance(queries, torch.Tensor) and queries.is_cuda and queries.is_contiguous() and queries.shape[1] == self.dim
                and queries.dtype in (torch.float32, torch.bfloat16) and queries.device.index == self.device):
            raise ValueError(f"queries must be a contiguous float32/bfloat16 CUDA tensor [nq, {self.dim}] on cuda:{self.device}")
        nq = queries.shape[0]
        if not (isinstance(hits_out, torch.Tensor) and hits_out.is_cuda and hits_out.is_contiguous() and hits_out.dtype == torch.int64
                and tuple(hits_out.shape) == (nq, k, 2)):
            raise ValueError("hits_out must be a contiguous CUDA int64 tensor [nq, k, 2]")
        nat.check(self._lib.b200_index_search_packed(self._h, ctypes.c_void_p(queries.data_ptr()),
                                                     0 if queries.dtype == torch.float32 else 1, nq, k, id_base,
                                                     ctypes.c_void_p(hits_out.data_ptr()), self._stream()))

    def search_l2(self, queries: Any, k: int = 10, id_base: int = 0):
        """``b200_index_search_l2``: the k nearest rows by squared Euclidean distance (FAISS-flat semantics,
        faiss_store.py:99-101,212-224): ``(ids[nq, k], squared distances[nq, k])`` ascending; fp32 queries, k <= 64."""
        import torch

        if isinstance(queries, torch.Tensor) and queries.is_cuda:
            if queries.device.index != self.device or not queries.is_contiguous() or queries.shape[1] != self.dim or \
                    queries.dtype != torch.float32:
                raise ValueError(f"device queries must be contiguous float32 [nq, {self.dim}] on cuda:{self.device}")
            nq = queries.shape[0]
            ids = torch.empty((nq, k), dtype=torch.int64, device=queries.device)
            dist = torch.empty((nq, k), dtype=torch.float32, device=queries.device)
            nat.check(self._lib.b200_index_search_l2(
                self._h, ctypes.c_void_p(queries.data_ptr()), 0, 1, nq, k, id_base, ctypes.c_void_p(ids.data_ptr()),
                ctypes.c_void_p(dist.data_ptr()), 1, self._stream()))
            return ids, dist
        q = np.ascontiguousarray(queries, dtype=np.float32)
        if q.ndim != 2 or q.shape[1] != self.dim:
            raise ValueError(f"queries must be [nq, {self.dim}]")
        ids = np.empty((q.shape[0], k), dtype=np.int64)
        dist = np.empty((q.shape[0], k), dtype=np.float32)
        nat.check(self._lib.b200_index_search_l2(self._h, ctypes.c_void_p(q.ctypes.data), 0, 0, q.shape[0], k, id_base,
                                                 ctypes.c_void_p(ids.ctypes.data), ctypes.c_void_p(dist.ctypes.data), 0,
                                                 self._stream()))
        return ids, dist

    def certificate_stats(self) -> tuple[int, int]:
        """``(queries checked, queries re-scored exhaustively)`` by the fused path since creation (synchronises)."""
        a, b = ctypes.c_int64(0), ctypes.c_int64(0)
        nat.check(self._lib.b200_index_certificate_stats(self._h, ctypes.byref(a), ctypes.byref(b)))
        return int(a.value), int(b.value)

    def search_static(self, queries: Any, ids_out: Any, scores_out: Any, id_base: int = 0) -> None:
        """``b200_index_search_static``: the same search for callers that may not allocate device memory
        on the query path (all HBM already owned by the host program's pool, or no tolerance for the
        device-wide sync inside cudaMalloc).  Nothing is allocated here either: the caller passes the
        CUDA query tensor ``[nq <= 128, dim]`` (fp32 or bf16) and pre-allocated CUDA outputs
        ``ids_out [nq, k] int64`` / ``scores_out [nq, k] float32``.  See the header for the shared-scratch rules."""
        import torch

        for t in (queries, ids_out, scores_out):
            if not isinstance(t, torch.Tensor) or not t.is_cuda or t.device.index != self.device or not t.is_contiguous():
                raise ValueError(f"search_static needs contiguous CUDA tensors on cuda:{self.device}")
        if queries.dtype not in (torch.float32, torch.bfloat16) or queries.shape[1] != self.dim:
            raise ValueError(f"queries must be float32 or bfloat16 [nq, {self.dim}]")
        nq, k = ids_out.shape
        if ids_out.dtype != torch.int64 or scores_out.dtype != torch.float32 or scores_out.shape != (nq, k) or queries.shape[0] != nq:
            raise ValueError("ids_out must be int64 [nq, k] and scores_out float32 [nq, k]")
        nat.check(self._lib.b200_index_search_static(
            self._h, ctypes.c_void_p(queries.data_ptr()), 0 if queries.dtype == torch.float32 else 1, nq, k, id_base,
            ctypes.c_void_p(ids_out.data_ptr()), ctypes.c_void_p(scores_out.data_ptr()), self._stream()))

    def get_rows(self, rows: Any) -> np.ndarray:
        r = np.ascontiguousarray(rows, dtype=np.int64)
        out = np.empty((len(r), self.dim), dtype=np.float32)
        if len(r):
            nat.check(self._lib.b200_index_get_rows(self._h, ctypes.c_void_p(r.ctypes.data), len(r),
                                                    ctypes.c_void_p(out.ctypes.data), 0, self._stream()))
        return out

    def remove_rows(self, rows: Any) -> None:
        r = np.ascontiguousarray(sorted(set(int(x) for x in rows)), dtype=np.int64)
        if len(r):
            nat.check(self._lib.b200_index_remove_rows(self._h, ctypes.c_void_p(r.ctypes.data), len(r), self._stream()))

    def clear(self) -> None:
        nat.check(self._lib.b200_index_clear(self._h))
