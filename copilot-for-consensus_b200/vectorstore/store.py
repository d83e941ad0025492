"""B200VectorStore -- drop-in for the reference's in-memory vector store driver.

Mirrors ``InMemoryVectorStore``
(/root/reference/adapters/copilot_vectorstore/copilot_vectorstore/inmemory.py:14-178): same seven
methods, same exception types and message fragments, same ordering (cosine score descending, ties
in insertion order).  Vectors live in HBM behind the C ABI (``b200_index_*``); the id <-> row map
and the metadata dicts stay here, as SURVEY.md section 8b lays out.

Deliberate differences, all documented in DESIGN.md:
  * a store has ONE dimension, fixed by the first insert; a later insert of another length raises
    ``ValueError("... dimension ...")`` (the reference accepts it and then fails inside ``query``);
  * ``top_k`` above 1024 and negative ``top_k`` raise instead of being served (up to 48 the fused
    select serves it; 49..1024 -- the orchestrator asks for 50 by default -- takes the exhaustive path);
  * ``delete`` is lazy: rows are compacted (order-preserving) before the next read.

Extensions (array in / array out, no per-element Python): ``add_embeddings_array``,
``query_batch``.

``metric="l2"`` gives the scoring of the reference's second in-process driver, ``FAISSVectorStore`` with a
flat index (faiss_store.py:99-101 ``IndexFlatL2``; :212-224): nearest by squared Euclidean distance,
``score = 1 / (1 + distance)``.  Same storage, same methods; only ``query`` / ``query_batch`` differ.
"""
from __future__ import annotations

import ctypes
import threading
from typing import Any, Sequence

import numpy as np

from .. import _native as nat
from .interface import SearchResult, VectorStore

MAX_TOP_K = 1024
MAX_TOP_K_L2 = 64


def _as_ptr(a: np.ndarray) -> ctypes.c_void_p:
    return ctypes.c_void_p(a.ctypes.data)


class B200VectorStore(VectorStore):
    """GPU-resident cosine top-k store behind the reference ``VectorStore`` interface."""

    def __init__(self, device: str | int = "cuda:0", upsert: bool = False, initial_capacity: int = 4096,
                 master_dtype: str = "float32", metric: str = "cosine", dimension: int | None = None):
        if metric not in ("cosine", "l2"):
            raise ValueError("metric must be 'cosine' or 'l2'")
        self._metric = metric
        self._fixed_dim = int(dimension) if dimension else None   # faiss_store.py:66: the flat index has a set dimension
        self._device = nat.device_index(device)
        self._upsert = bool(upsert)
        self._initial_capacity = int(initial_capacity)
        if master_dtype not in ("float32", "bfloat16"):
            raise ValueError("master_dtype must be 'float32' or 'bfloat16'")
        self._master = nat.MASTER_F32 if master_dtype == "float32" else nat.MASTER_BF16
        self._lib = nat.lib()  # fails loudly when the CUDA library is not built
        self._h: ctypes.c_void_p | None = None
        self._dim: int | None = None
        self._ids: list[str | None] = []        # row -> id (None = tombstone awaiting compaction)
        self._row_of: dict[str, int] = {}
        self._metadata: dict[str, dict[str, Any]] = {}
        self._dead: list[int] = []
        self._lock = threading.RLock()

    # -- construction helpers ----------------------------------------------------------------
    @classmethod
    def from_config(cls, config: Any = None, **kwargs: Any) -> "B200VectorStore":
        """inmemory.py:29-33 -- the inmemory driver config carries no fields."""
        _ = config
        return cls(**kwargs)

    def connect(self) -> bool:
        """Duck-typed hook the services call when present (embedding/main.py:317-327)."""
        return True

    def close(self) -> None:
        with self._lock:
            if self._h is not None:
                self._lib.b200_index_destroy(self._h)
                self._h = None

    def __del__(self) -> None:  # pragma: no cover - interpreter shutdown ordering
        try:
            self.close()
        except Exception:
            pass

    def _stream(self) -> ctypes.c_void_p:
        return ctypes.c_void_p(nat.current_stream_ptr(self._device))

    def _ensure_index(self, dim: int) -> None:
        if self._fixed_dim is not None and dim != self._fixed_dim:
            raise ValueError(f"Vector dimension ({dim}) doesn't match index dimension ({self._fixed_dim})")
        if self._h is None:
            h = ctypes.c_void_p()
            nat.check(self._lib.b200_index_create(dim, self._initial_capacity, self._device, self._master,
                                                  ctypes.byref(h)))
            self._h, self._dim = h, dim
        elif dim != self._dim:
            raise ValueError(f"Vector dimension ({dim}) doesn't match stored vectors dimension ({self._dim})")

    def _flush_deletes(self) -> None:
        if not self._dead:
            return
        rows = np.array(sorted(self._dead), dtype=np.int64)
        nat.check(self._lib.b200_index_remove_rows(self._h, _as_ptr(rows), len(rows), self._stream()))
        self._ids = [i for i in self._ids if i is not None]
        self._row_of = {i: r for r, i in enumerate(self._ids)}
        self._dead = []

    def _append_rows(self, ids: Sequence[str], arr: np.ndarray, metadatas: Sequence[dict[str, Any]]) -> None:
        if len(ids) == 0:
            return
        self._flush_deletes()
        self._ensure_index(arr.shape[1])
        nat.check(self._lib.b200_index_add(self._h, _as_ptr(arr), arr.shape[0], 0, self._stream()))
        base = len(self._ids)
        for j, (i, m) in enumerate(zip(ids, metadatas)):
            self._ids.append(i)
            self._row_of[i] = base + j
            self._metadata[i] = m.copy()

    def _host_rows(self, vectors: Any) -> np.ndarray:
        if self._master == nat.MASTER_BF16:
            import torch

            t = torch.as_tensor(np.asarray(vectors, dtype=np.float32)).to(torch.bfloat16)
            return t.view(torch.int16).numpy().view(np.uint16)
        return np.ascontiguousarray(vectors, dtype=np.float32)

    # -- VectorStore interface -----------------------------------------------------------------
    def add_embedding(self, id: str, vector: list[float], metadata: dict[str, Any]) -> None:
        """inmemory.py:35-53."""
        with self._lock:
            if id in self._row_of and not self._upsert:
                raise ValueError(f"ID '{id}' already exists in the vector store")
            if vector is None or len(vector) == 0:
                raise ValueError("Vector cannot be empty")
            arr = self._host_rows([vector])
            if self._dim is not None and arr.shape[1] != self._dim:
                raise ValueError(f"Vector dimension ({arr.shape[1]}) doesn't match stored vectors dimension ({self._dim})")
            if id in self._row_of:      # upsert: the old row goes only once the new one is known to be valid
                self.delete(id)
            self._append_rows([id], arr, [metadata])

    def add_embeddings(self, ids: list[str], vectors: list[list[float]], metadatas: list[dict[str, Any]]) -> None:
        """inmemory.py:55-76: all-or-nothing against ids already stored; a duplicate *inside* the batch
        raises after the rows before it went in, exactly as the reference's row loop does."""
        if not (len(ids) == len(vectors) == len(metadatas)):
            raise ValueError("ids, vectors, and metadatas must have the same length")
        with self._lock:
            if not self._upsert:
                for i in ids:
                    if i in self._row_of:
                        raise ValueError(f"ID '{i}' already exists in the vector store")
            n_ok, err = len(ids), None
            seen: set[str] = set()
            for j, (i, v) in enumerate(zip(ids, vectors)):
                if i in seen and not self._upsert:
                    n_ok, err = j, ValueError(f"ID '{i}' already exists in the vector store")
                    break
                if v is None or len(v) == 0:
                    n_ok, err = j, ValueError("Vector cannot be empty")
                    break
                seen.add(i)
            if self._upsert:
                for i in set(ids[:n_ok]):
                    if i in self._row_of:
                        self.delete(i)
                last = {i: j for j, i in enumerate(ids[:n_ok])}  # the last occurrence wins
                keep = sorted(last.values())
                ids_k = [ids[j] for j in keep]
                vec_k = [vectors[j] for j in keep]
                met_k = [metadatas[j] for j in keep]
            else:
                ids_k, vec_k, met_k = list(ids[:n_ok]), vectors[:n_ok], metadatas[:n_ok]
            if ids_k:
                lens = {len(v) for v in vec_k}
                if len(lens) != 1:
                    raise ValueError(f"Vector dimension mismatch inside the batch: lengths {sorted(lens)}")
                self._append_rows(ids_k, self._host_rows(vec_k), met_k)
            if err is not None:
                raise err

    def add_embeddings_array(self, ids: Sequence[str], vectors: Any, metadatas: Sequence[dict[str, Any]] | None = None) -> None:
        """Extension: rows as one ``[n, dim]`` numpy array / torch tensor (host or this GPU)."""
        n = len(ids)
        metadatas = metadatas if metadatas is not None else [{} for _ in range(n)]
        if not (n == len(vectors) == len(metadatas)):
            raise ValueError("ids, vectors, and metadatas must have the same length")
        with self._lock:
            if n == 0:
                return
            if self._upsert:
                # qdrant_store.py:287-297 semantics: stored ids are replaced; inside a batch the last row of an id wins
                last = {i: j for j, i in enumerate(ids)}
                if len(last) != n:
                    keep = sorted(last.values())
                    vectors = vectors[keep] if hasattr(vectors, "shape") else [vectors[j] for j in keep]
                    ids = [ids[j] for j in keep]
                    metadatas = [metadatas[j] for j in keep]
                    n = len(ids)
                for i in ids:
                    if i in self._row_of:
                        self.delete(i)
            else:
                if len(set(ids)) != n:
                    raise ValueError("duplicate ids inside the batch already exists in the vector store")
                for i in ids:
                    if i in self._row_of:
                        raise ValueError(f"ID '{i}' already exists in the vector store")
            dev_ptr = self._device_tensor_ptr(vectors, want_bf16=self._master == nat.MASTER_BF16)
            if dev_ptr is None:
                self._append_rows(ids, self._host_rows(vectors), metadatas)
                return
            self._flush_deletes()
            self._ensure_index(int(vectors.shape[1]))
            nat.check(self._lib.b200_index_add(self._h, ctypes.c_void_p(dev_ptr), n, 1, self._stream()))
            base = len(self._ids)
            self._ids.extend(ids)
            for j, (i, m) in enumerate(zip(ids, metadatas)):
                self._row_of[i] = base + j
                self._metadata[i] = m.copy()

    def _device_tensor_ptr(self, t: Any, want_bf16: bool) -> int | None:
        """data_ptr of a contiguous CUDA tensor on this store's GPU with the wanted dtype, else None."""
        try:
            import torch
        except Exception:  # pragma: no cover
            return None
        if not isinstance(t, torch.Tensor) or not t.is_cuda:
            return None
        if t.device.index != self._device:
            raise ValueError(f"tensor lives on {t.device}, the store on cuda:{self._device}")
        want = torch.bfloat16 if want_bf16 else torch.float32
        if t.dtype != want or not t.is_contiguous():
            raise ValueError(f"device tensors must be contiguous {want}")
        return int(t.data_ptr())

    def query(self, query_vector: list[float], top_k: int = 10) -> list[SearchResult]:
        """inmemory.py:78-128."""
        with self._lock:
            if not self._row_of:
                return []
            q = np.ascontiguousarray(query_vector, dtype=np.float32).reshape(-1)
            if q.shape[0] != self._dim:
                raise ValueError(
                    f"Query vector dimension ({q.shape[0]}) doesn't match stored vectors dimension ({self._dim})"
                )
            if top_k < 0:
                raise ValueError("top_k must be non-negative")
            k = min(int(top_k), len(self._row_of))
            if k == 0:
                return []
            rows, scores = self._search_host(q[None, :], k)
            rows, scores = rows[0], scores[0]
            if self._metric == "l2":     # faiss_store.py:220: score = 1.0 / (1.0 + float(dist)), a Python float
                scores = 1.0 / (1.0 + scores.astype(np.float64))
            rows = rows[rows >= 0]
            vecs = self._rows_to_host(rows)
            out = []
            for j, r in enumerate(rows):
                i = self._ids[int(r)]
                out.append(SearchResult(id=i, score=float(scores[j]), vector=vecs[j].tolist(),
                                        metadata=self._metadata[i].copy()))
            return out

    def _check_k(self, k: int) -> None:
        limit = MAX_TOP_K_L2 if self._metric == "l2" else MAX_TOP_K
        if k > limit:
            raise RuntimeError(f"top_k={k} is above the supported maximum of {limit}")

    def _search_host(self, q: np.ndarray, k: int) -> tuple[np.ndarray, np.ndarray]:
        self._check_k(k)
        self._flush_deletes()
        nq = q.shape[0]
        ids = np.empty((nq, k), dtype=np.int64)
        sc = np.empty((nq, k), dtype=np.float32)
        fn = self._lib.b200_index_search_l2 if self._metric == "l2" else self._lib.b200_index_search
        nat.check(fn(self._h, _as_ptr(q), 0, 0, nq, k, 0, _as_ptr(ids), _as_ptr(sc), 0, self._stream()))
        return ids, sc

    def _rows_to_host(self, rows: np.ndarray) -> np.ndarray:
        rows = np.ascontiguousarray(rows, dtype=np.int64)
        out = np.empty((len(rows), self._dim or 0), dtype=np.float32)
        if len(rows):
            nat.check(self._lib.b200_index_get_rows(self._h, _as_ptr(rows), len(rows), _as_ptr(out), 0, self._stream()))
        return out

    def query_batch(self, queries: Any, top_k: int = 10, id_base: int = 0):
        """Extension: ``[nq, dim]`` queries -> ``(rows[nq, top_k] int64, scores[nq, top_k] fp32)``.

        Rows are insertion-order indices (+ ``id_base``), -1 where the store holds fewer than
        ``top_k`` vectors; map them with :meth:`ids_for_rows`.  With ``metric="l2"`` the second array holds
        squared distances, ascending (``1 / (1 + d)`` is the score ``query`` reports).  A CUDA tensor in (fp32 or bf16, on
        this store's GPU) gives CUDA tensors out with no host round trip; numpy in gives numpy out."""
        with self._lock:
            if top_k < 0:
                raise ValueError("top_k must be non-negative")
            self._check_k(top_k)
            try:
                import torch
            except Exception:  # pragma: no cover
                torch = None
            if torch is not None and isinstance(queries, torch.Tensor) and queries.is_cuda:
                if queries.device.index != self._device:
                    raise ValueError(f"queries live on {queries.device}, the store on cuda:{self._device}")
                if queries.dtype not in (torch.float32, torch.bfloat16) or not queries.is_contiguous():
                    raise ValueError("device queries must be contiguous float32 or bfloat16")
                if self._metric == "l2" and queries.dtype != torch.float32:
                    raise ValueError("l2 queries must be float32")
                nq, d = queries.shape
                ids = torch.full((nq, top_k), -1, dtype=torch.int64, device=queries.device)
                sc = torch.zeros((nq, top_k), dtype=torch.float32, device=queries.device)
                if self._h is None or not self._row_of or top_k == 0 or nq == 0:
                    return ids, sc
                if d != self._dim:
                    raise ValueError(f"Query vector dimension ({d}) doesn't match stored vectors dimension ({self._dim})")
                self._flush_deletes()
                fn = self._lib.b200_index_search_l2 if self._metric == "l2" else self._lib.b200_index_search
                nat.check(fn(self._h, ctypes.c_void_p(queries.data_ptr()), 0 if queries.dtype == torch.float32 else 1, 1, nq,
                             top_k, id_base, ctypes.c_void_p(ids.data_ptr()), ctypes.c_void_p(sc.data_ptr()), 1, self._stream()))
                return ids, sc
            q = np.ascontiguousarray(queries, dtype=np.float32)
            if q.ndim != 2:
                raise ValueError("queries must be [nq, dim]")
            if self._h is None or not self._row_of or top_k == 0 or q.shape[0] == 0:
                return (np.full((q.shape[0], top_k), -1, np.int64), np.zeros((q.shape[0], top_k), np.float32))
            if q.shape[1] != self._dim:
                raise ValueError(f"Query vector dimension ({q.shape[1]}) doesn't match stored vectors dimension ({self._dim})")
            ids, sc = self._search_host(q, top_k)
            if id_base:
                ids = np.where(ids >= 0, ids + id_base, ids)
            return ids, sc

    def ids_for_rows(self, rows: Any) -> list[list[str | None]]:
        rows = np.asarray(rows.cpu() if hasattr(rows, "cpu") else rows)
        with self._lock:
            return [[self._ids[int(r)] if r >= 0 else None for r in row] for row in rows]

    def metadata_for_ids(self, ids: Sequence[str]) -> dict[str, dict[str, Any]]:
        """Copies of the stored metadata dicts (what ``SearchResult.metadata`` carries, inmemory.py:125)."""
        with self._lock:
            return {i: self._metadata[i].copy() for i in ids}

    def mean_vector_where(self, key: str, value: Any) -> list[float] | None:
        """Arithmetic mean (fp32) of the stored vectors whose metadata has ``metadata[key] == value``, in
        insertion order; None when there is none.  Rows are gathered on the GPU (``b200_index_get_rows`` into
        a device buffer) and averaged there; ``dim`` floats come back."""
        import torch

        with self._lock:
            ids = [i for i in self._ids if i is not None and self._metadata[i].get(key) == value]
            if not ids:
                return None
            self._flush_deletes()
            rows = np.ascontiguousarray([self._row_of[i] for i in ids], dtype=np.int64)
            with torch.cuda.device(self._device):
                out = torch.empty((len(rows), self._dim), dtype=torch.float32, device=f"cuda:{self._device}")
                nat.check(self._lib.b200_index_get_rows(self._h, _as_ptr(rows), len(rows), ctypes.c_void_p(out.data_ptr()), 1,
                                                        self._stream()))
                return out.mean(dim=0).cpu().tolist()

    def delete(self, id: str) -> None:
        """inmemory.py:130-143."""
        with self._lock:
            if id not in self._row_of:
                raise KeyError(f"ID '{id}' not found in vector store")
            r = self._row_of.pop(id)
            del self._metadata[id]
            self._ids[r] = None
            self._dead.append(r)

    def clear(self) -> None:
        """inmemory.py:145-148."""
        with self._lock:
            # the reference forgets the dimension too (its next insert may have any length), so the
            # device index is dropped rather than emptied
            self.close()
            self._dim = None
            self._ids, self._row_of, self._metadata, self._dead = [], {}, {}, []

    def count(self) -> int:
        """inmemory.py:150-156."""
        return len(self._row_of)

    def get(self, id: str) -> SearchResult:
        """inmemory.py:158-178."""
        with self._lock:
            if id not in self._row_of:
                raise KeyError(f"ID '{id}' not found in vector store")
            self._flush_deletes()
            v = self._rows_to_host(np.array([self._row_of[id]], dtype=np.int64))[0]
            return SearchResult(id=id, score=1.0, vector=v.tolist(), metadata=self._metadata[id].copy())

    @property
    def dimension(self) -> int | None:
        return self._dim if self._dim is not None else self._fixed_dim

    @property
    def metric(self) -> str:
        return self._metric

    def certificate_stats(self) -> tuple[int, int]:
        """``(queries checked, queries re-scored exhaustively)`` since the index was created: how often the
        fused select's exactness certificate (see ``b200_index_search`` in the header) sent a query to the
        exhaustive kernel.  Synchronises the device."""
        with self._lock:
            if self._h is None:
                return 0, 0
            a, b = ctypes.c_int64(0), ctypes.c_int64(0)
            nat.check(self._lib.b200_index_certificate_stats(self._h, ctypes.byref(a), ctypes.byref(b)))
            return int(a.value), int(b.value)

    @property
    def native_handle(self) -> ctypes.c_void_p | None:
        """The ``b200_index_t*`` (for the sharded multi-GPU wrapper and the benchmarks)."""
        return self._h
