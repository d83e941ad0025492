"""CPU oracles for the cosine top-k path -- TEST INFRASTRUCTURE, not product code.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` leg may import this module.  The product path never does.

Three restatements of ``InMemoryVectorStore.query``
(/root/reference/adapters/copilot_vectorstore/copilot_vectorstore/inmemory.py:78-128):

* :class:`InMemoryPort` -- a line-for-line behavioural port (dict of fp32 rows, one
  Python iteration per stored row, ``np.dot`` / ``np.linalg.norm``).  This is what the
  reference *runs*, so it is the CPU baseline that gets timed (``kind: "port"``).
* :func:`topk_numpy` -- the vectorised fp32 restatement SURVEY.md section 8c names
  (``Xn @ qn`` + stable argsort); fast, BLAS summation order.
* :func:`topk_canonical` -- the C library ``libsearch_oracle.so`` (search_oracle.c):
  the same arithmetic with ONE fixed summation order.  The CUDA rescoring kernel
  reproduces it bit for bit, which is what makes "ids identical" well-posed under
  fp32 near-ties.  This is the parity checker for the GPU path.

All three are pinned against the real reference by ``oracle/gen_golden.py`` and
``tests/test_oracle_search.py``.
"""
from __future__ import annotations

import ctypes
import os
import subprocess
from typing import Any

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libsearch_oracle.so")
_lib = None


def build(force: bool = False) -> str:
    """Compile search_oracle.c with the committed Makefile (building the checker is not using it)."""
    src = os.path.join(_HERE, "search_oracle.c")
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src):
        subprocess.check_call(["make", "-s", "-C", _HERE, "libsearch_oracle.so"])
    return _LIB_PATH


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        build()
        L = ctypes.CDLL(_LIB_PATH)
        f32p = ctypes.POINTER(ctypes.c_float)
        i64p = ctypes.POINTER(ctypes.c_int64)
        L.b200o_dot.restype = ctypes.c_float
        L.b200o_dot.argtypes = [f32p, f32p, ctypes.c_int]
        L.b200o_norm.restype = ctypes.c_float
        L.b200o_norm.argtypes = [f32p, ctypes.c_int]
        L.b200o_score.restype = ctypes.c_float
        L.b200o_score.argtypes = [f32p, f32p, ctypes.c_int]
        L.b200o_scores.restype = None
        L.b200o_scores.argtypes = [f32p, f32p, ctypes.c_int64, ctypes.c_int, f32p]
        L.b200o_topk.restype = ctypes.c_int
        L.b200o_topk.argtypes = [f32p, ctypes.c_int, f32p, ctypes.c_int64, ctypes.c_int,
                                 ctypes.c_int, i64p, f32p]
        L.b200o_l2_topk.restype = ctypes.c_int
        L.b200o_l2_topk.argtypes = [f32p, ctypes.c_int, f32p, ctypes.c_int64, ctypes.c_int,
                                    ctypes.c_int, i64p, f32p]
        L.b200o_round_bf16.restype = None
        L.b200o_round_bf16.argtypes = [f32p, ctypes.c_int64, f32p]
        _lib = L
    return _lib


def _f32(a) -> np.ndarray:
    return np.ascontiguousarray(a, dtype=np.float32)


def _p(a: np.ndarray, t=ctypes.c_float):
    return a.ctypes.data_as(ctypes.POINTER(t))


def dot_canonical(x, y) -> float:
    x, y = _f32(x), _f32(y)
    assert x.ndim == 1 and x.shape == y.shape
    return float(lib().b200o_dot(_p(x), _p(y), x.shape[0]))


def score_canonical(q, v) -> float:
    q, v = _f32(q), _f32(v)
    return float(lib().b200o_score(_p(q), _p(v), q.shape[0]))


def scores_canonical(q, rows) -> np.ndarray:
    q, rows = _f32(q), _f32(rows)
    n, d = rows.shape
    out = np.empty(n, dtype=np.float32)
    lib().b200o_scores(_p(q), _p(rows), n, d, _p(out))
    return out


def topk_canonical(queries, rows, k: int) -> tuple[np.ndarray, np.ndarray]:
    """(ids[nq,k] int64 row indices, -1 padded; scores[nq,k] fp32) under the canonical order."""
    queries, rows = _f32(queries), _f32(rows)
    if queries.ndim == 1:
        queries = queries[None, :]
    nq, d = queries.shape
    n = rows.shape[0]
    assert rows.ndim == 2 and (n == 0 or rows.shape[1] == d)
    ids = np.empty((nq, k), dtype=np.int64)
    sc = np.empty((nq, k), dtype=np.float32)
    rc = lib().b200o_topk(_p(queries), nq, _p(rows), n, d, k, _p(ids, ctypes.c_int64), _p(sc))
    if rc != 0:
        raise ValueError("b200o_topk: bad arguments")
    return ids, sc


def l2_topk_canonical(queries, rows, k: int) -> tuple[np.ndarray, np.ndarray]:
    """FAISS-flat mode (faiss_store.py:99-101,212-224): ``(ids[nq,k], squared L2 distances[nq,k])``,
    ascending distance, ties to the lower row, canonical fp32 summation order."""
    queries, rows = _f32(queries), _f32(rows)
    if queries.ndim == 1:
        queries = queries[None, :]
    nq, d = queries.shape
    n = rows.shape[0]
    ids = np.empty((nq, k), dtype=np.int64)
    dist = np.empty((nq, k), dtype=np.float32)
    rc = lib().b200o_l2_topk(_p(queries), nq, _p(rows), n, d, k, _p(ids, ctypes.c_int64), _p(dist))
    if rc != 0:
        raise ValueError("b200o_l2_topk: bad arguments")
    return ids, dist


def l2_topk_numpy(queries, rows, k: int) -> tuple[np.ndarray, np.ndarray]:
    """What ``faiss.IndexFlatL2.search`` defines (exact squared Euclidean distance, ascending), restated in
    float64 numpy.  faiss is not installable here, so this restatement -- not faiss -- pins the canonical one."""
    q = np.asarray(queries, dtype=np.float64)
    x = np.asarray(rows, dtype=np.float64)
    if q.ndim == 1:
        q = q[None, :]
    ids = np.full((q.shape[0], k), -1, dtype=np.int64)
    dist = np.zeros((q.shape[0], k), dtype=np.float64)
    for a in range(q.shape[0]):
        dd = ((x - q[a]) ** 2).sum(axis=1)
        order = np.argsort(dd, kind="stable")[:k]
        ids[a, : len(order)] = order
        dist[a, : len(order)] = dd[order]
    return ids, dist


def l2_score(dist) -> np.ndarray:
    """faiss_store.py:220: ``score = 1.0 / (1.0 + float(dist))`` (Python float = float64 of the fp32 distance)."""
    return 1.0 / (1.0 + np.asarray(dist, dtype=np.float32).astype(np.float64))


def round_bf16(a) -> np.ndarray:
    """fp32 -> bf16 (round to nearest even) -> fp32, so oracle and GPU see identical inputs."""
    a = _f32(a)
    out = np.empty_like(a)
    lib().b200o_round_bf16(_p(a.reshape(-1)), a.size, _p(out.reshape(-1)))
    return out


def topk_numpy(queries, rows, k: int) -> tuple[np.ndarray, np.ndarray]:
    """Vectorised fp32 restatement of inmemory.py:96-119 (BLAS summation order)."""
    queries, rows = _f32(queries), _f32(rows)
    if queries.ndim == 1:
        queries = queries[None, :]
    norms = np.sqrt(np.einsum("ij,ij->i", rows, rows, dtype=np.float32)).astype(np.float32)
    xn = rows / (norms + np.float32(1e-8))[:, None]
    ids = np.full((queries.shape[0], k), -1, dtype=np.int64)
    sc = np.zeros((queries.shape[0], k), dtype=np.float32)
    for a, q in enumerate(queries):
        qn = q / (np.float32(np.linalg.norm(q)) + np.float32(1e-8))
        s = xn @ qn
        order = np.argsort(-s, kind="stable")[:k]
        ids[a, : len(order)] = order
        sc[a, : len(order)] = s[order]
    return ids, sc


class PortSearchResult:
    """interface.py:11-25 -- id, score, vector, metadata."""

    __slots__ = ("id", "score", "vector", "metadata")

    def __init__(self, id: str, score: float, vector: list[float], metadata: dict[str, Any]):
        self.id, self.score, self.vector, self.metadata = id, score, vector, metadata

    def __repr__(self) -> str:  # pragma: no cover
        return f"PortSearchResult(id={self.id!r}, score={self.score!r})"


class InMemoryPort:
    """Behavioural port of the reference driver (inmemory.py:14-178), used as the timed CPU
    baseline on machines where /root/reference does not exist.  Same data structures (two
    dicts in insertion order), same per-row Python loop, same numpy calls, same errors."""

    def __init__(self) -> None:
        self._vectors: dict[str, np.ndarray] = {}
        self._metadata: dict[str, dict[str, Any]] = {}

    def add_embedding(self, id: str, vector, metadata: dict[str, Any]) -> None:  # inmemory.py:35-53
        if id in self._vectors:
            raise ValueError(f"ID '{id}' already exists in the vector store")
        if vector is None or len(vector) == 0:
            raise ValueError("Vector cannot be empty")
        self._vectors[id] = np.array(vector, dtype=np.float32)
        self._metadata[id] = metadata.copy()

    def add_embeddings(self, ids, vectors, metadatas) -> None:  # inmemory.py:55-76
        if not (len(ids) == len(vectors) == len(metadatas)):
            raise ValueError("ids, vectors, and metadatas must have the same length")
        for id in ids:
            if id in self._vectors:
                raise ValueError(f"ID '{id}' already exists in the vector store")
        for id, vector, metadata in zip(ids, vectors, metadatas):
            self.add_embedding(id, vector, metadata)

    def query(self, query_vector, top_k: int = 10) -> list[PortSearchResult]:  # inmemory.py:78-128
        if not self._vectors:
            return []
        q = np.array(query_vector, dtype=np.float32)
        first = next(iter(self._vectors.values()))
        if len(q) != len(first):
            raise ValueError(
                f"Query vector dimension ({len(q)}) doesn't match stored vectors dimension ({len(first)})"
            )
        qn = q / (np.linalg.norm(q) + 1e-8)
        sims = []
        for eid, v in self._vectors.items():
            vn = v / (np.linalg.norm(v) + 1e-8)
            sims.append((eid, float(np.dot(qn, vn))))
        sims.sort(key=lambda x: x[1], reverse=True)
        return [
            PortSearchResult(i, s, self._vectors[i].tolist(), self._metadata[i].copy())
            for i, s in sims[:top_k]
        ]

    def delete(self, id: str) -> None:  # inmemory.py:130-143
        if id not in self._vectors:
            raise KeyError(f"ID '{id}' not found in vector store")
        del self._vectors[id]
        del self._metadata[id]

    def clear(self) -> None:
        self._vectors.clear()
        self._metadata.clear()

    def count(self) -> int:
        return len(self._vectors)

    def get(self, id: str) -> PortSearchResult:  # inmemory.py:158-178
        if id not in self._vectors:
            raise KeyError(f"ID '{id}' not found in vector store")
        return PortSearchResult(id, 1.0, self._vectors[id].tolist(), self._metadata[id].copy())
