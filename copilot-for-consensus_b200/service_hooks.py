"""Service-side batching hooks: the callers either side of the hot path.

The reference's embedding service feeds its provider ONE text per call and its store 32 rows per call:

* ``EmbeddingService._generate_batch_embeddings`` -- embedding/app/service.py:373-419: for every chunk with
  text, ``vector = self.embedding_provider.embed(text)``, then a 15-key metadata dict;
* ``EmbeddingService._store_embeddings`` -- embedding/app/service.py:421-442: one
  ``vector_store.add_embeddings(ids, vectors, metadatas)`` per slice.

These functions build the SAME objects (same keys, same order, same skipping of text-less chunks, vectors
as ``list[float]``) but hand the provider the whole slice at once (``embed_batch``: one tokenizer pass, one
GPU batch bucketed by length) and hand the store an array (``add_embeddings_array``: one H2D copy, one ingest
kernel) when the adapters are the B200 ones.  With any other provider / store they fall back to the
reference's per-text / list calls, so a service patched to call them behaves identically against a
``Mock()`` -- batching is detected by ``isinstance``, never by ``hasattr`` (mocks auto-create attributes;
embedding/tests/test_service.py:62-67, 227).

A maintainer would replace the two method bodies with::

    return generate_batch_embeddings(self.embedding_provider, chunks, self.embedding_model, self.embedding_backend)
    store_embeddings(self.vector_store, embeddings, self.metrics_collector)

Query side:

* ``ReportingService.search_reports_by_topic`` -- reporting/app/service.py:819-864: ``embed(topic)`` ->
  ``query(vec, top_k=limit * 3)`` -> group the hits by ``metadata["thread_id"]`` (max / running mean / count)
  -> threads by max score.  :func:`rank_threads_by_topic` is that body up to ``sorted_threads``;
  :func:`rank_threads_by_topics` does it for many topics with ONE encoder batch and ONE batched search.
* ``SummarizationOrchestrator._retrieve_context`` -- orchestrator/app/service.py:559-565 -- never passes the
  ``query_vector`` that ``ThreadChunksSource.get_candidates`` (orchestrator/app/context_sources.py:57-68,102)
  is prepared to use, so "top-k relevance" falls back to the document store with a neutral score.
  :func:`thread_query_vector` supplies one: the mean of the thread's own stored embeddings, computed on the
  GPU from the store's rows; :func:`thread_candidates_query` wraps it into the ``query`` dict.
"""
from __future__ import annotations

from datetime import datetime, timezone
from typing import Any, Callable, Optional, Sequence

import numpy as np

from .embedding.provider import B200SentenceTransformerProvider
from .vectorstore.store import B200VectorStore


def _utc_now_z() -> str:
    # embedding/app/service.py:413 -- ISO-8601 UTC with a trailing "Z"
    return datetime.now(timezone.utc).isoformat().replace("+00:00", "Z")


def chunk_metadata(chunk: dict[str, Any], text: str, embedding_model: str, embedding_backend: str,
                   embedding_date: str) -> dict[str, Any]:
    """The metadata object of embedding/app/service.py:397-414, key for key."""
    meta = chunk.get("metadata", {})
    return {
        "chunk_id": chunk["_id"],
        "message_id": chunk.get("message_id", ""),
        "thread_id": chunk.get("thread_id", ""),
        "archive_id": chunk.get("archive_id", ""),
        "chunk_index": chunk.get("chunk_index", 0),
        "text": text,
        "sender": meta.get("sender", ""),
        "sender_name": meta.get("sender_name", ""),
        "date": meta.get("date", ""),
        "subject": meta.get("subject", ""),
        "draft_mentions": meta.get("draft_mentions", []),
        "token_count": chunk.get("token_count", 0),
        "embedding_model": embedding_model,
        "embedding_backend": embedding_backend,
        "embedding_date": embedding_date,
    }


def generate_batch_embeddings(provider: Any, chunks: Sequence[dict[str, Any]], embedding_model: str,
                              embedding_backend: str, now: Optional[Callable[[], str]] = None,
                              on_skip: Optional[Callable[[dict[str, Any]], None]] = None) -> list[dict[str, Any]]:
    """Drop-in body for ``_generate_batch_embeddings``: ``[{"id", "vector", "metadata"}, ...]`` for every
    chunk whose ``text`` is truthy, in input order.  ``now`` supplies the ``embedding_date`` string (one
    call per emitted chunk, like the reference); ``on_skip`` sees the chunks without text (the reference
    logs a warning there)."""
    now = now or _utc_now_z
    kept: list[tuple[dict[str, Any], str]] = []
    for chunk in chunks:
        text = chunk.get("text", "")
        if not text:                       # service.py:385-387: skipped, not an error
            if on_skip is not None:
                on_skip(chunk)
            continue
        kept.append((chunk, text))
    if not kept:
        return []
    if isinstance(provider, B200SentenceTransformerProvider):
        vectors = provider.embed_batch([t for _, t in kept])          # one batch through the encoder
    else:
        vectors = [provider.embed(t) for _, t in kept]                # the reference's call pattern
    out = []
    for (chunk, text), vector in zip(kept, vectors):
        out.append({"id": chunk["_id"], "vector": vector,
                    "metadata": chunk_metadata(chunk, text, embedding_model, embedding_backend, now())})
    return out


def store_embeddings(vector_store: Any, embeddings: Sequence[dict[str, Any]], metrics_collector: Any = None) -> int:
    """Drop-in body for ``_store_embeddings`` (service.py:421-442): one insert per slice; returns the number
    of rows stored.  The B200 store takes the vectors as one ``[n, dim]`` array."""
    if not embeddings:
        return 0
    ids = [e["id"] for e in embeddings]
    metadatas = [e["metadata"] for e in embeddings]
    if isinstance(vector_store, B200VectorStore):
        vector_store.add_embeddings_array(ids, np.asarray([e["vector"] for e in embeddings], dtype=np.float32), metadatas)
    else:
        vector_store.add_embeddings(ids, [e["vector"] for e in embeddings], metadatas)
    if metrics_collector:
        metrics_collector.increment("vector_store_documents_total", value=len(embeddings))
    return len(embeddings)


# ------------------------------------------------------------------------------------------------
# query side
# ------------------------------------------------------------------------------------------------
def _group_hits_by_thread(hits: Sequence[tuple[str, float, dict[str, Any]]], limit: int, min_score: float):
    """reporting/app/service.py:830-864, statement for statement: hits are (chunk id, score, metadata)."""
    thread_scores: dict[str, dict[str, Any]] = {}
    for chunk_id, score, metadata in hits:
        if score < min_score:
            continue
        thread_id = metadata.get("thread_id")
        if thread_id:
            if thread_id not in thread_scores:
                thread_scores[thread_id] = {"max_score": score, "avg_score": score, "chunk_count": 1, "chunks": [chunk_id]}
            else:
                t = thread_scores[thread_id]
                t["max_score"] = max(t["max_score"], score)
                t["avg_score"] = (t["avg_score"] * t["chunk_count"] + score) / (t["chunk_count"] + 1)
                t["chunk_count"] += 1
                t["chunks"].append(chunk_id)
    return sorted(thread_scores.items(), key=lambda x: x[1]["max_score"], reverse=True)[:limit]


def rank_threads_by_topic(provider: Any, vector_store: Any, topic: str, limit: int = 10,
                          min_score: float = 0.5) -> list[tuple[str, dict[str, Any]]]:
    """Drop-in for the first half of ``search_reports_by_topic`` (reporting/app/service.py:819-864): the
    ``sorted_threads`` list ``[(thread_id, {"max_score", "avg_score", "chunk_count", "chunks"}), ...]``.  Same
    error contract: provider / store failures come back as ``ValueError`` with the reference's messages."""
    if not vector_store or not provider:
        raise ValueError("Topic search requires vector store and embedding provider")
    try:
        topic_embedding = provider.embed(topic)
    except Exception as e:
        raise ValueError(f"Failed to generate topic embedding: {e}")
    try:
        results = vector_store.query(topic_embedding, top_k=limit * 3)
    except Exception as e:
        raise ValueError(f"Vector store query failed: {e}")
    return _group_hits_by_thread([(r.id, r.score, r.metadata) for r in results], limit, min_score)


def rank_threads_by_topics(provider: Any, vector_store: Any, topics: Sequence[str], limit: int = 10,
                           min_score: float = 0.5) -> list[list[tuple[str, dict[str, Any]]]]:
    """``rank_threads_by_topic`` for many topics: with the B200 adapters the topics go through the encoder as
    one batch and through the index as one batched search, embeddings never leaving the GPU; with any other
    adapters it is the per-topic loop.  Returns one ``sorted_threads`` list per topic."""
    if not vector_store or not provider:
        raise ValueError("Topic search requires vector store and embedding provider")
    if not (isinstance(provider, B200SentenceTransformerProvider) and isinstance(vector_store, B200VectorStore)):
        return [rank_threads_by_topic(provider, vector_store, t, limit, min_score) for t in topics]
    for t in topics:
        if t is None or not str(t).strip():      # what embed() raises, wrapped the way the service wraps it
            raise ValueError("Failed to generate topic embedding: Text cannot be empty or whitespace-only"
                             if t is not None else "Failed to generate topic embedding: Text cannot be None")
    if not topics:
        return []
    k = min(limit * 3, vector_store.count())
    if k == 0:
        return [[] for _ in topics]
    try:
        emb = provider.embed_array(list(topics), to_host=False)
    except Exception as e:
        raise ValueError(f"Failed to generate topic embedding: {e}")
    try:
        rows, scores = vector_store.query_batch(emb, top_k=k)
        rows, scores = rows.cpu().numpy(), scores.cpu().numpy()
        ids = vector_store.ids_for_rows(rows)
        metas = vector_store.metadata_for_ids([i for row in ids for i in row if i is not None])
    except Exception as e:
        raise ValueError(f"Vector store query failed: {e}")
    out = []
    for a in range(len(topics)):
        hits = [(i, float(scores[a, j]), metas[i]) for j, i in enumerate(ids[a]) if i is not None]
        out.append(_group_hits_by_thread(hits, limit, min_score))
    return out


def thread_query_vector(vector_store: Any, thread_id: str) -> Optional[list[float]]:
    """The query vector ``ThreadChunksSource.get_candidates`` never receives (orchestrator/app/service.py:559-565;
    context_sources.py:57-64 names "mean of thread embeddings" as the intended one): the arithmetic mean of the
    stored embeddings whose metadata carries this ``thread_id``, or None when the store holds none.  On the B200
    store the rows are gathered and averaged on the GPU; only the ``dim`` floats of the result cross to the host."""
    if isinstance(vector_store, B200VectorStore):
        return vector_store.mean_vector_where("thread_id", thread_id)
    vectors = getattr(vector_store, "_vectors", None)          # the reference's in-memory driver (inmemory.py:26-27)
    metadata = getattr(vector_store, "_metadata", None)
    if not isinstance(vectors, dict) or not isinstance(metadata, dict):
        return None
    rows = [vectors[i] for i, m in metadata.items() if m.get("thread_id") == thread_id]
    if not rows:
        return None
    return np.mean(np.asarray(rows, dtype=np.float32), axis=0, dtype=np.float32).tolist()


def thread_candidates_query(vector_store: Any, thread_id: str, top_k: int = 50,
                            min_score: Optional[float] = None) -> dict[str, Any]:
    """The ``query`` argument for ``context_source.get_candidates(thread_id=..., query=...)`` with a real
    ``query_vector`` in it (orchestrator/app/service.py:563-565 passes only ``top_k``).  Falls back to the
    reference's own dict when the store has no row for the thread."""
    q: dict[str, Any] = {"top_k": top_k}
    vec = thread_query_vector(vector_store, thread_id)
    if vec is not None:
        q["query_vector"] = vec
    if min_score is not None:
        q["min_score"] = min_score
    return q
