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
