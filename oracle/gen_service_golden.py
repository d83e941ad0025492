"""Generate tests/golden/service_batch_golden.json by running the REAL reference methods
``EmbeddingService._generate_batch_embeddings`` / ``_store_embeddings`` (embedding/app/service.py:373-442)
from /root/reference on a fixed list of chunk documents with a deterministic stand-in provider.

Run in the build container only (``python oracle/gen_service_golden.py``); the GPU box has no
/root/reference, so the output is committed and tests/test_service_hooks_cpu.py reads it.
``embedding_date`` is wall-clock in the reference; it is replaced by the placeholder "<now>" here and
the test injects the same placeholder through the hook's ``now`` argument."""
from __future__ import annotations

import glob
import json
import os
import sys
import types

REF = "/root/reference"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.dont_write_bytecode = True
for d in glob.glob(REF + "/adapters/*"):
    if os.path.isdir(d):
        sys.path.insert(0, d)
sys.path.insert(0, REF + "/embedding")

from app.service import EmbeddingService  # noqa: E402


def fake_vector(text: str) -> list[float]:
    """Deterministic, text-dependent 8-d vector (exactly representable values)."""
    h = sum((i + 1) * ord(c) for i, c in enumerate(text)) % 997
    return [float((h * (j + 3)) % 64) / 64.0 for j in range(8)]


class Provider:
    def embed(self, text):
        return fake_vector(text)


class Store:
    def __init__(self):
        self.calls = []

    def add_embeddings(self, ids, vectors, metadatas):
        self.calls.append({"ids": list(ids), "vectors": [list(v) for v in vectors], "metadatas": list(metadatas)})


CHUNKS = [
    {"_id": "c0", "text": "first chunk of the thread", "message_id": "m0", "thread_id": "t0", "archive_id": "a0",
     "chunk_index": 0, "token_count": 5,
     "metadata": {"sender": "a@example.org", "sender_name": "A", "date": "2024-01-01T00:00:00Z", "subject": "hello",
                  "draft_mentions": ["draft-ietf-quic-transport"]}},
    {"_id": "c1", "text": "", "message_id": "m0", "thread_id": "t0", "archive_id": "a0", "chunk_index": 1},
    {"_id": "c2", "text": "only the mandatory keys"},
    {"_id": "c3", "message_id": "m1"},
    {"_id": "c4", "text": "metadata without optional members", "metadata": {"sender": "b@example.org"},
     "chunk_index": 7, "token_count": 4, "thread_id": "t1"},
    {"_id": "c5", "text": "unicode éè 中文 text", "archive_id": "a9", "metadata": {}},
]


def main():
    svc = types.SimpleNamespace(embedding_provider=Provider(), embedding_model="all-MiniLM-L6-v2",
                                embedding_backend="sentencetransformers", vector_store=Store(), metrics_collector=None)
    out = EmbeddingService._generate_batch_embeddings(svc, CHUNKS)
    for e in out:
        assert e["metadata"]["embedding_date"].endswith("Z")
        e["metadata"]["embedding_date"] = "<now>"
    EmbeddingService._store_embeddings(svc, out)
    EmbeddingService._store_embeddings(svc, [])           # no call for an empty slice
    path = os.path.join(ROOT, "tests", "golden", "service_batch_golden.json")
    with open(path, "w", encoding="utf-8") as f:
        json.dump({"chunks": CHUNKS, "embeddings": out, "store_calls": svc.vector_store.calls,
                   "embedding_model": svc.embedding_model, "embedding_backend": svc.embedding_backend}, f, indent=1)
    print("wrote", path, len(out), "embeddings,", len(svc.vector_store.calls), "store call(s)")


if __name__ == "__main__":
    main()
