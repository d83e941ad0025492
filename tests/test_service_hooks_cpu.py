"""Service-side batching hooks against outputs of the REAL reference methods
``EmbeddingService._generate_batch_embeddings`` / ``_store_embeddings`` (embedding/app/service.py:373-442),
recorded by oracle/gen_service_golden.py into tests/golden/service_batch_golden.json."""
import json
import os
from unittest.mock import Mock

import numpy as np
import pytest

from conftest import GOLDEN, pkg


@pytest.fixture(scope="module")
def golden():
    with open(os.path.join(GOLDEN, "service_batch_golden.json"), encoding="utf-8") as f:
        return json.load(f)


def _fake_vector(text):
    # same stand-in as oracle/gen_service_golden.py
    h = sum((i + 1) * ord(c) for i, c in enumerate(text)) % 997
    return [float((h * (j + 3)) % 64) / 64.0 for j in range(8)]


class _Provider:
    def __init__(self):
        self.calls = []

    def embed(self, text):
        self.calls.append(text)
        return _fake_vector(text)


class _Store:
    def __init__(self):
        self.calls = []

    def add_embeddings(self, ids, vectors, metadatas):
        self.calls.append({"ids": list(ids), "vectors": [list(v) for v in vectors], "metadatas": list(metadatas)})


def test_generate_matches_reference_objects_key_for_key(golden):
    hooks = pkg("service_hooks")
    skipped = []
    prov = _Provider()
    out = hooks.generate_batch_embeddings(prov, golden["chunks"], golden["embedding_model"], golden["embedding_backend"],
                                          now=lambda: "<now>", on_skip=skipped.append)
    assert out == golden["embeddings"]
    # same key ORDER too (the reference builds dict literals; JSON keeps insertion order)
    for a, b in zip(out, golden["embeddings"]):
        assert list(a) == list(b) and list(a["metadata"]) == list(b["metadata"])
    assert [c["_id"] for c in skipped] == ["c1", "c3"]              # text-less chunks are skipped, not errors
    assert prov.calls == [e["metadata"]["text"] for e in golden["embeddings"]]   # one embed per kept chunk, in order


def test_default_date_is_utc_iso_z(golden):
    hooks = pkg("service_hooks")
    out = hooks.generate_batch_embeddings(_Provider(), golden["chunks"][:1], "m", "b")
    d = out[0]["metadata"]["embedding_date"]
    assert d.endswith("Z") and "+00:00" not in d and "T" in d


def test_store_matches_reference_call(golden):
    hooks = pkg("service_hooks")
    st = _Store()
    metrics = Mock()
    assert hooks.store_embeddings(st, golden["embeddings"], metrics) == len(golden["embeddings"])
    assert hooks.store_embeddings(st, [], metrics) == 0             # no call for an empty slice
    assert st.calls == golden["store_calls"]
    metrics.increment.assert_called_once_with("vector_store_documents_total", value=len(golden["embeddings"]))


def test_mock_provider_is_not_mistaken_for_a_batching_one(golden):
    """Mock() auto-creates embed_batch; the hook must route by isinstance (test_service.py:62-67, 227)."""
    hooks = pkg("service_hooks")
    m = Mock()
    m.embed.side_effect = _fake_vector
    out = hooks.generate_batch_embeddings(m, golden["chunks"], "all-MiniLM-L6-v2", "sentencetransformers", now=lambda: "<now>")
    assert out == golden["embeddings"]
    assert m.embed.call_count == len(golden["embeddings"]) and not m.embed_batch.called
    ms = Mock()
    hooks.store_embeddings(ms, out)
    assert ms.add_embeddings.call_count == 1 and not ms.add_embeddings_array.called


def test_b200_adapters_take_the_batched_calls(golden):
    """Instances of the B200 classes get ONE embed_batch / ONE add_embeddings_array per slice."""
    hooks = pkg("service_hooks")
    prov_cls = pkg("embedding.provider").B200SentenceTransformerProvider
    store_cls = pkg("vectorstore.store").B200VectorStore

    class P(prov_cls):
        def __init__(self):          # no device, no weights: only the routing is under test
            self.batches = []

        def embed(self, text):
            raise AssertionError("per-text call on the batched path")

        def embed_batch(self, texts):
            self.batches.append(list(texts))
            return [_fake_vector(t) for t in texts]

        def close(self):
            pass

        def __del__(self):
            pass

    class S(store_cls):
        def __init__(self):
            self.array_calls = []

        def add_embeddings(self, ids, vectors, metadatas):
            raise AssertionError("list call on the array path")

        def add_embeddings_array(self, ids, vectors, metadatas=None):
            self.array_calls.append((list(ids), np.asarray(vectors), list(metadatas)))

        def close(self):
            pass

        def __del__(self):
            pass

    p, s = P(), S()
    out = hooks.generate_batch_embeddings(p, golden["chunks"], golden["embedding_model"], golden["embedding_backend"],
                                          now=lambda: "<now>")
    assert out == golden["embeddings"] and len(p.batches) == 1 and len(p.batches[0]) == len(out)
    hooks.store_embeddings(s, out)
    assert len(s.array_calls) == 1
    ids, arr, metas = s.array_calls[0]
    assert ids == golden["store_calls"][0]["ids"] and metas == golden["store_calls"][0]["metadatas"]
    assert arr.dtype == np.float32 and arr.shape == (len(out), 8)
    np.testing.assert_array_equal(arr, np.asarray(golden["store_calls"][0]["vectors"], dtype=np.float32))
