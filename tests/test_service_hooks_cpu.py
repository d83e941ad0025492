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


# ---- query side: goldens recorded from the REAL reference methods by oracle/gen_query_golden.py
@pytest.fixture(scope="module")
def qgolden():
    with open(os.path.join(GOLDEN, "query_side_golden.json"), encoding="utf-8") as f:
        return json.load(f)


class _TableProvider:
    def __init__(self, table):
        self.table = table

    def embed(self, text):
        return self.table[text]


def _port_store(g):
    from oracle.search_oracle import InMemoryPort

    st = InMemoryPort()
    st.add_embeddings(g["ids"], g["vectors"], g["metadatas"])
    return st


def check_topic_golden(hooks, provider, store, g, atol):
    for key, want in g["search_reports_by_topic"].items():
        topic, limit, min_score = key.split("|")
        got = hooks.rank_threads_by_topic(provider, store, topic, limit=int(limit), min_score=float(min_score))
        assert [t for t, _ in got] == [w["thread_id"] for w in want], key
        for (_t, s), w in zip(got, want):
            assert s["chunk_count"] == w["matching_chunks"] == len(s["chunks"])
            assert abs(s["max_score"] - w["relevance_score"]) <= atol and abs(s["avg_score"] - w["avg_relevance_score"]) <= atol


def test_rank_threads_by_topic_equals_the_reference_method(qgolden):
    """reporting/app/service.py:819-864 through the hook, on the store port: same threads, same order, same max /
    running-mean / count as ReportingService.search_reports_by_topic returned in the build container."""
    hooks = pkg("service_hooks")
    check_topic_golden(hooks, _TableProvider(qgolden["topics"]), _port_store(qgolden), qgolden, atol=3e-7)
    # the batched form falls back to the per-topic loop for adapters that are not the B200 ones
    topics = list(qgolden["topics"])
    many = hooks.rank_threads_by_topics(_TableProvider(qgolden["topics"]), _port_store(qgolden), topics, limit=2, min_score=0.0)
    for t, got in zip(topics, many):
        assert [x for x, _ in got] == [w["thread_id"] for w in qgolden["search_reports_by_topic"][f"{t}|2|0.0"]]
    with pytest.raises(ValueError, match="Topic search requires vector store and embedding provider"):
        hooks.rank_threads_by_topic(None, _port_store(qgolden), "x")
    bad = Mock()
    bad.embed.side_effect = RuntimeError("boom")
    with pytest.raises(ValueError, match="Failed to generate topic embedding: boom"):
        hooks.rank_threads_by_topic(bad, _port_store(qgolden), "x")
    badst = Mock()
    badst.query.side_effect = RuntimeError("down")
    with pytest.raises(ValueError, match="Vector store query failed: down"):
        hooks.rank_threads_by_topic(_TableProvider({"x": [0.0] * 16}), badst, "x")


def test_thread_query_vector_is_the_mean_of_the_threads_rows(qgolden):
    """The vector the orchestrator never passes (orchestrator/app/service.py:559-565): with it, the reference's
    own ThreadChunksSource.get_candidates returned relevance-ordered chunks (golden) instead of neutral 0.5 scores."""
    hooks = pkg("service_hooks")
    st = _port_store(qgolden)
    for t, rec in qgolden["get_candidates"].items():
        q = hooks.thread_candidates_query(st, t, top_k=12, min_score=0.0)
        np.testing.assert_array_equal(np.float32(q["query_vector"]), np.float32(rec["query_vector"]))
        assert q["top_k"] == 12 and q["min_score"] == 0.0
        res = st.query(q["query_vector"], top_k=12)
        mine = [r.id for r in res if r.metadata.get("thread_id") in (None, t)]       # context_sources.py:160-166 thread filter
        assert mine == [c["_id"] for c in rec["candidates"]]
        assert all(c["similarity_score"] == 0.5 for c in rec["fallback"])    # what the services get today
    assert hooks.thread_query_vector(st, "no-such-thread") is None
    assert hooks.thread_candidates_query(st, "no-such-thread") == {"top_k": 50}
    assert hooks.thread_query_vector(Mock(), "tA") is None                   # unknown store type: no vector, reference fallback
