"""The service-side hooks with the real B200 adapters: one batch through the encoder and one array insert
must give the same store contents as the reference's per-text / per-row call pattern."""
import tempfile

import numpy as np
import pytest

from conftest import pkg

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")


def test_batched_hooks_equal_per_text_pattern():
    hooks, mio = pkg("service_hooks"), pkg("embedding.model_io")
    prov_mod, vs = pkg("embedding.provider"), pkg("vectorstore")
    with tempfile.TemporaryDirectory() as td:
        mio.write_random_model(td, "all-MiniLM-L6-v2", seed=2)
        prov = prov_mod.B200SentenceTransformerProvider(model_name=td, device="cuda:0")
    words = ["ba", "be", "bi", "bo", "bu", "ca", "ce", "ci", "co", "cu"]
    chunks = [{"_id": f"c{i}", "text": " ".join(words[(i * 7 + j) % 10] for j in range(3 + (i * 13) % 90)),
               "thread_id": f"t{i % 3}", "chunk_index": i, "metadata": {"subject": f"s{i}"}} for i in range(40)]
    chunks.insert(5, {"_id": "empty", "text": ""})
    embs = hooks.generate_batch_embeddings(prov, chunks, "all-MiniLM-L6-v2", "sentencetransformers", now=lambda: "<now>")
    assert [e["id"] for e in embs] == [c["_id"] for c in chunks if c.get("text")]
    # the reference's pattern: one embed() per text (embedding/app/service.py:391)
    per_text = np.array([prov.embed(c["text"]) for c in chunks if c.get("text")], dtype=np.float32)
    got = np.array([e["vector"] for e in embs], dtype=np.float32)
    assert all(isinstance(e["vector"], list) and isinstance(e["vector"][0], float) for e in embs)
    np.testing.assert_allclose(got, per_text, atol=2e-3)     # same weights; only the padding length of the batch differs

    st_a, st_b = vs.B200VectorStore(device="cuda:0"), vs.B200VectorStore(device="cuda:0")
    assert hooks.store_embeddings(st_a, embs) == len(embs)
    st_b.add_embeddings([e["id"] for e in embs], [e["vector"] for e in embs], [e["metadata"] for e in embs])
    assert st_a.count() == st_b.count() == len(embs)
    for e in embs[::7]:
        ra, rb = st_a.query(e["vector"], top_k=5), st_b.query(e["vector"], top_k=5)
        assert [r.id for r in ra] == [r.id for r in rb] and ra[0].id == e["id"]
        assert [r.score for r in ra] == [r.score for r in rb]
        assert ra[0].metadata == e["metadata"] and ra[0].metadata["embedding_date"] == "<now>"
    with pytest.raises(ValueError, match="already exists"):
        hooks.store_embeddings(st_a, embs[:2])               # the in-memory slot raises on an existing id
    prov.close()


def test_query_side_hooks_on_the_b200_store_match_the_reference_goldens():
    """Row f3 with the real B200 store: topic ranking and the thread-mean query vector give what the reference
    methods gave (tests/golden/query_side_golden.json, recorded by oracle/gen_query_golden.py)."""
    import json
    import os

    from conftest import GOLDEN
    from test_service_hooks_cpu import _TableProvider, check_topic_golden

    hooks, vs = pkg("service_hooks"), pkg("vectorstore")
    g = json.load(open(os.path.join(GOLDEN, "query_side_golden.json"), encoding="utf-8"))
    st = vs.B200VectorStore(device="cuda:0")
    st.add_embeddings(g["ids"], g["vectors"], g["metadatas"])
    check_topic_golden(hooks, _TableProvider(g["topics"]), st, g, atol=3e-7)
    for t, rec in g["get_candidates"].items():
        q = hooks.thread_candidates_query(st, t, top_k=12, min_score=0.0)
        np.testing.assert_allclose(q["query_vector"], rec["query_vector"], rtol=0, atol=1e-6)   # GPU mean: other summation order
        res = st.query(rec["query_vector"], top_k=12)
        assert [r.id for r in res if r.metadata.get("thread_id") in (None, t)] == [c["_id"] for c in rec["candidates"]]
        np.testing.assert_allclose([r.score for r in res if r.metadata.get("thread_id") in (None, t)],
                                   [c["similarity_score"] for c in rec["candidates"]], rtol=0, atol=3e-7)
    assert hooks.thread_query_vector(st, "no-such-thread") is None
    st.close()


def test_batched_topic_search_equals_the_per_topic_loop_and_upsert_is_honoured():
    hooks, mio = pkg("service_hooks"), pkg("embedding.model_io")
    prov_mod, vs = pkg("embedding.provider"), pkg("vectorstore")
    with tempfile.TemporaryDirectory() as td:
        mio.write_random_model(td, "all-MiniLM-L6-v2", seed=4)
        prov = prov_mod.B200SentenceTransformerProvider(model_name=td, device="cuda:0")
    words = ["ba", "be", "bi", "bo", "bu", "ca", "ce", "ci", "co", "cu", "da", "de"]
    chunks = [{"_id": f"c{i}", "text": " ".join(words[(i * 5 + j * (1 + i % 3)) % 12] for j in range(4 + (i * 11) % 60)),
               "thread_id": f"t{i % 7}"} for i in range(120)]
    embs = hooks.generate_batch_embeddings(prov, chunks, "m", "b", now=lambda: "<now>")
    st = vs.B200VectorStore(device="cuda:0", upsert=True)
    hooks.store_embeddings(st, embs)
    hooks.store_embeddings(st, embs[:10])            # upsert=True: re-embedding a chunk replaces it
    assert st.count() == 120
    assert st.ids_for_rows(np.arange(110, 120)[None])[0] == [e["id"] for e in embs[:10]]   # replaced rows go to the end
    topics = [c["text"] for c in chunks[::17]] + ["bo bu ca"]
    many = hooks.rank_threads_by_topics(prov, st, topics, limit=4, min_score=0.2)
    for t, got in zip(topics, many):
        one = hooks.rank_threads_by_topic(prov, st, t, limit=4, min_score=0.2)
        assert [x for x, _ in got] == [x for x, _ in one]
        for (_a, sa), (_b, sb) in zip(got, one):
            assert sa["chunks"] == sb["chunks"] and abs(sa["max_score"] - sb["max_score"]) <= 2e-3
    with pytest.raises(ValueError, match="Failed to generate topic embedding"):
        hooks.rank_threads_by_topics(prov, st, ["ok", "  "])
    prov.close()
    st.close()
