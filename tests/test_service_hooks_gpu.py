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
