"""Generate tests/golden/query_side_golden.json by running the REAL reference methods on fixed inputs:

* ``ReportingService.search_reports_by_topic`` (reporting/app/service.py:797-898) with the reference's own
  ``InMemoryVectorStore`` and a deterministic stand-in provider / document store;
* ``ThreadChunksSource.get_candidates`` (orchestrator/app/context_sources.py:40-186) with a ``query_vector``
  equal to the mean of the thread's stored embeddings -- the vector the orchestrator never passes
  (orchestrator/app/service.py:559-565).

Run in the build container only (``python oracle/gen_query_golden.py``); the GPU box has no /root/reference, so
the output is committed and tests/test_service_hooks_cpu.py / test_service_hooks_gpu.py read it."""
from __future__ import annotations

import glob
import importlib.util
import json
import os
import sys
import types

import numpy as np

REF = "/root/reference"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.dont_write_bytecode = True
for d in glob.glob(REF + "/adapters/*"):
    if os.path.isdir(d):
        sys.path.insert(0, d)


def _load(name: str, path: str, package_dir: str):
    """Import <service>/app/<module>.py under a unique package name (every service calls its package ``app``)."""
    pkg = types.ModuleType(name.split(".")[0])
    pkg.__path__ = [package_dir]
    sys.modules[name.split(".")[0]] = pkg
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


DIM = 16
THREADS = {"tA": 9, "tB": 7, "tC": 5, "tD": 3}


def corpus():
    """Chunks of four threads: each thread is a cluster around its own direction, so topic and thread-mean queries
    have an unambiguous answer; values are multiples of 1/256 so every implementation sees identical inputs."""
    rng = np.random.default_rng(42)
    centers = {t: rng.standard_normal(DIM) for t in THREADS}
    ids, vecs, metas = [], [], []
    for t, n in THREADS.items():
        for j in range(n):
            v = centers[t] + 0.35 * rng.standard_normal(DIM)
            v = np.round(v * 256) / 256
            ids.append(f"{t}-c{j}")
            vecs.append(v.astype(np.float32).tolist())
            metas.append({"thread_id": t, "chunk_id": f"{t}-c{j}", "chunk_index": j})
    # chunks without a thread id are ignored by the grouping (service.py:841)
    ids.append("orphan")
    vecs.append((np.round(centers["tA"] * 256) / 256).astype(np.float32).tolist())
    metas.append({"chunk_id": "orphan"})
    topics = {f"topic about {t}": (np.round((centers[t] + 0.2 * rng.standard_normal(DIM)) * 256) / 256).astype(np.float32).tolist()
              for t in THREADS}
    return ids, vecs, metas, topics


class Provider:
    def __init__(self, table):
        self.table = table

    def embed(self, text):
        return self.table[text]


class DocStore:
    """query_documents for the two collections the methods touch."""

    def __init__(self, chunks):
        self.chunks = chunks

    def query_documents(self, collection, filter_dict=None, limit=100, **_):
        filter_dict = filter_dict or {}
        if collection == "summaries":
            return [{"thread_id": filter_dict["thread_id"], "summary": "s-" + filter_dict["thread_id"]}]
        if collection == "chunks":
            if "_id" in filter_dict:
                want = set(filter_dict["_id"]["$in"])
                return [dict(c) for c in self.chunks if c["_id"] in want][:limit]
            return [dict(c) for c in self.chunks if c.get("thread_id") == filter_dict.get("thread_id")][:limit]
        return []


def main():
    from copilot_vectorstore.inmemory import InMemoryVectorStore

    rep = _load("reporting_app.service", REF + "/reporting/app/service.py", REF + "/reporting/app")
    ids, vecs, metas, topics = corpus()
    store = InMemoryVectorStore()
    store.add_embeddings(ids, vecs, metas)
    chunks = [{"_id": i, "thread_id": m.get("thread_id"), "text": "text of " + i} for i, m in zip(ids, metas)]
    docs = DocStore(chunks)

    svc = types.SimpleNamespace(vector_store=store, embedding_provider=Provider(topics), document_store=docs)
    svc.get_thread_summary = lambda tid: rep.ReportingService.get_thread_summary(svc, tid)
    svc._get_thread_by_id = lambda tid: None
    svc._get_archive_by_id = lambda aid: None
    topic_out = {}
    for topic in topics:
        for limit, min_score in [(10, 0.5), (2, 0.0), (3, 0.8)]:
            r = rep.ReportingService.search_reports_by_topic(svc, topic, limit=limit, min_score=min_score)
            topic_out[f"{topic}|{limit}|{min_score}"] = [
                {"thread_id": x["thread_id"], "relevance_score": x["relevance_score"],
                 "avg_relevance_score": x["avg_relevance_score"], "matching_chunks": x["matching_chunks"]} for x in r]

    ctx = _load("orchestrator_app.context_sources", REF + "/orchestrator/app/context_sources.py", REF + "/orchestrator/app")
    src = ctx.ThreadChunksSource(document_store=docs, vector_store=store)
    cand_out = {}
    for t in THREADS:
        rows = np.asarray([v for v, m in zip(vecs, metas) if m.get("thread_id") == t], dtype=np.float32)
        qv = np.mean(rows, axis=0, dtype=np.float32).tolist()
        got = src.get_candidates(t, {"query_vector": qv, "top_k": 12, "min_score": 0.0})
        cand_out[t] = {"query_vector": qv, "candidates": [{"_id": c["_id"], "similarity_score": c["similarity_score"]} for c in got]}
        fallback = src.get_candidates(t, {"top_k": 12})
        cand_out[t]["fallback"] = [{"_id": c["_id"], "similarity_score": c["similarity_score"]} for c in fallback]

    path = os.path.join(ROOT, "tests", "golden", "query_side_golden.json")
    with open(path, "w", encoding="utf-8") as f:
        json.dump({"ids": ids, "vectors": vecs, "metadatas": metas, "topics": topics, "search_reports_by_topic": topic_out,
                   "get_candidates": cand_out}, f, indent=1)
    print("wrote", path, len(topic_out), "topic searches,", len(cand_out), "threads")


if __name__ == "__main__":
    main()
