"""Generate tests/golden/*.npz by running the REAL reference from /root/reference.

Run in the build container only (``python oracle/gen_golden.py``); /root/reference does not exist
on the GPU box, so the outputs are committed as small fixtures and every test reads those.

search_golden.npz  -- the real ``InMemoryVectorStore`` (inmemory.py:14-178) on:
    rand   : 600 x 384 bf16-rounded N(0,1) rows, 8 queries, top-10
    ties   : exact duplicate rows at known indices (insertion-order tie-break, inmemory.py:118)
    unit3  : the vectors of the reference's own unit tests (test_inmemory.py:70-91, 180-192) and
             of embedding/tests/test_integration.py:242-254
    odd    : dimensions that are not multiples of 32 (3, 5, 33, 100), raw fp32 rows
encode_golden.npz  -- two pins for the fp32 encode restatement (see oracle/encode_oracle.py):
    hf_*   : installed transformers.BertModel (fp32) + masked mean + L2 normalise, ragged batch
    ref_*  : the reference's in-tree HuggingFaceEmbeddingProvider.embed (huggingface_provider.py:71-103)
             on a locally written random-init model dir, plus the token ids its tokenizer produced
"""
from __future__ import annotations

import importlib
import os
import sys
import tempfile

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"
sys.path.insert(0, ROOT)
sys.dont_write_bytecode = True
for sub in ("adapters/copilot_config", "adapters/copilot_vectorstore", "adapters/copilot_embedding"):
    sys.path.insert(0, os.path.join(REF, sub))

from oracle import encode_oracle as eo  # noqa: E402
from oracle import search_oracle as so  # noqa: E402

pkg = importlib.import_module("copilot-for-consensus_b200.embedding.model_io")


def bf16_bits(a: np.ndarray) -> np.ndarray:
    r = so.round_bf16(a)
    return (r.view(np.uint32) >> 16).astype(np.uint16)


def bits_to_f32(b: np.ndarray) -> np.ndarray:
    return (b.astype(np.uint32) << 16).view(np.float32)


def ref_query(rows: np.ndarray, queries: np.ndarray, k: int):
    from copilot_vectorstore.inmemory import InMemoryVectorStore

    st = InMemoryVectorStore()
    st.add_embeddings([f"r{i}" for i in range(len(rows))], [r.tolist() for r in rows], [{} for _ in rows])
    ids = np.full((len(queries), k), -1, dtype=np.int64)
    sc = np.zeros((len(queries), k), dtype=np.float32)
    for a, q in enumerate(queries):
        res = st.query(q.tolist(), top_k=k)
        for j, r in enumerate(res):
            ids[a, j] = int(r.id[1:])
            sc[a, j] = r.score
    return ids, sc


def gen_search(out: str) -> None:
    rng = np.random.default_rng(1)
    d = {}
    rows = bits_to_f32(bf16_bits(rng.standard_normal((600, 384), dtype=np.float32)))
    qs = bits_to_f32(bf16_bits(np.random.default_rng(2).standard_normal((8, 384), dtype=np.float32)))
    ids, sc = ref_query(rows, qs, 10)
    d.update(rand_rows_bf16=bf16_bits(rows), rand_queries_bf16=bf16_bits(qs), rand_ids=ids, rand_scores=sc)

    # exact duplicates: rows 5, 17, 40, 41 identical; rows 3 and 50 identical; query near both
    t = bits_to_f32(bf16_bits(rng.standard_normal((64, 384), dtype=np.float32)))
    t[17] = t[5]; t[40] = t[5]; t[41] = t[5]; t[50] = t[3]
    tq = np.stack([t[5], t[3], t[5] + t[3]]).astype(np.float32)
    ids, sc = ref_query(t, tq, 10)
    d.update(ties_rows=t, ties_queries=tq, ties_ids=ids, ties_scores=sc)

    # the reference's own unit-test vectors
    u = np.array([[1, 0, 0], [0, 1, 0], [0.5, 0.5, 0]], dtype=np.float32)
    uq = np.array([[0.9, 0.1, 0.0], [1, 0, 0], [0, 0, 1]], dtype=np.float32)
    ids, sc = ref_query(u, uq, 3)
    d.update(unit3_rows=u, unit3_queries=uq, unit3_ids=ids, unit3_scores=sc)
    w = np.full((5, 384), 0.1, dtype=np.float32) * np.arange(1, 6, dtype=np.float32)[:, None]
    w[3, :7] = -0.3
    ids, sc = ref_query(w, np.full((1, 384), 0.1, dtype=np.float32), 2)
    d.update(integ_rows=w, integ_ids=ids, integ_scores=sc)

    for dim in (3, 5, 33, 100):
        r = rng.standard_normal((100, dim), dtype=np.float32)
        r[7] = 0.0  # a zero vector scores 0.0 (0 / (0 + 1e-8))
        q = rng.standard_normal((4, dim), dtype=np.float32)
        ids, sc = ref_query(r, q, 10)
        d.update({f"odd{dim}_rows": r, f"odd{dim}_queries": q, f"odd{dim}_ids": ids, f"odd{dim}_scores": sc})
    np.savez_compressed(out, **d)
    print("wrote", out, {k: v.shape for k, v in d.items()})


def gen_encode(out: str) -> None:
    from transformers import BertConfig, BertModel

    cfg = eo.MINILM_L6
    w = eo.make_weights(cfg, seed=0)
    d = {}
    ids, mask, lens = eo.synthetic_tokens(4, 64, cfg.vocab_size, seed=0, ragged=True)
    hf = BertModel(BertConfig(**pkg.config_to_hf(pkg.KNOWN_ARCHS["all-MiniLM-L6-v2"])), add_pooling_layer=False).eval()
    missing, unexpected = hf.load_state_dict(w, strict=False)
    assert not unexpected and all("position_ids" in m for m in missing), (missing, unexpected)
    with torch.no_grad():
        h = hf(input_ids=ids.long(), attention_mask=mask.long()).last_hidden_state
    emb = eo.pool_normalize(h, mask)
    d.update(hf_input_ids=ids.numpy(), hf_mask=mask.numpy(), hf_embeddings=emb.numpy(),
             hf_last_hidden_row0=h[0, : int(lens[0])].numpy())

    # the reference's own in-tree transformer provider on a local random-init model dir
    from copilot_embedding.huggingface_provider import HuggingFaceEmbeddingProvider

    texts = [
        "The working group reached rough consensus on the QUIC draft; see RFC 9000, section 4.2!",
        "naive cafe resume -- mailing-list thread about DNSSEC key rollover (re: draft-ietf-dnsop)",
        "ok",
    ]
    with tempfile.TemporaryDirectory() as td:
        pkg.write_random_model(td, "all-MiniLM-L6-v2", seed=0)
        prov = HuggingFaceEmbeddingProvider(model_name=td, device="cpu", max_length=256)
        outs, toks = [], []
        for t in texts:
            v = np.array(prov.embed(t), dtype=np.float32)
            outs.append(v / max(np.linalg.norm(v), 1e-12))
            toks.append(prov.tokenizer(t, truncation=True, max_length=256)["input_ids"])
    L = max(len(t) for t in toks)
    tok = np.zeros((len(toks), L), dtype=np.int32)
    for i, t in enumerate(toks):
        tok[i, : len(t)] = t
    d.update(ref_texts=np.array(texts), ref_token_ids=tok, ref_token_lens=np.array([len(t) for t in toks], dtype=np.int32),
             ref_embeddings=np.stack(outs))
    np.savez_compressed(out, **d)
    print("wrote", out, {k: getattr(v, "shape", None) for k, v in d.items()})


if __name__ == "__main__":
    gdir = os.path.join(ROOT, "tests", "golden")
    os.makedirs(gdir, exist_ok=True)
    gen_search(os.path.join(gdir, "search_golden.npz"))
    gen_encode(os.path.join(gdir, "encode_golden.npz"))
