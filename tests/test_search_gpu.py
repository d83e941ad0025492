"""Parity of the CUDA cosine top-k path with the oracle and with the reference's golden outputs.

Everything here calls through the C ABI (ctypes -> libb200_retrieval.so), either directly or via
the B200VectorStore wrapper.  Bar: ids identical to the canonical oracle (ties to the lower row),
scores bit-identical to it, and within 3e-7 of the real reference's fp32 scores (the reference's
own BLAS summation order is not reproducible below that; see oracle/search_oracle.c)."""
import ctypes

import numpy as np
import pytest

from conftest import bits_to_f32, pkg
from oracle import search_oracle as so

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")


@pytest.fixture(scope="module")
def vs():
    return pkg("vectorstore")


def _store_with(vs, rows, **kw):
    st = vs.B200VectorStore(device="cuda:0", **kw)
    st.add_embeddings([f"r{i}" for i in range(len(rows))], [r.tolist() for r in rows], [{"i": i} for i in range(len(rows))])
    return st


def _assert_matches_oracle(rows_ids, scores, queries, rows, k):
    oi, os_ = so.topk_canonical(queries, rows, k)
    np.testing.assert_array_equal(rows_ids, oi)
    np.testing.assert_array_equal(scores.view(np.uint32), os_.view(np.uint32))  # bit-exact fp32


GOLD = ["rand", "ties", "unit3", "integ", "odd3", "odd5", "odd33", "odd100"]


@pytest.mark.parametrize("name", GOLD)
def test_reference_golden_vectors(vs, search_golden, name):
    g = search_golden
    if name == "rand":
        rows, qs = bits_to_f32(g["rand_rows_bf16"]), bits_to_f32(g["rand_queries_bf16"])
    elif name == "integ":
        rows, qs = g["integ_rows"], np.full((1, 384), 0.1, np.float32)
    else:
        rows, qs = g[f"{name}_rows"], g[f"{name}_queries"]
    ref_ids, ref_sc = g[f"{name}_ids"], g[f"{name}_scores"]
    k = ref_ids.shape[1]
    st = _store_with(vs, rows)
    for a, q in enumerate(qs):
        res = st.query(q.tolist(), top_k=k)
        got_ids = np.array([int(r.id[1:]) for r in res])
        got_sc = np.array([r.score for r in res], dtype=np.float32)
        n = len(res)
        assert n == min(k, len(rows))
        np.testing.assert_allclose(got_sc, ref_sc[a, :n], rtol=0, atol=3e-7)
        for j in range(n):  # ids equal wherever the reference itself is not inside its summation noise
            if got_ids[j] != ref_ids[a, j]:
                assert (np.abs(ref_sc[a] - ref_sc[a, j]) <= 3e-7).sum() > 1, (name, a, j)
        # SearchResult carries the stored vector and a metadata copy (inmemory.py:122-126)
        np.testing.assert_array_equal(np.array(res[0].vector, np.float32), rows[got_ids[0]])
        assert res[0].metadata == {"i": int(got_ids[0])}
    ids, sc = st.query_batch(qs, top_k=k)
    _assert_matches_oracle(ids, sc, qs, rows, k)


def test_ties_resolve_to_insertion_order(vs, search_golden):
    g = search_golden
    st = _store_with(vs, g["ties_rows"])
    res = st.query(g["ties_queries"][0].tolist(), top_k=4)
    assert [r.id for r in res] == ["r5", "r17", "r40", "r41"]
    ids, _ = st.query_batch(g["ties_queries"], top_k=10)
    np.testing.assert_array_equal(ids, g["ties_ids"])


@pytest.mark.parametrize("n,d,nq,k", [(20000, 384, 300, 10), (5000, 384, 129, 24), (3000, 384, 5, 48),
                                      (777, 200, 64, 10), (1, 384, 3, 10), (7, 64, 2, 10), (4097, 96, 130, 10),
                                      # dimensions beyond the resident query tile (streamed mode): mpnet-class 768,
                                      # OpenAI-class 1536, an odd one, the largest accepted
                                      (6000, 768, 200, 10), (3000, 1536, 129, 10), (900, 1000, 5, 48), (600, 4096, 3, 10)])
def test_random_corpora_match_oracle_bit_exact(vs, n, d, nq, k):
    rng = np.random.default_rng(n + d)
    rows = so.round_bf16(rng.standard_normal((n, d), dtype=np.float32))
    qs = so.round_bf16(rng.standard_normal((nq, d), dtype=np.float32))
    st = vs.B200VectorStore(device="cuda:0")
    st.add_embeddings_array([f"r{i}" for i in range(n)], rows)
    ids, sc = st.query_batch(qs, top_k=k)
    _assert_matches_oracle(ids, sc, qs, rows, k)


@pytest.mark.parametrize("n,d,nq,k,master", [(20000, 384, 7, 50, "float32"), (9000, 384, 3, 200, "float32"),
                                             (5000, 96, 2, 1024, "float32"), (30, 384, 2, 50, "float32"),
                                             (6000, 768, 4, 64, "bfloat16")])
def test_large_top_k_exhaustive_path_matches_oracle_bit_exact(vs, n, d, nq, k, master):
    """top_k above the fused select's 48 (orchestrator/app/context_sources.py:66 asks for 50): every row is
    scored exactly; ids, order and fp32 scores equal the oracle's, short lists are padded with -1."""
    rng = np.random.default_rng(n + k)
    rows = so.round_bf16(rng.standard_normal((n, d), dtype=np.float32))
    rows[11] = rows[3]                      # exact duplicates: the lower row wins the tie
    rows[n - 1] = rows[3]
    qs = so.round_bf16(rng.standard_normal((nq, d), dtype=np.float32))
    qs[0] = rows[3]
    st = vs.B200VectorStore(device="cuda:0", master_dtype=master)
    st.add_embeddings_array([f"r{i}" for i in range(n)], rows)
    ids, sc = st.query_batch(qs, top_k=k)
    _assert_matches_oracle(ids, sc, qs, rows, k)
    assert list(ids[0, :3]) == [3, 11, n - 1]
    res = st.query(qs[1].tolist(), top_k=k)             # the reference-shaped call
    assert len(res) == min(k, n) and [r.id for r in res] == [f"r{i}" for i in ids[1, :min(k, n)]]


def test_fp32_inputs_drop_in_mode(vs):
    """Raw fp32 vectors (not bf16-representable): candidates come from the bf16 search copy, the
    final order and scores from the fp32 master copy, so the result still equals the oracle."""
    rng = np.random.default_rng(11)
    rows = rng.standard_normal((6000, 384), dtype=np.float32) * rng.uniform(0.01, 30, (6000, 1)).astype(np.float32)
    rows[123] = 0.0                                   # zero vector scores exactly 0.0
    qs = rng.standard_normal((40, 384), dtype=np.float32)
    st = vs.B200VectorStore(device="cuda:0")
    st.add_embeddings_array([f"r{i}" for i in range(len(rows))], rows)
    ids, sc = st.query_batch(qs, top_k=10)
    _assert_matches_oracle(ids, sc, qs, rows, 10)


def test_bf16_master_and_device_tensors(vs):
    rng = np.random.default_rng(5)
    rows = so.round_bf16(rng.standard_normal((9000, 384), dtype=np.float32))
    qs = so.round_bf16(rng.standard_normal((200, 384), dtype=np.float32))
    st = vs.B200VectorStore(device="cuda:0", master_dtype="bfloat16")
    st.add_embeddings_array([f"r{i}" for i in range(len(rows))], torch.from_numpy(rows).cuda().bfloat16().contiguous())
    ids, sc = st.query_batch(torch.from_numpy(qs).cuda().bfloat16().contiguous(), top_k=10, id_base=1000)
    assert ids.is_cuda and sc.is_cuda
    torch.cuda.synchronize()
    _assert_matches_oracle(ids.cpu().numpy() - 1000, sc.cpu().numpy(), qs, rows, 10)
    ids2, sc2 = st.query_batch(torch.from_numpy(qs).cuda().contiguous(), top_k=10)   # fp32 queries, same values
    torch.cuda.synchronize()
    np.testing.assert_array_equal(ids2.cpu().numpy(), ids.cpu().numpy() - 1000)


def test_delete_compacts_in_insertion_order(vs):
    rng = np.random.default_rng(9)
    rows = so.round_bf16(rng.standard_normal((3000, 128), dtype=np.float32))
    qs = so.round_bf16(rng.standard_normal((16, 128), dtype=np.float32))
    st = vs.B200VectorStore(device="cuda:0")
    names = [f"r{i}" for i in range(len(rows))]
    st.add_embeddings_array(names, rows)
    gone = sorted(set(rng.integers(0, 3000, 400).tolist()) | {0, 1, 2999})
    for i in gone:
        st.delete(f"r{i}")
    with pytest.raises(KeyError, match="not found"):
        st.delete("r0")
    keep = np.array([i for i in range(3000) if i not in set(gone)])
    assert st.count() == len(keep)
    ids, sc = st.query_batch(qs, top_k=10)
    oi, os_ = so.topk_canonical(qs, rows[keep], 10)
    np.testing.assert_array_equal(ids, oi)
    np.testing.assert_array_equal(sc.view(np.uint32), os_.view(np.uint32))
    assert st.ids_for_rows(ids[:1])[0] == [f"r{keep[j]}" for j in oi[0]]
    np.testing.assert_array_equal(np.array(st.get(f"r{keep[5]}").vector, np.float32), rows[keep[5]])
    # re-adding a deleted id appends it at the end of insertion order
    st.add_embedding("r0", rows[0].tolist(), {})
    assert st.ids_for_rows([[st.count() - 1]])[0] == ["r0"]


def test_error_contract_and_edge_cases(vs):
    """Exception types / fragments of test_inmemory.py:27-41,55-68,107-131,146-178."""
    st = vs.B200VectorStore(device="cuda:0")
    assert st.query([1.0, 0.0, 0.0]) == [] and st.count() == 0
    st.add_embedding("a", [1.0, 0.0, 0.0], {"k": [1]})
    with pytest.raises(ValueError, match="already exists"):
        st.add_embedding("a", [1.0, 0.0, 0.0], {})
    with pytest.raises(ValueError, match="cannot be empty"):
        st.add_embedding("b", [], {})
    with pytest.raises(ValueError, match="same length"):
        st.add_embeddings(["x", "y"], [[1.0, 0, 0]], [{}, {}])
    with pytest.raises(ValueError, match="dimension"):
        st.query([1.0, 0.0])
    with pytest.raises(KeyError, match="not found"):
        st.get("zz")
    got = st.get("a")
    assert got.score == 1.0 and got.vector == [1.0, 0.0, 0.0]
    got.metadata["k"] = "changed"                      # metadata is a copy (test_inmemory.py:166-178)
    assert st.get("a").metadata == {"k": [1]}
    st.add_embeddings(["b", "c"], [[0.0, 1.0, 0.0], [0.5, 0.5, 0.0]], [{}, {}])
    res = st.query([0.9, 0.1, 0.0], top_k=2)
    assert [r.id for r in res] == ["a", "c"]           # test_inmemory.py:70-105
    assert len(st.query([0.9, 0.1, 0.0], top_k=10)) == 3 and st.query([0.9, 0.1, 0.0], top_k=0) == []
    assert abs(st.query([1.0, 0.0, 0.0], top_k=1)[0].score - 1.0) < 1e-3
    assert st.query([0.0, 0.0, 1.0], top_k=3)[0].score < 0.1
    with pytest.raises(ValueError, match="already exists"):   # duplicate inside a batch: earlier rows stay
        st.add_embeddings(["d", "e", "d"], [[1.0, 1, 1]] * 3, [{}] * 3)
    assert st.count() == 5
    st.clear()
    assert st.count() == 0 and st.query([1.0, 0.0, 0.0]) == []
    st.add_embedding("z", [1.0, 2.0], {})              # a cleared store accepts a new dimension
    assert st.query([1.0, 2.0])[0].id == "z"
    with pytest.raises(RuntimeError, match="maximum"):
        st.query_batch(np.ones((1, 2), np.float32), top_k=1025)


def test_factory_dispatch(vs):
    cfgm = pkg("config")
    st = vs.create_vector_store(cfgm.AdapterConfig_VectorStore(vector_store_type="InMemory", driver=cfgm.DriverConfig_VectorStore_Inmemory()))
    assert isinstance(st, vs.B200VectorStore) and isinstance(st, vs.VectorStore)
    with pytest.raises(ValueError, match="vector_store config is required"):
        vs.create_vector_store(None)
    with pytest.raises(ValueError, match="Unknown vector_store driver"):
        vs.create_vector_store(cfgm.AdapterConfig_VectorStore(vector_store_type="nope", driver=None))
    with pytest.raises(TypeError, match="Expected inmemory config"):
        vs.create_vector_store(cfgm.AdapterConfig_VectorStore(vector_store_type="inmemory", driver=object()))


def test_merge_topk_abi():
    nat = pkg("_native")
    L = nat.lib()
    rng = np.random.default_rng(3)
    nq, nl, kin, kout = 37, 8, 10, 10
    sc = np.sort(rng.standard_normal((nl, nq, kin)).astype(np.float32), axis=2)[:, :, ::-1].copy()
    sc[1] = sc[0]                                        # equal scores across lists -> lower id wins
    ids = rng.permutation(nl * nq * kin).reshape(nl, nq, kin).astype(np.int64)
    ids[3, :, 7:] = -1                                   # short lists are -1 padded
    d_ids, d_sc = torch.from_numpy(ids).cuda(), torch.from_numpy(sc).cuda()
    o_ids = torch.empty((nq, kout), dtype=torch.int64, device="cuda")
    o_sc = torch.empty((nq, kout), dtype=torch.float32, device="cuda")
    nat.check(L.b200_merge_topk(ctypes.c_void_p(d_ids.data_ptr()), ctypes.c_void_p(d_sc.data_ptr()), nq, nl, kin, kout,
                                ctypes.c_void_p(o_ids.data_ptr()), ctypes.c_void_p(o_sc.data_ptr()), 0, None))
    torch.cuda.synchronize()
    for q in range(nq):
        cand = [(-float(sc[l, q, e]), int(ids[l, q, e])) for l in range(nl) for e in range(kin) if ids[l, q, e] >= 0]
        cand.sort()
        want = cand[:kout]
        assert [int(x) for x in o_ids[q].cpu()] == [c[1] for c in want]
        np.testing.assert_array_equal(o_sc[q].cpu().numpy(), np.array([-c[0] for c in want], np.float32))


@pytest.mark.slow
def test_million_row_corpus_sampled_oracle_and_properties(vs):
    """BASELINE.json configs[1] search half: 512 queries over 1M x 384 bf16.  The oracle checks a
    sample of queries exhaustively; every query is checked through size-independent properties."""
    n, d, nq, k = 1_000_000, 384, 512, 10
    g = torch.Generator(device="cuda").manual_seed(1)
    rows = torch.randn((n, d), generator=g, device="cuda", dtype=torch.float32).bfloat16().contiguous()
    st = vs.B200VectorStore(device="cuda:0", master_dtype="bfloat16", initial_capacity=n)
    st.add_embeddings_array([str(i) for i in range(n)], rows)
    pick = torch.randint(0, n, (nq,), generator=g, device="cuda")
    qs = rows[pick].contiguous()                         # queries are corpus rows: rank 1 is the row itself
    qs[1::2] += torch.randn((nq // 2, d), generator=g, device="cuda").bfloat16() * 0.5
    ids, sc = st.query_batch(qs, top_k=k)
    torch.cuda.synchronize()
    ids_h, sc_h = ids.cpu().numpy(), sc.cpu().numpy()
    assert (ids_h >= 0).all() and (ids_h < n).all()
    assert (np.diff(sc_h, axis=1) <= 0).all()                                   # sorted descending
    assert all(len(set(r)) == k for r in ids_h)                                   # no duplicates
    np.testing.assert_array_equal(ids_h[0::2, 0], pick.cpu().numpy()[0::2])       # self-match first
    np.testing.assert_allclose(sc_h[0::2, 0], 1.0, atol=1e-6)
    rows_h = rows.float().cpu().numpy()
    sample = np.arange(0, nq, 16)
    oi, os_ = so.topk_canonical(qs.float().cpu().numpy()[sample], rows_h, k)
    np.testing.assert_array_equal(ids_h[sample], oi)
    np.testing.assert_array_equal(sc_h[sample].view(np.uint32), os_.view(np.uint32))


@pytest.mark.parametrize("opts", [dict(sim_drain=1), dict(sim_drain=2), dict(sim_pair=0), dict(sim_pair_kp=16, sim_drain=2),
                                  dict(sim_pair_kp=20, sim_drain=2), dict(sim_bound=0), dict(sim_bound=1, sim_drain=2),
                                  dict(sim_nsplit=3, sim_drain=2), dict(sim_refresh_mask=3)])
def test_similarity_kernel_variants_match_oracle(vs, opts):
    """Every switch of the similarity kernel (one / two drain groups, CTA pairs or single CTAs, 16 / 20 kept candidates,
    cross-CTA bounds on / off, forced slice counts, lazier bound refresh) must leave ids and fp32 score bits untouched:
    300 independent queries (three query tiles -> the CTA-pair kernel with a padded tile) over 70 000 rows, a tail tile."""
    nat = pkg("_native")
    rng = np.random.default_rng(42)
    rows = rng.standard_normal((70_013, 384)).astype(np.float32)
    qs = rng.standard_normal((300, 384)).astype(np.float32)
    ix = vs.B200Index(384, device="cuda:0", master_dtype="float32")
    ix.add(rows)
    oi, os_ = so.topk_canonical(qs, rows, 10)
    with nat.options(**opts):
        ids, sc = ix.search(qs, 10)
    np.testing.assert_array_equal(ids, oi)
    np.testing.assert_array_equal(np.ascontiguousarray(sc).view(np.uint32), np.ascontiguousarray(os_).view(np.uint32))
    ix.close()
