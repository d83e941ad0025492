"""Exactness of the fused select is certified per query, not observed (include/b200_retrieval.h,
"Exactness" under b200_index_search): adversarial corpora where the 16-bit tensor-core ranking cannot tell
rows apart must still return the oracle's ids and fp32 score bits, through the exhaustive re-run; and on
ordinary data the certificate must pass nearly always (otherwise the re-run would be the real cost).
Also here: the FAISS-flat (squared L2) mode against its oracle, and two searches on two streams on one handle."""
import numpy as np
import pytest

from conftest import pkg
from oracle import search_oracle as so

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")


@pytest.fixture(scope="module")
def vs():
    return pkg("vectorstore")


def _bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def _near_duplicates(n_dup, d, seed, spread):
    """n_dup rows within `spread` (relative) of one direction: their half-precision images collide or differ
    by one ulp in a few elements, so the 16-bit scores tie or rank them arbitrarily, while the fp32 scores differ."""
    rng = np.random.default_rng(seed)
    base = rng.standard_normal(d).astype(np.float32)
    return (base[None, :] * (1.0 + spread * rng.standard_normal((n_dup, d)))).astype(np.float32)


@pytest.mark.parametrize("k,n_dup,spread", [(10, 64, 2e-5), (10, 200, 1e-4), (24, 96, 2e-5), (48, 128, 5e-5)])
def test_near_duplicate_rows_are_exact(vs, k, n_dup, spread):
    d, n_bg = 384, 30000
    rng = np.random.default_rng(k)
    rows = rng.standard_normal((n_bg, d)).astype(np.float32)
    dup = _near_duplicates(n_dup, d, 7, spread)
    pos = np.sort(rng.choice(n_bg, n_dup, replace=False))
    rows[pos] = dup                                        # scattered over every corpus slice
    qs = np.concatenate([dup[:3] * 1.7, _near_duplicates(3, d, 7, spread), rng.standard_normal((2, d)).astype(np.float32)])
    ix = vs.B200Index(d, device="cuda:0", master_dtype="float32")
    ix.add(rows)
    nat = pkg("_native")
    with nat.options(search_certify=0):
        ids_u, sc_u = ix.search(qs, k)
    ids, sc = ix.search(qs, k)
    oi, os_ = so.topk_canonical(qs, rows, k)
    np.testing.assert_array_equal(ids, oi)
    np.testing.assert_array_equal(_bits(sc), _bits(os_))
    checked, redone = ix.certificate_stats()
    assert checked == len(qs) and redone >= 6, (checked, redone)       # the six near-duplicate queries went to the exact kernel
    # the point of the certificate: without it the same kernels return a wrong neighbour on this corpus
    if n_dup > 64:
        assert (ids_u != oi).any(), "the uncertified path was expected to miss on this corpus"
    ix.close()


@pytest.mark.parametrize("screen", [1, 0])
@pytest.mark.parametrize("k", [10, 48])
def test_near_duplicates_packed_into_few_blocks(vs, k, screen):
    """The exhaustive kernels merge one sorted candidate list per CTA: when several of the true top-k sit in
    the rows of ONE CTA (runs of consecutive near-duplicates, at every position of the grid), every entry of
    that list has to come through, not only its head.  Both passes: screened + full (search_screen=1) and
    full alone (search_screen=0)."""
    d, n_bg = 384, 60000
    rng = np.random.default_rng(100 + k)
    rows = rng.standard_normal((n_bg, d)).astype(np.float32)
    dup = _near_duplicates(96, d, 11, 3e-5)
    # runs of 3..9 consecutive rows scattered over the corpus: 96 rows in 16 runs
    starts = np.sort(rng.choice(n_bg // 16 - 1, 16, replace=False)) * 16
    at, pos = 0, []
    for i, s0 in enumerate(starts):
        n = 3 + (i % 7)
        n = min(n, 96 - at)
        pos.extend(range(int(s0), int(s0) + n)); at += n
    pos.extend(range(n_bg - (96 - at), n_bg))            # whatever is left goes to the last rows
    pos = np.array(pos[:96])
    rows[pos] = dup
    qs = np.concatenate([dup[:4] * 0.3, rng.standard_normal((1, d)).astype(np.float32)])
    ix = vs.B200Index(d, device="cuda:0", master_dtype="float32")
    ix.add(rows)
    nat = pkg("_native")
    with nat.options(search_screen=screen):
        ids, sc = ix.search(qs, k)
    oi, os_ = so.topk_canonical(qs, rows, k)
    np.testing.assert_array_equal(ids, oi)
    np.testing.assert_array_equal(_bits(sc), _bits(os_))
    checked, redone = ix.certificate_stats()
    assert checked == len(qs) and redone >= 4, (checked, redone)
    ix.close()


def test_screened_pass_settles_small_clusters(vs):
    """A cluster smaller than 64 - k leaves the 64th screen value far below the k-th score: the screened pass
    alone must give the oracle's answer (bit for bit), also with a bf16 master copy."""
    d, n_bg, k = 384, 50000, 10
    rng = np.random.default_rng(5)
    rows = rng.standard_normal((n_bg, d)).astype(np.float32)
    dup = _near_duplicates(30, d, 3, 2e-5)
    rows[np.sort(rng.choice(n_bg, 30, replace=False))] = dup
    rows = torch.from_numpy(rows).bfloat16().float().numpy()           # exactly representable in both master types
    qs = np.concatenate([dup[:5] * 2.0, rng.standard_normal((3, d)).astype(np.float32)])
    oi, os_ = so.topk_canonical(qs, rows, k)
    for master in ("float32", "bfloat16"):
        ix = vs.B200Index(d, device="cuda:0", master_dtype=master)
        ix.add(rows)
        ids, sc = ix.search(qs, k)
        np.testing.assert_array_equal(ids, oi)
        np.testing.assert_array_equal(_bits(sc), _bits(os_))
        ix.close()


def test_exact_duplicates_and_zero_rows(vs):
    """All-equal scores (every row identical; zero rows score 0 against everything): ties resolve to the lower row."""
    d = 384
    rng = np.random.default_rng(3)
    v = rng.standard_normal(d).astype(np.float32)
    rows = np.tile(v, (5000, 1))
    rows[1000:1100] = 0.0
    qs = np.stack([v, -v, np.zeros(d, np.float32)])
    ix = vs.B200Index(d, device="cuda:0", master_dtype="float32")
    ix.add(rows)
    ids, sc = ix.search(qs, 10)
    oi, os_ = so.topk_canonical(qs, rows, 10)
    np.testing.assert_array_equal(ids, oi)
    np.testing.assert_array_equal(_bits(sc), _bits(os_))
    ix.close()


def test_certificate_passes_on_ordinary_data(vs):
    d, n, nq = 384, 200000, 2048
    g = torch.Generator(device="cuda").manual_seed(5)
    rows = torch.randn((n, d), generator=g, device="cuda").bfloat16().contiguous()
    qs = torch.randn((nq, d), generator=g, device="cuda").bfloat16().contiguous()
    ix = vs.B200Index(d, device="cuda:0", master_dtype="bfloat16", capacity=n)
    ix.add(rows)
    ids, sc = ix.search(qs, 10)
    checked, redone = ix.certificate_stats()
    assert checked == nq
    assert redone <= nq // 100, f"{redone} of {nq} queries failed the certificate on random data"
    sel = np.arange(0, nq, 97)
    oi, os_ = so.topk_canonical(qs[sel].float().cpu().numpy(), rows.float().cpu().numpy(), 10)
    np.testing.assert_array_equal(ids[sel].cpu().numpy(), oi)
    np.testing.assert_array_equal(_bits(sc[sel].cpu().numpy()), _bits(os_))
    ix.close()


@pytest.mark.parametrize("n,d,nq,k", [(20000, 384, 9, 10), (3000, 100, 4, 64), (5, 384, 2, 10), (70000, 768, 3, 5)])
def test_l2_mode_matches_oracle(vs, n, d, nq, k):
    rng = np.random.default_rng(n)
    rows = rng.standard_normal((n, d)).astype(np.float32)
    rows[n // 2] = rows[0]                                   # an exact tie: the lower row must come first
    qs = rng.standard_normal((nq, d)).astype(np.float32)
    qs[0] = rows[0] + 0.01
    ix = vs.B200Index(d, device="cuda:0", master_dtype="float32")
    ix.add(rows)
    ids, dist = ix.search_l2(qs, k)
    oi, od = so.l2_topk_canonical(qs, rows, k)
    np.testing.assert_array_equal(ids, oi)
    np.testing.assert_array_equal(_bits(dist), _bits(od))
    ids_t, dist_t = ix.search_l2(torch.from_numpy(qs).cuda(), k)
    np.testing.assert_array_equal(ids_t.cpu().numpy(), oi)
    ix.close()


def test_l2_store_scores_like_the_faiss_driver(vs):
    """faiss_store.py:212-224: ascending distance, score = 1 / (1 + dist); same ids / metadata / vector contract."""
    rng = np.random.default_rng(0)
    rows = rng.standard_normal((500, 48)).astype(np.float32)
    st = vs.B200VectorStore(device="cuda:0", metric="l2", dimension=48)
    st.add_embeddings([f"r{i}" for i in range(500)], [r.tolist() for r in rows], [{"i": i} for i in range(500)])
    q = rows[17] + 0.05
    res = st.query(q.tolist(), top_k=5)
    oi, od = so.l2_topk_canonical(q[None], rows, 5)
    assert [r.id for r in res] == [f"r{i}" for i in oi[0]]
    np.testing.assert_allclose([r.score for r in res], so.l2_score(od[0]), rtol=0, atol=0)
    assert res[0].id == "r17" and res[0].metadata == {"i": 17}
    np.testing.assert_array_equal(np.array(res[0].vector, np.float32), rows[17])
    with pytest.raises(ValueError, match="dimension"):
        st.add_embedding("bad", [1.0] * 47, {})
    with pytest.raises(ValueError, match="dimension"):
        st.query([0.0] * 47)
    st.close()


def test_two_streams_on_one_index_do_not_race(vs):
    """Device-pointer searches return while their kernels are queued; the handle orders calls that arrive on
    different streams with an event (include/b200_retrieval.h, conventions).  Both results must be oracle-exact."""
    d, n = 384, 300000
    g = torch.Generator(device="cuda").manual_seed(9)
    rows = torch.randn((n, d), generator=g, device="cuda").bfloat16().contiguous()
    qa = torch.randn((700, d), generator=g, device="cuda").bfloat16().contiguous()
    qb = torch.randn((130, d), generator=g, device="cuda").bfloat16().contiguous()
    ix = vs.B200Index(d, device="cuda:0", master_dtype="bfloat16", capacity=n)
    ix.add(rows)
    torch.cuda.synchronize()
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    outs = []
    for rep in range(3):
        with torch.cuda.stream(s1):
            ra = ix.search(qa, 10)
        with torch.cuda.stream(s2):
            rb = ix.search(qb, 24)
        outs.append((ra, rb))
    torch.cuda.synchronize()
    rows_h = rows.float().cpu().numpy()
    sel_a, sel_b = np.arange(0, 700, 53), np.arange(0, 130, 17)
    oa = so.topk_canonical(qa[sel_a].float().cpu().numpy(), rows_h, 10)
    ob = so.topk_canonical(qb[sel_b].float().cpu().numpy(), rows_h, 24)
    for ra, rb in outs:
        np.testing.assert_array_equal(ra[0][sel_a].cpu().numpy(), oa[0])
        np.testing.assert_array_equal(_bits(ra[1][sel_a].cpu().numpy()), _bits(oa[1]))
        np.testing.assert_array_equal(rb[0][sel_b].cpu().numpy(), ob[0])
        np.testing.assert_array_equal(_bits(rb[1][sel_b].cpu().numpy()), _bits(ob[1]))
    ix.close()
