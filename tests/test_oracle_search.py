"""The search oracles against golden outputs of the REAL reference (oracle/gen_golden.py ran
InMemoryVectorStore from /root/reference; inmemory.py:78-128)."""
import numpy as np
import pytest

from conftest import bits_to_f32
from oracle import search_oracle as so

CASES = ["rand", "ties", "unit3", "odd3", "odd5", "odd33", "odd100"]


def _case(g, name):
    if name == "rand":
        return bits_to_f32(g["rand_rows_bf16"]), bits_to_f32(g["rand_queries_bf16"]), g["rand_ids"], g["rand_scores"]
    return g[f"{name}_rows"], g[f"{name}_queries"], g[f"{name}_ids"], g[f"{name}_scores"]


def _assert_ids_equal_up_to_near_ties(ids, ref_ids, ref_sc, tag=""):
    """ids identical wherever the reference's own neighbouring scores are separated by more than
    the BLAS-summation-order noise (the reference is itself not reproducible below that)."""
    for a in range(ids.shape[0]):
        for j in range(ids.shape[1]):
            if ids[a, j] != ref_ids[a, j]:
                near = np.abs(ref_sc[a] - ref_sc[a, j]) <= 3e-7
                assert near.sum() > 1, (tag, a, j)


@pytest.mark.parametrize("name", CASES)
def test_canonical_oracle_matches_reference_golden(search_golden, name):
    rows, qs, ref_ids, ref_sc = _case(search_golden, name)
    k = ref_ids.shape[1]
    ids, sc = so.topk_canonical(qs, rows, k)
    # scores: same arithmetic, different summation order -> a few fp32 ulps
    np.testing.assert_allclose(sc, ref_sc, rtol=0, atol=3e-7)
    _assert_ids_equal_up_to_near_ties(ids, ref_ids, ref_sc, name)


def test_golden_cases_have_no_accidental_near_ties(search_golden):
    """So that the comparison above is exact-id on the random cases."""
    for name in ["rand", "odd33", "odd100"]:
        rows, qs, ref_ids, ref_sc = _case(search_golden, name)
        ids, _ = so.topk_canonical(qs, rows, ref_ids.shape[1])
        np.testing.assert_array_equal(ids, ref_ids)


def test_tie_break_is_insertion_order(search_golden):
    g = search_golden
    ids, sc = so.topk_canonical(g["ties_queries"], g["ties_rows"], 10)
    np.testing.assert_array_equal(ids, g["ties_ids"])
    assert list(ids[0, :4]) == [5, 17, 40, 41]          # four identical rows, earliest first
    assert list(ids[1, :2]) == [3, 50]


def test_integration_fixture_vectors(search_golden):
    """embedding/tests/test_integration.py:242-254 -- query [0.1]*384, top_k=2.  Four of the five
    rows are positive multiples of the query, i.e. cosine 1 up to rounding: a genuine near-tie,
    so only the scores (and membership in the tied set) are comparable."""
    g = search_golden
    ids, sc = so.topk_canonical(np.full((1, 384), 0.1, np.float32), g["integ_rows"], 2)
    np.testing.assert_allclose(sc, g["integ_scores"], atol=3e-7)
    assert set(ids[0]) <= {0, 1, 2, 4} and 3 not in ids[0]


def test_numpy_restatement_and_port_agree_with_golden(search_golden):
    rows, qs, ref_ids, ref_sc = _case(search_golden, "rand")
    ids, sc = so.topk_numpy(qs, rows, 10)
    np.testing.assert_array_equal(ids, ref_ids)
    np.testing.assert_allclose(sc, ref_sc, atol=3e-7)
    port = so.InMemoryPort()
    port.add_embeddings([f"r{i}" for i in range(len(rows))], [r.tolist() for r in rows], [{} for _ in rows])
    for a in range(2):
        res = port.query(qs[a].tolist(), top_k=10)
        assert [int(r.id[1:]) for r in res] == list(ref_ids[a])
        np.testing.assert_allclose([r.score for r in res], ref_sc[a], atol=1e-7)


def test_port_error_contract():
    """Exception types and message fragments pinned by test_inmemory.py:27-41,55-60,107-131."""
    p = so.InMemoryPort()
    assert p.query([1.0, 0.0]) == []
    p.add_embedding("a", [1.0, 0.0, 0.0], {"x": 1})
    with pytest.raises(ValueError, match="already exists"):
        p.add_embedding("a", [1.0, 0.0, 0.0], {})
    with pytest.raises(ValueError, match="cannot be empty"):
        p.add_embedding("b", [], {})
    with pytest.raises(ValueError, match="same length"):
        p.add_embeddings(["x", "y"], [[1.0]], [{}, {}])
    with pytest.raises(ValueError, match="dimension"):
        p.query([1.0, 0.0])
    with pytest.raises(KeyError, match="not found"):
        p.delete("zz")
    assert p.get("a").score == 1.0 and p.count() == 1


def test_canonical_edge_cases():
    rows = np.zeros((0, 8), np.float32)
    ids, sc = so.topk_canonical(np.ones((2, 8), np.float32), rows, 3)
    assert (ids == -1).all() and (sc == 0).all()
    rows = np.eye(4, 8, dtype=np.float32)
    ids, sc = so.topk_canonical(rows[1], rows, 10)          # k > n: -1 padded
    assert list(ids[0, :4]) == [1, 0, 2, 3] and (ids[0, 4:] == -1).all()
    assert sc[0, 0] == pytest.approx(1.0, abs=1e-6)
    # zero query / zero row score exactly 0.0, as 0/(0+1e-8) does in the reference
    assert so.score_canonical(np.zeros(8, np.float32), rows[0]) == 0.0


def test_round_bf16_matches_torch():
    import torch

    x = np.random.default_rng(0).standard_normal(10000).astype(np.float32)
    np.testing.assert_array_equal(so.round_bf16(x), torch.from_numpy(x).bfloat16().float().numpy())
