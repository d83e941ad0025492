"""Randomised differential test of the search against the canonical oracle on CLUSTERED corpora: many groups of
near-duplicate rows of every size (2 .. 150), so that a large share of the queries fails the candidate-sufficiency
certificate and travels through the screened and the full exhaustive pass, with several failing queries per launch
(the passes reuse their list buffers between consecutive queries).  Bit-exact ids and score bits, every case."""
import numpy as np
import pytest

from conftest import pkg
from oracle import search_oracle as so

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")


def _clustered(rng, n_rows, d, n_clusters, spread):
    rows = rng.standard_normal((n_rows, d)).astype(np.float32)
    centres = []
    at = rng.permutation(n_rows)
    used = 0
    for c in range(n_clusters):
        size = int(rng.choice([2, 3, 7, 20, 40, 70, 150]))
        if used + size > n_rows // 2:
            break
        base = rng.standard_normal(d).astype(np.float32) * float(rng.uniform(0.2, 5.0))
        members = base[None, :] * (1.0 + spread * rng.standard_normal((size, d)).astype(np.float32))
        where = at[used:used + size]
        if c % 3 == 0:                                   # every third cluster: one run of consecutive rows
            s0 = int(rng.integers(0, n_rows - size))
            where = np.arange(s0, s0 + size)
        rows[where] = members
        used += size
        centres.append(base)
    return rows, np.stack(centres)


CASES = [
    # rows,   d,  k, master,     nq, spread, seed
    (40000, 384, 10, "float32", 24, 2e-5, 1),
    (40000, 384, 10, "bfloat16", 24, 1e-3, 2),
    (25000, 384, 48, "float32", 12, 5e-5, 3),
    (25000, 384, 24, "bfloat16", 12, 2e-3, 4),
    (30000, 768, 10, "float32", 16, 2e-5, 5),
    (30000, 64, 5, "float32", 40, 1e-5, 6),
    (20000, 100, 10, "float32", 16, 2e-5, 7),            # d % 8 != 0: no screened pass, generic row loop
    (50000, 384, 1, "float32", 200, 2e-5, 8),            # two query tiles: the CTA-pair kernel
    (9000, 384, 10, "float32", 300, 1e-4, 9),
    (30000, 128, 64, "float32", 6, 2e-5, 10),            # k > 48: the sorting exhaustive path
]


@pytest.mark.parametrize("n_rows,d,k,master,nq,spread,seed", CASES)
def test_clustered_corpus_matches_oracle(n_rows, d, k, master, nq, spread, seed):
    vs, nat = pkg("vectorstore"), pkg("_native")
    rng = np.random.default_rng(seed)
    rows, centres = _clustered(rng, n_rows, d, 60, spread)
    if master == "bfloat16":
        rows = torch.from_numpy(rows).bfloat16().float().numpy()
    pick = rng.integers(0, len(centres), nq)
    qs = centres[pick] * (1.0 + spread * rng.standard_normal((nq, d)).astype(np.float32))
    qs[::5] = rng.standard_normal((len(qs[::5]), d)).astype(np.float32)       # some ordinary queries in between
    oi, os_ = so.topk_canonical(qs, rows, k)
    for screen in (1, 0):
        ix = vs.B200Index(d, device="cuda:0", master_dtype=master)
        for c0 in range(0, n_rows, 16384):                                   # several ingest calls
            ix.add(rows[c0:c0 + 16384])
        with nat.options(search_screen=screen):
            ids, sc = ix.search(qs, k)
        np.testing.assert_array_equal(ids, oi, err_msg=f"screen={screen}")
        np.testing.assert_array_equal(np.ascontiguousarray(sc).view(np.uint32), np.ascontiguousarray(os_).view(np.uint32))
        checked, redone = ix.certificate_stats()
        if k <= 48:
            assert checked == nq
        ix.close()
