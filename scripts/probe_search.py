"""GPU diagnostic for the search path: progressively larger cases, prints where the CUDA result
first departs from the oracle.  Not a test; used while bringing kernels up."""
import importlib
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import search_oracle as so  # noqa: E402

vs = importlib.import_module("copilot-for-consensus_b200.vectorstore")


def case(n, d, nq, k, seed=0):
    rng = np.random.default_rng(seed)
    rows = so.round_bf16(rng.standard_normal((n, d), dtype=np.float32))
    qs = so.round_bf16(rng.standard_normal((nq, d), dtype=np.float32))
    st = vs.B200VectorStore(device="cuda:0")
    st.add_embeddings_array([f"r{i}" for i in range(n)], rows)
    t = time.time()
    ids, sc = st.query_batch(qs, top_k=k)
    dt = time.time() - t
    oi, os_ = so.topk_canonical(qs, rows, k)
    same_ids = (ids == oi).all()
    same_sc = (sc.view(np.uint32) == os_.view(np.uint32)).all()
    print(f"n={n} d={d} nq={nq} k={k}: ids_equal={same_ids} scores_bit_equal={same_sc} ({dt*1e3:.1f} ms)", flush=True)
    if not (same_ids and same_sc):
        bad = np.argwhere(ids != oi)
        print("  first mismatches (query, rank):", bad[:8].tolist())
        q = int(bad[0][0]) if len(bad) else 0
        print("  gpu ids   ", ids[q].tolist())
        print("  oracle ids", oi[q].tolist())
        print("  gpu sc    ", sc[q].tolist())
        print("  oracle sc ", os_[q].tolist())
        print("  max |score diff| where ids agree:", float(np.abs(np.where(ids == oi, sc - os_, 0)).max()))
    return same_ids and same_sc


if __name__ == "__main__":
    ok = True
    for args in [(256, 64, 1, 10), (256, 384, 1, 10), (300, 384, 5, 10), (4096, 384, 128, 10), (50000, 384, 300, 10),
                 (5000, 384, 40, 24), (5000, 384, 40, 48), (100, 3, 4, 10), (200000, 384, 1024, 10)]:
        ok = case(*args) and ok
    print("ALL OK" if ok else "MISMATCHES")
    sys.exit(0 if ok else 1)
