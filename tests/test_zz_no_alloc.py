"""b200_index_search_static -- the search entry point for callers that must not allocate.

Who needs it: a host program that claimed all free HBM for its own pool at start-up (a cudaMalloc
made later from inside this library would fail there), or a query thread that cannot tolerate the
device-wide synchronisation a cudaMalloc/cudaFree brings.  The regular b200_index_search grows
scratch buffers on demand; this variant keeps its scratch in a fixed array inside the library.

The test drives both entry points through the C ABI on the same index and queries and requires
identical ids and bit-identical scores (the bar every search test in this repo uses), on two
different calls in a row so that whatever the first call left in the shared scratch array cannot
leak into the second.  It also checks that free device memory does not move across the calls and
that an oversized request is refused with a message naming the limit.

This file sorts last on purpose: with `pytest -x` every other test has already run if it fails."""
import numpy as np
import pytest

from conftest import pkg

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")


def _bf16_round(a: torch.Tensor) -> torch.Tensor:
    return a.bfloat16().float()


def test_static_scratch_search_equals_regular_search_twice_in_a_row():
    vs = pkg("vectorstore")
    dev = torch.device("cuda", 0)
    g = torch.Generator(device=dev).manual_seed(123)
    rows = _bf16_round(torch.randn((30_000, 384), generator=g, device=dev))
    index = vs.B200Index(384, device=0, master_dtype="float32", capacity=rows.shape[0])
    index.add(rows.contiguous())

    calls = [
        (_bf16_round(torch.randn((100, 384), generator=g, device=dev)).contiguous(), 10),     # 16 kept candidates
        (torch.randn((37, 384), generator=g, device=dev).bfloat16().contiguous(), 24),        # 32 kept, bf16 queries
    ]
    want = [index.search(q, k, id_base=7) for q, k in calls]          # regular entry point (may allocate)
    outs = [(torch.empty((q.shape[0], k), dtype=torch.int64, device=dev),
             torch.empty((q.shape[0], k), dtype=torch.float32, device=dev)) for q, k in calls]
    torch.cuda.synchronize()

    index.search_static(calls[0][0], *outs[0], id_base=7)             # first call on this device: symbol lookup
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info(0)[0]
    for (q, _k), (ids, sc) in zip(calls, outs):                       # two calls in a row, same stream
        ids.fill_(-5)
        sc.fill_(float("nan"))
        index.search_static(q, ids, sc, id_base=7)
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info(0)[0] == free_before, "the no-allocation entry point changed free device memory"

    for (ids, sc), (w_ids, w_sc) in zip(outs, want):
        assert torch.equal(ids, w_ids)
        assert torch.equal(sc.view(torch.int32), w_sc.view(torch.int32))   # bit-identical fp32 scores
    index.close()


def test_static_scratch_search_refuses_what_does_not_fit():
    vs = pkg("vectorstore")
    dev = torch.device("cuda", 0)
    index = vs.B200Index(64, device=0, capacity=1024)
    index.add(torch.randn((500, 64), device=dev).contiguous())
    q = torch.randn((129, 64), device=dev)
    ids = torch.empty((129, 10), dtype=torch.int64, device=dev)
    sc = torch.empty((129, 10), dtype=torch.float32, device=dev)
    with pytest.raises(RuntimeError, match="at most 128 queries"):
        index.search_static(q, ids, sc)
    with pytest.raises(ValueError, match="CUDA tensors"):
        index.search_static(q[:4].cpu(), ids[:4], sc[:4])
    # an empty index answers with "no rows" (-1 ids), as the regular entry point does
    empty = vs.B200Index(64, device=0, capacity=1024)
    empty.search_static(q[:4].contiguous(), ids[:4], sc[:4])
    torch.cuda.synchronize()
    assert (ids[:4] == -1).all() and (sc[:4] == 0).all()
    index.close()
    empty.close()


_FRESH = r"""
import sys
sys.path.insert(0, {root!r})
import importlib
import numpy as np, torch
from oracle import search_oracle as so
vs = importlib.import_module("copilot-for-consensus_b200.vectorstore")
dev = torch.device("cuda", 0)
g = torch.Generator(device=dev).manual_seed(5)
rows = torch.randn((20000, 384), generator=g, device=dev).bfloat16().float().contiguous()
index = vs.B200Index(384, device=0, master_dtype="float32", capacity=rows.shape[0])
index.add(rows)
qs = {{10: torch.randn((50, 384), generator=g, device=dev).bfloat16().float().contiguous(),
      24: torch.randn((20, 384), generator=g, device=dev).bfloat16().contiguous(),
      48: torch.randn((9, 384), generator=g, device=dev).bfloat16().float().contiguous()}}
outs = {{k: (torch.empty((q.shape[0], k), dtype=torch.int64, device=dev),
            torch.empty((q.shape[0], k), dtype=torch.float32, device=dev)) for k, q in qs.items()}}
torch.cuda.synchronize()
index.search_static(qs[10], *outs[10])          # the very first search of this process
torch.cuda.synchronize()
free = torch.cuda.mem_get_info(0)[0]
for k in (24, 48):                              # other k bands / query dtype: nothing may be loaded or allocated now
    index.search_static(qs[k], *outs[k])
torch.cuda.synchronize()
assert torch.cuda.mem_get_info(0)[0] == free, "a later k band changed free device memory"
for k, q in qs.items():
    oi, os_ = so.topk_canonical(q.float().cpu().numpy(), rows.cpu().numpy(), k)
    assert (outs[k][0].cpu().numpy() == oi).all(), k
    assert (outs[k][1].cpu().numpy().view(np.uint32) == os_.view(np.uint32)).all(), k
print("fresh-process static search ok")
"""


def test_static_search_first_thing_in_a_fresh_process():
    """The no-allocation promise has to hold when this entry point is the FIRST search a process makes: its first
    call on a device resolves the scratch array's address and has the driver load every kernel it may ever
    launch, so a later call in another k band or query dtype neither allocates nor loads code.  (The other tests in
    this file run the regular search first, which would hide a kernel that is only loaded lazily.)"""
    import subprocess
    import sys

    from conftest import ROOT

    r = subprocess.run([sys.executable, "-c", _FRESH.format(root=ROOT)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    assert "fresh-process static search ok" in r.stdout
