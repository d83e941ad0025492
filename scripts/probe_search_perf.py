"""GPU diagnostic: per-kernel time of the search path over a 10M x 384 bf16 corpus for several
query-batch sizes (library-internal CUDA-event timing).  Not a test."""
import importlib
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
P = "copilot-for-consensus_b200"
nat = importlib.import_module(P + "._native")
vs = importlib.import_module(P + ".vectorstore")

n = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
dev = torch.device("cuda", 0)
g = torch.Generator(device=dev).manual_seed(1)
idx = vs.B200Index(384, device=0, master_dtype="bfloat16", capacity=n)
for c0 in range(0, n, 1 << 20):
    m = min(1 << 20, n - c0)
    idx.add(torch.randn((m, 384), generator=g, device=dev).bfloat16().contiguous())
settings = [("", "")] if len(sys.argv) < 3 else [tuple(x.split(":")) for x in sys.argv[2].split(",")]
for refresh, bound in settings:
  if refresh != "":
    os.environ["B200_SIM_REFRESH"] = refresh
    os.environ["B200_SIM_BOUND"] = bound
    print(f"== refresh_mask={refresh} bound_mode={bound} (1=buckets 2=single-CTA 3=both)")
  for nq in (1, 32, 128, 512, 4096):
    q = torch.randn((nq, 384), generator=g, device=dev).bfloat16().contiguous()
    for _ in range(2):
        idx.search(q, 10)
    torch.cuda.synchronize()
    nat.profile(True)
    for _ in range(3):
        idx.search(q, 10)
    prof = nat.profile_read()
    nat.profile(False)
    line = "  ".join(f"{k}={v['ms']/v['launches']:.3f}ms" for k, v in prof.items())
    hbm_ms = n * 768 / 6555.2e9 * 1e3
    tc_ms = nq * n * 768 / 1388.2e12 * 1e3
    print(f"nq={nq:5d}: {line}   roofline hbm {hbm_ms:.3f} tensor {tc_ms:.3f} ms", flush=True)
