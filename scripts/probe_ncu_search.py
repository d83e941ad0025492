"""Search kernels at the BASELINE.json configs[2] shape (10 M x 384 bf16 rows): one warm-up and one profiled search at
128 queries (HBM-bound, single-tile kernel) and at 4096 queries (tensor-bound, CTA-pair kernel); for ncu."""
import importlib, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
vs = importlib.import_module("copilot-for-consensus_b200.vectorstore")
dev = torch.device("cuda", 0)
g = torch.Generator(device=dev).manual_seed(1)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
ix = vs.B200Index(384, device=0, master_dtype="bfloat16", capacity=n)
for c0 in range(0, n, 1 << 20):
    ix.add(torch.randn((min(1 << 20, n - c0), 384), generator=g, device=dev).bfloat16().contiguous())
for nq in (128, 4096):
    q = torch.randn((nq, 384), generator=g, device=dev).bfloat16().contiguous()
    for _ in range(2):
        ix.search(q, 10)
    torch.cuda.synchronize()
print("done", ix.certificate_stats())
