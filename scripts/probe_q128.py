import importlib, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
vs = importlib.import_module("copilot-for-consensus_b200.vectorstore")
n = int(sys.argv[1]) if len(sys.argv) > 1 else 4_000_000
nq = int(sys.argv[2]) if len(sys.argv) > 2 else 128
dev = torch.device("cuda", 0)
g = torch.Generator(device=dev).manual_seed(1)
idx = vs.B200Index(384, device=0, master_dtype="bfloat16", capacity=n)
for c0 in range(0, n, 1 << 20):
    m = min(1 << 20, n - c0)
    idx.add(torch.randn((m, 384), generator=g, device=dev).bfloat16().contiguous())
q = torch.randn((nq, 384), generator=g, device=dev).bfloat16().contiguous()
for _ in range(4):
    idx.search(q, 10)
torch.cuda.synchronize()
print("done")
