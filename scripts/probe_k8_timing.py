"""Per-tile clocks of the first drain warp of CTA 0 of the similarity kernel (library built with EXTRA=-DB200_K8_TIMING)."""
import ctypes, importlib, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
P = "copilot-for-consensus_b200"
vs = importlib.import_module(P + ".vectorstore"); nat = importlib.import_module(P + "._native")
L = nat.lib()
dev = torch.device("cuda", 0)
g = torch.Generator(device=dev).manual_seed(1)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_250_000
ix = vs.B200Index(384, device=0, master_dtype="bfloat16", capacity=n)
for c0 in range(0, n, 1 << 20):
    ix.add(torch.randn((min(1 << 20, n - c0), 384), generator=g, device=dev).bfloat16().contiguous())
buf = (ctypes.c_ulonglong * 128)()
for nq in ([int(x) for x in sys.argv[2].split(',')] if len(sys.argv) > 2 else (1, 32, 128)):
    q = torch.randn((nq, 384), generator=g, device=dev).bfloat16().contiguous()
    for _ in range(3): ix.search(q, 10)
    L.b200_debug_k8_timing(buf)

    scan = [buf[t] for t in range(36)]; wait = [buf[64 + t] for t in range(36)]
    print("nq", nq, "scan clk per tile:", scan)
    print("nq", nq, "wait clk per tile:", wait, "sum scan", sum(scan), "sum wait", sum(wait), flush=True)
