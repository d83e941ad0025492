import ctypes, importlib, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
P = "copilot-for-consensus_b200"
nat = importlib.import_module(P + "._native")
nat.LIB_PATH = os.path.join(ROOT, P, "libb200_stats.so")
vs = importlib.import_module(P + ".vectorstore")
L = nat.lib()
L.b200_debug_sim_stats.argtypes = [ctypes.c_void_p, ctypes.c_int]
n = 10_000_000
dev = torch.device("cuda", 0)
g = torch.Generator(device=dev).manual_seed(1)
idx = vs.B200Index(384, device=0, master_dtype="bfloat16", capacity=n)
for c0 in range(0, n, 1 << 20):
    m = min(1 << 20, n - c0)
    idx.add(torch.randn((m, 384), generator=g, device=dev).bfloat16().contiguous())
buf = (ctypes.c_ulonglong * 4)()
for nq in (1, 32, 128, 4096):
    q = torch.randn((nq, 384), generator=g, device=dev).bfloat16().contiguous()
    idx.search(q, 10); torch.cuda.synchronize()
    L.b200_debug_sim_stats(None, 1)
    idx.search(q, 10); torch.cuda.synchronize()
    L.b200_debug_sim_stats(buf, 0)
    print(f"nq={nq}: chunks={buf[3]} slow-path entries={buf[0]} hit columns={buf[1]} lane inserts={buf[2]}  per query inserts={buf[2]/nq:.0f}", flush=True)
