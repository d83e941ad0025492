import ctypes, importlib, os, sys, tempfile, time
import torch
sys.path.insert(0, "/root/repo")
P = "copilot-for-consensus_b200"
mio = importlib.import_module(P + ".embedding.model_io")
prov_mod = importlib.import_module(P + ".embedding.provider")
nat = importlib.import_module(P + "._native")
with tempfile.TemporaryDirectory() as td:
    mio.write_random_model(td, "all-MiniLM-L6-v2", seed=0)
    prov = prov_mod.B200SentenceTransformerProvider(model_name=td, device="cuda:0")
B, S = 512, 256
ids = torch.randint(1000, 30522, (B, S), dtype=torch.int32, device="cuda")
lens = torch.full((B,), S, dtype=torch.int32, device="cuda")
L = nat.lib()
buf = (ctypes.c_ulonglong * 8)()
if hasattr(L, "b200_debug_gemm_timing"):
    L.b200_debug_gemm_timing(buf)      # timing build: sets the experiment flag from the environment
for _ in range(3):
    out = prov.embed_tokens(ids, lens)
torch.cuda.synchronize()
nat.profile(True) if hasattr(nat, "profile") else None
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20):
    out = prov.embed_tokens(ids, lens)
e1.record(); torch.cuda.synchronize()
print("encode ms:", e0.elapsed_time(e1) / 20)
