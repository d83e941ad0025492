"""Where the GEMM MMA issuer spends its clocks (library built with EXTRA=-DB200_GEMM_TIMING)."""
import ctypes, importlib, os, sys, tempfile
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
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
for _ in range(3):
    out = prov.embed_tokens(ids, lens)
torch.cuda.synchronize()
L = nat.lib()
buf = (ctypes.c_ulonglong * 8)()
L.b200_debug_gemm_timing(buf)
out = prov.embed_tokens(ids, lens)
L.b200_debug_gemm_timing(buf)
n = max(1, buf[3])
print("tiles of pair 0 (all GEMMs of one encode):", buf[3])
print("per tile: wait accumulator %.0f  wait operands %.0f  issue %.0f  total %.0f clk" % (buf[0] / n, buf[1] / n, buf[2] / n, buf[4] / n))
