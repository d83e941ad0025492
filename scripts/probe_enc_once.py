import importlib, os, sys, tempfile
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
P = "copilot-for-consensus_b200"
mio = importlib.import_module(P + ".embedding.model_io")
prov_mod = importlib.import_module(P + ".embedding.provider")
with tempfile.TemporaryDirectory() as td:
    mio.write_random_model(td, "all-MiniLM-L6-v2", seed=0)
    prov = prov_mod.B200SentenceTransformerProvider(model_name=td, device="cuda:0")
B, S = 512, 256
ids = torch.randint(1000, 30522, (B, S), dtype=torch.int32, device="cuda")
lens = torch.full((B,), S, dtype=torch.int32, device="cuda")
for _ in range(3):
    out = prov.embed_tokens(ids, lens)
torch.cuda.synchronize()
print("ok", float(out.abs().sum()))
# per-phase clocks of the attention softmax warps (library built with EXTRA=-DB200_ATT_TIMING)
import ctypes
nat = importlib.import_module(P + "._native")
L = nat.lib()
if hasattr(L, "b200_debug_att_timing"):
    buf = (ctypes.c_ulonglong * 16)()
    L.b200_debug_att_timing(buf)
    out = prov.embed_tokens(ids, lens)
    L.b200_debug_att_timing(buf)
    n = max(1, buf[7])
    names = ["reference(next)", "wait scores", "exact max", "wait turn", "exp pass", "wait output", "output"]
    tot = sum(buf[i] for i in range(7))
    print("units(per sampling lane)", buf[7], "clk/unit", tot / n)
    for i, nm in enumerate(names):
        print(f"  {nm:16s} {buf[i]/n:9.1f} clk  {100.0*buf[i]/tot:5.1f}%")
    print("  MMA warp: P V issue %.1f clk per unit" % (buf[13] / max(1, 2 * n)))
    print("  exp pass per warp of group 0:", " ".join("%.0f" % (buf[8 + w] / n) for w in range(8)))
