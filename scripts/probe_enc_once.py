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
    n = max(1, buf[5])
    names = ["wait_S", "pass1", "pass2", "wait_O", "output", None, "wait_turn"]
    tot = sum(buf[i] for i in range(7) if i != 5)
    print("units(per sampling lane)", buf[5], "clk/unit", tot / n)
    for i, nm in enumerate(names):
        if nm is None:
            continue
        print(f"  {nm:8s} {buf[i]/n:9.1f} clk  {100.0*buf[i]/tot:5.1f}%")
    print("  pass 2 per unit: ld-wait %.1f  probs+st issue %.1f  tail (st wait, arrive) %.1f" % (buf[8]/n, buf[9]/n, buf[10]/n))
    print("  output: O ld %.1f  pair barrier %.1f ; MMA warp PV issue %.1f" % (buf[11]/n, buf[12]/n, buf[13]/n))
    print("  output: reciprocal %.1f" % (buf[14]/n))
