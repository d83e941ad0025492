"""bert-base-class encoder (768-d, 12 layers, 12 heads x 64, seq 384), one batch of 128: per-kernel breakdown."""
import importlib, json, os, sys, tempfile
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
P = "copilot-for-consensus_b200"
mio = importlib.import_module(P + ".embedding.model_io"); prov_mod = importlib.import_module(P + ".embedding.provider")
nat = importlib.import_module(P + "._native")
with tempfile.TemporaryDirectory() as td:
    mio.write_random_model(td, "bert-base-class", seed=0)
    prov = prov_mod.B200SentenceTransformerProvider(model_name=td, device="cuda:0")
B, S = 128, 384
ids = torch.randint(1000, 30522, (B, S), dtype=torch.int32, device="cuda")
lens = torch.full((B,), S, dtype=torch.int32, device="cuda")
for _ in range(3): prov.embed_tokens(ids, lens)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10): prov.embed_tokens(ids, lens)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 10
nat.profile(True)
for _ in range(3): prov.embed_tokens(ids, lens)
torch.cuda.synchronize()
pr = nat.profile_read(); nat.profile(False)
tot = sum(v["ms"] for v in pr.values())
print(json.dumps({"encode_ms": round(ms, 3), "chunks_per_s": round(B / ms * 1e3), "tflops": round(B * 70.67e9 / (ms * 1e-3) / 1e12, 1)}))
for k, v in pr.items():
    print(json.dumps({"kernel": k, "launches_per_batch": v["launches"] / 3, "ms_per_launch": round(v["ms"] / v["launches"], 4),
                      "share": round(v["ms"] / tot, 3), "tflops": round(v["flops"] / (v["ms"] * 1e9), 1) if v["ms"] else None}))
