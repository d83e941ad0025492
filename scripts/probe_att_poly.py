"""Encode time and attention kernel time against the share of exponentials computed on the FMA pipe (option att_poly)."""
import importlib, json, os, sys, tempfile
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
P = "copilot-for-consensus_b200"
mio = importlib.import_module(P + ".embedding.model_io"); prov_mod = importlib.import_module(P + ".embedding.provider")
nat = importlib.import_module(P + "._native")
with tempfile.TemporaryDirectory() as td:
    mio.write_random_model(td, "all-MiniLM-L6-v2", seed=0)
    prov = prov_mod.B200SentenceTransformerProvider(model_name=td, device="cuda:0")
B, S = 512, 256
ids = torch.randint(1000, 30522, (B, S), dtype=torch.int32, device="cuda")
lens = torch.full((B,), S, dtype=torch.int32, device="cuda")
for rep in range(2):
    for n in (0, 1, 2, 3, 4):
        with nat.options(att_poly=n):
            for _ in range(3): prov.embed_tokens(ids, lens)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(20): prov.embed_tokens(ids, lens)
            e1.record(); torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / 20
            nat.profile(True)
            for _ in range(3): prov.embed_tokens(ids, lens)
            torch.cuda.synchronize()
            pr = nat.profile_read(); nat.profile(False)
        print(json.dumps({"att_poly": n, "encode_ms": round(ms, 4), "attention_ms_per_layer": round(pr["attention"]["ms"] / pr["attention"]["launches"], 4)}), flush=True)
