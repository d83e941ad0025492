"""mpnet-base-class encoder (768-d, 12 layers, 12 heads x 64, seq 384), one batch of 128: for ncu launch lists."""
import importlib, os, sys, tempfile
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
P = "copilot-for-consensus_b200"
mio = importlib.import_module(P + ".embedding.model_io")
prov_mod = importlib.import_module(P + ".embedding.provider")
with tempfile.TemporaryDirectory() as td:
    mio.write_random_model(td, "bert-base-class", seed=0)
    prov = prov_mod.B200SentenceTransformerProvider(model_name=td, device="cuda:0")
B, S = 128, 384
ids = torch.randint(1000, 30522, (B, S), dtype=torch.int32, device="cuda")
lens = torch.full((B,), S, dtype=torch.int32, device="cuda")
for _ in range(2):
    out = prov.embed_tokens(ids, lens)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5):
    out = prov.embed_tokens(ids, lens)
e1.record(); torch.cuda.synchronize()
print("mpnet-class encode ms per batch of 128 x 384:", e0.elapsed_time(e1) / 5)
