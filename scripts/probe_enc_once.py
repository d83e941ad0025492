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
