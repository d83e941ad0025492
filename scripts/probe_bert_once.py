"""Two bert-base-class encodes (128 x 384), for profiler captures: ncu -s / -c pick kernels of the second one."""
import importlib, os, sys, tempfile
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
P = "copilot-for-consensus_b200"
mio = importlib.import_module(P + ".embedding.model_io"); prov_mod = importlib.import_module(P + ".embedding.provider")
with tempfile.TemporaryDirectory() as td:
    mio.write_random_model(td, "bert-base-class", seed=0)
    prov = prov_mod.B200SentenceTransformerProvider(model_name=td, device="cuda:0")
ids = torch.randint(1000, 30522, (128, 384), dtype=torch.int32, device="cuda")
lens = torch.full((128,), 384, dtype=torch.int32, device="cuda")
for _ in range(2):
    out = prov.embed_tokens(ids, lens)
torch.cuda.synchronize()
print("ok", float(out.abs().sum()))
