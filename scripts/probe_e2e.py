"""Where the end-to-end (host buffers) step spends its time beyond the kernels."""
import importlib, os, sys, tempfile, time
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
P = "copilot-for-consensus_b200"
mio = importlib.import_module(P + ".embedding.model_io")
prov_mod = importlib.import_module(P + ".embedding.provider")
vs = importlib.import_module(P + ".vectorstore")
with tempfile.TemporaryDirectory() as td:
    mio.write_random_model(td, "all-MiniLM-L6-v2", seed=0)
    prov = prov_mod.B200SentenceTransformerProvider(model_name=td, device="cuda:0")
B, S, N = 512, 256, 1_000_000
g = torch.Generator(device="cuda").manual_seed(0)
index = vs.B200Index(384, device=0, master_dtype="bfloat16", capacity=N)
index.add(torch.randn((N, 384), generator=g, device="cuda").bfloat16().contiguous())
ids = torch.randint(1000, 30522, (B, S), dtype=torch.int32).pin_memory()
lens = torch.full((B,), S, dtype=torch.int32).pin_memory()
d_ids, d_lens = ids.cuda(), lens.cuda()

def t(fn, n=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n):
        r = fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / n * 1e3, r

ms_dev, emb_d = t(lambda: prov.embed_tokens(d_ids, d_lens))
ms_host, emb_h = t(lambda: prov.embed_tokens(ids.numpy(), lens.numpy()))
ms_sdev, _ = t(lambda: index.search(emb_d, 10))
ms_shost, _ = t(lambda: index.search(emb_h, 10))
print(f"encode: device tensors {ms_dev:.3f} ms, host arrays {ms_host:.3f} ms (+{ms_host - ms_dev:.3f})")
print(f"search: device tensors {ms_sdev:.3f} ms, host arrays {ms_shost:.3f} ms (+{ms_shost - ms_sdev:.3f})")
