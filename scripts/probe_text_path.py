"""Text in -> vectors out: tokenizer throughput (Python restatement vs the native ASCII path) and the
whole embed_array(texts) call on the GPU box."""
import importlib, os, random, sys, tempfile, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
P = "copilot-for-consensus_b200"
mio = importlib.import_module(P + ".embedding.model_io")
prov_mod = importlib.import_module(P + ".embedding.provider")
with tempfile.TemporaryDirectory() as td:
    mio.write_random_model(td, "all-MiniLM-L6-v2", seed=0)
    prov = prov_mod.B200SentenceTransformerProvider(model_name=td, device="cuda:0")
tok = prov.tokenizer
random.seed(0)
words = [v for v in tok.vocab if v.isalpha() and v.isascii() and len(v) > 2][:5000]
texts = [" ".join(random.choice(words) for _ in range(random.randrange(60, 200))) for _ in range(8192)]
t0 = time.perf_counter(); a = [tok.encode(t, 256) for t in texts[:400]]; dt = time.perf_counter() - t0
print(f"python tokenizer: {400 / dt:.0f} texts/s (1 thread)")
for th in (1, 4, 0):
    tok.encode_arrays(texts[:64], 256, n_threads=th)
    best = 1e9
    for _ in range(3):
        t0 = time.perf_counter(); ids, lens = tok.encode_arrays(texts, 256, n_threads=th); best = min(best, time.perf_counter() - t0)
    print(f"native tokenizer threads={th or os.cpu_count()}: {len(texts) / best:.0f} texts/s, {lens.sum() / best / 1e6:.1f} M tokens/s")
assert [ids[i, : lens[i]].tolist() for i in range(400)] == a
prov.embed_array(texts[:512])
best = 1e9
for _ in range(3):
    t0 = time.perf_counter(); out = prov.embed_array(texts); best = min(best, time.perf_counter() - t0)
print(f"embed_array(8192 texts): {len(texts) / best:.0f} texts/s end to end (tokenise + encode + copy back), mean {lens.mean():.0f} tokens")
print("cores", os.cpu_count())
