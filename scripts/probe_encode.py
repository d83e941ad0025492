"""GPU diagnostic for the encoder: one-layer model, every stage compared with an fp32 torch
restatement through b200_encoder_peek; then full models against the oracle.  Not a test."""
import ctypes
import importlib
import math
import os
import sys
import tempfile
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import encode_oracle as eo  # noqa: E402

P = "copilot-for-consensus_b200"
mio = importlib.import_module(P + ".embedding.model_io")
prov_mod = importlib.import_module(P + ".embedding.provider")
nat = importlib.import_module(P + "._native")


def stage_trace(cfg, w, ids, mask):
    """one layer, returns dict of fp32 intermediates"""
    B, S = ids.shape
    H, nh, dh = cfg.hidden, cfg.heads, cfg.head_dim
    x0 = (w["embeddings.word_embeddings.weight"][ids.long()] + w["embeddings.position_embeddings.weight"][:S][None]
          + w["embeddings.token_type_embeddings.weight"][0][None, None])
    x0 = eo._layer_norm(x0, w["embeddings.LayerNorm.weight"], w["embeddings.LayerNorm.bias"], cfg.ln_eps)
    p = "encoder.layer.0."
    lin = lambda t, n: t @ w[p + n + ".weight"].T + w[p + n + ".bias"]
    q, k, v = lin(x0, "attention.self.query"), lin(x0, "attention.self.key"), lin(x0, "attention.self.value")
    qkv = torch.cat([q, k, v], dim=-1)
    qh, kh, vh = (t.view(B, S, nh, dh).transpose(1, 2) for t in (q, k, v))
    add = (1.0 - mask.float())[:, None, None, :] * torch.finfo(torch.float32).min
    ctx = (torch.softmax(qh @ kh.transpose(-1, -2) / math.sqrt(dh) + add, -1) @ vh).transpose(1, 2).reshape(B, S, H)
    y1 = lin(ctx, "attention.output.dense") + x0
    x1 = eo._layer_norm(y1, w[p + "attention.output.LayerNorm.weight"], w[p + "attention.output.LayerNorm.bias"], cfg.ln_eps)
    ffn = eo._gelu_erf(lin(x1, "intermediate.dense"))
    y2 = lin(ffn, "output.dense") + x1
    x2 = eo._layer_norm(y2, w[p + "output.LayerNorm.weight"], w[p + "output.LayerNorm.bias"], cfg.ln_eps)
    return {"qkv": qkv, "ctx": ctx, "ffn": ffn, "y": y2, "x": x2}


def peek(prov, which, shape):
    n = int(np.prod(shape))
    out = np.empty(n, dtype=np.float32)
    nat.check(nat.lib().b200_encoder_peek(prov._h, which, ctypes.c_void_p(out.ctypes.data), n))
    return torch.from_numpy(out.reshape(shape))


def one_layer(hidden, heads, inter, S, B=3):
    cfg = mio.EncoderConfig(vocab_size=2000, hidden=hidden, layers=1, heads=heads, intermediate=inter, max_pos=max(S, 64),
                            max_seq_length=S)
    ocfg = eo.EncoderConfig(vocab_size=2000, hidden=hidden, layers=1, heads=heads, intermediate=inter, max_pos=max(S, 64))
    with tempfile.TemporaryDirectory() as td:
        mio.write_random_model(td, cfg, seed=5)
        prov = prov_mod.B200SentenceTransformerProvider(model_name=td, device="cuda:0")
    w = eo.round_weights_bf16(eo.make_weights(ocfg, seed=5))
    ids, mask, lens = eo.synthetic_tokens(B, S, 2000, seed=2, ragged=True)
    ids = ids.clamp(max=1999)
    out = prov.embed_tokens(ids.numpy(), lens.numpy())
    tr = stage_trace(ocfg, w, ids, mask)
    print(f"--- one layer hidden={hidden} heads={heads} inter={inter} S={S}")
    valid = mask.bool()
    for which, name, width in [(1, "qkv", 3 * hidden), (2, "ctx", hidden), (4, "ffn", inter), (3, "y", hidden), (0, "x", hidden)]:
        got = peek(prov, which, (B, S, width))
        ref = tr[name]
        err = (got - ref).abs()
        e_valid = err[valid].max().item()
        print(f"  {name:4s} max|err| valid rows {e_valid:.4e}  (ref absmax {ref[valid].abs().max().item():.3f})"
              f"  finite={bool(torch.isfinite(got).all())}")
    ref_emb = eo.pool_normalize(tr["x"], mask).numpy()
    cos = (out * ref_emb).sum(1)
    print(f"  embedding: min cos {cos.min():.6f}  max|err| {np.abs(out - ref_emb).max():.3e}")
    prov.close()


def full(arch, B, S, ragged):
    cfg = mio.KNOWN_ARCHS[arch]
    ocfg = eo.EncoderConfig(hidden=cfg.hidden, layers=cfg.layers, heads=cfg.heads, intermediate=cfg.intermediate)
    with tempfile.TemporaryDirectory() as td:
        mio.write_random_model(td, arch, seed=0)
        prov = prov_mod.B200SentenceTransformerProvider(model_name=td, device="cuda:0")
    w = eo.make_weights(ocfg, seed=0)
    ids, mask, lens = eo.synthetic_tokens(B, S, cfg.vocab_size, seed=1, ragged=ragged)
    t = time.time()
    out = prov.embed_tokens(ids.numpy(), lens.numpy())
    dt = time.time() - t
    ref = eo.encode(ocfg, w, ids, mask).numpy()
    cos = (out * ref).sum(1)
    print(f"--- {arch} B={B} S={S} ragged={ragged}: min cos {cos.min():.6f} max|err| {np.abs(out - ref).max():.3e}"
          f" ({dt*1e3:.1f} ms incl. first-call setup)")
    prov.close()


if __name__ == "__main__":
    one_layer(128, 4, 256, 64)
    one_layer(384, 12, 1536, 256)
    one_layer(256, 4, 512, 128)      # head dim 64
    full("all-MiniLM-L6-v2", 8, 256, False)
    full("all-MiniLM-L6-v2", 16, 256, True)
    full("bert-base-class", 2, 384, True)
