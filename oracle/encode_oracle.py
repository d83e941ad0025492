"""fp32 CPU oracle for the chunk-embedding path -- TEST INFRASTRUCTURE, not product code.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` leg may import this module.

What it restates.  The reference's embed hot path is
``SentenceTransformerEmbeddingProvider.embed`` ->
``self.model.encode(text, convert_to_numpy=True)``
(/root/reference/adapters/copilot_embedding/copilot_embedding/sentence_transformer_provider.py:76-94).
The arithmetic behind ``encode`` lives in the third-party, UN-VENDORED and UNPINNED
dependency ``sentence-transformers>=2.0.0`` (adapters/copilot_embedding/setup.py:47-49), which is
not installed in this image and cannot be.  Its published pipeline for all-MiniLM-L6-v2 is
Transformer(BertModel) -> Pooling(mean, attention-mask weighted) -> Normalize(L2):

  tokens -> word + position + token-type(0) embeddings -> LayerNorm(eps 1e-12)
  L x { h = LN(h + Wo.softmax(QK^T/sqrt(dh) + mask).V) ; h = LN(h + W2.gelu_erf(W1.h)) }   (post-LN)
  pooled = sum_t(h_t * m_t) / max(sum_t m_t, 1e-9) ;  out = pooled / max(||pooled||_2, 1e-12)

PARITY PIN: the reference's own tests mock the library at this boundary
(adapters/copilot_embedding/tests/test_sentence_transformer_provider.py:28-34,64-78), so
there is no golden vector for ``encode`` in the reference: parity with sentence-transformers
itself is UNPINNED.  What *is* pinned (oracle/gen_golden.py -> tests/golden/encode_*.npz):
  (1) this restatement == installed ``transformers.BertModel`` (fp32, eager+sdpa) fed the same
      weights, then masked-mean + F.normalize, to fp32 round-off;
  (2) this restatement == the reference's in-tree ``HuggingFaceEmbeddingProvider.embed``
      (huggingface_provider.py:71-103: tokenizer -> AutoModel -> last_hidden_state.mean(dim=1)),
      L2-normalised, for single un-padded texts, on a locally saved random-init model dir.

Everything is explicit torch fp32 on CPU; no transformers import at run time.
"""
from __future__ import annotations

import math
from dataclasses import dataclass

import torch


@dataclass(frozen=True)
class EncoderConfig:
    vocab_size: int = 30522
    hidden: int = 384
    layers: int = 6
    heads: int = 12
    intermediate: int = 1536
    max_pos: int = 512
    type_vocab: int = 2
    ln_eps: float = 1e-12

    @property
    def head_dim(self) -> int:
        return self.hidden // self.heads


MINILM_L6 = EncoderConfig()
MPNET_BASE_CLASS = EncoderConfig(hidden=768, layers=12, heads=12, intermediate=3072)


def make_weights(cfg: EncoderConfig, seed: int = 0) -> dict[str, torch.Tensor]:
    """Seeded random-init weights under HF ``BertModel`` state_dict names.

    Matrices ~ N(0, 0.02) as HF's initializer_range; biases ~ N(0, 0.02) and LayerNorm
    gamma = 1 + N(0, 0.1), beta = N(0, 0.1) instead of HF's zeros/ones so that a dropped bias
    or affine term in a fused epilogue shows up in parity tests.
    """
    g = torch.Generator(device="cpu").manual_seed(seed)

    def n(*shape, std=0.02, mean=0.0):
        return torch.randn(*shape, generator=g, dtype=torch.float32) * std + mean

    H, I = cfg.hidden, cfg.intermediate
    w: dict[str, torch.Tensor] = {
        "embeddings.word_embeddings.weight": n(cfg.vocab_size, H),
        "embeddings.position_embeddings.weight": n(cfg.max_pos, H),
        "embeddings.token_type_embeddings.weight": n(cfg.type_vocab, H),
        "embeddings.LayerNorm.weight": n(H, std=0.1, mean=1.0),
        "embeddings.LayerNorm.bias": n(H, std=0.1),
    }
    for l in range(cfg.layers):
        p = f"encoder.layer.{l}."
        for name in ("query", "key", "value"):
            w[p + f"attention.self.{name}.weight"] = n(H, H)
            w[p + f"attention.self.{name}.bias"] = n(H)
        w[p + "attention.output.dense.weight"] = n(H, H)
        w[p + "attention.output.dense.bias"] = n(H)
        w[p + "attention.output.LayerNorm.weight"] = n(H, std=0.1, mean=1.0)
        w[p + "attention.output.LayerNorm.bias"] = n(H, std=0.1)
        w[p + "intermediate.dense.weight"] = n(I, H)
        w[p + "intermediate.dense.bias"] = n(I)
        w[p + "output.dense.weight"] = n(H, I)
        w[p + "output.dense.bias"] = n(H)
        w[p + "output.LayerNorm.weight"] = n(H, std=0.1, mean=1.0)
        w[p + "output.LayerNorm.bias"] = n(H, std=0.1)
    return w


def round_weights_bf16(w: dict[str, torch.Tensor]) -> dict[str, torch.Tensor]:
    """bf16-round the GEMM operands + embedding tables (what the GPU path stores); biases and
    LayerNorm parameters stay fp32 on the GPU, so they stay fp32 here."""
    out = {}
    for k, v in w.items():
        is_matrix = v.ndim == 2
        out[k] = v.to(torch.bfloat16).to(torch.float32) if is_matrix else v.clone()
    return out


def _layer_norm(x: torch.Tensor, g: torch.Tensor, b: torch.Tensor, eps: float) -> torch.Tensor:
    mu = x.mean(dim=-1, keepdim=True)
    var = ((x - mu) ** 2).mean(dim=-1, keepdim=True)
    return (x - mu) / torch.sqrt(var + eps) * g + b


def _gelu_erf(x: torch.Tensor) -> torch.Tensor:
    return x * 0.5 * (1.0 + torch.erf(x / math.sqrt(2.0)))


@torch.no_grad()
def hidden_states(cfg: EncoderConfig, w: dict[str, torch.Tensor], input_ids: torch.Tensor,
                  attention_mask: torch.Tensor, return_all: bool = False):
    """BertModel(add_pooling_layer=False).last_hidden_state, fp32.  ids/mask: [B,S] integer."""
    ids = input_ids.long()
    B, S = ids.shape
    H, nh, dh = cfg.hidden, cfg.heads, cfg.head_dim
    pos = torch.arange(S)
    x = (w["embeddings.word_embeddings.weight"][ids]
         + w["embeddings.position_embeddings.weight"][pos][None]
         + w["embeddings.token_type_embeddings.weight"][0][None, None])
    x = _layer_norm(x, w["embeddings.LayerNorm.weight"], w["embeddings.LayerNorm.bias"], cfg.ln_eps)
    add_mask = (1.0 - attention_mask.to(torch.float32))[:, None, None, :] * torch.finfo(torch.float32).min
    trace = [x]
    for l in range(cfg.layers):
        p = f"encoder.layer.{l}."

        def lin(t, name):
            return t @ w[p + name + ".weight"].T + w[p + name + ".bias"]

        q = lin(x, "attention.self.query").view(B, S, nh, dh).transpose(1, 2)
        k = lin(x, "attention.self.key").view(B, S, nh, dh).transpose(1, 2)
        v = lin(x, "attention.self.value").view(B, S, nh, dh).transpose(1, 2)
        sc = q @ k.transpose(-1, -2) / math.sqrt(dh) + add_mask
        ctx = (torch.softmax(sc, dim=-1) @ v).transpose(1, 2).reshape(B, S, H)
        x = _layer_norm(lin(ctx, "attention.output.dense") + x,
                        w[p + "attention.output.LayerNorm.weight"], w[p + "attention.output.LayerNorm.bias"],
                        cfg.ln_eps)
        h = _gelu_erf(lin(x, "intermediate.dense"))
        x = _layer_norm(lin(h, "output.dense") + x,
                        w[p + "output.LayerNorm.weight"], w[p + "output.LayerNorm.bias"], cfg.ln_eps)
        trace.append(x)
    return trace if return_all else x


@torch.no_grad()
def pool_normalize(h: torch.Tensor, attention_mask: torch.Tensor) -> torch.Tensor:
    """sentence-transformers Pooling(mean) + Normalize: masked mean, then x / max(||x||, 1e-12)."""
    m = attention_mask.to(torch.float32)[:, :, None]
    pooled = (h * m).sum(dim=1) / m.sum(dim=1).clamp(min=1e-9)
    return pooled / pooled.norm(p=2, dim=1, keepdim=True).clamp(min=1e-12)


@torch.no_grad()
def encode(cfg: EncoderConfig, w: dict[str, torch.Tensor], input_ids: torch.Tensor,
           attention_mask: torch.Tensor) -> torch.Tensor:
    """ids/mask [B,S] -> L2-normalised sentence embeddings [B,H], fp32 on CPU."""
    return pool_normalize(hidden_states(cfg, w, input_ids, attention_mask), attention_mask)


@torch.no_grad()
def encode_fast(cfg: EncoderConfig, w: dict[str, torch.Tensor], input_ids: torch.Tensor,
                attention_mask: torch.Tensor) -> torch.Tensor:
    """Same arithmetic as :func:`encode`, written with the fused torch.nn.functional ops that
    transformers' BertModel (sdpa attention) and therefore sentence-transformers run on CPU.  This
    is the variant bench.py TIMES as the reference's CPU encode; tests pin it to :func:`encode`."""
    import torch.nn.functional as F

    ids = input_ids.long()
    B, S = ids.shape
    H, nh, dh = cfg.hidden, cfg.heads, cfg.head_dim
    x = (F.embedding(ids, w["embeddings.word_embeddings.weight"])
         + w["embeddings.position_embeddings.weight"][:S][None]
         + w["embeddings.token_type_embeddings.weight"][0][None, None])
    x = F.layer_norm(x, (H,), w["embeddings.LayerNorm.weight"], w["embeddings.LayerNorm.bias"], cfg.ln_eps)
    keep = attention_mask.bool()[:, None, None, :]
    for l in range(cfg.layers):
        p = f"encoder.layer.{l}."
        q = F.linear(x, w[p + "attention.self.query.weight"], w[p + "attention.self.query.bias"]).view(B, S, nh, dh).transpose(1, 2)
        k = F.linear(x, w[p + "attention.self.key.weight"], w[p + "attention.self.key.bias"]).view(B, S, nh, dh).transpose(1, 2)
        v = F.linear(x, w[p + "attention.self.value.weight"], w[p + "attention.self.value.bias"]).view(B, S, nh, dh).transpose(1, 2)
        ctx = F.scaled_dot_product_attention(q, k, v, attn_mask=keep).transpose(1, 2).reshape(B, S, H)
        x = F.layer_norm(F.linear(ctx, w[p + "attention.output.dense.weight"], w[p + "attention.output.dense.bias"]) + x,
                         (H,), w[p + "attention.output.LayerNorm.weight"], w[p + "attention.output.LayerNorm.bias"], cfg.ln_eps)
        h = F.gelu(F.linear(x, w[p + "intermediate.dense.weight"], w[p + "intermediate.dense.bias"]))
        x = F.layer_norm(F.linear(h, w[p + "output.dense.weight"], w[p + "output.dense.bias"]) + x,
                         (H,), w[p + "output.LayerNorm.weight"], w[p + "output.LayerNorm.bias"], cfg.ln_eps)
    return pool_normalize(x, attention_mask)


def synthetic_tokens(batch: int, seq: int, vocab: int = 30522, seed: int = 0, ragged: bool = False):
    """SURVEY.md section 8d synthetic chunks: ids ~ U{1000..vocab-1}, [CLS]=101 first, [SEP]=102
    last valid token; ``ragged`` draws lengths ~ U{32..seq} and pads with id 0 / mask 0."""
    g = torch.Generator(device="cpu").manual_seed(seed)
    ids = torch.randint(1000, vocab, (batch, seq), generator=g, dtype=torch.int32)
    if ragged:
        lens = torch.randint(min(32, seq), seq + 1, (batch,), generator=g, dtype=torch.int32)
    else:
        lens = torch.full((batch,), seq, dtype=torch.int32)
    ar = torch.arange(seq, dtype=torch.int32)[None, :]
    mask = (ar < lens[:, None]).to(torch.int32)
    ids = ids * mask
    ids[:, 0] = 101
    ids[torch.arange(batch), (lens - 1).long()] = 102
    return ids, mask, lens
