"""Model directories for the B200 embedding provider: config, vocab, weights.

A model directory is the HuggingFace layout sentence-transformers itself reads
(``config.json`` + ``vocab.txt`` + ``model.safetensors`` or ``pytorch_model.bin``, optionally
under ``0_Transformer/`` with ``sentence_bert_config.json`` carrying ``max_seq_length``).
``model_name`` is resolved the way ``SentenceTransformer(model_name, cache_folder=...)`` is called by
the reference (sentence_transformer_provider.py:50): a local path, else
``<cache_dir>/<model_name>`` / ``<cache_dir>/sentence-transformers_<model_name>``.

There is no network on the target boxes, so ``write_random_model`` lays down a random-init
model of the named architecture (BASELINE.json: "random-init weights of the same architecture")
with a synthetic WordPiece vocabulary; the parity oracle reads the very same files.
"""
from __future__ import annotations

import json
import os
from dataclasses import asdict, dataclass

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
    max_seq_length: int = 256

    @property
    def head_dim(self) -> int:
        return self.hidden // self.heads


#: architectures BASELINE.json names.  configs[3] is "mpnet-base-class 768-d encoder (12 layers, seq 384)": the SHAPE
#: of all-mpnet-base-v2 on the BERT architecture this path implements.  The real all-mpnet-base-v2 checkpoint is an
#: MPNetModel (relative position bias, other state_dict keys, another tokenizer) and is refused by load_model_dir.
KNOWN_ARCHS = {
    "all-MiniLM-L6-v2": EncoderConfig(),
    "bert-base-class": EncoderConfig(hidden=768, layers=12, heads=12, intermediate=3072, max_seq_length=384),
}


def config_from_hf(d: dict, max_seq_length: int | None = None) -> EncoderConfig:
    mt = d.get("model_type", "bert")
    if mt != "bert":
        raise ValueError(f"model_type {mt!r} is not supported: this provider implements the BERT architecture only "
                         "(MPNet, RoBERTa, DistilBERT... differ in position handling, parameter names or tokenizer)")
    if d.get("hidden_act", "gelu") != "gelu":
        raise ValueError(f"unsupported hidden_act {d.get('hidden_act')!r}: only exact-erf gelu is implemented")
    if d.get("position_embedding_type", "absolute") != "absolute":
        raise ValueError("only absolute position embeddings are implemented")
    return EncoderConfig(
        vocab_size=int(d["vocab_size"]), hidden=int(d["hidden_size"]), layers=int(d["num_hidden_layers"]),
        heads=int(d["num_attention_heads"]), intermediate=int(d["intermediate_size"]),
        max_pos=int(d.get("max_position_embeddings", 512)), type_vocab=int(d.get("type_vocab_size", 2)),
        ln_eps=float(d.get("layer_norm_eps", 1e-12)),
        max_seq_length=int(max_seq_length or min(256, int(d.get("max_position_embeddings", 512)))),
    )


def config_to_hf(cfg: EncoderConfig) -> dict:
    return {
        "architectures": ["BertModel"], "model_type": "bert", "vocab_size": cfg.vocab_size,
        "hidden_size": cfg.hidden, "num_hidden_layers": cfg.layers, "num_attention_heads": cfg.heads,
        "intermediate_size": cfg.intermediate, "max_position_embeddings": cfg.max_pos,
        "type_vocab_size": cfg.type_vocab, "layer_norm_eps": cfg.ln_eps, "hidden_act": "gelu",
        "hidden_dropout_prob": 0.0, "attention_probs_dropout_prob": 0.0,
        "initializer_range": 0.02, "pad_token_id": 0, "position_embedding_type": "absolute",
    }


def synthetic_vocab(size: int = 30522) -> list[str]:
    """Deterministic BERT-shaped vocabulary: specials at their usual ids ([PAD]=0, [UNK]=100,
    [CLS]=101, [SEP]=102, [MASK]=103), every printable ASCII character as a word-start and as a
    ``##`` continuation (so any ASCII text tokenises without [UNK]), then consonant-vowel syllable
    pieces up to ``size``."""
    vocab = ["[PAD]"] + [f"[unused{i}]" for i in range(99)] + ["[UNK]", "[CLS]", "[SEP]", "[MASK]"]
    vocab += [f"[unused{i}]" for i in range(99, 99 + 896)]  # ids 104..999, as in bert-base-uncased
    seen = set(vocab)

    def push(tok: str) -> None:
        if tok not in seen and len(vocab) < size:
            seen.add(tok)
            vocab.append(tok)

    chars = [chr(c) for c in range(33, 127) if not ("A" <= chr(c) <= "Z")]
    for c in chars:
        push(c)
    for c in chars:
        push("##" + c)
    cons, vows = "bcdfghjklmnpqrstvwxyz", "aeiou"
    syll = [c + v for c in cons for v in vows] + [v + c for v in vows for c in cons]
    for s in syll:
        push(s)
        push("##" + s)
    for a in syll:
        for b in syll:
            if len(vocab) >= size:
                break
            push(a + b)
            push("##" + a + b)
    i = 0
    while len(vocab) < size:
        push(f"w{i}")
        i += 1
    return vocab[:size]


def random_weights(cfg: EncoderConfig, seed: int = 0) -> dict[str, torch.Tensor]:
    """Seeded fp32 weights under HF ``BertModel`` state_dict names.  Matrices N(0, 0.02); biases
    N(0, 0.02); LayerNorm gamma 1+N(0, 0.1), beta N(0, 0.1) (non-trivial on purpose so fused
    bias / affine epilogues are exercised).  Same recipe as oracle/encode_oracle.make_weights."""
    g = torch.Generator(device="cpu").manual_seed(seed)

    def n(*shape, std=0.02, mean=0.0):
        return torch.randn(*shape, generator=g, dtype=torch.float32) * std + mean

    H, I = cfg.hidden, cfg.intermediate
    w = {
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


def write_random_model(path: str, arch: str | EncoderConfig = "all-MiniLM-L6-v2", seed: int = 0) -> str:
    """Create a loadable random-init model directory (config.json, vocab.txt, model.safetensors)."""
    cfg = KNOWN_ARCHS[arch] if isinstance(arch, str) else arch
    os.makedirs(path, exist_ok=True)
    with open(os.path.join(path, "config.json"), "w") as f:
        json.dump(config_to_hf(cfg), f, indent=1)
    with open(os.path.join(path, "sentence_bert_config.json"), "w") as f:
        json.dump({"max_seq_length": cfg.max_seq_length, "do_lower_case": False}, f)
    with open(os.path.join(path, "vocab.txt"), "w", encoding="utf-8") as f:
        f.write("\n".join(synthetic_vocab(cfg.vocab_size)) + "\n")
    with open(os.path.join(path, "tokenizer_config.json"), "w") as f:
        json.dump({"do_lower_case": True, "tokenizer_class": "BertTokenizer", "model_max_length": cfg.max_pos}, f)
    from safetensors.torch import save_file

    save_file({k: v.contiguous() for k, v in random_weights(cfg, seed).items()},
              os.path.join(path, "model.safetensors"))
    with open(os.path.join(path, "b200_random_init.json"), "w") as f:
        json.dump({"seed": seed, "config": asdict(cfg)}, f)
    # the sentence-transformers module stack of all-MiniLM-L6-v2: Transformer -> mean Pooling -> Normalize
    with open(os.path.join(path, "modules.json"), "w") as f:
        json.dump([{"idx": 0, "name": "0", "path": "", "type": "sentence_transformers.models.Transformer"},
                   {"idx": 1, "name": "1", "path": "1_Pooling", "type": "sentence_transformers.models.Pooling"},
                   {"idx": 2, "name": "2", "path": "2_Normalize", "type": "sentence_transformers.models.Normalize"}], f, indent=1)
    os.makedirs(os.path.join(path, "1_Pooling"), exist_ok=True)
    with open(os.path.join(path, "1_Pooling", "config.json"), "w") as f:
        json.dump({"word_embedding_dimension": cfg.hidden, "pooling_mode_cls_token": False, "pooling_mode_mean_tokens": True,
                   "pooling_mode_max_tokens": False, "pooling_mode_mean_sqrt_len_tokens": False}, f, indent=1)
    return path


def check_module_stack(path: str) -> None:
    """``SentenceTransformer.encode`` runs whatever ``modules.json`` lists.  The kernels implement exactly one
    stack -- Transformer -> masked-mean Pooling -> Normalize (all-MiniLM-L6-v2's) -- so anything else (CLS or max
    pooling, no Normalize stage, a Dense head) is refused instead of silently returning other embeddings than the
    library would.  A bare HuggingFace directory without ``modules.json`` is what sentence-transformers wraps in
    Transformer + mean Pooling; there the Normalize stage is this provider's documented addition."""
    mj = os.path.join(path, "modules.json")
    if not os.path.isfile(mj):
        return
    with open(mj) as f:
        mods = json.load(f)
    kinds = [str(m.get("type", "")).rsplit(".", 1)[-1] for m in mods]
    if kinds != ["Transformer", "Pooling", "Normalize"]:
        raise ValueError(f"unsupported sentence-transformers module stack {kinds}: only Transformer -> Pooling(mean) -> "
                         "Normalize is implemented")
    pool_dir = os.path.join(path, str(mods[1].get("path", "1_Pooling")))
    pc = os.path.join(pool_dir, "config.json")
    if os.path.isfile(pc):
        with open(pc) as f:
            cfg = json.load(f)
        modes = {k: bool(v) for k, v in cfg.items() if k.startswith("pooling_mode_")}
        on = sorted(k for k, v in modes.items() if v)
        if on != ["pooling_mode_mean_tokens"]:
            raise ValueError(f"unsupported pooling {on}: only pooling_mode_mean_tokens is implemented")


def resolve_model_dir(model_name: str, cache_dir: str | None) -> str | None:
    """Where ``SentenceTransformer(model_name, cache_folder=cache_dir)`` would find local files."""
    cands = [model_name]
    roots = [cache_dir] if cache_dir else []   # no environment lookups: the reference pins adapters to that
    for root in roots:
        cands += [
            os.path.join(root, model_name),
            os.path.join(root, "sentence-transformers_" + model_name.replace("/", "_")),
            os.path.join(root, model_name.replace("/", "_")),
        ]
    for c in cands:
        for sub in ("", "0_Transformer"):
            d = os.path.join(c, sub) if sub else c
            if os.path.isfile(os.path.join(d, "config.json")):
                return c
    return None


def load_model_dir(path: str) -> tuple[EncoderConfig, dict[str, torch.Tensor], list[str], bool]:
    """-> (config, fp32 state dict with the ``bert.`` prefix stripped, vocab, do_lower_case)."""
    check_module_stack(path)
    tdir = os.path.join(path, "0_Transformer") if os.path.isfile(os.path.join(path, "0_Transformer", "config.json")) else path
    with open(os.path.join(tdir, "config.json")) as f:
        hf = json.load(f)
    max_seq = None
    for d in (path, tdir):
        p = os.path.join(d, "sentence_bert_config.json")
        if os.path.isfile(p):
            with open(p) as f:
                max_seq = json.load(f).get("max_seq_length")
            break
    cfg = config_from_hf(hf, max_seq)
    st_path, bin_path = os.path.join(tdir, "model.safetensors"), os.path.join(tdir, "pytorch_model.bin")
    if os.path.isfile(st_path):
        from safetensors.torch import load_file

        sd = load_file(st_path)
    elif os.path.isfile(bin_path):
        sd = torch.load(bin_path, map_location="cpu", weights_only=True)
    else:
        raise FileNotFoundError(f"no model.safetensors or pytorch_model.bin under {tdir}")
    sd = {(k[5:] if k.startswith("bert.") else k): v.to(torch.float32) for k, v in sd.items()}
    do_lower = True
    tc = os.path.join(tdir, "tokenizer_config.json")
    if os.path.isfile(tc):
        with open(tc) as f:
            do_lower = bool(json.load(f).get("do_lower_case", True))
    with open(os.path.join(tdir, "vocab.txt"), encoding="utf-8") as f:
        vocab = [line.rstrip("\n") for line in f]
    return cfg, sd, vocab, do_lower
