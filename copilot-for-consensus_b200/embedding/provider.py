"""B200SentenceTransformerProvider -- drop-in for the reference's SentenceTransformer driver on CUDA.

Mirrors ``SentenceTransformerEmbeddingProvider``
(/root/reference/adapters/copilot_embedding/copilot_embedding/sentence_transformer_provider.py:19-94):
same constructor arguments and attributes (``model_name``, ``device``, ``cache_dir``), same
``from_config``, same ``embed(text) -> list[float]`` with the same ``ValueError`` messages.  What
``self.model.encode(text, convert_to_numpy=True)`` did (tokenise -> BERT -> masked mean -> L2
normalise) is split here into host WordPiece (wordpiece.py) and the CUDA encoder behind the C ABI
(``b200_encode``).

Extensions for throughput (the reference interface is one text per call):
``embed_batch(texts)``, ``embed_array(texts)`` and ``embed_tokens(input_ids, lengths)``.
``dimension`` short-circuits the service's start-up probe (embedding/main.py:290-296).
"""
from __future__ import annotations

import ctypes
import logging
import threading
from typing import Any, Sequence

import numpy as np

from .. import _native as nat
from . import model_io
from .base import EmbeddingProvider
from .wordpiece import WordPieceTokenizer

logger = logging.getLogger(__name__)


class B200SentenceTransformerProvider(EmbeddingProvider):
    """BERT-class sentence encoder on one B200, behind the reference's provider interface."""

    #: lets a patched service detect the batch API with isinstance/attribute checks that a Mock
    #: provider would not satisfy by accident
    supports_batch = True

    def __init__(self, model_name: str = "all-MiniLM-L6-v2", device: str = "cuda:0", cache_dir: str | None = None,
                 max_tokens_per_call: int = 512 * 256):
        self.model_name = model_name
        self.device = device
        self.cache_dir = cache_dir
        self._dev = nat.device_index(device)  # refuses anything that is not CUDA: there is no CPU path here
        self._lib = nat.lib()
        path = model_io.resolve_model_dir(model_name, cache_dir)
        if path is None:
            raise RuntimeError(
                f"model '{model_name}' was not found locally (cache_dir={cache_dir!r}) and this deployment has no "
                "network; point model_name/cache_dir at a HuggingFace-layout directory, or create a random-init one "
                "with copilot-for-consensus_b200.embedding.model_io.write_random_model()"
            )
        logger.info(f"Loading BERT-class model from {path} onto {device}")
        cfg, sd, vocab, do_lower = model_io.load_model_dir(path)
        self.config = cfg
        self.dimension = cfg.hidden
        self.max_seq_length = min(cfg.max_seq_length, cfg.max_pos)
        self.tokenizer = WordPieceTokenizer(vocab, do_lower_case=do_lower)
        self._max_tokens = int(max_tokens_per_call)
        self._lock = threading.Lock()
        self._h = ctypes.c_void_p()
        c = nat.EncoderConfigC(cfg.vocab_size, cfg.hidden, cfg.layers, cfg.heads, cfg.intermediate, cfg.max_pos,
                               cfg.type_vocab, cfg.ln_eps)
        nat.check(self._lib.b200_encoder_create(ctypes.byref(c), self._dev, ctypes.byref(self._h)))
        for name, t in sd.items():
            if name.startswith("pooler.") or name.endswith("position_ids"):
                continue
            a = np.ascontiguousarray(t.detach().cpu().numpy(), dtype=np.float32)
            nat.check(self._lib.b200_encoder_load_tensor(self._h, name.encode(), ctypes.c_void_p(a.ctypes.data), a.size))
        nat.check(self._lib.b200_encoder_finalize(self._h))
        logger.info("BERT-class model loaded")

    @classmethod
    def from_config(cls, driver_config: Any) -> "B200SentenceTransformerProvider":
        """sentence_transformer_provider.py:53-74."""
        return cls(model_name=str(driver_config.model_name), device=str(driver_config.device),
                   cache_dir=driver_config.cache_dir)

    def close(self) -> None:
        if getattr(self, "_h", None) is not None and self._h:
            self._lib.b200_encoder_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self) -> None:  # pragma: no cover
        try:
            self.close()
        except Exception:
            pass

    # -- the reference interface -----------------------------------------------------------------
    def embed(self, text: str) -> list[float]:
        """sentence_transformer_provider.py:76-94."""
        if text is None:
            raise ValueError("Text cannot be None")
        if not text.strip():
            raise ValueError("Text cannot be empty or whitespace-only")
        return self.embed_array([text])[0].tolist()

    # -- batch extensions ----------------------------------------------------------------------------
    def embed_batch(self, texts: Sequence[str]) -> list[list[float]]:
        """One embedding per text, same validation as :meth:`embed` applied to every element."""
        return self.embed_array(texts).tolist()

    def embed_array(self, texts: Sequence[str], to_host: bool = True):
        """``[len(texts), dimension]`` fp32.  Texts are bucketed by token length so that short chunks
        do not pay for the longest one's padding.  ``to_host=False`` leaves the result on the GPU (a CUDA
        tensor), e.g. to feed ``B200VectorStore.query_batch`` without a host round trip."""
        for t in texts:
            if t is None:
                raise ValueError("Text cannot be None")
            if not t.strip():
                raise ValueError("Text cannot be empty or whitespace-only")
        n = len(texts)
        if to_host:
            out = nat.host_out((n, self.dimension), np.float32, self._dev)
        else:
            import torch

            out = torch.empty((n, self.dimension), dtype=torch.float32, device=f"cuda:{self._dev}")
        if n == 0:
            return out
        all_ids, all_lens = self.tokenizer.encode_arrays(texts, self.max_seq_length)
        order = np.argsort(all_lens, kind="stable")
        sorted_lens = all_lens[order]
        s_cap = -(-self.config.max_pos // 16) * 16
        i = 0
        while i < n:
            # grow the bucket while (padded length x rows) stays inside the per-call token budget
            j = i
            while j < n:
                S = -(-int(sorted_lens[j]) // 64) * 64
                if (j - i + 1) * S > self._max_tokens and j > i:
                    break
                j += 1
            sel = order[i:j]
            S = min(-(-int(sorted_lens[j - 1]) // 64) * 64, s_cap)
            ids = np.zeros((len(sel), S), dtype=np.int32)
            w = min(S, all_ids.shape[1])
            ids[:, :w] = all_ids[sel, :w]
            lens = np.ascontiguousarray(all_lens[sel])
            if to_host:
                out[sel] = self._encode_host(ids, lens)
            else:
                dev = out.device
                emb = self.embed_tokens(torch.from_numpy(ids).to(dev), torch.from_numpy(lens).to(dev))
                out[torch.from_numpy(np.ascontiguousarray(sel)).to(dev)] = emb
            i = j
        return out

    def _stream(self) -> ctypes.c_void_p:
        return ctypes.c_void_p(nat.current_stream_ptr(self._dev))

    def _encode_host(self, ids: np.ndarray, lens: np.ndarray) -> np.ndarray:
        B, S = ids.shape
        out = nat.host_out((B, self.dimension), np.float32, self._dev)
        with self._lock:
            nat.check(self._lib.b200_encode(self._h, ctypes.c_void_p(ids.ctypes.data), ctypes.c_void_p(lens.ctypes.data),
                                            B, S, 0, ctypes.c_void_p(out.ctypes.data), 0, self._stream()))
        return out

    def embed_tokens(self, input_ids: Any, lengths: Any):
        """Token ids ``[B, S]`` int32 (0-padded, S a multiple of 16) + valid lengths ``[B]`` int32 ->
        embeddings ``[B, dimension]`` fp32.  CUDA tensors in give a CUDA tensor out with no host
        round trip (asynchronous on torch's current stream); numpy in gives numpy out."""
        try:
            import torch
        except Exception:  # pragma: no cover
            torch = None
        if torch is not None and isinstance(input_ids, torch.Tensor) and input_ids.is_cuda:
            if input_ids.device.index != self._dev or not lengths.is_cuda:
                raise ValueError(f"token tensors must live on cuda:{self._dev}")
            if input_ids.dtype != torch.int32 or lengths.dtype != torch.int32:
                raise ValueError("input_ids and lengths must be int32")
            if not input_ids.is_contiguous() or not lengths.is_contiguous():
                raise ValueError("token tensors must be contiguous")
            B, S = input_ids.shape
            out = torch.empty((B, self.dimension), dtype=torch.float32, device=input_ids.device)
            with self._lock:
                nat.check(self._lib.b200_encode(self._h, ctypes.c_void_p(input_ids.data_ptr()),
                                                ctypes.c_void_p(lengths.data_ptr()), B, S, 1,
                                                ctypes.c_void_p(out.data_ptr()), 1, self._stream()))
            return out
        ids = np.ascontiguousarray(input_ids, dtype=np.int32)
        lens = np.ascontiguousarray(lengths, dtype=np.int32)
        if ids.ndim != 2 or lens.shape != (ids.shape[0],):
            raise ValueError("input_ids must be [B, S] and lengths [B]")
        return self._encode_host(ids, lens)
