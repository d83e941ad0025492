"""Host-side BERT WordPiece tokenisation (stays on the CPU by design -- BASELINE.json north_star).

The reference never tokenises itself: ``SentenceTransformer.encode`` does it inside the
third-party library (sentence_transformer_provider.py:93).  For all-MiniLM-L6-v2 that is the
uncased BERT recipe: clean -> lower-case + strip accents -> split on whitespace/punctuation/CJK
-> greedy longest-match WordPiece with "##" continuations -> ``[CLS] ... [SEP]``, truncated to
``max_seq_length`` (256 for all-MiniLM-L6-v2, 384 for all-mpnet-base-v2).

``tests/test_wordpiece.py`` checks this implementation token-for-token against the installed
``transformers.BertTokenizer`` on the same vocab.
"""
from __future__ import annotations

import unicodedata
from typing import Iterable, Sequence

import numpy as np

PAD, UNK, CLS, SEP, MASK = "[PAD]", "[UNK]", "[CLS]", "[SEP]", "[MASK]"


def _is_whitespace(ch: str) -> bool:
    return ch in " \t\n\r" or unicodedata.category(ch) == "Zs"


def _is_control(ch: str) -> bool:
    if ch in "\t\n\r":
        return False
    return unicodedata.category(ch).startswith("C")


def _is_punctuation(ch: str) -> bool:
    cp = ord(ch)
    if 33 <= cp <= 47 or 58 <= cp <= 64 or 91 <= cp <= 96 or 123 <= cp <= 126:
        return True
    return unicodedata.category(ch).startswith("P")


def _is_cjk(cp: int) -> bool:
    return (
        0x4E00 <= cp <= 0x9FFF or 0x3400 <= cp <= 0x4DBF or 0x20000 <= cp <= 0x2A6DF
        or 0x2A700 <= cp <= 0x2B73F or 0x2B740 <= cp <= 0x2B81F or 0x2B820 <= cp <= 0x2CEAF
        or 0xF900 <= cp <= 0xFAFF or 0x2F800 <= cp <= 0x2FA1F
    )


class WordPieceTokenizer:
    """Uncased BERT tokenizer over a ``vocab.txt`` (one piece per line, line number = id)."""

    def __init__(self, vocab: Sequence[str] | dict[str, int], do_lower_case: bool = True,
                 max_chars_per_word: int = 100):
        if not isinstance(vocab, dict):
            vocab = {tok: i for i, tok in enumerate(vocab)}
        self.vocab: dict[str, int] = vocab
        self.do_lower_case = do_lower_case
        self.max_chars_per_word = max_chars_per_word
        for special in (PAD, UNK, CLS, SEP):
            if special not in vocab:
                raise ValueError(f"vocab is missing the special token {special}")
        self.pad_id, self.unk_id = vocab[PAD], vocab[UNK]
        self.cls_id, self.sep_id = vocab[CLS], vocab[SEP]
        self._never_split = {PAD, UNK, CLS, SEP, MASK}
        self._cache: dict[str, tuple[int, ...]] = {}

    @classmethod
    def from_file(cls, path: str, do_lower_case: bool = True) -> "WordPieceTokenizer":
        with open(path, encoding="utf-8") as f:
            return cls([line.rstrip("\n") for line in f], do_lower_case=do_lower_case)

    # -- basic tokenisation -------------------------------------------------------------------
    def _clean(self, text: str) -> str:
        out = []
        for ch in text:
            cp = ord(ch)
            if cp == 0 or cp == 0xFFFD or _is_control(ch):
                continue
            if _is_whitespace(ch):
                out.append(" ")
            elif _is_cjk(cp):
                out.append(f" {ch} ")
            else:
                out.append(ch)
        return "".join(out)

    def _split_word(self, word: str) -> list[str]:
        if word in self._never_split:
            return [word]
        if self.do_lower_case:
            word = word.lower()
            word = "".join(c for c in unicodedata.normalize("NFD", word) if unicodedata.category(c) != "Mn")
        pieces: list[str] = []
        cur: list[str] = []
        for ch in word:
            if _is_punctuation(ch):
                if cur:
                    pieces.append("".join(cur))
                    cur = []
                pieces.append(ch)
            else:
                cur.append(ch)
        if cur:
            pieces.append("".join(cur))
        return pieces

    # -- wordpiece ----------------------------------------------------------------------------
    def _wordpiece(self, word: str) -> tuple[int, ...]:
        hit = self._cache.get(word)
        if hit is not None:
            return hit
        if len(word) > self.max_chars_per_word:
            ids: tuple[int, ...] = (self.unk_id,)
        else:
            out: list[int] = []
            start, n = 0, len(word)
            while start < n:
                end = n
                found = -1
                while start < end:
                    sub = word[start:end]
                    if start > 0:
                        sub = "##" + sub
                    tid = self.vocab.get(sub)
                    if tid is not None:
                        found = tid
                        break
                    end -= 1
                if found < 0:
                    out = [self.unk_id]
                    break
                out.append(found)
                start = end
            ids = tuple(out)
        if len(self._cache) < 1_000_000:
            self._cache[word] = ids
        return ids

    def tokenize_to_ids(self, text: str) -> list[int]:
        """Piece ids without [CLS]/[SEP]."""
        text = unicodedata.normalize("NFC", self._clean(text))
        ids: list[int] = []
        for raw in text.split():
            for w in self._split_word(raw):
                ids.extend(self._wordpiece(w))
        return ids

    def encode(self, text: str, max_length: int) -> list[int]:
        """``[CLS] pieces... [SEP]`` truncated so the total length is at most ``max_length``."""
        body = self.tokenize_to_ids(text)[: max(0, max_length - 2)]
        return [self.cls_id, *body, self.sep_id]

    # -- many texts at once: native ASCII fast path + this implementation for the rest -------------
    def _native_handle(self):
        """Handle of the C++ tokenizer over the same vocabulary (csrc/wordpiece.cu), or None when the
        shared library cannot be loaded.  Built lazily, once."""
        if getattr(self, "_native_tried", False):
            return self._native
        self._native_tried, self._native = True, None
        try:
            import ctypes

            from .. import _native as nat

            lib = nat.lib()
            by_id = sorted(self.vocab.items(), key=lambda kv: kv[1])
            if [i for _, i in by_id] != list(range(len(by_id))):
                return None                                  # ids are not 0..n-1: keep the Python path
            enc = [t.encode("utf-8") for t, _ in by_id]
            offs = np.zeros(len(enc) + 1, dtype=np.int64)
            np.cumsum([len(b) for b in enc], out=offs[1:])
            blob = b"".join(enc)
            h = ctypes.c_void_p()
            nat.check(lib.b200_wordpiece_create(ctypes.c_char_p(blob), ctypes.c_void_p(offs.ctypes.data), len(enc),
                                                1 if self.do_lower_case else 0, ctypes.byref(h)))
            self._native = (lib, h)
        except Exception:                                    # no library on this machine: Python only
            self._native = None
        return self._native

    def encode_arrays(self, texts: Sequence[str], max_length: int, n_threads: int = 0) -> tuple[np.ndarray, np.ndarray]:
        """``(ids[n, max_length] int32 padded with [PAD], lengths[n] int32)`` -- :meth:`encode` for every text
        without a Python object per token.  ASCII texts go through the multi-threaded C++ tokenizer
        (``b200_wordpiece_encode``), texts with any non-ASCII character through :meth:`encode`; the result is
        the same either way (tests/test_wordpiece.py compares the two on the same inputs)."""
        n = len(texts)
        ids = np.full((n, max_length), self.pad_id, dtype=np.int32)
        lens = np.zeros(n, dtype=np.int32)
        nat_h = self._native_handle() if n >= 1 and max_length >= 2 and self.max_chars_per_word == 100 else None
        if nat_h is None:
            todo = range(n)
        else:
            import ctypes

            from .. import _native as nat

            lib, h = nat_h
            enc = [t.encode("utf-8", "surrogatepass") for t in texts]
            offs = np.zeros(n + 1, dtype=np.int64)
            np.cumsum([len(b) for b in enc], out=offs[1:])
            blob = b"".join(enc)
            nat.check(lib.b200_wordpiece_encode(h, ctypes.c_char_p(blob), ctypes.c_void_p(offs.ctypes.data), n, max_length,
                                                ctypes.c_void_p(ids.ctypes.data), ctypes.c_void_p(lens.ctypes.data),
                                                n_threads))
            todo = np.flatnonzero(lens < 0).tolist()          # texts with non-ASCII characters
        for i in todo:
            row = self.encode(texts[i], max_length)
            ids[i, : len(row)] = row
            ids[i, len(row):] = self.pad_id
            lens[i] = len(row)
        return ids, lens

    def encode_many(self, texts: Sequence[str], max_length: int, n_threads: int = 0) -> list[list[int]]:
        """:meth:`encode` for every text, as lists (see :meth:`encode_arrays`)."""
        if max_length < 2:
            return [self.encode(t, max_length) for t in texts]
        ids, lens = self.encode_arrays(texts, max_length, n_threads)
        return [ids[i, : int(lens[i])].tolist() for i in range(len(texts))]

    def __del__(self):
        nat_h = getattr(self, "_native", None)
        if nat_h:
            try:
                nat_h[0].b200_wordpiece_destroy(nat_h[1])
            except Exception:
                pass

    def encode_batch(self, texts: Iterable[str], max_length: int, pad_to_multiple: int = 1
                     ) -> tuple[np.ndarray, np.ndarray]:
        """Padded ``(input_ids[B,S] int32, lengths[B] int32)``; S = longest row rounded up."""
        texts = list(texts)
        if max_length < 2:
            rows = [self.encode(t, max_length) for t in texts]
            full = np.full((len(rows), max(2, max_length)), self.pad_id, dtype=np.int32)
            lens = np.array([len(r) for r in rows], dtype=np.int32)
            for i, r in enumerate(rows):
                full[i, : len(r)] = r
        else:
            full, lens = self.encode_arrays(texts, max_length)
        longest = int(lens.max()) if len(texts) else 1
        S = -(-longest // pad_to_multiple) * pad_to_multiple
        ids = np.full((len(texts), S), self.pad_id, dtype=np.int32)
        w = min(S, full.shape[1])
        ids[:, :w] = full[:, :w]
        return ids, lens
