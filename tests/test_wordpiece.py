"""Host WordPiece against transformers.BertTokenizer on the same vocabulary, and against the
token ids the reference's HuggingFaceEmbeddingProvider tokenizer produced (golden)."""
import numpy as np
import pytest

from conftest import pkg

TEXTS = [
    "The working group reached rough consensus on the QUIC draft; see RFC 9000, section 4.2!",
    "naïve café résumé — mailing-list thread about DNSSEC key rollover (re: draft-ietf-dnsop)",
    "  multiple   spaces\tand\nnewlines  ",
    "UPPER lower MiXeD 12345 3.14159 a_b-c/d\\e",
    "汉字 mixed with ascii and emoji \U0001f600 ok",
    "x" * 150 + " short",
    "[CLS] literal specials [SEP] [MASK] [unused5]",
    "ba##be ##ba don't can't it's",
    "",
]


@pytest.fixture(scope="module")
def vocab():
    return pkg("embedding.model_io").synthetic_vocab(30522)


def test_synthetic_vocab_shape(vocab):
    assert len(vocab) == 30522 and len(set(vocab)) == 30522
    assert vocab[0] == "[PAD]" and vocab[100] == "[UNK]" and vocab[101] == "[CLS]"
    assert vocab[102] == "[SEP]" and vocab[103] == "[MASK]"


def test_matches_transformers_bert_tokenizer(vocab, tmp_path):
    transformers = pytest.importorskip("transformers")
    (tmp_path / "vocab.txt").write_text("\n".join(vocab) + "\n", encoding="utf-8")
    (tmp_path / "tokenizer_config.json").write_text('{"do_lower_case": true, "tokenizer_class": "BertTokenizer"}')
    hf = transformers.AutoTokenizer.from_pretrained(str(tmp_path))
    wp = pkg("embedding.wordpiece").WordPieceTokenizer(vocab)
    for t in TEXTS:
        want = hf(t, truncation=True, max_length=64)["input_ids"]
        assert wp.encode(t, 64) == want, t


def test_matches_reference_tokenizer_golden(vocab, encode_golden):
    g = encode_golden
    wp = pkg("embedding.wordpiece").WordPieceTokenizer(vocab)
    for i, t in enumerate(g["ref_texts"]):
        n = int(g["ref_token_lens"][i])
        assert wp.encode(str(t), 256) == list(g["ref_token_ids"][i, :n])


def test_encode_batch_padding(vocab):
    wp = pkg("embedding.wordpiece").WordPieceTokenizer(vocab)
    ids, lens = wp.encode_batch(["ok", "ba be bi bo bu ca ce"], max_length=256, pad_to_multiple=8)
    assert ids.dtype == np.int32 and ids.shape[1] % 8 == 0
    assert ids[0, 0] == 101 and ids[0, lens[0] - 1] == 102 and (ids[0, lens[0]:] == 0).all()
    ids, lens = wp.encode_batch(["ba " * 500], max_length=16)
    assert ids.shape == (1, 16) and ids[0, -1] == 102


ASCII_TEXTS = [
    "The working group reached rough consensus on the QUIC draft; see RFC 9000, section 4.2!",
    "  multiple   spaces\tand\nnewlines\r\n  ",
    "UPPER lower MiXeD 12345 3.14159 a_b-c/d\\e ~tilde `tick` {brace} |pipe| ^caret",
    "x" * 150 + " short " + "y" * 100 + " " + "z" * 101,
    "[CLS] literal specials [SEP] [MASK] [unused5] [PAD] [UNK] [cls]",
    "ba##be ##ba don't can't it's a.b.c ... !!! ???",
    "control\x01chars\x7finside\x00words and\x0bvertical\x0ctabs\x1f",
    "",
    "   ",
    "a",
    "word " * 400,
]


def test_native_tokenizer_equals_python_and_transformers(vocab, tmp_path):
    """csrc/wordpiece.cu (ASCII fast path, threads) against the Python restatement and against
    transformers.BertTokenizer, token for token; non-ASCII texts fall back to Python inside encode_many."""
    wpm = pkg("embedding.wordpiece")
    wp = wpm.WordPieceTokenizer(vocab)
    assert wp._native_handle() is not None, "the shared library must load (the tokenizer needs no GPU)"
    texts = ASCII_TEXTS + TEXTS
    for max_len in (2, 3, 16, 64, 256):
        want = [wp.encode(t, max_len) for t in texts]
        for threads in (1, 0, 5):
            assert wp.encode_many(texts, max_len, n_threads=threads) == want, (max_len, threads)
    transformers = pytest.importorskip("transformers")
    (tmp_path / "vocab.txt").write_text("\n".join(vocab) + "\n", encoding="utf-8")
    (tmp_path / "tokenizer_config.json").write_text('{"do_lower_case": true, "tokenizer_class": "BertTokenizer"}')
    hf = transformers.AutoTokenizer.from_pretrained(str(tmp_path))
    got = wp.encode_many(ASCII_TEXTS, 64)
    for t, g in zip(ASCII_TEXTS, got):
        assert g == hf(t, truncation=True, max_length=64)["input_ids"], t


def test_native_tokenizer_random_ascii_and_batch_layout(vocab):
    import random

    wp = pkg("embedding.wordpiece").WordPieceTokenizer(vocab)
    rnd = random.Random(3)
    alphabet = "abcdefghijklmnopqrstuvwxyzABCDEFGHIJKLMNOPQRSTUVWXYZ0123456789 .,;:!?-_'\"()[]#/\\\t\n\x02"
    words = [v for v in vocab if v.isalpha() and v.isascii() and len(v) > 1][:3000]
    texts = []
    for i in range(300):
        if i % 3 == 0:
            texts.append("".join(rnd.choice(alphabet) for _ in range(rnd.randrange(0, 400))))
        else:
            texts.append(" ".join(rnd.choice(words) for _ in range(rnd.randrange(1, 300))))
    want = [wp.encode(t, 128) for t in texts]
    assert wp.encode_many(texts, 128) == want
    ids, lens = wp.encode_batch(texts, 128, pad_to_multiple=16)
    assert ids.shape[1] % 16 == 0 and ids.dtype == np.int32 and lens.tolist() == [len(r) for r in want]
    for i, r in enumerate(want):
        assert ids[i, : len(r)].tolist() == r and (ids[i, len(r):] == wp.pad_id).all()


def test_native_tokenizer_unicode_bmp_equals_python(vocab):
    """Non-ASCII text of the Basic Multilingual Plane goes through the generated tables
    (csrc/unicode_tables.h, derived from Python's unicodedata by scripts/gen_unicode_tables.py); anything the
    tables do not cover (combining marks, astral characters, final-sigma contexts) comes back as -1 and is
    tokenised by the Python implementation -- either way encode_many must equal encode."""
    import ctypes
    import random

    wp = pkg("embedding.wordpiece").WordPieceTokenizer(vocab)
    lib, h = wp._native_handle()
    samples = [
        "naïve café résumé — mailing-list thread about DNSSEC key rollover (re: draft-ietf-dnsop)",
        "“smart quotes” and ‘single’ … ellipsis – en dash — em dash • bullet § section ¶",
        "Ångström Øresund Łódź Dvořák İstanbul ıspanak ẞ straße ǅ ǆ Ǆ ﬁ ﬂ",
        "汉字 mixed with ascii 漢字かな交じり文 カタカナ ひらがな ｆｕｌｌｗｉｄｔｈ！",
        "한국어 텍스트 토큰화 테스트",
        "Ελληνικά κείμενο ΑΒΓ δοκιμή",          # no capital sigma: stays on the native path
        "ΣΟΦΟΣ λόγος",                          # capital sigma: Python's context-sensitive lower()
        "русский текст Ёжик ПРИВЕТ мир",
        "zero​width nbsp line para　ideographic space﻿bom",
        "é combining acute and å ring",          # combining marks -> Python
        "emoji \U0001f600 and astral \U00020000 ideograph",    # beyond the BMP -> Python
        "תל אביב عربى نص हिन्दी पाठ ไทย ข้อความ",
        "mixed[CLS]inside and [SEP] alone №5 ½ ² ™ © ® ° ± × ÷",
    ]
    rnd = random.Random(11)
    pools = [list(range(0x20, 0x7F)), list(range(0xA0, 0x250)), list(range(0x370, 0x530)), list(range(0x2000, 0x2070)),
             list(range(0x3000, 0x3100)), list(range(0x4E00, 0x4E80)), list(range(0xAC00, 0xAC80)), list(range(0xFF00, 0xFFF0)),
             list(range(0x1E00, 0x2000)), list(range(0x2100, 0x2200))]
    for _ in range(400):
        n = rnd.randrange(1, 120)
        pool = rnd.choice(pools) + pools[0] + [0x20] * 40
        samples.append("".join(chr(rnd.choice(pool)) for _ in range(n)))
    # every BMP character once, in words of 7, so that every table entry is exercised
    cps = [cp for cp in range(0x80, 0x10000) if not 0xD800 <= cp <= 0xDFFF]
    for i in range(0, len(cps), 700):
        chunk = cps[i:i + 700]
        samples.append(" ".join("a" + "".join(chr(c) for c in chunk[j:j + 7]) + "b" for j in range(0, len(chunk), 7)))
    for max_len in (16, 256, 1024):
        want = [wp.encode(t, max_len) for t in samples]
        assert wp.encode_many(samples, max_len) == want
    # the native path really handles BMP text itself (no silent fallback for the common cases)
    n = 9
    enc = [t.encode("utf-8") for t in samples[:n]]
    offs = np.zeros(n + 1, dtype=np.int64)
    np.cumsum([len(b) for b in enc], out=offs[1:])
    ids = np.empty((n, 64), dtype=np.int32)
    lens = np.empty(n, dtype=np.int32)
    assert lib.b200_wordpiece_encode(h, ctypes.c_char_p(b"".join(enc)), ctypes.c_void_p(offs.ctypes.data), n, 64,
                                     ctypes.c_void_p(ids.ctypes.data), ctypes.c_void_p(lens.ctypes.data), 1) == 0
    assert (lens[[0, 1, 2, 3, 4, 5, 7, 8]] > 0).all() and lens[6] == -1
