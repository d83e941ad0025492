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
