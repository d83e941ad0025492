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
