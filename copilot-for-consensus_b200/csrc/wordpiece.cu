// wordpiece.cu -- host-side BERT WordPiece for ASCII text, many texts at a time on several threads.
// Pure host code (compiled by nvcc's host compiler so that it lands in the same shared library).
//
// What `SentenceTransformer.encode` does before the encoder runs (SURVEY.md section 8, a4'): clean ->
// lower-case + strip accents -> split on whitespace / punctuation / CJK -> greedy longest-match WordPiece
// with "##" continuations -> [CLS] ... [SEP], truncated.  embedding/wordpiece.py restates that in Python
// (token-for-token equal to transformers.BertTokenizer, tests/test_wordpiece.py) at ~600 texts/s per
// core -- 250x slower than the encoder it feeds.  For text made of ASCII bytes only, the Unicode steps
// (NFC / NFD, accent stripping, category lookups, CJK spacing) are identities or simple byte tests, so
// this file implements exactly the remaining steps; a text with any byte >= 0x80 is reported back
// (length -1) and the caller sends it through the Python implementation.
#include <stdint.h>

#include <atomic>
#include <cstring>
#include <string>
#include <string_view>
#include <thread>
#include <unordered_map>
#include <vector>

#include "host_util.h"

namespace {

struct WordPiece {
  std::vector<std::string> pieces;                              // owns the vocabulary strings
  std::unordered_map<std::string_view, int32_t> vocab;          // views into `pieces`
  int32_t pad_id = 0, unk_id = 0, cls_id = 0, sep_id = 0;
  bool lower = true;
  int max_chars_per_word = 100;
};

inline bool is_space(unsigned char c) { return c == ' ' || c == '\t' || c == '\n' || c == '\r'; }
// Unicode category C* among ASCII: the C0 controls (other than \t \n \r, handled before) and DEL
inline bool is_control(unsigned char c) { return c < 32 || c == 127; }
inline bool is_punct(unsigned char c) {
  return (c >= 33 && c <= 47) || (c >= 58 && c <= 64) || (c >= 91 && c <= 96) || (c >= 123 && c <= 126);
}
inline bool never_split(std::string_view w) {
  return w == "[PAD]" || w == "[UNK]" || w == "[CLS]" || w == "[SEP]" || w == "[MASK]";
}

// greedy longest-match-first; a word with an unmatchable remainder becomes a single [UNK]
void wordpiece_word(const WordPiece& wp, std::string_view word, std::string& scratch, std::vector<int32_t>& out) {
  if ((int)word.size() > wp.max_chars_per_word) {
    out.push_back(wp.unk_id);
    return;
  }
  const size_t mark = out.size();
  size_t start = 0;
  const size_t n = word.size();
  while (start < n) {
    size_t end = n;
    int32_t found = -1;
    while (start < end) {
      std::string_view sub;
      if (start > 0) {
        scratch.assign("##");
        scratch.append(word.data() + start, end - start);
        sub = scratch;
      } else {
        sub = word.substr(0, end);
      }
      auto it = wp.vocab.find(sub);
      if (it != wp.vocab.end()) {
        found = it->second;
        break;
      }
      --end;
    }
    if (found < 0) {
      out.resize(mark);
      out.push_back(wp.unk_id);
      return;
    }
    out.push_back(found);
    start = end;
  }
}

struct ThreadState {
  std::unordered_map<std::string, std::vector<int32_t>> cache;   // piece (lower-cased, punctuation-free run) -> ids
  std::string word, piece, scratch;
  std::vector<int32_t> body;
};

// pieces of one cleaned, whitespace-delimited word appended to ts.body
void encode_word(const WordPiece& wp, ThreadState& ts) {
  std::string& w = ts.word;
  if (w.empty()) return;
  if (never_split(w)) {                        // special tokens pass through as written (wordpiece.py:_split_word)
    wordpiece_word(wp, w, ts.scratch, ts.body);
    return;
  }
  if (wp.lower)
    for (char& c : w)
      if (c >= 'A' && c <= 'Z') c = (char)(c + 32);
  auto flush = [&](std::string& p) {
    if (p.empty()) return;
    auto it = ts.cache.find(p);
    if (it == ts.cache.end()) {
      std::vector<int32_t> ids;
      wordpiece_word(wp, p, ts.scratch, ids);
      if (ts.cache.size() < 200000) it = ts.cache.emplace(p, std::move(ids)).first;
      else {
        ts.body.insert(ts.body.end(), ids.begin(), ids.end());
        p.clear();
        return;
      }
    }
    ts.body.insert(ts.body.end(), it->second.begin(), it->second.end());
    p.clear();
  };
  std::string& p = ts.piece;
  p.clear();
  for (char c : w) {
    if (is_punct((unsigned char)c)) {          // every punctuation character is a piece of its own
      flush(p);
      p.assign(1, c);
      flush(p);
    } else {
      p.push_back(c);
    }
  }
  flush(p);
}

// returns the number of ids written ([CLS] body [SEP], at most max_length), or -1 for non-ASCII text
int encode_text(const WordPiece& wp, ThreadState& ts, const char* text, size_t len, int max_length, int32_t* out) {
  for (size_t i = 0; i < len; ++i)
    if ((unsigned char)text[i] >= 0x80) return -1;
  ts.body.clear();
  ts.word.clear();
  const size_t body_max = (size_t)(max_length > 2 ? max_length - 2 : 0);
  for (size_t i = 0; i <= len; ++i) {
    const unsigned char c = i < len ? (unsigned char)text[i] : (unsigned char)' ';
    if (is_space(c)) {
      encode_word(wp, ts);
      ts.word.clear();
      if (ts.body.size() >= body_max) break;   // the rest would be truncated away
    } else if (c == 0 || is_control(c)) {
      continue;                                // dropped by the cleaning step, also inside a word
    } else {
      ts.word.push_back((char)c);
    }
  }
  const size_t nb = ts.body.size() < body_max ? ts.body.size() : body_max;
  int n = 0;
  if (max_length >= 1) out[n++] = wp.cls_id;
  for (size_t i = 0; i < nb && n < max_length; ++i) out[n++] = ts.body[i];
  if (n < max_length || max_length < 1) {
    if (max_length >= 1) out[n++] = wp.sep_id;
  }
  return n;
}

}  // namespace

struct b200_wordpiece {
  WordPiece wp;
};

extern "C" {

int b200_wordpiece_create(const char* vocab_blob, const int64_t* offsets, int32_t n_vocab, int do_lower_case,
                          b200_wordpiece_t** out) {
  if (!vocab_blob || !offsets || n_vocab <= 0 || !out) B200_FAIL(B200_ERR_INVALID, "bad vocabulary arguments");
  auto* h = new b200_wordpiece();
  h->wp.lower = do_lower_case != 0;
  h->wp.pieces.reserve((size_t)n_vocab);
  for (int32_t i = 0; i < n_vocab; ++i)
    h->wp.pieces.emplace_back(vocab_blob + offsets[i], (size_t)(offsets[i + 1] - offsets[i]));
  h->wp.vocab.reserve((size_t)n_vocab * 2);
  for (int32_t i = 0; i < n_vocab; ++i) h->wp.vocab.emplace(std::string_view(h->wp.pieces[(size_t)i]), i);   // first id wins
  const char* names[4] = {"[PAD]", "[UNK]", "[CLS]", "[SEP]"};
  int32_t* ids[4] = {&h->wp.pad_id, &h->wp.unk_id, &h->wp.cls_id, &h->wp.sep_id};
  for (int k = 0; k < 4; ++k) {
    auto it = h->wp.vocab.find(std::string_view(names[k]));
    if (it == h->wp.vocab.end()) {
      delete h;
      B200_FAIL(B200_ERR_INVALID, "vocab is missing the special token %s", names[k]);
    }
    *ids[k] = it->second;
  }
  *out = h;
  return B200_OK;
}

int b200_wordpiece_destroy(b200_wordpiece_t* h) {
  delete h;
  return B200_OK;
}

int b200_wordpiece_encode(const b200_wordpiece_t* h, const char* texts_blob, const int64_t* offsets, int64_t n_texts,
                          int32_t max_length, int32_t* ids, int32_t* lengths, int32_t n_threads) {
  if (!h) B200_FAIL(B200_ERR_INVALID, "tokenizer is NULL");
  if (n_texts < 0 || max_length < 2) B200_FAIL(B200_ERR_INVALID, "n_texts must be >= 0 and max_length >= 2");
  if (n_texts == 0) return B200_OK;
  if (!texts_blob || !offsets || !ids || !lengths) B200_FAIL(B200_ERR_INVALID, "NULL buffer");
  int nt = n_threads > 0 ? n_threads : (int)std::thread::hardware_concurrency();
  if (nt < 1) nt = 1;
  if ((int64_t)nt > (n_texts + 15) / 16) nt = (int)((n_texts + 15) / 16);     // at least 16 texts per thread
  std::atomic<int64_t> next{0};
  auto work = [&]() {
    ThreadState ts;
    for (;;) {
      const int64_t i0 = next.fetch_add(16);
      if (i0 >= n_texts) break;
      const int64_t i1 = i0 + 16 < n_texts ? i0 + 16 : n_texts;
      for (int64_t i = i0; i < i1; ++i) {
        int32_t* row = ids + (size_t)i * max_length;
        const int n = encode_text(h->wp, ts, texts_blob + offsets[i], (size_t)(offsets[i + 1] - offsets[i]), max_length, row);
        lengths[i] = n;
        for (int j = n < 0 ? 0 : n; j < max_length; ++j) row[j] = h->wp.pad_id;
      }
    }
  };
  if (nt == 1) {
    work();
  } else {
    std::vector<std::thread> pool;
    pool.reserve((size_t)nt);
    for (int t = 0; t < nt; ++t) pool.emplace_back(work);
    for (auto& t : pool) t.join();
  }
  return B200_OK;
}

}  // extern "C"
