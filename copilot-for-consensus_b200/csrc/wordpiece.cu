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

#include <mutex>

#include "host_util.h"
#include "unicode_tables.h"

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

// ---- Basic Multilingual Plane tables (generated from Python's unicodedata, see unicode_tables.h) ----
uint8_t g_cls[65536];            // 0 ordinary, 1 whitespace, 2 dropped, 3 punctuation, 4 CJK, 5 "use the Python path"
int16_t g_fold_idx[65536];       // index into kFolds, -1 = the character folds to itself
std::once_flag g_tables_once;

void init_tables() {
  std::call_once(g_tables_once, [] {
    size_t cp = 0;
    for (int r = 0; r < b200_unicode::kNumClassRuns; ++r)
      for (uint16_t k = 0; k < b200_unicode::kClassRuns[r][0]; ++k) g_cls[cp++] = (uint8_t)b200_unicode::kClassRuns[r][1];
    for (size_t i = 0; i < 65536; ++i) g_fold_idx[i] = -1;
    for (int i = 0; i < b200_unicode::kNumFolds; ++i) g_fold_idx[b200_unicode::kFolds[i].cp] = (int16_t)i;
  });
}

inline void utf8_append(std::string& s, uint32_t cp) {     // cp <= 0xFFFF
  if (cp < 0x80) s.push_back((char)cp);
  else if (cp < 0x800) {
    s.push_back((char)(0xC0 | (cp >> 6)));
    s.push_back((char)(0x80 | (cp & 0x3F)));
  } else {
    s.push_back((char)(0xE0 | (cp >> 12)));
    s.push_back((char)(0x80 | ((cp >> 6) & 0x3F)));
    s.push_back((char)(0x80 | (cp & 0x3F)));
  }
}

// greedy longest-match-first; a word with an unmatchable remainder becomes a single [UNK]
void wordpiece_word(const WordPiece& wp, std::string_view word, std::string& scratch, std::vector<int32_t>& out) {
  // character boundaries (UTF-8): the match shrinks one CHARACTER at a time, the length limit counts characters
  size_t bounds[104];
  size_t nchar = 0;
  for (size_t i = 0; i < word.size(); ++i)
    if (((unsigned char)word[i] & 0xC0) != 0x80) {
      if (nchar > (size_t)wp.max_chars_per_word) break;
      bounds[nchar++] = i;
    }
  if ((int)nchar > wp.max_chars_per_word) {
    out.push_back(wp.unk_id);
    return;
  }
  bounds[nchar] = word.size();
  const size_t mark = out.size();
  size_t start = 0;                          // in characters
  while (start < nchar) {
    size_t end = nchar;
    int32_t found = -1;
    while (start < end) {
      std::string_view sub;
      if (start > 0) {
        scratch.assign("##");
        scratch.append(word.data() + bounds[start], bounds[end] - bounds[start]);
        sub = scratch;
      } else {
        sub = word.substr(0, bounds[end]);
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
  std::vector<uint16_t> uword;                                     // code points of the current word (BMP text)
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

// ---- text with non-ASCII characters of the Basic Multilingual Plane (uncased models) ----
// pieces of one cleaned word given as code points: fold (lower-case, NFD, drop accents), split at
// punctuation, WordPiece each run
void flush_piece(const WordPiece& wp, ThreadState& ts) {
  std::string& p = ts.piece;
  if (p.empty()) return;
  auto it = ts.cache.find(p);
  if (it == ts.cache.end()) {
    std::vector<int32_t> ids;
    wordpiece_word(wp, p, ts.scratch, ids);
    if (ts.cache.size() >= 200000) {
      ts.body.insert(ts.body.end(), ids.begin(), ids.end());
      p.clear();
      return;
    }
    it = ts.cache.emplace(p, std::move(ids)).first;
  }
  ts.body.insert(ts.body.end(), it->second.begin(), it->second.end());
  p.clear();
}

void encode_word_unicode(const WordPiece& wp, ThreadState& ts) {
  std::vector<uint16_t>& w = ts.uword;
  if (w.empty()) return;
  bool ascii = true;
  for (uint16_t c : w) ascii = ascii && c < 0x80;
  if (ascii) {                                 // includes the special tokens, which must not be folded
    ts.word.assign(w.begin(), w.end());
    encode_word(wp, ts);
    return;
  }
  ts.piece.clear();
  auto emit = [&](uint32_t o) {
    if (g_cls[o] == 3) {                       // punctuation (tested on the folded character): a piece of its own
      flush_piece(wp, ts);
      utf8_append(ts.piece, o);
      flush_piece(wp, ts);
    } else {
      utf8_append(ts.piece, o);
    }
  };
  for (uint16_t c : w) {
    if (c >= 0xAC00 && c <= 0xD7A3) {          // Hangul syllable: NFD is arithmetic (L V [T] jamo)
      const uint32_t si = c - 0xAC00u;
      emit(0x1100 + si / 588);
      emit(0x1161 + (si % 588) / 28);
      if (si % 28) emit(0x11A7 + si % 28);
      continue;
    }
    const int16_t fi = g_fold_idx[c];
    if (fi < 0) {
      emit(c);
    } else {
      const b200_unicode::Fold& f = b200_unicode::kFolds[fi];
      for (int k = 0; k < f.n; ++k) emit(f.out[k]);
    }
  }
  flush_piece(wp, ts);
}

// returns the ids written, or -1 when the text must go to the Python implementation
int encode_text_unicode(const WordPiece& wp, ThreadState& ts, const char* text, size_t len, int max_length, int32_t* out) {
  ts.body.clear();
  ts.uword.clear();
  const size_t body_max = (size_t)(max_length > 2 ? max_length - 2 : 0);
  size_t i = 0;
  bool stop = false;
  while (i <= len && !stop) {
    uint32_t cp;
    if (i == len) {
      cp = ' ';
      ++i;
    } else {
      const unsigned char b0 = (unsigned char)text[i];
      if (b0 < 0x80) {
        cp = b0;
        i += 1;
      } else if ((b0 & 0xE0) == 0xC0 && i + 1 < len && ((unsigned char)text[i + 1] & 0xC0) == 0x80) {
        cp = ((uint32_t)(b0 & 0x1F) << 6) | ((unsigned char)text[i + 1] & 0x3F);
        if (cp < 0x80) return -1;              // overlong
        i += 2;
      } else if ((b0 & 0xF0) == 0xE0 && i + 2 < len && ((unsigned char)text[i + 1] & 0xC0) == 0x80 &&
                 ((unsigned char)text[i + 2] & 0xC0) == 0x80) {
        cp = ((uint32_t)(b0 & 0x0F) << 12) | (((uint32_t)(unsigned char)text[i + 1] & 0x3F) << 6) |
             ((unsigned char)text[i + 2] & 0x3F);
        if (cp < 0x800) return -1;             // overlong
        i += 3;
      } else {
        return -1;                             // beyond the BMP (4-byte sequences) or malformed: Python decides
      }
    }
    const uint8_t c = g_cls[cp];
    if (c == 5) return -1;
    if (c == 2) continue;
    if (c == 1 || c == 4) {
      encode_word_unicode(wp, ts);
      ts.uword.clear();
      if (c == 4) {                            // an ideograph is a word of its own
        ts.uword.push_back((uint16_t)cp);
        encode_word_unicode(wp, ts);
        ts.uword.clear();
      }
      if (ts.body.size() >= body_max) stop = true;
    } else {
      ts.uword.push_back((uint16_t)cp);
    }
  }
  const size_t nb = ts.body.size() < body_max ? ts.body.size() : body_max;
  int n = 0;
  out[n++] = wp.cls_id;
  for (size_t k = 0; k < nb; ++k) out[n++] = ts.body[k];
  out[n++] = wp.sep_id;
  return n;
}

// returns the number of ids written ([CLS] body [SEP], at most max_length), or -1 for text the Python
// implementation has to handle (cased models with non-ASCII text, characters outside the tables)
int encode_text(const WordPiece& wp, ThreadState& ts, const char* text, size_t len, int max_length, int32_t* out) {
  for (size_t i = 0; i < len; ++i)
    if ((unsigned char)text[i] >= 0x80) return wp.lower ? encode_text_unicode(wp, ts, text, len, max_length, out) : -1;
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
  init_tables();
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
