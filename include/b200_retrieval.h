/*
 * b200_retrieval.h -- C ABI of the B200-native retrieval hot path.
 *
 * This is the boundary a maintainer of Copilot-for-Consensus binds (ctypes) to
 * replace the arithmetic behind two adapter drivers, citations relative to the reference tree:
 *
 *   adapters/copilot_embedding/copilot_embedding/sentence_transformer_provider.py:76-94
 *       SentenceTransformerEmbeddingProvider.embed -> self.model.encode(...)
 *   adapters/copilot_vectorstore/copilot_vectorstore/inmemory.py:35-178
 *       InMemoryVectorStore.add_embedding(s) / query / delete / clear / count / get
 *
 * Conventions
 *   - plain C types only; no exceptions cross the boundary.  Every function returns 0 on success
 *     or a negative b200_status; b200_last_error() gives the calling thread's last message.
 *   - handles are opaque and own their device memory; buffers passed in are borrowed for the call.
 *   - `*_on_device` flags say whether a pointer is a device pointer on the handle's GPU (1) or a
 *     host pointer (0; pinned memory makes the copies asynchronous).
 *   - `stream` is a cudaStream_t passed as void* (NULL = the legacy default stream).  Calls are
 *     asynchronous with respect to the host when every buffer is on the device; calls with host
 *     output buffers synchronise the stream before returning.
 *   - a handle may be used from several threads and streams.  Calls on one handle are serialised on the
 *     host by the handle's mutex and on the GPU by an event the handle keeps: a call that arrives on a
 *     different stream than the previous one first makes its stream wait for that call's last kernel
 *     (every handle owns one set of scratch buffers).  b200_index_search_static is the exception: see there.
 *   - the library never reads environment variables (the reference pins its adapters to that:
 *     adapters/copilot_vectorstore/tests/test_factory.py:75-107, scripts/check_no_runtime_env_vars.py:6-16).
 *   - there is no CPU fallback: without a usable sm_100 device every call fails with
 *     B200_ERR_CUDA / B200_ERR_UNSUPPORTED.
 */
#ifndef B200_RETRIEVAL_H
#define B200_RETRIEVAL_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum b200_status {
  B200_OK = 0,
  B200_ERR_INVALID = -1,     /* bad argument (maps to ValueError) */
  B200_ERR_DIMENSION = -2,   /* vector dimension mismatch (ValueError "dimension", inmemory.py:100-104) */
  B200_ERR_NOT_FOUND = -3,   /* row out of range (KeyError "not found", inmemory.py:139,170) */
  B200_ERR_CUDA = -4,        /* CUDA runtime / driver failure (RuntimeError) */
  B200_ERR_UNSUPPORTED = -5, /* shape outside what the kernels implement (RuntimeError) */
  B200_ERR_NOMEM = -6        /* device allocation failed (MemoryError) */
} b200_status;

typedef struct b200_index b200_index_t;
typedef struct b200_encoder b200_encoder_t;
typedef struct b200_wordpiece b200_wordpiece_t;

/* ---- library ------------------------------------------------------------------------------ */
const char* b200_last_error(void);
int b200_version(void);
/* number of visible sm_100 devices; 0 when there is none (then nothing else will work) */
int b200_device_count(void);
/* Process-wide switches between kernel variants, for A/B timing and for the tests that drive every
 * alternative path against the oracle (no reference counterpart).  Names and defaults:
 *   pdl 0 | gemm_pair 1 | gemm_ast 1 | attention_tc 1 | att_exact_max 0 | att_turns 1 | att_poly 1 (0..4) |
 *   fuse_ln 1 | fuse_ffn 1 | serpentine 1 | sim_pair 1 | sim_bound 3 | sim_refresh_mask 0 | search_certify 1 | sim_pair_kp 0 | sim_nsplit 0 | search_screen 1 | att_norms_fused 1 | defer_ln 1 | sim_drain 0
 * Unknown names and out-of-range values are refused with B200_ERR_INVALID. */
int b200_set_option(const char* name, int value);
int b200_get_option(const char* name, int* value);

/* ---- vector index: replaces InMemoryVectorStore's storage + query arithmetic ---------------
 * Rows are identified by their position in insertion order (row 0 = first inserted), which is
 * exactly the reference's tie-break order (stable sort at inmemory.py:118).  The id <-> row map
 * and the metadata dicts stay in the Python wrapper. */

enum {
  B200_INDEX_MASTER_F32 = 0,  /* keep the caller's fp32 rows (drop-in mode; SearchResult.vector, inmemory.py:124) */
  B200_INDEX_MASTER_BF16 = 1  /* rows are given as bf16 and kept as bf16 (BASELINE.json "bf16 corpus") */
};

/* dim: vector dimension (1..4096; up to 384 the query tile stays resident in shared memory, beyond that it is
 * streamed); capacity: initial row capacity, grows on demand.  Besides the caller's rows (the master copy) the
 * index keeps a unit-normalised 16-bit copy (IEEE half) that the tensor-core pass streams. */
int b200_index_create(int dim, int64_t capacity, int device, int master_dtype, b200_index_t** out);
int b200_index_destroy(b200_index_t* idx);
/* inmemory.py:145-148 clear() */
int b200_index_clear(b200_index_t* idx);
/* inmemory.py:150-156 count() */
int64_t b200_index_count(const b200_index_t* idx);
int b200_index_dim(const b200_index_t* idx);
int b200_index_reserve(b200_index_t* idx, int64_t capacity);

/* inmemory.py:35-76 add_embedding(s): append n rows [n, dim] (fp32 for MASTER_F32 indexes, bf16 bit
 * patterns for MASTER_BF16).  Builds the unit-normalised bf16 search copy (kernel K11). */
int b200_index_add(b200_index_t* idx, const void* rows, int64_t n, int rows_on_device, void* stream);

/* inmemory.py:130-143 delete(): remove rows (ascending, unique) and close the gaps so that the
 * remaining rows keep their relative (insertion) order. */
int b200_index_remove_rows(b200_index_t* idx, const int64_t* rows_host, int64_t n, void* stream);

/* inmemory.py:124,174 SearchResult.vector / get(): copy n stored rows (as fp32) to out[n, dim]. */
int b200_index_get_rows(b200_index_t* idx, const int64_t* rows_host, int64_t n, float* out, int out_on_device,
                        void* stream);

/* inmemory.py:78-128 query(), batched: for each of nq raw query vectors return the k best rows by
 * cosine similarity, scores computed as the reference does (fp32, eps 1e-8), ordered by
 * (score descending, row ascending).  ids[nq*k] = id_base + row, -1 where fewer than k rows exist;
 * scores[nq*k].  q_dtype: 0 = fp32, 1 = bf16.  k <= 1024: up to 48 the fused
 * select serves it; 49..1024 (orchestrator/app/context_sources.py:66 asks for 50) scores every row exactly, one
 * query at a time -- same ids, order and scores, HBM-bound on the master copy.
 *
 * Exactness.  The reference scores EVERY row exactly (inmemory.py:111-119).  The fused path ranks rows by a
 * 16-bit tensor-core score, keeps KP = 20 (16 in batches of more than 128 queries over fewer than 2.5 M rows) / 32 / 64 candidates per query (k <= 10 / 24 / 48) and rescores those
 * exactly.  It then proves, per query, that no other row can belong to the exact top-k: every row outside the
 * candidates has a 16-bit score <= A (the KP-th best), its exact score can exceed that by at most
 * eps = r_v + r_q (1 + r_v) + slop (r_v: largest rounding distance of any stored row, tracked at insert;
 * r_q: this query's rounding distance; slop: fp32 evaluation error), and the k-th exact score must lie strictly
 * above A + eps.  A query that fails the test -- near-duplicate rows, ties -- is re-scored against every row in
 * the same call (no host round trip: the kernels are always launched and return at once when nothing failed):
 * first by a screened pass (one cheap fp32 score per row, the reference's arithmetic for the best 64 rows, final
 * when the 64th lies clearly below the k-th), then, for clusters of more than 64 - k near-equal rows, by the
 * reference's arithmetic on every row.  So the ids, their order and the fp32 score bits are those of the
 * exhaustive computation, always; the cost of an uncertified query is one or two passes over the master copy. */
int b200_index_search(b200_index_t* idx, const void* queries, int q_dtype, int queries_on_device, int64_t nq,
                      int k, int64_t id_base, int64_t* ids, float* scores, int out_on_device, void* stream);

/* Cumulative counters of the certificate above (no reference counterpart): queries checked by the fused path
 * since the index was created, and how many of them had to be re-scored exhaustively.  Synchronises the device. */
int b200_index_certificate_stats(b200_index_t* idx, int64_t* checked, int64_t* recomputed);

/* adapters/copilot_vectorstore/copilot_vectorstore/faiss_store.py:99-101 (IndexFlatL2), :212-224 (search):
 * the k nearest rows by squared Euclidean distance, ascending, ties to the lower row; dist[nq*k] holds the
 * squared distances (the store turns them into scores 1/(1+d), faiss_store.py:220).  Exact brute force over the
 * master copy -- what a flat FAISS index is -- one HBM pass per query; fp32 queries, k <= 64. */
int b200_index_search_l2(b200_index_t* idx, const void* queries, int q_dtype, int queries_on_device, int64_t nq,
                         int k, int64_t id_base, int64_t* ids, float* dist, int out_on_device, void* stream);

/* b200_index_search for callers that must not trigger a device allocation (this entry point has no
 * counterpart in the reference).  Meant for processes that gave all free HBM to their own pool at
 * start-up, where a cudaMalloc issued from inside a library fails, and for query threads that cannot
 * afford the device-wide synchronisation cudaMalloc/cudaFree imply.  Same result as
 * b200_index_search, with these differences:
 *   - it calls no cudaMalloc / cudaFree / allocating library routine.  All scratch lives in a fixed
 *     4.7 MB array inside the library image, resident on every GPU that loads the library until exit.
 *   - limits: nq <= 128 per call, index dimension <= 384, k <= 48; larger requests are refused with
 *     B200_ERR_UNSUPPORTED and a message naming the limit.
 *   - queries, ids and scores must be device pointers on the index's GPU.
 *   - every call on a GPU shares that one array (across all indexes and threads).  The caller never
 *     has to clear it -- each call zeroes the 34 KB that must start at zero and rewrites the rest
 *     before reading it -- but two calls whose kernels overlap on the GPU (different streams) race on
 *     it: keep such calls on one stream or order them yourself (the per-handle stream ordering of the
 *     other entry points does not cover this array, which all indexes share).
 *   - the first call on each device looks up the array's device address (cudaGetSymbolAddress) and
 *     has the driver load the code of every kernel this entry point can launch (all k bands, both query
 *     dtypes), so later calls load nothing; make that first call before any stream capture begins on
 *     that device.
 *   - the exactness certificate of b200_index_search applies here too, with its bookkeeping and the
 *     exhaustive kernel's scratch inside the same array. */
int b200_index_search_static(b200_index_t* idx, const void* queries_dev, int q_dtype, int64_t nq, int k,
                             int64_t id_base, int64_t* ids_dev, float* scores_dev, void* stream);

/* Multi-GPU search without glue kernels (no reference counterpart: the reference is single-process).
 * b200_index_search_packed is b200_index_search writing each hit as ONE 16-byte record
 *   { uint64 orderable(score) << 32 (0 = no hit), int64 id_base + row (-1 = no hit) }
 * into hits_dev[nq*k] -- which may be the send buffer of an all-gather.  b200_merge_hits takes the gathered
 * buffer, laid out [nlists][nq_per_list][k_in] (list l = shard l's records), and writes the k_out best of
 * queries [q_first, q_first + nq) under (score desc, id asc) as ids_out[nq*k_out] / scores_out[nq*k_out].
 * Everything is a device pointer; between a shard's search and the merge there is the collective and nothing else. */
int b200_index_search_packed(b200_index_t* idx, const void* queries_dev, int q_dtype, int64_t nq, int k,
                             int64_t id_base, void* hits_dev, void* stream);
int b200_merge_hits(const void* hits_in, int nlists, int64_t nq_per_list, int k_in, int64_t q_first, int64_t nq,
                    int k_out, int64_t* ids_out, float* scores_out, int device, void* stream);

/* The same exchange without the collective (no reference counterpart): every rank's receive buffer and flag array are
 * mapped into every other rank (NVLink peer memory, e.g. torch symmetric memory).  b200_index_search_exchange runs the
 * search and has its last kernels STORE this shard's hit records straight into slot `rank` of every rank's buffer
 *   peer_hits[i] -> rank i's buffer for this search, laid out [world][nq][k] records; peer_flags[i] -> rank i's uint32[world]
 * then publishes `epoch` into flag `rank` of every rank (release, system scope).  b200_merge_hits_wait is b200_merge_hits
 * that first waits until all `nlists` flags of THIS rank have reached `epoch`.  Use two receive buffers alternately
 * (epoch parity) and epochs that grow by one per search: a rank can then run at most one search ahead of the slowest,
 * into the other buffer, and no further synchronisation is needed.  world <= 8. */
int b200_index_search_exchange(b200_index_t* idx, const void* queries_dev, int q_dtype, int64_t nq, int k, int64_t id_base,
                               void* const* peer_hits, void* const* peer_flags, int world, int rank, uint32_t epoch,
                               void* stream);
int b200_merge_hits_wait(const void* hits_local, const void* flags_local, uint32_t epoch, int nlists, int64_t nq_per_list,
                         int k_in, int64_t q_first, int64_t nq, int k_out, int64_t* ids_out, float* scores_out, int device,
                         void* stream);

/* Multi-GPU candidate merge (no reference counterpart: the reference is single-process).  After an
 * all-gather of per-shard results laid out [nlists][nq][k_in], keep the best k_out per query under
 * the same (score desc, id asc) order.  All pointers are device pointers. */
int b200_merge_topk(const int64_t* ids_in, const float* scores_in, int64_t nq, int nlists, int k_in, int k_out,
                    int64_t* ids_out, float* scores_out, int device, void* stream);

/* ---- encoder: replaces SentenceTransformer.encode for BERT-class models ---------------------
 * sentence_transformer_provider.py:93 -> tokenise (host, not here) -> BertModel -> masked mean
 * pool -> L2 normalise.  Weights are handed over by HuggingFace state_dict key. */

typedef struct b200_encoder_config {
  int32_t vocab_size;
  int32_t hidden;       /* 384 (MiniLM-L6) or 768 (mpnet-base class) */
  int32_t layers;
  int32_t heads;
  int32_t intermediate;
  int32_t max_pos;
  int32_t type_vocab;
  float ln_eps;
} b200_encoder_config_t;

int b200_encoder_create(const b200_encoder_config_t* cfg, int device, b200_encoder_t** out);
int b200_encoder_destroy(b200_encoder_t* enc);
/* name: HF BertModel state_dict key (e.g. "encoder.layer.0.attention.self.query.weight");
 * data: fp32 host buffer of numel elements. */
int b200_encoder_load_tensor(b200_encoder_t* enc, const char* name, const float* data, int64_t numel);
/* checks that every tensor arrived, packs Q/K/V, converts GEMM operands to bf16 */
int b200_encoder_finalize(b200_encoder_t* enc);
int b200_encoder_hidden(const b200_encoder_t* enc);
/* input_ids[B*S] int32 (0-padded), lengths[B] int32 valid tokens per row (attention mask = first
 * lengths[b] tokens); out[B*hidden] fp32 L2-normalised embeddings. */
int b200_encode(b200_encoder_t* enc, const int32_t* input_ids, const int32_t* lengths, int batch, int seq,
                int inputs_on_device, float* out, int out_on_device, void* stream);

/* Diagnostics (no reference counterpart): copy the first numel elements of an internal bf16
 * activation buffer, as left by the last b200_encode call, to a host fp32 buffer.
 * which: 0 = x (hidden state), 1 = qkv, 2 = ctx (attention output), 3 = y (pre-LayerNorm), 4 = ffn. */
int b200_encoder_peek(b200_encoder_t* enc, int which, float* out_host, int64_t numel);

/* ---- host-side WordPiece ---------------------------------------------------------------------------
 * Replaces the tokenisation `SentenceTransformer.encode` performs inside
 * sentence_transformer_provider.py:93 (uncased BERT recipe: clean, lower-case + strip accents, split on
 * whitespace / punctuation / CJK, greedy longest-match WordPiece, [CLS] ... [SEP], truncation) for ASCII
 * text and, for uncased models, for text of the Basic Multilingual Plane (character classes and folded
 * forms come from tables generated out of Python's unicodedata: csrc/unicode_tables.h).  No CUDA involved.
 *   vocab_blob / offsets[n_vocab + 1]: the pieces of vocab.txt concatenated (UTF-8), id = line number.
 *   texts_blob / offsets[n_texts + 1]: the texts concatenated (UTF-8).
 *   ids[n_texts * max_length] (padded with [PAD]), lengths[n_texts]; lengths[i] = -1 marks a text the
 *   tables do not cover (combining marks, characters beyond U+FFFF, GREEK CAPITAL SIGMA, malformed UTF-8,
 *   any non-ASCII text of a cased model), which the caller must tokenise with the full implementation
 *   (copilot-for-consensus_b200/embedding/wordpiece.py).  n_threads <= 0: one per hardware thread. */
int b200_wordpiece_create(const char* vocab_blob, const int64_t* offsets, int32_t n_vocab, int do_lower_case,
                          b200_wordpiece_t** out);
int b200_wordpiece_destroy(b200_wordpiece_t* h);
int b200_wordpiece_encode(const b200_wordpiece_t* h, const char* texts_blob, const int64_t* offsets, int64_t n_texts,
                          int32_t max_length, int32_t* ids, int32_t* lengths, int32_t n_threads);

/* Optional per-kernel timing (no reference counterpart), used by bench.py for the roofline line.
 * While enabled, every kernel launch is bracketed by CUDA events on its stream.  profile_read
 * synchronises the device and writes one line per kernel name into buf:
 *   "<name> <launches> <total_ms> <algorithmic_flops> <algorithmic_bytes>\n"
 * and returns the number of characters written. */
int b200_profile_enable(int on);
int b200_profile_reset(void);
int64_t b200_profile_read(char* buf, int64_t cap);

#ifdef __cplusplus
}
#endif
#endif /* B200_RETRIEVAL_H */
