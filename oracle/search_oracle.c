/*
 * search_oracle.c -- TEST INFRASTRUCTURE, not product code.
 *
 * CPU restatement of the reference's in-memory cosine top-k
 * (adapters/copilot_vectorstore/copilot_vectorstore/inmemory.py:78-128):
 *
 *   q_hat = q / (||q|| + 1e-8)                       inmemory.py:108
 *   for every stored v, in insertion order:          inmemory.py:111
 *       v_hat = v / (||v|| + 1e-8)                   inmemory.py:113
 *       s     = dot(q_hat, v_hat)        (fp32)      inmemory.py:114
 *   stable sort by s descending, take [:top_k]       inmemory.py:118-119
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
 * reference leg may load this library; the product path never does.
 *
 * What is fixed here that the reference leaves open: the reference's three
 * reductions (||q||^2, ||v||^2, q_hat.v_hat) go through BLAS sdot, whose
 * summation order depends on the BLAS build and the CPU.  To make "ids
 * identical" a well-posed statement this restatement -- and the CUDA rescoring
 * kernel that must agree with it bit for bit -- uses ONE canonical order:
 *
 *   element i is accumulated into partial (i & 31) with a fused multiply-add,
 *   elements of one partial in ascending i; the 32 partials are then folded
 *   by the halving tree  p[l] += p[l+16], p[l] += p[l+8], ... p[0] += p[1].
 *
 * Everything else is the reference's arithmetic verbatim: fp32 throughout,
 * IEEE sqrt and divide, epsilon 1e-8 added in fp32 (NumPy >= 2 scalar rules;
 * numpy 2.3.5 is what this image carries), ties broken by the lower row index
 * (= earlier insertion, what Python's stable sort does at inmemory.py:118).
 *
 * Pinned against the real reference by oracle/gen_golden.py ->
 * tests/golden/search_*.npz and tests/test_oracle_search.py.
 *
 * Build: see oracle/Makefile (gcc -O3 -mavx2 -mfma -ffp-contract=off -fopenmp).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define B200O_LANES 32

/* canonical-order fp32 dot product (see header) */
float b200o_dot(const float *x, const float *y, int d)
{
    float p[B200O_LANES];
    for (int l = 0; l < B200O_LANES; ++l) p[l] = 0.0f;
    int i = 0;
    for (; i + B200O_LANES <= d; i += B200O_LANES)
        for (int l = 0; l < B200O_LANES; ++l)
            p[l] = __builtin_fmaf(x[i + l], y[i + l], p[l]);
    for (int l = 0; i + l < d; ++l)
        p[l] = __builtin_fmaf(x[i + l], y[i + l], p[l]);
    for (int s = B200O_LANES / 2; s >= 1; s >>= 1)
        for (int l = 0; l < s; ++l) p[l] = p[l] + p[l + s];
    return p[0];
}

/* ||x||_2 as np.linalg.norm computes it for a 1-D fp32 array: sqrt(dot(x,x)) */
float b200o_norm(const float *x, int d) { return sqrtf(b200o_dot(x, x, d)); }

/* x / (||x|| + 1e-8), element-wise fp32 (inmemory.py:108,113) */
void b200o_normalize(const float *x, int d, float *out)
{
    const float den = b200o_norm(x, d) + 1e-8f;
    for (int i = 0; i < d; ++i) out[i] = x[i] / den;
}

/* the reference score of one (query,row) pair, both given raw */
float b200o_score(const float *q, const float *v, int d)
{
    float qh[d > 0 ? d : 1], vh[d > 0 ? d : 1];
    b200o_normalize(q, d, qh);
    b200o_normalize(v, d, vh);
    return b200o_dot(qh, vh, d);
}

/* scores of one raw query against n raw rows */
void b200o_scores(const float *q, const float *rows, int64_t n, int d, float *out)
{
    float *qh = (float *)malloc(sizeof(float) * (size_t)(d > 0 ? d : 1));
    b200o_normalize(q, d, qh);
#pragma omp parallel
    {
        float *vh = (float *)malloc(sizeof(float) * (size_t)(d > 0 ? d : 1));
#pragma omp for schedule(static)
        for (int64_t r = 0; r < n; ++r) {
            b200o_normalize(rows + r * (int64_t)d, d, vh);
            out[r] = b200o_dot(qh, vh, d);
        }
        free(vh);
    }
    free(qh);
}

/* one sorted candidate list: score descending, row ascending on equal score */
typedef struct { float s; int64_t r; } cand_t;

static inline int cand_before(float s, int64_t r, const cand_t *c)
{
    return (s > c->s) || (s == c->s && r < c->r);
}

static inline void cand_insert(cand_t *list, int *cnt, int k, float s, int64_t r)
{
    int n = *cnt;
    if (n == k) {
        if (!cand_before(s, r, &list[k - 1])) return;
        n = k - 1;
    }
    int j = n;
    while (j > 0 && cand_before(s, r, &list[j - 1])) { list[j] = list[j - 1]; --j; }
    list[j].s = s; list[j].r = r;
    *cnt = n + 1;
}

/*
 * Batched top-k: nq raw queries against n raw rows, k results per query.
 * ids[nq*k] (-1 padded when n < k), scores[nq*k] (0 padded).
 * Rows are normalised once (v_hat does not depend on the query), walked in
 * cache-sized blocks against every query; each thread keeps its own lists and
 * the lists are merged under the same (score desc, row asc) order, so the
 * result does not depend on the thread count.
 */
int b200o_topk(const float *queries, int nq, const float *rows, int64_t n, int d,
               int k, int64_t *ids, float *scores)
{
    if (nq < 0 || n < 0 || d <= 0 || k < 0) return -1;
    for (int64_t i = 0; i < (int64_t)nq * k; ++i) { ids[i] = -1; scores[i] = 0.0f; }
    if (nq == 0 || n == 0 || k == 0) return 0;

    float *qh = (float *)malloc(sizeof(float) * (size_t)nq * d);
    for (int a = 0; a < nq; ++a) b200o_normalize(queries + (int64_t)a * d, d, qh + (int64_t)a * d);

    int nthreads = 1;
#ifdef _OPENMP
    nthreads = omp_get_max_threads();
#endif
    cand_t *lists = (cand_t *)malloc(sizeof(cand_t) * (size_t)nthreads * nq * k);
    int *cnts = (int *)calloc((size_t)nthreads * nq, sizeof(int));
    const int64_t BLK = 128;

#pragma omp parallel
    {
        int t = 0;
#ifdef _OPENMP
        t = omp_get_thread_num();
#endif
        cand_t *my = lists + (size_t)t * nq * k;
        int *mc = cnts + (size_t)t * nq;
        float *vh = (float *)malloc(sizeof(float) * (size_t)BLK * d);
#pragma omp for schedule(dynamic, 4)
        for (int64_t b0 = 0; b0 < n; b0 += BLK) {
            int64_t bn = n - b0 < BLK ? n - b0 : BLK;
            for (int64_t r = 0; r < bn; ++r)
                b200o_normalize(rows + (b0 + r) * (int64_t)d, d, vh + r * d);
            for (int a = 0; a < nq; ++a) {
                const float *qa = qh + (int64_t)a * d;
                cand_t *la = my + (size_t)a * k;
                for (int64_t r = 0; r < bn; ++r) {
                    float s = b200o_dot(qa, vh + r * d, d);
                    if (mc[a] < k || cand_before(s, b0 + r, &la[k - 1]))
                        cand_insert(la, &mc[a], k, s, b0 + r);
                }
            }
        }
        free(vh);
    }

    for (int a = 0; a < nq; ++a) {
        cand_t *out = lists + (size_t)a * k; /* thread 0's list doubles as the result */
        int oc = cnts[a];
        for (int t = 1; t < nthreads; ++t) {
            cand_t *lt = lists + ((size_t)t * nq + a) * k;
            int ct = cnts[(size_t)t * nq + a];
            for (int j = 0; j < ct; ++j) cand_insert(out, &oc, k, lt[j].s, lt[j].r);
        }
        for (int j = 0; j < oc; ++j) { ids[(int64_t)a * k + j] = out[j].r; scores[(int64_t)a * k + j] = out[j].s; }
    }
    free(lists); free(cnts); free(qh);
    return 0;
}

/*
 * FAISS-flat mode (adapters/copilot_vectorstore/copilot_vectorstore/faiss_store.py:99-101
 * IndexFlatL2, :212-224 search + score conversion): squared Euclidean distance, ascending,
 * score = 1 / (1 + dist).  faiss itself is not installable here, so parity
 * with faiss's own summation order is UNPINNED; pinned are the reference's call site semantics
 * (exact flat search, ascending distance, 1/(1+d)) against a float64 numpy restatement
 * (tests/test_oracle_search.py).  Canonical order as above: element i -> partial (i & 31) by
 * fma(t, t, p) with t = q[i] - v[i] rounded to fp32; halving tree; ties -> lower row.
 */
float b200o_l2(const float *q, const float *v, int d)
{
    float p[B200O_LANES];
    for (int l = 0; l < B200O_LANES; ++l) p[l] = 0.0f;
    for (int i = 0; i < d; ++i) {
        const float t = q[i] - v[i];
        p[i & (B200O_LANES - 1)] = __builtin_fmaf(t, t, p[i & (B200O_LANES - 1)]);
    }
    for (int s = B200O_LANES / 2; s >= 1; s >>= 1)
        for (int l = 0; l < s; ++l) p[l] = p[l] + p[l + s];
    return p[0];
}

/* nq queries x n rows -> k nearest by squared L2; ids -1 padded, dist 0 padded */
int b200o_l2_topk(const float *queries, int nq, const float *rows, int64_t n, int d,
                  int k, int64_t *ids, float *dist)
{
    if (nq < 0 || n < 0 || d <= 0 || k < 0) return -1;
    for (int64_t i = 0; i < (int64_t)nq * k; ++i) { ids[i] = -1; dist[i] = 0.0f; }
    if (nq == 0 || n == 0 || k == 0) return 0;
#pragma omp parallel for schedule(dynamic, 1)
    for (int a = 0; a < nq; ++a) {
        cand_t *list = (cand_t *)malloc(sizeof(cand_t) * (size_t)k);
        int cnt = 0;
        const float *qa = queries + (int64_t)a * d;
        for (int64_t r = 0; r < n; ++r) {
            /* "better first" = smaller distance: order by -dist descending, row ascending */
            const float s = -b200o_l2(qa, rows + r * (int64_t)d, d);
            if (cnt < k || cand_before(s, r, &list[k - 1])) cand_insert(list, &cnt, k, s, r);
        }
        for (int j = 0; j < cnt; ++j) { ids[(int64_t)a * k + j] = list[j].r; dist[(int64_t)a * k + j] = -list[j].s; }
        free(list);
    }
    return 0;
}

/* bf16 (round-to-nearest-even) helpers so tests can build "identical inputs" */
static inline uint16_t f32_to_bf16_rne(float f)
{
    uint32_t u; memcpy(&u, &f, 4);
    if ((u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((u >> 16) | 0x40); /* NaN */
    u += 0x7fffu + ((u >> 16) & 1u);
    return (uint16_t)(u >> 16);
}

void b200o_round_bf16(const float *in, int64_t n, float *out)
{
    for (int64_t i = 0; i < n; ++i) {
        uint32_t u = (uint32_t)f32_to_bf16_rne(in[i]) << 16;
        memcpy(&out[i], &u, 4);
    }
}
