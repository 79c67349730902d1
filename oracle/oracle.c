/*
 * oracle.c -- multi-threaded C restatement of the CPU path (TEST / BASELINE INFRASTRUCTURE ONLY).
 *
 * Same Spark 3.5.5 semantics as oracle/spark_oracle.py (which is pinned on the golden vectors and is the
 * statement of record); tests/test_oracle_c.py proves the two agree.  bench.py times THIS file on the
 * GPU box's host cores as `cpu_baseline` (kind "port": the reference's own implementation is the JVM
 * inside pyspark==3.5.5, which is not available offline).  Nothing under
 * fraud-detection-spark-kafka-llm_b200/ links or calls it.
 *
 * Restated rules (reference call sites in brackets):
 *   clean      lower + regexp_replace("[^a-zA-Z ]", "")            [utils/agent_api.py:143-144]
 *   tokenise   Java String.split("\\s") after toLowerCase          [fraud_detection_spark.py:50]
 *   stop words exact match against the (case-folded) list          [fraud_detection_spark.py:51]
 *   hashingTF  murmur3_x86_32(seed 42) + nonNegativeMod, counts    [stages/2_HashingTF_*]
 *   idf        value * idf[index]                                  [fraud_detection_spark.py:53]
 *   classify   LR / DT / RF / GBT as ProbabilisticClassificationModel [agent_api.py:158-167, spark.py:59-74]
 */
#include <math.h>
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef struct {
    /* stop words: open-addressed set of (offset, len) into sw_bytes, keyed by murmur3 */
    const uint8_t* sw_bytes;
    const int32_t* sw_offs;
    int32_t n_sw;
    int32_t* sw_table; /* size sw_cap, -1 empty */
    int32_t sw_cap;
    int32_t num_features;
    int32_t binary;
    const double* idf; /* may be NULL */
    /* classifier */
    int32_t kind; /* 0 none, 1 lr, 2 dt, 3 rf, 4 gbt */
    const double* coef;
    double intercept, raw_threshold;
    int32_t n_trees;
    const int32_t* tree_offs;
    const int32_t* feature;
    const double* threshold;
    const int32_t* left;
    const int32_t* right;
    const double* leaf; /* [n_nodes,2] or [n_nodes] */
    const double* tree_w;
} oracle_model;

static inline uint32_t rotl32(uint32_t x, int r) { return (x << r) | (x >> (32 - r)); }

static uint32_t murmur3_32(const uint8_t* p, int n, uint32_t seed) {
    uint32_t h = seed;
    const int nb = n / 4;
    for (int i = 0; i < nb; ++i) {
        uint32_t k;
        memcpy(&k, p + 4 * i, 4);
        k *= 0xcc9e2d51u;
        k = rotl32(k, 15);
        k *= 0x1b873593u;
        h ^= k;
        h = rotl32(h, 13);
        h = h * 5u + 0xe6546b64u;
    }
    uint32_t k = 0;
    for (int i = 0; i < (n & 3); ++i) k |= (uint32_t)p[4 * nb + i] << (8 * i);
    if (n & 3) {
        k *= 0xcc9e2d51u;
        k = rotl32(k, 15);
        k *= 0x1b873593u;
        h ^= k;
    }
    h ^= (uint32_t)n;
    h ^= h >> 16;
    h *= 0x85ebca6bu;
    h ^= h >> 13;
    h *= 0xc2b2ae35u;
    h ^= h >> 16;
    return h;
}

static inline int32_t nonneg_mod(int32_t x, int32_t m) {
    int32_t r = x % m;
    return r < 0 ? r + m : r;
}

uint32_t oracle_murmur3(const uint8_t* p, int32_t n) { return murmur3_32(p, n, 42u); }

static int is_stop(const oracle_model* m, const uint8_t* t, int len, uint32_t h) {
    if (m->n_sw == 0) return 0;
    uint32_t s = h & (uint32_t)(m->sw_cap - 1);
    for (;;) {
        const int32_t w = m->sw_table[s];
        if (w < 0) return 0;
        const int wl = m->sw_offs[w + 1] - m->sw_offs[w];
        if (wl == len && memcmp(m->sw_bytes + m->sw_offs[w], t, (size_t)len) == 0) return 1;
        s = (s + 1) & (uint32_t)(m->sw_cap - 1);
    }
}

static int cmp_i32(const void* a, const void* b) {
    const int32_t x = *(const int32_t*)a, y = *(const int32_t*)b;
    return (x > y) - (x < y);
}

typedef struct {
    uint8_t* clean;
    size_t clean_cap;
    int32_t* idx;
    size_t idx_cap;
    int32_t* uidx;
    double* uval;
} scratch_t;

static void reserve(scratch_t* s, size_t bytes) {
    if (bytes + 1 > s->clean_cap) {
        s->clean_cap = (bytes + 1) * 2;
        s->clean = (uint8_t*)realloc(s->clean, s->clean_cap);
    }
    if (bytes + 2 > s->idx_cap) {
        s->idx_cap = (bytes + 2) * 2;
        s->idx = (int32_t*)realloc(s->idx, s->idx_cap * sizeof(int32_t));
        s->uidx = (int32_t*)realloc(s->uidx, s->idx_cap * sizeof(int32_t));
        s->uval = (double*)realloc(s->uval, s->idx_cap * sizeof(double));
    }
}

/* one dialogue -> ascending (index, tf-idf) pairs in s->uidx / s->uval; returns nnz */
static int featurize_row(const oracle_model* m, const uint8_t* raw, int64_t n, int do_clean, scratch_t* s) {
    reserve(s, (size_t)n);
    uint8_t* c = s->clean;
    int64_t cn = 0;
    if (do_clean) {
        for (int64_t p = 0; p < n; ++p) {
            const uint8_t b = raw[p];
            if (b >= 'A' && b <= 'Z') c[cn++] = b | 0x20;
            else if ((b >= 'a' && b <= 'z') || b == ' ') c[cn++] = b;
            else if (b == 0xB0 && p >= 1 && raw[p - 1] == 0xC4) c[cn++] = 'i';
            else if (b == 0xAA && p >= 2 && raw[p - 1] == 0x84 && raw[p - 2] == 0xE2) c[cn++] = 'k';
        }
    } else {
        for (int64_t p = 0; p < n; ++p) {
            const uint8_t b = raw[p];
            c[cn++] = (b >= 'A' && b <= 'Z') ? (b | 0x20) : b;
        }
    }
    /* Java split("\\s"): every ws char is a boundary; trailing empties dropped; no match -> whole string */
    int nidx = 0;
    int64_t start = 0;
    int matched = 0;
    int64_t last_nonempty_end = -1; /* number of idx entries up to and including the last non-empty token */
    int kept_upto = 0;
    for (int64_t p = 0; p <= cn; ++p) {
        const int ws = p < cn && (c[p] == ' ' || (c[p] >= 0x09 && c[p] <= 0x0D));
        if (p < cn && !ws) continue;
        if (ws) matched = 1;
        const int len = (int)(p - start);
        const uint32_t h = murmur3_32(c + start, len, 42u);
        /* record token; stop-word filtering happens here, trailing-empty trimming below */
        if (!is_stop(m, c + start, len, h)) s->idx[nidx++] = nonneg_mod((int32_t)h, m->num_features);
        else s->idx[nidx++] = -1;
        if (len > 0) {
            last_nonempty_end = nidx;
        }
        start = p + 1;
    }
    if (matched) kept_upto = last_nonempty_end < 0 ? 0 : (int)last_nonempty_end;
    else kept_upto = nidx; /* exactly one token */
    int k = 0;
    for (int i = 0; i < kept_upto; ++i)
        if (s->idx[i] >= 0) s->idx[k++] = s->idx[i];
    qsort(s->idx, (size_t)k, sizeof(int32_t), cmp_i32);
    int nnz = 0;
    for (int i = 0; i < k;) {
        int j = i;
        while (j < k && s->idx[j] == s->idx[i]) ++j;
        const double tf = m->binary ? 1.0 : (double)(j - i);
        s->uidx[nnz] = s->idx[i];
        s->uval[nnz] = m->idf ? tf * m->idf[s->idx[i]] : tf;
        ++nnz;
        i = j;
    }
    return nnz;
}

static double sparse_get(const int32_t* idx, const double* val, int n, int f) {
    int lo = 0, hi = n;
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (idx[mid] < f) lo = mid + 1;
        else hi = mid;
    }
    return (lo < n && idx[lo] == f) ? val[lo] : 0.0;
}

static int walk(const oracle_model* m, int t, const int32_t* idx, const double* val, int n) {
    const int base = m->tree_offs[t];
    int nd = 0;
    while (m->feature[base + nd] >= 0) {
        const double x = sparse_get(idx, val, n, m->feature[base + nd]);
        nd = x <= m->threshold[base + nd] ? m->left[base + nd] : m->right[base + nd];
    }
    return base + nd;
}

static void score_row(const oracle_model* m, const int32_t* idx, const double* val, int n, double* pred, double* prob,
                      double* raw) {
    if (m->kind == 1) {
        double s = 0.0;
        for (int i = 0; i < n; ++i) {
            const double c = m->coef[idx[i]];
            if (c != 0.0) s = s + val[i] * c;
        }
        const double mg = s + m->intercept;
        raw[0] = -mg;
        raw[1] = mg;
        prob[0] = 1.0 / (1.0 + exp(mg));
        prob[1] = 1.0 - prob[0];
        *pred = mg > m->raw_threshold ? 1.0 : 0.0;
    } else if (m->kind == 2) {
        const int lf = walk(m, 0, idx, val, n);
        raw[0] = m->leaf[2 * lf];
        raw[1] = m->leaf[2 * lf + 1];
        const double tot = raw[0] + raw[1];
        prob[0] = tot != 0.0 ? raw[0] / tot : raw[0];
        prob[1] = tot != 0.0 ? raw[1] / tot : raw[1];
        *pred = raw[1] > raw[0] ? 1.0 : 0.0;
    } else if (m->kind == 3) {
        double v0 = 0.0, v1 = 0.0;
        for (int t = 0; t < m->n_trees; ++t) {
            const int lf = walk(m, t, idx, val, n);
            const double c0 = m->leaf[2 * lf], c1 = m->leaf[2 * lf + 1], tot = c0 + c1;
            if (tot != 0.0) {
                v0 += c0 / tot;
                v1 += c1 / tot;
            }
        }
        raw[0] = v0;
        raw[1] = v1;
        const double tot = v0 + v1;
        prob[0] = tot != 0.0 ? v0 / tot : v0;
        prob[1] = tot != 0.0 ? v1 / tot : v1;
        *pred = v1 > v0 ? 1.0 : 0.0;
    } else {
        double mg = 0.0;
        for (int t = 0; t < m->n_trees; ++t) mg = mg + m->leaf[walk(m, t, idx, val, n)] * m->tree_w[t];
        raw[0] = -mg;
        raw[1] = mg;
        prob[0] = 1.0 / (1.0 + exp(-2.0 * raw[0]));
        prob[1] = 1.0 - prob[0];
        *pred = raw[1] > raw[0] ? 1.0 : 0.0;
    }
}

typedef struct {
    const oracle_model* m;
    const uint8_t* text;
    const int64_t* offs;
    int64_t r0, r1;
    int do_clean;
    double *pred, *prob, *raw;
    int32_t* nnz_out;        /* featurize: per-row nnz (pass 1) */
    const int64_t* rowptr;   /* featurize pass 2 */
    int32_t* out_idx;
    double* out_val;
} job_t;

static void* worker(void* arg) {
    job_t* j = (job_t*)arg;
    scratch_t s;
    memset(&s, 0, sizeof(s));
    for (int64_t r = j->r0; r < j->r1; ++r) {
        const int nnz = featurize_row(j->m, j->text + j->offs[r], j->offs[r + 1] - j->offs[r], j->do_clean, &s);
        if (j->pred) score_row(j->m, s.uidx, s.uval, nnz, j->pred + r, j->prob + 2 * r, j->raw + 2 * r);
        if (j->nnz_out) j->nnz_out[r] = nnz;
        if (j->rowptr) {
            memcpy(j->out_idx + j->rowptr[r], s.uidx, (size_t)nnz * sizeof(int32_t));
            memcpy(j->out_val + j->rowptr[r], s.uval, (size_t)nnz * sizeof(double));
        }
    }
    free(s.clean);
    free(s.idx);
    free(s.uidx);
    free(s.uval);
    return NULL;
}

static void run_jobs(job_t proto, int64_t n_rows, int n_threads) {
    if (n_threads < 1) n_threads = 1;
    if (n_threads > 256) n_threads = 256;
    pthread_t th[256];
    job_t jobs[256];
    for (int t = 0; t < n_threads; ++t) {
        jobs[t] = proto;
        jobs[t].r0 = n_rows * t / n_threads;
        jobs[t].r1 = n_rows * (t + 1) / n_threads;
        pthread_create(&th[t], NULL, worker, &jobs[t]);
    }
    for (int t = 0; t < n_threads; ++t) pthread_join(th[t], NULL);
}

/* build the stop-word table inside the model struct (caller passes a zeroed table of size cap = pow2 >= 2n) */
void oracle_build_stopwords(oracle_model* m) {
    for (int i = 0; i < m->sw_cap; ++i) m->sw_table[i] = -1;
    for (int w = 0; w < m->n_sw; ++w) {
        const int len = m->sw_offs[w + 1] - m->sw_offs[w];
        uint32_t s = murmur3_32(m->sw_bytes + m->sw_offs[w], len, 42u) & (uint32_t)(m->sw_cap - 1);
        while (m->sw_table[s] >= 0) s = (s + 1) & (uint32_t)(m->sw_cap - 1);
        m->sw_table[s] = w;
    }
}

void oracle_transform(const oracle_model* m, const uint8_t* text, const int64_t* offs, int64_t n_rows, int do_clean,
                      double* pred, double* prob, double* raw, int n_threads) {
    job_t j;
    memset(&j, 0, sizeof(j));
    j.m = m;
    j.text = text;
    j.offs = offs;
    j.do_clean = do_clean;
    j.pred = pred;
    j.prob = prob;
    j.raw = raw;
    run_jobs(j, n_rows, n_threads);
}

/* pass 1: per-row nnz; pass 2 (rowptr given): fill idx/val */
void oracle_featurize(const oracle_model* m, const uint8_t* text, const int64_t* offs, int64_t n_rows, int do_clean,
                      int32_t* nnz_out, const int64_t* rowptr, int32_t* out_idx, double* out_val, int n_threads) {
    job_t j;
    memset(&j, 0, sizeof(j));
    j.m = m;
    j.text = text;
    j.offs = offs;
    j.do_clean = do_clean;
    j.nnz_out = nnz_out;
    j.rowptr = rowptr;
    j.out_idx = out_idx;
    j.out_val = out_val;
    run_jobs(j, n_rows, n_threads);
}
