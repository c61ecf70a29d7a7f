/*
 * oracle.c — CPU restatement of the reference's RBAC-filtered kNN hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under vectorsearch-rbac_b200/ may import, link or call this
 * file; it is the checker for tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs.  The product path is CUDA-only and fails loudly without its extension.
 *
 * Parity status: PINNED.  tests/test_oracle_golden.py checks these functions against every
 * known-answer vector the reference holds for this path (pgvector/test/expected/vector_type.out:
 * 373-429, ivfflat_vector.out:3-20,39-50,86-111, t/003_ivfflat_vector_build_recall.pl:104-116)
 * and against oracle/_ref/libpgvref.so, which oracle/Makefile compiles from the reference's own
 * pgvector/src/vector.c:549-563,596-606 where the file lies under /root/reference.
 *
 * What is restated, and from where (paths relative to the reference checkout):
 *   distance loops .......... pgvector/src/vector.c:549-563 (L2 squared), :596-606 (inner product)
 *   SQL operators ........... vector.c:568-578 `<->` = sqrt((double) sum);  :626-636 `<#>` = -(double) sum
 *   index ordering proc ..... vector.c:584-594 (double) squared sum (ivfflat FUNCTION 1, sql/vector.sql:292-303)
 *   ORDER BY .. LIMIT k ..... PostgreSQL SeqScan + top-N sort (PG core, not in the checkout) driven by
 *                             controller/baseline/prefilter/prefilter_role.py:162-172  -> bounded heap
 *   merge / dedup ........... prefilter_role.py:174-189; controller/dynamic_partition/search.py:347-364
 *   ground truth ............ basic_benchmark/common_function.py:699-714 (no dedup), :1114-1124 (dedup)
 *   IVFFlat scan ............ pgvector/src/ivfscan.c:36-107 (GetScanLists), :112-176 (GetScanItems),
 *                             :257-266 (probe clamps), :375-381 (iterative refill)
 *   IVFFlat assignment ...... pgvector/src/ivfbuild.c:141-198 (nearest centre, first minimum wins)
 *   recall .................. common_function.py:1154-1160
 *
 * Summation order.  "strict" = one fp32 accumulator, left to right, multiply and add rounded
 * separately: what the reference's own build script produces (compile_pgvector.sh:7, -g -O0).
 * This file is compiled with -ffp-contract=off and without -ffast-math so -O2 keeps that order.
 * Ties: PostgreSQL leaves the order of equal distances unspecified; the oracle (and the engine)
 * break them by ascending tuple id so results are reproducible.
 */
#include <math.h>
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define ORC_METRIC_L2 0 /* `<->`  */
#define ORC_METRIC_IP 1 /* `<#>`  */

typedef float (*orc_distfn)(int dim, float *a, float *b);

/* vector.c:549-563 */
float orc_l2sq(int dim, float *a, float *b)
{
    float acc = 0.0f;
    for (int i = 0; i < dim; i++) {
        float d = a[i] - b[i];
        acc += d * d;
    }
    return acc;
}

/* vector.c:596-606 */
float orc_ip(int dim, float *a, float *b)
{
    float acc = 0.0f;
    for (int i = 0; i < dim; i++)
        acc += a[i] * b[i];
    return acc;
}

/* vector.c:568-578 / :626-636: what the SQL operator returns (float8). */
double orc_operator_value(int metric, float sum)
{
    if (metric == ORC_METRIC_L2)
        return sqrt((double) sum);
    return (double) -sum;
}

/* vector.c:584-594 / :626-636: what an index orders by (opclass FUNCTION 1). */
double orc_order_value(int metric, float sum)
{
    if (metric == ORC_METRIC_L2)
        return (double) sum;
    return (double) -sum;
}

/* vector.c:61-67 CheckDims: 0 ok, -2 = "different vector dimensions". */
int orc_check_dims(int dim_a, int dim_b) { return dim_a == dim_b ? 0 : -2; }

/* vector.c:102-113 CheckElement: NaN / Inf are rejected on input. */
int orc_check_finite(const float *x, size_t n)
{
    for (size_t i = 0; i < n; i++)
        if (isnan(x[i]) || isinf(x[i]))
            return -3;
    return 0;
}

/* ------------------------------------------------------------------------------------------- */
/* Bounded top-k: a max-heap on (order value, tid).                                             */

typedef struct {
    double key;   /* order value */
    uint64_t tid;
    uint32_t row;
    float sum;    /* raw fp32 sum */
} orc_item;

static inline int item_less(const orc_item *a, const orc_item *b)
{
    if (a->key != b->key)
        return a->key < b->key;
    return a->tid < b->tid;
}

static void heap_sift_down(orc_item *h, int n, int i)
{
    for (;;) {
        int l = 2 * i + 1, r = l + 1, m = i;
        if (l < n && item_less(&h[m], &h[l])) m = l;
        if (r < n && item_less(&h[m], &h[r])) m = r;
        if (m == i) return;
        orc_item t = h[i]; h[i] = h[m]; h[m] = t;
        i = m;
    }
}

static void heap_sift_up(orc_item *h, int i)
{
    while (i > 0) {
        int p = (i - 1) / 2;
        if (!item_less(&h[p], &h[i])) return;
        orc_item t = h[i]; h[i] = h[p]; h[p] = t;
        i = p;
    }
}

static int item_cmp(const void *pa, const void *pb)
{
    const orc_item *a = (const orc_item *) pa, *b = (const orc_item *) pb;
    if (item_less(a, b)) return -1;
    if (item_less(b, a)) return 1;
    return 0;
}

/*
 * SELECT .. ORDER BY distance LIMIT k over one table (prefilter_role.py:162-172): every row of
 * `rows` (NULL = rows 0..nrows-1) is scored with `fn`, the k smallest are kept.  Returns count.
 */
static int scan_topk(const float *X, int dim, const uint64_t *tid, const uint32_t *rows, size_t nrows,
                     const float *q, int metric, int k, orc_distfn fn, orc_item *out /* k */)
{
    int n = 0;
    for (size_t i = 0; i < nrows; i++) {
        uint32_t r = rows ? rows[i] : (uint32_t) i;
        float s = fn(dim, (float *) (X + (size_t) r * dim), (float *) q);
        orc_item it = { orc_order_value(metric, s), tid ? tid[r] : r, r, s };
        if (n < k) {
            out[n] = it;
            heap_sift_up(out, n);
            n++;
        } else if (item_less(&it, &out[0])) {
            out[0] = it;
            heap_sift_down(out, n, 0);
        }
    }
    qsort(out, (size_t) n, sizeof(orc_item), item_cmp);
    return n;
}

static orc_distfn pick_fn(int metric, orc_distfn l2, orc_distfn ip)
{
    if (metric == ORC_METRIC_L2) return l2 ? l2 : orc_l2sq;
    return ip ? ip : orc_ip;
}

/* ------------------------------------------------------------------------------------------- */
/* Batched filtered search over partitions.                                                     */

typedef struct {
    const float *X; int dim; const uint64_t *tid; size_t nrows;
    /* partitions: CSR over row indices; partition p = part_rows[part_off[p] .. part_off[p+1]) */
    const uint64_t *part_off; const uint32_t *part_rows;
    /* queries: CSR over partition ids */
    size_t nq; const float *Q; const uint64_t *qpart_off; const uint32_t *qpart;
    int metric; int k; int dedup; /* dedup: prefilter_role.py:183-189 / common_function.py:1114-1121 */
    int limit_per_partition;      /* 1 = LIMIT k per partition (search); 0 = no LIMIT (ground truth, common_function.py:699-705) */
    orc_distfn fn;
    uint64_t *out_tid; uint32_t *out_row; double *out_dist; float *out_sum; int32_t *out_count;
    size_t q_begin, q_end;
} orc_job;

static void run_queries(orc_job *j)
{
    int k = j->k;
    orc_item *tmp = NULL;
    size_t tmp_cap = 0;
    for (size_t qi = j->q_begin; qi < j->q_end; qi++) {
        const float *q = j->Q + qi * (size_t) j->dim;
        size_t np = (size_t) (j->qpart_off[qi + 1] - j->qpart_off[qi]);
        size_t need = np * (size_t) k + 1;
        if (need > tmp_cap) {
            free(tmp);
            tmp = (orc_item *) malloc(need * sizeof(orc_item));
            tmp_cap = need;
        }
        size_t n = 0;
        for (size_t pi = 0; pi < np; pi++) {
            uint32_t p = j->qpart[j->qpart_off[qi] + pi];
            const uint32_t *rows = j->part_rows ? j->part_rows + j->part_off[p] : NULL;
            size_t nr = j->part_rows ? (size_t) (j->part_off[p + 1] - j->part_off[p]) : j->nrows;
            /* the ground-truth variant has no LIMIT per partition; keeping k per partition is
             * equivalent for the final [:k] unless duplicates across partitions are kept, in which
             * case k per partition still suffices (each copy is within its partition's top-k). */
            n += (size_t) scan_topk(j->X, j->dim, j->tid, rows, nr, q, j->metric, k, j->fn, tmp + n);
        }
        size_t m = 0;
        if (np > 1) {
            /* all_results.sort(key=distance) then first-seen dedup on (document_id, block_id) */
            qsort(tmp, n, sizeof(orc_item), item_cmp);
            for (size_t i = 0; i < n && m < (size_t) k; i++) {
                if (j->dedup && m > 0 && tmp[m - 1].tid == tmp[i].tid) continue;
                tmp[m++] = tmp[i];
            }
        } else {
            m = n;
        }
        for (size_t i = 0; i < (size_t) k; i++) {
            size_t o = qi * (size_t) k + i;
            if (i < m) {
                j->out_tid[o] = tmp[i].tid;
                if (j->out_row) j->out_row[o] = tmp[i].row;
                j->out_dist[o] = orc_operator_value(j->metric, tmp[i].sum);
                if (j->out_sum) j->out_sum[o] = tmp[i].sum;
            } else {
                j->out_tid[o] = UINT64_MAX;
                if (j->out_row) j->out_row[o] = UINT32_MAX;
                j->out_dist[o] = INFINITY;
                if (j->out_sum) j->out_sum[o] = INFINITY;
            }
        }
        j->out_count[qi] = (int32_t) m;
    }
    free(tmp);
}

static void *job_thread(void *arg) { run_queries((orc_job *) arg); return NULL; }

/*
 * nq independent filtered searches (prefilter_role.py:81-195 per query).  part_rows == NULL means
 * "every partition is the whole table" (unfiltered scan).  fn_l2 / fn_ip may point into
 * oracle/_ref/libpgvref.so (the reference's own loops) or be NULL for the port above.
 */
int orc_search_batch(const float *X, int dim, const uint64_t *tid, size_t nrows,
                     const uint64_t *part_off, const uint32_t *part_rows,
                     size_t nq, const float *Q, const uint64_t *qpart_off, const uint32_t *qpart,
                     int metric, int k, int dedup, orc_distfn fn_l2, orc_distfn fn_ip, int nthreads,
                     uint64_t *out_tid, uint32_t *out_row, double *out_dist, float *out_sum,
                     int32_t *out_count)
{
    if (!X || !Q || dim <= 0 || k <= 0 || !out_tid || !out_dist || !out_count) return -1;
    orc_job base = { X, dim, tid, nrows, part_off, part_rows, nq, Q, qpart_off, qpart, metric, k,
                     dedup, 1, pick_fn(metric, fn_l2, fn_ip), out_tid, out_row, out_dist, out_sum,
                     out_count, 0, nq };
    if (nthreads <= 1 || nq < 2) {
        run_queries(&base);
        return 0;
    }
    if ((size_t) nthreads > nq) nthreads = (int) nq;
    pthread_t *th = (pthread_t *) malloc(sizeof(pthread_t) * (size_t) nthreads);
    orc_job *jobs = (orc_job *) malloc(sizeof(orc_job) * (size_t) nthreads);
    for (int t = 0; t < nthreads; t++) {
        jobs[t] = base;
        jobs[t].q_begin = nq * (size_t) t / (size_t) nthreads;
        jobs[t].q_end = nq * (size_t) (t + 1) / (size_t) nthreads;
        pthread_create(&th[t], NULL, job_thread, &jobs[t]);
    }
    for (int t = 0; t < nthreads; t++) pthread_join(th[t], NULL);
    free(th);
    free(jobs);
    return 0;
}

/* ------------------------------------------------------------------------------------------- */
/* IVFFlat                                                                                       */

/* ivfbuild.c:141-198 AddTupleToSort: nearest centre by the opclass distance, first minimum wins
 * (`if (distance < minDistance)` with centres visited in order). */
int orc_ivf_assign(const float *X, int dim, size_t nrows, const float *centers, int lists, int metric,
                   uint32_t *out_list)
{
    orc_distfn fn = pick_fn(metric, NULL, NULL);
    for (size_t r = 0; r < nrows; r++) {
        double best = 1.7976931348623157e308; /* DBL_MAX, ivfbuild.c:150 */
        uint32_t bi = 0;
        for (int c = 0; c < lists; c++) {
            double d = orc_order_value(metric, fn(dim, (float *) (X + r * (size_t) dim),
                                                  (float *) (centers + (size_t) c * dim)));
            if (d < best) { best = d; bi = (uint32_t) c; }
        }
        out_list[r] = bi;
    }
    return 0;
}

typedef struct { double d; int list; } orc_listd;

static int listd_cmp(const void *a, const void *b)
{
    const orc_listd *x = (const orc_listd *) a, *y = (const orc_listd *) b;
    if (x->d < y->d) return -1;
    if (x->d > y->d) return 1;
    return (x->list > y->list) - (x->list < y->list);
}

/*
 * ivfscan.c:36-107 GetScanLists: distance from q to every centre, keep the max_probes nearest,
 * nearest first.  (pgvector's pairing heap leaves ties between equal centre distances in an
 * unspecified order; ascending list number here.)  null_query: all distances 0 (ivfscan.c:181-185).
 */
int orc_ivf_order_lists(const float *centers, int lists, int dim, const float *q, int metric,
                        int max_probes, int null_query, int32_t *out_lists, double *out_d)
{
    orc_distfn fn = pick_fn(metric, NULL, NULL);
    orc_listd *ld = (orc_listd *) malloc(sizeof(orc_listd) * (size_t) lists);
    for (int c = 0; c < lists; c++) {
        ld[c].list = c;
        ld[c].d = null_query ? 0.0
                             : orc_order_value(metric, fn(dim, (float *) (centers + (size_t) c * dim), (float *) q));
    }
    qsort(ld, (size_t) lists, sizeof(orc_listd), listd_cmp);
    if (max_probes > lists) max_probes = lists; /* ivfscan.c:265-266 */
    for (int i = 0; i < max_probes; i++) {
        out_lists[i] = ld[i].list;
        if (out_d) out_d[i] = ld[i].d;
    }
    free(ld);
    return max_probes;
}

/*
 * ivfscan.c:112-176 GetScanItems for one probe batch: every tuple of lists
 * probe_lists[0..nprobe) is scored and the whole batch is sorted (tuplesort, not top-k).  Rows
 * failing `allowed` (NULL = all pass) are dropped — in PostgreSQL the executor applies the WHERE /
 * RLS qual to what the index returns, which yields the same rows in the same
 * order.  Writes at most `cap` entries; returns the number of qualifying tuples in the batch.
 */
long orc_ivf_scan_batch(const float *X, int dim, const uint64_t *tid, const uint64_t *list_off,
                        const uint32_t *list_rows, const int32_t *probe_lists, int nprobe,
                        const uint8_t *allowed, const float *q, int metric, int null_query, long cap,
                        uint64_t *out_tid, uint32_t *out_row, double *out_dist)
{
    orc_distfn fn = pick_fn(metric, NULL, NULL);
    size_t total = 0;
    for (int p = 0; p < nprobe; p++)
        total += (size_t) (list_off[probe_lists[p] + 1] - list_off[probe_lists[p]]);
    orc_item *it = (orc_item *) malloc(sizeof(orc_item) * (total + 1));
    size_t n = 0;
    for (int p = 0; p < nprobe; p++) {
        int l = probe_lists[p];
        for (uint64_t i = list_off[l]; i < list_off[l + 1]; i++) {
            uint32_t r = list_rows[i];
            if (allowed && !allowed[r]) continue;
            float s = null_query ? 0.0f : fn(dim, (float *) (X + (size_t) r * dim), (float *) q);
            it[n].key = null_query ? 0.0 : orc_order_value(metric, s);
            it[n].tid = tid ? tid[r] : r;
            it[n].row = r;
            it[n].sum = s;
            n++;
        }
    }
    qsort(it, n, sizeof(orc_item), item_cmp);
    for (size_t i = 0; i < n && (long) i < cap; i++) {
        out_tid[i] = it[i].tid;
        if (out_row) out_row[i] = it[i].row;
        out_dist[i] = null_query ? 0.0 : orc_operator_value(metric, it[i].sum);
    }
    free(it);
    return (long) n;
}

/* common_function.py:1154-1160: |true ∩ predicted| / |true| over tuple ids. */
double orc_recall(const uint64_t *truth, int nt, const uint64_t *pred, int np)
{
    if (nt <= 0) return 0.0;
    int hit = 0;
    for (int i = 0; i < nt; i++) {
        int dup = 0;
        for (int j = 0; j < i; j++) if (truth[j] == truth[i]) { dup = 1; break; }
        if (dup) { continue; }
        for (int j = 0; j < np; j++) if (pred[j] == truth[i]) { hit++; break; }
    }
    int uniq = 0;
    for (int i = 0; i < nt; i++) {
        int dup = 0;
        for (int j = 0; j < i; j++) if (truth[j] == truth[i]) { dup = 1; break; }
        uniq += !dup;
    }
    return (double) hit / (double) uniq;
}
