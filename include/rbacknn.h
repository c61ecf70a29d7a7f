/*
 * rbacknn.h — C-ABI of the B200-native RBAC-filtered kNN engine (librbacknn.so).
 *
 * Drop-in boundary for ONE hot path of rjzhb/VectorSearch-RBAC: the role / user permission
 * prefilter fused with L2 / inner-product scoring and top-k selection that the reference runs
 * as pgvector exact scans and IVFFlat index scans over role-partitioned tables and in its
 * ground-truth computation.  Every entry point below names the reference interface it stands in
 * for (paths relative to the reference checkout).
 *
 * Conventions
 *   - plain C, no torch / C++ types in any signature; all sizes are explicit.
 *   - the caller owns every host buffer it passes; the library copies what it keeps.
 *   - no exceptions / longjmp cross this boundary: functions return 0 (RK_OK) or a negative
 *     rk_status and leave a message for rk_last_error() (thread-local).  The reference reports the
 *     same conditions through ereport(ERROR) (pgvector/src/vector.c:61-67, 102-113).
 *   - there is NO CPU fallback: without a CUDA device rk_open() fails with RK_ERR_NO_DEVICE.
 *   - an rk_index is immutable after rk_finalize(); any number of rk_scan handles (one per host
 *     thread) may search it concurrently, like PostgreSQL backends each owning one IndexScanDesc
 *     (pgvector/src/ivfflat.c:203 amcanparallel=false, one scan == one thread).
 */
#ifndef RBACKNN_H
#define RBACKNN_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif
#if defined(__GNUC__)
#pragma GCC visibility push(default)
#endif

#define RK_ABI_VERSION 3

typedef struct rk_index rk_index;
typedef struct rk_scan rk_scan;

typedef enum rk_status {
    RK_OK = 0,
    RK_ERR_INVALID = -1,      /* bad argument (NULL, k out of range, ...)                         */
    RK_ERR_DIM = -2,          /* "different vector dimensions %d and %d"  (vector.c:61-67)        */
    RK_ERR_NONFINITE = -3,    /* NaN / Inf element in a vector            (vector.c:102-113)      */
    RK_ERR_STATE = -4,        /* call order violated (e.g. search before rk_finalize)             */
    RK_ERR_NO_DEVICE = -5,    /* no CUDA device / wrong architecture: the engine has no CPU path  */
    RK_ERR_CUDA = -6,         /* a CUDA runtime call failed; message carries cudaGetErrorString   */
    RK_ERR_NOMEM = -7,
    RK_ERR_UNSUPPORTED = -8,  /* e.g. dim > RK_MAX_DIM                                            */
    RK_ERR_NO_ORDER = -9      /* "cannot scan ivfflat index without order" (ivfscan.c:360-361)    */
} rk_status;

/* Distance operators.  RK_METRIC_L2 is SQL `<->` (l2_distance, vector.c:568-578: sqrt((double)
 * fp32 sum)); RK_METRIC_IP is SQL `<#>` (vector_negative_inner_product, vector.c:626-636:
 * -(double) fp32 sum).  Index scans order by opclass FUNCTION 1, i.e. the squared L2 sum widened to
 * double (vector.c:584-594) — the same order as `<->`. */
typedef enum rk_metric { RK_METRIC_L2 = 0, RK_METRIC_IP = 1 } rk_metric;

/* How the permission prefilter is applied.
 *   RK_FILTER_PARTITION  layout-faithful: user -> role set -> partition ids -> scan each partition,
 *                        merge + dedup on tuple id (controller/baseline/prefilter/prefilter_role.py:
 *                        120-189; prefilter_combination_role.py:122-161; controller/dynamic_partition/
 *                        search.py:44-111,347-364).
 *   RK_FILTER_BITMASK    one pass over the row store with the per-document role mask tested against
 *                        the user's role mask inside the scan kernel — the RLS predicate of
 *                        controller/baseline/pg_row_security/row_level_security.py:55-65.
 *   RK_FILTER_NONE       no permission filter (plain pgvector scan). */
typedef enum rk_filter { RK_FILTER_PARTITION = 0, RK_FILTER_BITMASK = 1, RK_FILTER_NONE = 2 } rk_filter;

/* Which materialised layout a partition-filtered scan uses.
 *   RK_LAYOUT_ROLE         one scan per role of the user, merged (prefilter_role.py:124-189)
 *   RK_LAYOUT_COMBINATION  the single table of the user's role set (prefilter_combination_role.py:122-161)
 *   RK_LAYOUT_DYNAMIC      CombRolePartitions lookup, RLS predicate inside impure partitions
 *                          (dynamic_partition/search.py:44-111; load_result_to_database.py:591-622)
 *   RK_LAYOUT_AUTO         dynamic if a role-set map was given, else combination, else role */
typedef enum rk_layout { RK_LAYOUT_AUTO = 0, RK_LAYOUT_ROLE = 1, RK_LAYOUT_COMBINATION = 2, RK_LAYOUT_DYNAMIC = 3 } rk_layout;

/* Scoring engine of a batched search.
 *   RK_ENGINE_SCAN    bandwidth-bound CUDA-core scan, one pass over the visible rows per query
 *                     (what a single pgvector scan does; exact fp32).
 *   RK_ENGINE_TENSOR  tcgen05 tensor cores: the whole query batch against the row store as a dense
 *                     contraction (|q|^2 - 2 q.x + |x|^2 or -q.x) with the permission mask in the
 *                     epilogue; nominates candidates in TF32, re-scores them in exact fp32 and falls
 *                     back to the scan for any query whose answer it cannot prove.  Results are
 *                     identical to RK_ENGINE_SCAN.  Exact scans only (probes = 0), k <= 384.
 *   RK_ENGINE_AUTO    RK_ENGINE_TENSOR for batches of 32 or more exact scans it supports (role,
 *                     combination, bitmask or no filter), RK_ENGINE_SCAN otherwise. */
typedef enum rk_engine { RK_ENGINE_AUTO = 0, RK_ENGINE_SCAN = 1, RK_ENGINE_TENSOR = 2 } rk_engine;

#define RK_MAX_DIM 2048   /* pgvector IVFFLAT_MAX_DIM is 2000 (pgvector/src/ivfflat.h:24)          */
#define RK_MAX_K 1024     /* top-k limit of the GPU selection network                               */
#define RK_MAX_LISTS 32768 /* IVFFLAT_MAX_LISTS (pgvector/src/ivfflat.h:44)                         */

typedef struct rk_config {
    int32_t abi_version;  /* RK_ABI_VERSION */
    int32_t dim;          /* vector dimension, 1..RK_MAX_DIM (vector.h:11-17 `int16 dim`)           */
    int32_t metric;       /* rk_metric */
    int32_t device;       /* CUDA device ordinal */
    int32_t k_max;        /* largest k any scan will ask for, 1..RK_MAX_K */
    int32_t reserved[11];
} rk_config;

/* ---- lifecycle ------------------------------------------------------------------------------ */

/* Stands in for opening the relation + index (RelationGetIndexScan / IvfflatGetMetaPageInfo,
 * pgvector/src/ivfscan.c:252-255).  Returns NULL on failure (see rk_last_error()). */
rk_index *rk_open(const rk_config *cfg);
void rk_close(rk_index *ix);
const char *rk_last_error(void);
int rk_abi_version(void);
/* Tuning / test switches of one index (stand-ins for pgvector's GUCs, pgvector/src/ivfflat.c:33-54, for engine
 * internals): "tensor_plan" (0 cheapest | 1 dense | 2 per user mask | 3 per document class), "tensor_pair" (-1 by row
 * length | 0 one-CTA kernel | 1 CTA-pair kernel), "tensor_bf16" (-1 by row length | 0 TF32 | 1 bf16 nomination),
 * "tensor_sample", "tensor_sdiv", "tensor_smin", "tensor_tpu", "tensor_host_timing", "ivf_seg_factor",
 * "ivf_probe_single", "ivf_assign_tc" (1 = k-means / build assignment on the tensor cores, 0 = fp32 kernel), "ivf_kmeanspp" (1 = k-means++ seeding,
 * 0 = first sampled rows), "ivf_kmeans_iters", "ivf_sample_identity", "single_fused",
 * "single_zero_copy", "class_order" (before rk_finalize: 1 = order the documents, and with them the rows, by (role mask, id)
 * when role masks repeat, 2 = always, 0 = by id), "tensor_kp_exact" (nominations beyond k while nomination is exact, < 0 = off),
 * "tensor_round_mark" (buffers above this many entries are compacted in every compaction round), "tensor_loose",
 * "tensor_kp_extra", "tensor_kp_bf16", "tensor_qtau", "seg_cost_x10".  Defaults come from the RK_* environment variables of the same names
 * (upper case), which are read once, in rk_open — never on the per-call path. */
int rk_set_option(rk_index *ix, const char *name, int64_t value);

/* ---- loading: what PostgreSQL's heap + catalog hold today ------------------------------------ */

/* Append n rows of table `documentblocks` (controller/initialize_main_tables.py:52-63):
 * x[n*dim] row-major fp32 (struct Vector payload, pgvector/src/vector.h:11-17), tid[n] the tuple
 * id handed back by scans (block_id, or a packed ItemPointer), doc[n] the row's document_id.
 * Rejects NaN/Inf like vector_in (vector.c:102-113). */
int rk_load_rows(rk_index *ix, const float *x, const uint64_t *tid, const uint32_t *doc, size_t n);

/* Same, with x[n*dim] already resident in HBM on the index's device (tid / doc stay host arrays).
 * The NaN/Inf check runs as a kernel. */
int rk_load_rows_device(rk_index *ix, const float *d_x, const uint64_t *tid, const uint32_t *doc, size_t n);

/* Optional: size the HBM row store once for n rows before a chunked rk_load_rows sequence. */
int rk_reserve_rows(rk_index *ix, size_t n);

/* PermissionAssignment(role_id, document_id) pairs (initialize_main_tables.py:66-72). */
int rk_set_doc_roles(rk_index *ix, const uint32_t *doc, const uint32_t *role, size_t n);
/* UserRoles(user_id, role_id) pairs (initialize_main_tables.py:33-39). */
int rk_set_user_roles(rk_index *ix, const uint32_t *user, const uint32_t *role, size_t n);

/* One table documentblocks_role_<r> per role: rows whose document is granted to r
 * (controller/baseline/prefilter/initialize_partitions.py:281-311). Partition id == role id. */
int rk_build_role_partitions(rk_index *ix);
/* One table per distinct sorted role set of a user, holding the de-duplicated union
 * (initialize_partitions.py:549-555, 610-700). */
int rk_build_combination_partitions(rk_index *ix);
/* Dynamic-partition layout: documentblocks_partition_<p> (controller/dynamic_partition/
 * load_result_to_database.py:207-219) and CombRolePartitions(comb_role INT[], partition_id)
 * (:293-299).  `roles` must be sorted ascending like search.py:44-52 does. */
int rk_define_partition(rk_index *ix, uint32_t part_id, const uint32_t *docs, size_t n);
int rk_map_roleset_to_partitions(rk_index *ix, const uint32_t *roles, size_t nroles,
                                 const uint32_t *parts, size_t nparts);

/* Freeze the index: rows are ordered by document, uploaded to HBM, partitions become extent
 * tables over the single row store, per-document and per-user role masks are built. */
int rk_finalize(rk_index *ix);

/* ---- IVFFlat (pgvector/src/ivfbuild.c, ivfkmeans.c, ivfscan.c) ------------------------------- */

/* CREATE INDEX ... USING ivfflat (vector vector_l2_ops|vector_ip_ops) WITH (lists = L):
 * sample max(10000, 50*L) rows (ivfbuild.c:404-441), k-means (ivfkmeans.c), assign every row to
 * its nearest centre (ivfbuild.c:141-198), store rows list-contiguous (ivfbuild.c:251-310). */
int rk_build_ivf(rk_index *ix, int lists, uint64_t seed);
/* Same index from caller-supplied centres[lists*dim] (assignment + packing only). */
int rk_build_ivf_from_centers(rk_index *ix, int lists, const float *centers);
/* Same index from centres AND the row -> list assignment of an index that already exists (row_list[n] by load order, values
 * < lists), e.g. what rk_ivfpages_read() found on disk: the lists are scanned as stored (ivfscan.c:128-168), nothing is
 * re-assigned. */
int rk_build_ivf_from_lists(rk_index *ix, int lists, const float *centers, const uint32_t *row_list);
/* Introspection for tests: copies out centres and the row->list assignment. */
int rk_ivf_info(const rk_index *ix, int *lists, float *centers, uint32_t *row_list /* by load order */);

/* ---- the on-disk form of a pgvector ivfflat index --------------------------
 * Reader for the main fork of an ivfflat index relation (base/<db>/<relfilenode>, segments concatenated): BLCKSZ 8192,
 * little-endian, IVFFLAT_VERSION 1.  Host code, needs no device.  Replaces the page walks of pgvector/src/ivfscan.c:45-101
 * (meta page ivfflat.h:222-230 -> list pages, IvfflatListData ivfflat.h:240-246, chained by IvfflatPageOpaqueData.nextblkno
 * ivfflat.h:232-238) and ivfscan.c:128-168 (a list's entry pages from startPage; index tuples formed at ivfbuild.c:240):
 * an AM shim reads the index file once, loads the rows and builds the same lists in HBM with
 * rk_load_rows + rk_build_ivf_from_lists.  `rel` must stay valid until rk_ivfpages_close.  Corrupt or foreign pages are
 * an error of the call (RK_ERR_INVALID / RK_ERR_UNSUPPORTED with the block and item in rk_last_error()), never a wild read. */
typedef struct rk_ivfpages rk_ivfpages;
int rk_ivfpages_open(const void *rel, size_t nbytes, rk_ivfpages **out);
/* unused_items: line pointers that are not LP_NORMAL (skipped, as PageGetItem callers after a vacuum would not see them) */
int rk_ivfpages_info(const rk_ivfpages *p, int32_t *dim, int32_t *lists, uint64_t *tuples, uint64_t *unused_items);
/* Any output may be NULL.  centers[lists*dim] in list order (list l = l-th item along the list pages); tuples come out list
 * by list in page order: list_start[lists+1], row_list[tuples] (the list of each tuple), x[tuples*dim], tid[tuples] = heap
 * tuple id packed as (block << 16) | offset — the packing INTEGRATION.md §2 unpacks into xs_heaptid. */
int rk_ivfpages_read(const rk_ivfpages *p, float *centers, uint32_t *list_start, uint32_t *row_list, float *x, uint64_t *tid);
void rk_ivfpages_close(rk_ivfpages *p);

/* ---- scans: the IndexAmRoutine contract (pgvector/src/ivfflat.c:236-240, ivfflat.h:335-338) --- */

typedef struct rk_scan_opts {
    int32_t k;              /* LIMIT pushed into the scan; 1..k_max                                 */
    int32_t probes;         /* ivfflat.probes; 0 = exact scan (SeqScan + top-N sort).  probes > 0 searches ONE
                               global IVFFlat index over all rows with the permission test fused into the list scan
                               (filter BITMASK semantics), whatever `layout` says: the reference's per-partition
                               indexes (one ivfflat per role table, initialize_partitions.py:325-336) would choose
                               lists per partition, so approximate results differ from its partitioned IVF arms
                               while exact results (probes = lists) are identical                              */
    int32_t max_probes;     /* ivfflat.max_probes; > probes turns on iterative_scan=relaxed_order   */
    int32_t filter;         /* rk_filter */
    int32_t null_query;     /* ORDER BY vec <-> NULL: every distance is 0 (ivfscan.c:181-200)       */
    int32_t layout;         /* rk_layout: which partition layout RK_FILTER_PARTITION walks          */
    int32_t keep_duplicates;/* 1 = ground-truth merge without dedup (common_function.py:713-714)    */
    int32_t engine;         /* rk_engine: which kernels score the batch                             */
} rk_scan_opts;

/* ivfflatbeginscan (ivfscan.c:241-315): allocates scan state (result ring, device scratch). */
rk_scan *rk_beginscan(rk_index *ix);
/* ivfflatrescan (ivfscan.c:320-334) + the first-call work of ivfflatgettuple (:355-373):
 * binds the ORDER BY operand `q` (borrowed for this call only) and the user whose permissions
 * filter the rows, and runs the scan on the GPU.  An unknown user has no roles and gets an empty
 * result (prefilter_role.py:120-124). */
int rk_rescan(rk_scan *sc, const float *q, int32_t q_dim, uint32_t user_id, const rk_scan_opts *o);
/* ivfflatgettuple (ivfscan.c:339-389): next tuple in ascending distance.  1 = *tid / *distance
 * set (distance is what `vector <-> q` / `vector <#> q` evaluates to, float8); 0 = exhausted
 * (after the last probe batch in iterative mode, ivfscan.c:375-381); <0 = error.
 *   Index scans (probes > 0) stream EVERY tuple of the probed lists the user may see, in order, like the reference's
 * tuplesort (ivfscan.c:375-388): opts.k is only how many tuples one GPU pass brings back; when the ring is drained and
 * the batch held more, the list scan is re-run with the lower bound (key, tid) > the last pair returned, and only when
 * the batch is exhausted does the scan move to the next `probes` lists.
 *   Exact scans (probes = 0) stand for `ORDER BY distance LIMIT k` (prefilter_role.py:162-172): opts.k IS the limit and
 * the scan ends after k tuples. */
int rk_gettuple(rk_scan *sc, uint64_t *tid, uint32_t *doc, double *distance);
/* Up to max_n consecutive rk_gettuple results in one call (what a harness-side caller wants: `cur.fetchall()` of
 * prefilter_role.py:172); arrays may be NULL.  Returns the number of tuples written, 0 = exhausted, <0 = error. */
int rk_gettuple_batch(rk_scan *sc, int max_n, uint64_t *tid, uint32_t *doc, double *distance);
/* ivfflatendscan (ivfscan.c:394-406). */
void rk_endscan(rk_scan *sc);

/* Throughput entry point: nq independent scans in one persistent launch.  Equivalent to nq
 * (rk_rescan + k x rk_gettuple) sequences; replaces the per-query loop of
 * basic_benchmark/common_function.py:1366-1382 and the ground truth of :671-759 / :859-1139.
 *   Q[nq*dim], user_ids[nq]            host buffers
 *   out_tid[nq*k], out_doc[nq*k] (may be NULL), out_dist[nq*k], out_count[nq]   host buffers
 * Rows past out_count[i] are filled with tid = UINT64_MAX, distance = +Inf. */
int rk_search_batch(rk_index *ix, size_t nq, const float *Q, const uint32_t *user_ids,
                    const rk_scan_opts *o, uint64_t *out_tid, uint32_t *out_doc, double *out_dist,
                    int32_t *out_count);

/* Same, with the query matrix and outputs already resident in HBM (device pointers); user_ids
 * stays a host array (it drives the host-side plan).  Asynchronous on the index stream.  The NaN / Inf check of the
 * queries (vector.c:102-113) runs as a kernel and is reported by the next rk_synchronize (RK_ERR_NONFINITE).
 * The rk_search_batch* / rk_merge_shards entry points of one index share a workspace and serialise on a mutex; for
 * concurrent searches from several host threads use one rk_scan handle per thread. */
int rk_search_batch_device(rk_index *ix, size_t nq, const float *dQ, const uint32_t *user_ids,
                           const rk_scan_opts *o, uint64_t *d_out_tid, uint32_t *d_out_doc,
                           double *d_out_dist, int32_t *d_out_count);

/* ---- a search that allocates nothing ------------------------------------------------------------------------------
 * No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid allocation).
 * Exact, unfiltered (RK_FILTER_NONE) top-k of nq device-resident queries over all rows of the index, with the same kernels,
 * arithmetic and tie rule as rk_search_batch_device(filter = RK_FILTER_NONE, engine = RK_ENGINE_SCAN) — results are
 * identical — but the scan's work list and candidate lists live in a file-scope static device array (about 4.9 MB, held by
 * every GPU that loads the library until the process exits) and the work list is written on the device: the call makes no
 * cudaMalloc / cudaFree / cudaHostAlloc and no host<->device copy, only three kernel launches on the index stream
 * (asynchronous).  Limits: nq <= RK_NOALLOC_MAX_NQ, k <= RK_NOALLOC_MAX_K (and <= k_max), dim % 4 == 0; larger sizes are
 * rejected with RK_ERR_INVALID.  All calls share the array, which keeps the previous call's contents (every word a call
 * reads is rewritten first, so calls in stream order are independent); calls that OVERLAP on the GPU — other streams, other
 * indexes on the same device, several host threads — race on it and must be serialised by the caller.  Inputs and outputs
 * are the caller's device buffers: dQ [nq][dim], d_out_tid / d_out_doc (may be NULL) / d_out_dist [nq][k], d_out_count [nq]. */
#define RK_NOALLOC_MAX_NQ 16
#define RK_NOALLOC_MAX_K 128
int rk_search_rows_noalloc(rk_index *ix, size_t nq, const float *dQ, int32_t k, uint64_t *d_out_tid, uint32_t *d_out_doc,
                           double *d_out_dist, int32_t *d_out_count);

/* ---- multi-GPU shard merge (new; the reference has no collective) ------------- */

/* Local shard result in exchange form: per query k (fp32 ordering key, tid, doc) triples sorted
 * ascending, padded with (+Inf, UINT64_MAX).  d_key/d_tid/d_doc are device buffers [nq*k]. */
int rk_search_batch_shard(rk_index *ix, size_t nq, const float *dQ, const uint32_t *user_ids,
                          const rk_scan_opts *o, float *d_key, uint64_t *d_tid, uint32_t *d_doc);
/* IVFFlat list choice alone (GetScanLists, pgvector/src/ivfscan.c:36-107): for each of nq device-resident queries the
 * min(o->probes, lists) nearest lists, nearest first, as list numbers into the device buffer d_lists [nq][that many].
 * With centres replicated on every rank of a sharded index each rank chooses the lists of nq/G
 * queries, the choices are all-gathered (4 * probes bytes per query) and rk_search_batch_shard_probed scans them:
 * the same lists every rank would have chosen itself, at 1/G of the centre scoring per rank. */
int rk_ivf_probe_device(rk_index *ix, size_t nq, const float *dQ, const rk_scan_opts *o, uint32_t *d_lists);
/* rk_search_batch_shard with the list choice supplied (d_lists from rk_ivf_probe_device, same o->probes). */
int rk_search_batch_shard_probed(rk_index *ix, size_t nq, const float *dQ, const uint32_t *user_ids,
                                 const rk_scan_opts *o, const uint32_t *d_lists, float *d_key, uint64_t *d_tid,
                                 uint32_t *d_doc);
/* G-way select over all-gathered shard results; writes the final [nq][k] like
 * rk_search_batch_device.  Device pointers.  Shard j's block of each array starts
 * shard_stride_bytes * j bytes after shard 0's (0 = dense [G][nq][k] arrays), so one all-gather of
 * a packed per-rank buffer {key[nq*k] | doc[nq*k] | tid[nq*k]} can feed it directly. */
int rk_merge_shards(rk_index *ix, size_t nq, int k, int n_shards, size_t shard_stride_bytes,
                    const float *d_key, const uint64_t *d_tid, const uint32_t *d_doc,
                    uint64_t *d_out_tid, uint32_t *d_out_doc, double *d_out_dist, int32_t *d_out_count);

/* ---- peer exchange: the shard merge over NVLink peer memory, no collective library on the data path -------------
 * Every rank owns a ring of `depth` receive slots ([world] packed shard results {key f32 | doc u32 | tid u64}[nq * k] each)
 * with an arrival and an acknowledgement flag per (slot, rank).  rk_search_batch_shard_push searches this rank's shard
 * into its own block of the next slot and a small kernel stores that block into every peer's slot and raises the peers'
 * arrival flags; rk_merge_shards_pull waits (in the kernel, bounded) for all peers' flags of the oldest pushed step,
 * merges, and acknowledges the slot.  Ranks meet only through those flags: the merge of step s may run on another stream
 * under the scan of step s + 1, and a rank may run up to `depth` steps ahead of the slowest.  All ranks must make the
 * same sequence of push / pull calls.  One process per GPU: exchange rk_exchange_get_handle's 64 bytes (a CUDA IPC
 * handle) and rk_exchange_connect_ipc them; several GPUs in one process: rk_exchange_connect_ptr(rk_exchange_base). */
typedef struct rk_exchange rk_exchange;
int rk_exchange_create(rk_index *ix, size_t nq_max, int k_max, int world, int rank, int depth, rk_exchange **out);
int rk_exchange_destroy(rk_exchange *ex);
int rk_exchange_get_handle(rk_exchange *ex, void *handle64);
void *rk_exchange_base(rk_exchange *ex);
size_t rk_exchange_block_bytes(rk_exchange *ex);
int rk_exchange_connect_ipc(rk_exchange *ex, int peer_rank, const void *handle64);
int rk_exchange_connect_ptr(rk_exchange *ex, int peer_rank, void *peer_base, int peer_device);
/* d_lists: NULL, or the IVFFlat list choice as for rk_search_batch_shard_probed.  Asynchronous on the index stream. */
int rk_search_batch_shard_push(rk_index *ix, rk_exchange *ex, size_t nq, const float *dQ, const uint32_t *user_ids,
                               const rk_scan_opts *o, const uint32_t *d_lists);
/* cuda_stream: NULL = the index stream; another stream must be ordered after the matching push by the caller (an event). */
int rk_merge_shards_pull(rk_index *ix, rk_exchange *ex, size_t nq, int k, void *cuda_stream, uint64_t *d_out_tid,
                         uint32_t *d_out_doc, double *d_out_dist, int32_t *d_out_count);
/* Synchronises cuda_stream (NULL = the index stream); RK_ERR_CUDA if a wait for a peer timed out since creation. */
int rk_exchange_status(rk_exchange *ex, void *cuda_stream);

/* ---- several GPUs of one box behind one handle (one process; SURVEY.md §8b `n_gpus`) --------------------------------
 * The nearest thing in the reference is FAISS's IndexShards: host threads over sub-indexes and a CPU merge
 * (logical_partition_benchmark/benchmark/faiss/faiss/IndexShards.cpp:197-256).  rk_multi keeps one rk_index per device
 * (document d on shard d mod n_gpus), one host thread per device and the peer exchange above between them; the calls
 * mirror the single-index ones and results are identical to one index holding all rows.  devices[n_gpus]: CUDA ordinals
 * (an ordinal may repeat: several shards on one GPU, for tests).  nq_max / cfg->k_max size the exchange rings. */
typedef struct rk_multi rk_multi;
rk_multi *rk_multi_open(const rk_config *cfg, const int32_t *devices, int32_t n_gpus, size_t nq_max);
void rk_multi_close(rk_multi *m);
int rk_multi_n_gpus(const rk_multi *m);
rk_index *rk_multi_shard(rk_multi *m, int g);           /* the per-device index (statistics, options) */
int rk_multi_load_rows(rk_multi *m, const float *x, const uint64_t *tid, const uint32_t *doc, size_t n);
int rk_multi_set_doc_roles(rk_multi *m, const uint32_t *doc, const uint32_t *role, size_t n);
int rk_multi_set_user_roles(rk_multi *m, const uint32_t *user, const uint32_t *role, size_t n);
int rk_multi_build_role_partitions(rk_multi *m);
int rk_multi_build_combination_partitions(rk_multi *m);
int rk_multi_define_partition(rk_multi *m, uint32_t part_id, const uint32_t *docs, size_t n);
int rk_multi_map_roleset_to_partitions(rk_multi *m, const uint32_t *roles, size_t nroles, const uint32_t *parts, size_t nparts);
int rk_multi_finalize(rk_multi *m);
/* Shard 0 trains the centres on its stripe, the other shards assign their rows to them. */
int rk_multi_build_ivf(rk_multi *m, int lists, uint64_t seed);
/* rk_search_batch over all shards: host buffers in and out, synchronous. */
int rk_multi_search_batch(rk_multi *m, size_t nq, const float *Q, const uint32_t *user_ids, const rk_scan_opts *o,
                          uint64_t *out_tid, uint32_t *out_doc, double *out_dist, int32_t *out_count);

/* ---- plumbing ------------------------------------------------------------------------------- */

/* Run on the caller's CUDA stream (a cudaStream_t), e.g. torch's current stream. NULL = own stream. */
int rk_set_stream(rk_index *ix, void *cuda_stream);
int rk_synchronize(rk_index *ix);

typedef struct rk_stats {
    uint64_t rows;             /* rows in the store                                                */
    uint64_t docs;
    uint64_t partitions;
    uint64_t device_bytes;     /* HBM held by the index                                            */
    /* last batch */
    uint64_t launches;         /* kernels launched by the last search call                         */
    uint64_t segments;         /* work-list segments of the last scan launch                       */
    uint64_t rows_scanned;     /* rows whose vectors the plan makes the kernel read                */
    uint64_t algo_bytes;       /* rows_scanned*4d + mask bytes + 12k per query    */
    float scan_ms;             /* CUDA-event time of the dominant scan kernel (timing enabled)     */
    float select_ms;           /* CUDA-event time of the merge/select kernel                       */
    float probe_ms;            /* IVF centre scoring + list choice                                 */
    float reserved_f;
    uint64_t total_launches;   /* kernels launched since rk_open                                   */
    uint64_t tensor_redo;      /* RK_ENGINE_TENSOR: queries of the last batch re-run on the scan   */
    uint64_t tensor_exact;     /* RK_ENGINE_TENSOR: 1 = TF32 was exact for the last batch          */
    uint64_t tensor_plan;      /* last batch: 0 = not on the tensor engine, 1 = dense, 2 = per user mask,
                                  3 = per document class (atoms); +16 when the sampling pass ran, +32 CTA-pair kernel,
                                  +64 bf16 nomination                                                */
    uint64_t tensor_flop;      /* last tensor batch: 2 * dim * (query, row) pairs scored (main pass) */
    uint64_t tensor_plan_bytes;/* last tensor batch: bytes the plan stages (visited tiles x query tiles) */
    uint64_t segment_kinds;    /* last scan launch: bit i set = the work list held segments of kind i (0 extent table,
                                  2 row range, 3 row range with a per-row permission test, 4 class-ordered IVF lists) */
    uint64_t single_fused;     /* last call: 1 = a single query ran as ONE fused launch (scan + select)              */
    uint64_t ivf_assign_tc_rows;       /* IVFFlat builds so far: row assignments made through the tensor-core path ...  */
    uint64_t ivf_assign_fallback_rows; /* ... and how many of them its proof refused (re-assigned by the fp32 kernel)   */
} rk_stats;
int rk_get_stats(rk_index *ix, rk_stats *out);
/* Record CUDA events around the kernels of each search call (costs two events per kernel). */
int rk_enable_timing(rk_index *ix, int on);
/* Durations (ms) of the most recent scan-kernel launches made with timing enabled, newest last;
 * up to RK_TIMING_RING are kept.  Synchronizes the index stream.  Returns how many were written. */
#define RK_TIMING_RING 64
int rk_get_scan_times(rk_index *ix, float *out_ms, int max_n);

#if defined(__GNUC__)
#pragma GCC visibility pop
#endif
#ifdef __cplusplus
}
#endif
#endif /* RBACKNN_H */
