// rk_host.h — host-side state of librbacknn (internal; the public surface is include/rbacknn.h).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <atomic>
#include <cstring>
#include <map>
#include <mutex>
#include <string>
#include <unordered_map>
#include <utility>
#include <vector>

#include "../../include/rbacknn.h"
#include "rk_kernels.h"

int rk_fail(int code, const char *fmt, ...);

// true when no element is NaN or +-Inf (exponent bits all ones).  Branch-free blocks so the compiler vectorises it: the
// check sits on the host path of every call (pgvector rejects such vectors in vector_in, vector.c:102-113).
static inline bool rk_all_finite(const float *x, size_t n)
{
    size_t i = 0;
    while (i < n) {
        const size_t e = n - i > 8192 ? i + 8192 : n;
        uint32_t bad = 0;
        for (; i < e; i++) {
            uint32_t v;
            memcpy(&v, x + i, 4);
            bad |= (uint32_t) ((v & 0x7f800000u) == 0x7f800000u);
        }
        if (bad) return false;
    }
    return true;
}

struct DevBuf {
    void *p = nullptr;
    size_t cap = 0;
    int ensure(size_t bytes);   // grows (contents NOT preserved)
    void release();
    template <typename T> T *as() const { return reinterpret_cast<T *>(p); }
};

struct PinBuf {
    void *p = nullptr;
    size_t cap = 0;
    int ensure(size_t bytes);
    void release();
    template <typename T> T *as() const { return reinterpret_cast<T *>(p); }
};

// One cached single-query plan (fused path): the work list and the user's role mask, resident in HBM.  It depends on the
// user's role set, the filter and the layout only — not on the query vector or k — so a pgvector-shaped scan
// (one query per rk_rescan) uploads nothing but its query.
struct FusedPlan {
    DevBuf buf;                 // [segments][role mask]
    uint32_t nseg = 0;
    size_t off_qmask = 0, off_xrow = 0, off_xvstart = 0;
    bool private_extents = false;   // the plan carries its own extent table: only the documents the role set may see
    uint64_t rows_scanned = 0, algo_bytes = 0;   // algo_bytes without the 12k output term
};

// Everything one search context needs: a stream, the uploaded plan, per-segment candidate lists
// and result buffers.  The index owns one (rk_search_batch*), every rk_scan owns another.
struct Workspace {
    int device = 0;
    cudaStream_t stream = nullptr, own_stream = nullptr;
    cudaEvent_t ev[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    DevBuf segs, order, ctl /* [0]=nseg [1]=counter */, cand_key, cand_tid, cand_row, cand_count, qseg, qmask, dq;
    DevBuf out_tid, out_doc, out_dist, out_count;
    DevBuf qflag;                        // [1] u32: set by the device-side NaN / Inf check of device-resident queries
    bool qflag_armed = false;            // a check was launched since the flag was last read
    uint32_t *h_qflag = nullptr;         // pinned
    DevBuf probe_key, probe_list;        // IVF: per-query selected lists
    const uint32_t *ext_probe = nullptr; // IVF: list choice supplied by the caller (rk_search_batch_shard_probed), [nq][probes]
    DevBuf ex_key, ex_tid, ex_doc;       // exchange-form results
    // fused single-query path
    std::unordered_map<uint64_t, FusedPlan> fused_plans;
    DevBuf fz_ctl, fz_key, fz_row, fz_pack;   // control words, flat hand-over list, packed outputs [count | tid | dist | doc | key]
    PinBuf h_fz;                              // pinned staging: query in, packed outputs back
    // zero-copy variant: one MAPPED pinned buffer [flag | packed outputs | query] the kernel reads / writes directly
    unsigned char *h_zc = nullptr, *d_zc = nullptr;
    size_t zc_cap = 0;
    uint32_t zc_seq = 0;
    unsigned long long *h_dbg = nullptr;      // pinned [sm + 1][8] phase stamps (option single_debug)
    DevBuf d_dbg;
    PinBuf h_plan_ring[2], h_out;        // plan staging alternates so the host may run one batch ahead
    cudaEvent_t ring_a[RK_TIMING_RING] = {}, ring_b[RK_TIMING_RING] = {};   // scan-kernel event pairs
    uint64_t ring_n = 0;
    cudaEvent_t ev_plan[2] = {nullptr, nullptr};
    bool ev_plan_used[2] = {false, false};
    int plan_idx = 0;
    bool timing = false;
    // lower bound of a streaming scan (rk_gettuple past the k-th tuple): only tuples with (key, tid) > (lb_key, lb_tid) are kept
    bool lb_on = false;
    uint32_t lb_key = 0;
    uint64_t lb_tid = 0;
    // stats of the last call
    uint64_t launches = 0, segments = 0, rows_scanned = 0, algo_bytes = 0;
    uint32_t seg_kinds = 0;               // bit per RK_SEG_* kind in the last scan launch's work list
    bool single_fused = false;            // the last call ran as one fused launch
    bool plan_single_partition = false;   // last exact plan: no query walks more than one partition (no tuple can repeat)
    bool timed_scan = false, timed_select = false, timed_probe = false;
    int init(int device);
    void destroy();
};

struct PartInfo {
    uint32_t ext_off = 0;     // index of first extent in the global extent arrays
    uint32_t n_ext = 0;
    uint64_t rows = 0;
    std::vector<uint32_t> docs;   // dense document indices, ascending
    std::vector<uint32_t> vstart; // [n_ext + 1] virtual row offset of each extent, last = rows
};

// The extents of a partition one role mask may see, as maximal runs of neighbouring extents [lo, hi) (indices inside the
// partition).  `few` = at most RK_FEW_RUNS of them (documents ordered by role mask: rk_finalize) — only then are lo / hi filled.
#define RK_FEW_RUNS 64
struct VisRuns {
    bool few = false;
    std::vector<uint32_t> lo, hi;
};

struct UserInfo {
    std::vector<uint32_t> roles;        // raw role ids, ascending, unique
    std::vector<uint64_t> mask;         // [mask_words]
    uint64_t allowed_rows = 0;          // rows the RLS predicate lets this user see
    uint64_t allowed_docs = 0;
    uint32_t mask_id = 0;               // users with the same role mask share an id (dense, assigned by rk_finalize)
};

struct IvfIndex {
    bool built = false;
    uint64_t assign_tc_rows = 0, assign_fallback_rows = 0;   // rk_stats: tensor-core assignments and the rows their proof refused
    int lists = 0;
    DevBuf d_x, d_tid, d_rowdoc;        // list-sorted copy of the row store
    DevBuf d_centers, d_center_tid, d_list_start;
    std::vector<uint32_t> list_start;   // [lists+1]
    std::vector<uint32_t> row_list;     // by finalized row order
    std::vector<float> centers;         // [lists][dim]
    uint32_t max_list_len = 0;
    int chunks_per_list = 1;
    // rows of every list ordered by document class (distinct role mask) when there are at most 128 classes
    int ncls = 0;                                   // 0 = lists keep row order, masks are tested per row
    DevBuf d_cls_off;                               // [lists * (ncls + 1) + 1]
    std::vector<std::vector<uint64_t>> cls_mask;    // role mask of each class
    std::vector<uint64_t> qcls_by_mask_id;          // [n_mask_ids][2] visible classes per user mask id (filled by the build)
};

// State of the batched tensor-core engine (rk_tensor.cu / rk_bscan.cu); used by the index's own
// workspace only (rk_search_batch*).
struct TensorState {
    bool ready = false;
    DevBuf d_xnorm, d_rowmask, d_rowmask2, d_probe;
    bool x_integer = false;
    float x_absmax = 0.f, xnorm_max = 0.f;
    DevBuf d_plan, d_qp, d_tiles;
    DevBuf d_xb, d_qpb;                  // bf16 nomination: shadow copy of the row store [rows][dim_b] (built on first use), of the gathered queries
    bool bf16_off = false;               // the bf16 nomination's proof refused too many queries of a batch: this index nominates in TF32 from now on
    int dim_b = 0;                       // elements per bf16 row (dim_pad rounded up to 8: 16-byte rows for the tensor map)
    std::vector<uint32_t> h_tiles;                                              // tile lists of all cached role masks
    std::map<std::vector<uint64_t>, std::pair<uint32_t, uint32_t>> tile_cache;  // role mask -> (offset, count) in h_tiles
    std::vector<std::pair<uint32_t, uint32_t>> tile_by_mask_id;                 // the same, by UserInfo::mask_id (count ~0u = not built yet)
    struct Atom { std::vector<uint64_t> mask; uint32_t tl_off = 0, tl_n = 0; uint64_t rows = 0; };
    std::vector<Atom> atoms;                                                    // document classes by role mask (empty: too many)
    bool tiles_dirty = false;                                                   // h_tiles grew since the last upload
    int last_plan = 0;
    uint64_t last_plan_bytes = 0;
    DevBuf d_redo;                       // queries the proof refused: their indices, vectors and the sub-batch's outputs (one block)
    DevBuf d_units, d_qnorm, d_scratch, d_cand_key, d_cand_row, d_cand_count, d_sel_key, d_sel_row, d_sel_count, d_tau_key, d_tau_count, d_flags;
    PinBuf h_flags;
    double last_flop = 0.0;
    uint32_t last_redo = 0;
    bool last_exact = false;
};

// Tuning / test switches.  Defaults come from the RK_* environment variables, read ONCE in rk_open (never on the per-call
// path); rk_set_option changes them per index afterwards.
struct RkTuning {
    int tensor_tpu = 0;          // RK_TENSOR_TPU: tiles per unit (0 = planner's choice)
    int tensor_pair = -1;        // RK_TENSOR_PAIR: -1 = by row length, 0 = one-CTA kernel, 1 = CTA-pair kernel
    int tensor_plan = 0;         // RK_TENSOR_PLAN: 0 = cheapest, 1 = dense, 2 = grouped, 3 = atoms
    int tensor_sample = 1;       // RK_TENSOR_SAMPLE: 0 disables the sampling pass
    int tensor_sdiv = 0;         // RK_TENSOR_SDIV: 0 = default (400)
    int tensor_smin = 1024;      // RK_TENSOR_SMIN
    int tensor_host_timing = 0;  // RK_TENSOR_HOST_TIMING
    int tensor_qtau = 1;         // RK_TENSOR_QTAU: 0 = every unit of the tensor engine keeps its own threshold
    int tensor_kp_exact = 28;    // RK_TENSOR_KP_EXACT: nominations beyond k while nomination is EXACT (integer data, the previous batch proved it); < 0 = off
    int tensor_loose = 128;      // RK_TENSOR_LOOSE: entries beyond kp a mid-unit compaction may keep
    int tensor_round_mark = 512; // RK_TENSOR_ROUND_MARK: buffers above this many entries are compacted whenever a compaction round runs (trigger: 768)
    int class_order = 1;         // RK_CLASS_ORDER: 1 = rk_finalize orders the documents (= the rows) by (role mask, id) when masks repeat (at least 4 documents per distinct mask on average), 2 = always, 0 = by id
    int tensor_kp_extra = 0;     // RK_TENSOR_KP_EXTRA: added to the nominations per query (kp = max(2k, k + 128) + ...; at most 512)
    int tensor_kp_bf16 = 64;     // RK_TENSOR_KP_BF16: added on top when the nomination runs in bf16 (C2: 9 refused queries per 1000 -> 1)
    int tensor_bf16 = -1;        // RK_TENSOR_BF16: -1 = bf16 on the CTA-pair kernel while the proof accepts it, 0 = TF32 nomination, 1 = bf16 always
    int ivf_seg_factor = 4;      // RK_IVF_SEG_FACTOR
    int ivf_kmeanspp = 1;        // RK_IVF_KMEANSPP: 1 = k-means++ seeding (ivfkmeans.c:21-89), 0 = the first `lists` sampled rows
    int ivf_kmeans_iters = 25;   // RK_IVF_KMEANS_ITERS: Lloyd iterations at most (0 = the seeds become the centres; tests)
    int ivf_sample_identity = 0; // RK_IVF_SAMPLE_IDENTITY: the sample is the first rows in index order, unshuffled (tests)
    int ivf_assign_tc = 1;       // RK_IVF_ASSIGN_TC: 0 = the fp32 assignment kernel for k-means and the final assignment
    int ivf_probe_single = 0;    // RK_IVF_PROBE_SINGLE
    int single_fused = 1;        // RK_SINGLE_FUSED: 0 = single queries take the two-launch scan + select path
    int seg_cost_x10 = 80;       // RK_SEG_COST_X10: fixed cost of one segment in 0.1 us (measured: 7.8 us on the 1/8 stripe of C3), for the segment sizing (segments per CTA = sqrt(T / 2o))
    int single_debug = 0;        // RK_SINGLE_DEBUG: print the fused launch's phase timeline (globaltimer stamps) to stderr
    int single_segs = 0;         // RK_SINGLE_SEGS: segments per CTA of a fused single-query plan (0 = chosen from the scan length)
    int single_seed = 0;         // RK_SINGLE_SEED: candidates a CTA of a long fused scan holds before its first compaction (0 = off)
    int single_seed_min = 2048;  // RK_SINGLE_SEED_MIN: ... for plans with at least this many rows per CTA
    int single_zero_copy = 1;    // RK_SINGLE_ZERO_COPY: 0 = cudaMemcpy in / out + stream sync instead of mapped host memory + flag
};

struct rk_index {
    rk_config cfg{};
    RkTuning tune;
    std::mutex batch_mu;         // rk_search_batch* / rk_merge_shards share ix->ws: one caller at a time
    int dim = 0, dim_pad = 0, metric = 0, device = 0, k_max = 0, sm_count = 148;
    int tile_rows = 0, nstages = 0;
    size_t scan_smem = 0;
    bool finalized = false;
    // staging
    size_t nrows = 0;
    DevBuf d_x;                                   // [nrows][dim_pad]
    std::vector<uint64_t> h_tid;
    std::vector<uint32_t> h_doc;                  // raw document ids (by row)
    std::vector<std::pair<uint32_t, uint32_t>> doc_roles, user_roles;
    std::map<uint32_t, std::vector<uint32_t>> part_defs;
    std::map<std::vector<uint32_t>, std::vector<uint32_t>> roleset_parts;
    bool want_role_parts = false, want_combo_parts = false;
    // finalized layout
    std::vector<uint32_t> load_to_row;            // load order -> finalized row
    std::vector<uint32_t> doc_ids;                // dense -> raw document id (dense order: by (role mask, raw id), see rk_finalize)
    std::vector<uint32_t> doc_row_start;          // [docs+1]
    std::vector<uint32_t> h_rowdoc;               // dense doc of each finalized row
    std::vector<uint32_t> role_ids;               // dense -> raw role id (ascending)
    int mask_words = 1;
    std::vector<uint64_t> doc_mask;               // [docs][mask_words]
    std::unordered_map<uint32_t, UserInfo> users;
    std::vector<PartInfo> parts;                  // [0] = ALL documents
    std::unordered_map<uint32_t, uint32_t> role_part;                    // raw role -> parts index
    std::map<std::vector<uint32_t>, uint32_t> combo_part;                // role set -> parts index
    std::unordered_map<uint32_t, uint32_t> dyn_part;                     // part_id -> parts index
    uint32_t n_mask_ids = 0;                      // distinct user role masks (UserInfo::mask_id < n_mask_ids)
    std::map<std::pair<std::vector<uint32_t>, uint32_t>, bool> pure_cache;  // (role set, part) -> pure?
    std::map<std::pair<std::vector<uint32_t>, uint32_t>, uint64_t> visible_cache;  // (role set, part) -> rows of the part the role set sees
    std::map<std::pair<uint32_t, uint32_t>, VisRuns> runs_cache;   // (user mask id, part) -> the visible extents as runs (rk_search.cu: visible_runs)
    std::mutex mu;
    DevBuf d_tid, d_rowdoc, d_doc_ids, d_doc_mask, d_ext_row, d_ext_vstart, d_ext_mask, d_part_ext_off;
    uint64_t device_bytes = 0;
    IvfIndex ivf;
    TensorState tensor;
    Workspace ws;
    std::atomic<uint64_t> total_launches{0};
    float last_scan_ms = 0, last_select_ms = 0, last_probe_ms = 0;
};

struct rk_scan {
    rk_index *ix = nullptr;
    Workspace ws;
    std::vector<float> q;
    uint32_t user = 0;
    rk_scan_opts opts{};
    bool bound = false;
    // result ring of the current probe batch
    std::vector<uint64_t> r_tid;
    std::vector<uint32_t> r_doc;
    std::vector<double> r_dist;
    std::vector<float> r_key;   // fp32 ordering values of the ring (the resume bound of a streaming scan)
    int r_count = 0, r_pos = 0;
    bool batch_more = false;    // the last fetch returned k tuples: the batch may hold more
    int batch_offset = 0, batch_probes = 0;   // probe offset / count of the batch being streamed
    int probes_done = 0;        // lists consumed so far (iterative scan)
    int max_probes_eff = 0, probes_eff = 0;
};

// rk_tensor.cu
int rk_tensor_supported(const rk_index *ix, const rk_scan_opts *o);
int rk_tensor_search(rk_index *ix, Workspace &ws, size_t nq, const float *dQ, const uint32_t *user_ids, const rk_scan_opts *o,
                     const RkSelectParams &out, std::vector<uint32_t> &redo);

// rk_ivf.cu
int rk_ivf_search(rk_index *ix, Workspace &ws, size_t nq, const float *dQ, const uint32_t *user_ids,
                  const rk_scan_opts *o, int probe_offset, const RkSelectParams &out);
