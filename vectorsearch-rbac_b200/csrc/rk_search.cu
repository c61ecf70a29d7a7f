// rk_search.cu — query planning and launches: (query vector, user) batches -> work list -> scan
// kernel -> select kernel; plus the pgvector-shaped scan iterator.
//
// Mirrors, on the host, what the harness does around each SQL statement:
//   user -> roles                 prefilter_role.py:120-124 (unknown user -> no roles -> empty result)
//   roles -> role partitions      prefilter_role.py:124-172
//   roles -> combination table    prefilter_combination_role.py:122-133
//   roles -> dynamic partitions   controller/dynamic_partition/search.py:44-60, RLS unless pure
//                                 (load_result_to_database.py:591-622)
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>

#include "rk_host.h"

#define RK_CUDA(call)                                                                                    \
    do {                                                                                                 \
        cudaError_t e__ = (call);                                                                        \
        if (e__ != cudaSuccess)                                                                          \
            return rk_fail(RK_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), __FILE__, __LINE__); \
    } while (0)

static int validate_opts(const rk_index *ix, const rk_scan_opts *o)
{
    if (!o) return rk_fail(RK_ERR_INVALID, "scan options are NULL");
    if (o->k < 1 || o->k > ix->k_max) return rk_fail(RK_ERR_INVALID, "k = %d outside 1..k_max (%d)", o->k, ix->k_max);
    if (o->probes < 0 || o->probes > RK_MAX_LISTS) return rk_fail(RK_ERR_INVALID, "probes = %d outside 0..%d", o->probes, RK_MAX_LISTS);
    if (o->filter < RK_FILTER_PARTITION || o->filter > RK_FILTER_NONE) return rk_fail(RK_ERR_INVALID, "unknown filter %d", o->filter);
    if (o->engine < RK_ENGINE_AUTO || o->engine > RK_ENGINE_TENSOR) return rk_fail(RK_ERR_INVALID, "unknown engine %d", o->engine);
    if (o->probes > 0 && !ix->ivf.built) return rk_fail(RK_ERR_STATE, "probes > 0 but no IVFFlat index was built");
    return 0;
}

static bool partition_pure(rk_index *ix, const UserInfo &u, uint32_t pidx)
{
    // a partition is pure for a role set when every document in it is visible to that role set
    // (load_result_to_database.py:591-604 skips the RLS policy for those)
    std::lock_guard<std::mutex> g(ix->mu);
    auto key = std::make_pair(u.roles, pidx);
    auto it = ix->pure_cache.find(key);
    if (it != ix->pure_cache.end()) return it->second;
    bool pure = true;
    const int W = ix->mask_words;
    for (uint32_t d : ix->parts[pidx].docs) {
        uint64_t acc = 0;
        for (int w = 0; w < W; w++) acc |= ix->doc_mask[(size_t) d * W + w] & u.mask[w];
        if (!acc) { pure = false; break; }
    }
    ix->pure_cache[key] = pure;
    return pure;
}

// Rows of a partition a role set may see (what a row-filtered scan of an impure partition reads: SURVEY.md §8(d) counts
// only those towards the algorithmic bytes).
static uint64_t partition_visible_rows(rk_index *ix, const UserInfo &u, uint32_t pidx)
{
    std::lock_guard<std::mutex> g(ix->mu);
    auto key = std::make_pair(u.roles, pidx);
    auto it = ix->visible_cache.find(key);
    if (it != ix->visible_cache.end()) return it->second;
    const int W = ix->mask_words;
    uint64_t rows = 0;
    for (uint32_t d : ix->parts[pidx].docs) {
        uint64_t acc = 0;
        for (int w = 0; w < W; w++) acc |= ix->doc_mask[(size_t) d * W + w] & u.mask[w];
        if (acc) rows += ix->doc_row_start[d + 1] - ix->doc_row_start[d];
    }
    ix->visible_cache[key] = rows;
    return rows;
}

// Which extents of partition pidx does this user's role mask see?  Cached per (mask id, partition).
static const VisRuns *visible_runs(rk_index *ix, const UserInfo &u, uint32_t pidx)
{
    std::lock_guard<std::mutex> g(ix->mu);
    auto key = std::make_pair(u.mask_id, pidx);
    auto it = ix->runs_cache.find(key);
    if (it != ix->runs_cache.end()) return &it->second;
    VisRuns vr;
    vr.few = true;
    const int W = ix->mask_words;
    const PartInfo &P = ix->parts[pidx];
    bool open = false;
    for (uint32_t e = 0; e < P.n_ext && vr.few; e++) {
        const uint32_t d = P.docs[e];
        uint64_t acc = 0;
        for (int w = 0; w < W; w++) acc |= ix->doc_mask[(size_t) d * W + w] & u.mask[w];
        if (acc) {
            if (!open) {
                if (vr.lo.size() == RK_FEW_RUNS) { vr.few = false; break; }
                vr.lo.push_back(e); vr.hi.push_back(e + 1); open = true;
            } else vr.hi.back() = e + 1;
        } else open = false;
    }
    if (!vr.few) { vr.lo.clear(); vr.hi.clear(); }
    return &(ix->runs_cache[key] = std::move(vr));
}

// rows_est: rows the segments cover (sizing, scheduling); rows_algo: rows that count as algorithmic traffic; runs: the
// visible extents of a masked partition when they form a few runs (segments are then cut inside the runs, need no mask test
// and skip what lies between them)
struct PlanItem { uint32_t part; uint32_t flags; uint64_t rows_est; uint64_t rows_algo; const VisRuns *runs = nullptr; };

// Which partitions does this user's scan walk?  Returns <0 on error.
static int plan_user(rk_index *ix, const UserInfo *u, const rk_scan_opts *o, std::vector<PlanItem> &items)
{
    items.clear();
    if (o->filter == RK_FILTER_NONE) {
        items.push_back({0u, 0u, ix->parts[0].rows, ix->parts[0].rows});
        return 0;
    }
    if (!u || u->roles.empty()) return 0;   // no roles -> nothing visible
    if (o->filter == RK_FILTER_BITMASK) {
        if (u->allowed_rows) {
            items.push_back({0u, RK_SEGF_MASKED, u->allowed_rows, u->allowed_rows});
            const VisRuns *vr = visible_runs(ix, *u, 0u);
            if (vr->few) items.back().runs = vr;
        }
        return 0;
    }
    int layout = o->layout;
    if (layout == RK_LAYOUT_AUTO) {
        if (!ix->roleset_parts.empty()) layout = RK_LAYOUT_DYNAMIC;
        else if (ix->want_combo_parts) layout = RK_LAYOUT_COMBINATION;
        else if (ix->want_role_parts) layout = RK_LAYOUT_ROLE;
        else return rk_fail(RK_ERR_STATE, "partition filter requested but no partition layout was built");
    }
    if (layout == RK_LAYOUT_ROLE) {
        if (!ix->want_role_parts) return rk_fail(RK_ERR_STATE, "role partitions were not built");
        for (uint32_t r : u->roles) {
            auto it = ix->role_part.find(r);
            if (it != ix->role_part.end() && ix->parts[it->second].rows)
                items.push_back({it->second, 0u, ix->parts[it->second].rows, ix->parts[it->second].rows});
        }
    } else if (layout == RK_LAYOUT_COMBINATION) {
        if (!ix->want_combo_parts) return rk_fail(RK_ERR_STATE, "combination partitions were not built");
        auto it = ix->combo_part.find(u->roles);
        if (it != ix->combo_part.end() && ix->parts[it->second].rows)
            items.push_back({it->second, 0u, ix->parts[it->second].rows, ix->parts[it->second].rows});
    } else if (layout == RK_LAYOUT_DYNAMIC) {
        auto it = ix->roleset_parts.find(u->roles);
        if (it == ix->roleset_parts.end()) return 0;   // no CombRolePartitions row -> no partitions (search.py:54-60)
        for (uint32_t pid : it->second) {
            uint32_t pidx = ix->dyn_part[pid];
            if (!ix->parts[pidx].rows) continue;
            bool pure = partition_pure(ix, *u, pidx);
            items.push_back({pidx, pure ? 0u : RK_SEGF_MASKED, ix->parts[pidx].rows, pure ? ix->parts[pidx].rows : partition_visible_rows(ix, *u, pidx)});
            if (!pure) {
                const VisRuns *vr = visible_runs(ix, *u, pidx);
                if (vr->few) { items.back().runs = vr; items.back().rows_est = items.back().rows_algo; }
            }
        }
    } else {
        return rk_fail(RK_ERR_INVALID, "unknown layout %d", layout);
    }
    return 0;
}

// Cuts one partition (extent table P) into segments of about `target` rows, appended to segs.
static void segments_of_part(const PartInfo &P, uint32_t flags, uint64_t rows_est, uint32_t q, uint64_t target, bool merge_ok,
                             std::vector<RkSegment> &segs)
{
    uint64_t nchunks = (rows_est + target - 1) / target;
    if (nchunks < 1) nchunks = 1;
    if (nchunks > P.rows) nchunks = P.rows;
    for (uint64_t c = 0; c < nchunks; c++) {
        RkSegment s;
        s.query = q; s.kind = RK_SEG_EXTENTS;
        s.flags = (uint16_t) (flags | (merge_ok ? RK_SEGF_MERGE_OK : 0u));
        s.begin = P.rows * c / nchunks;
        s.end = P.rows * (c + 1) / nchunks;
        if (s.end <= s.begin) continue;
        // extents [lo, hi) of the partition intersect virtual rows [begin, end)
        const auto vb = P.vstart.begin(), ve = P.vstart.end();
        const uint32_t lo = (uint32_t) (std::upper_bound(vb, ve, (uint32_t) s.begin) - vb) - 1;
        const uint32_t hi = (uint32_t) (std::lower_bound(vb, ve, (uint32_t) s.end) - vb);
        s.ext_lo = P.ext_off + lo;
        s.ext_hi = P.ext_off + hi;
        segs.push_back(s);
    }
}

// Cuts the partitions one query walks into segments of about `target` rows (appended to segs); returns the extent /
// mask bytes the plan makes the producer read (SURVEY.md §8(d) "M").
static uint64_t append_segments(const rk_index *ix, const std::vector<PlanItem> &items, uint32_t q, uint64_t target,
                                std::vector<RkSegment> &segs)
{
    const int W = ix->mask_words;
    uint64_t mask_bytes = 0;
    for (auto &pi : items) {
        const PartInfo &P = ix->parts[pi.part];
        if (pi.runs) {
            // the visible extents are a few runs: cut each run on its own — every row of such a segment is visible, so no mask
            // test (and no mask bytes), nothing between the runs is walked, and the segments are equal in VISIBLE rows
            for (size_t r = 0; r < pi.runs->lo.size(); r++) {
                const uint64_t v0 = P.vstart[pi.runs->lo[r]], v1 = P.vstart[pi.runs->hi[r]];
                uint64_t nch = (v1 - v0 + target / 2) / target;
                if (nch < 1) nch = 1;
                for (uint64_t c = 0; c < nch; c++) {
                    RkSegment s;
                    s.query = q; s.kind = RK_SEG_EXTENTS;
                    s.flags = (uint16_t) ((pi.flags & ~RK_SEGF_MASKED) | (items.size() == 1 ? RK_SEGF_MERGE_OK : 0u));
                    s.begin = v0 + (v1 - v0) * c / nch;
                    s.end = v0 + (v1 - v0) * (c + 1) / nch;
                    if (s.end <= s.begin) continue;
                    const auto vb = P.vstart.begin(), ve = P.vstart.end();
                    s.ext_lo = P.ext_off + (uint32_t) (std::upper_bound(vb, ve, (uint32_t) s.begin) - vb) - 1;
                    s.ext_hi = P.ext_off + (uint32_t) (std::lower_bound(vb, ve, (uint32_t) s.end) - vb);
                    segs.push_back(s);
                }
            }
            continue;
        }
        segments_of_part(P, pi.flags, pi.rows_est, q, target, items.size() == 1, segs);
        if (pi.flags & RK_SEGF_MASKED) mask_bytes += (uint64_t) P.n_ext * (8ull * W + 8);
    }
    return mask_bytes;
}

// Build the work list of an exact (probes == 0) batch into ws.h_plan and upload it.
static int build_exact_plan(rk_index *ix, Workspace &ws, size_t nq, const uint32_t *user_ids, const rk_scan_opts *o,
                            uint32_t *out_nseg)
{
    const int W = ix->mask_words;
    std::vector<std::vector<PlanItem>> per_q(nq);
    std::vector<const UserInfo *> uinfo(nq, nullptr);
    uint64_t total_rows = 0, algo_rows = 0;
    for (size_t q = 0; q < nq; q++) {
        auto it = ix->users.find(user_ids ? user_ids[q] : 0u);
        uinfo[q] = it == ix->users.end() ? nullptr : &it->second;
        int rc = plan_user(ix, uinfo[q], o, per_q[q]);
        if (rc < 0) return rc;
        for (auto &pi : per_q[q]) { total_rows += pi.rows_est; algo_rows += pi.rows_algo; }
    }
    ws.plan_single_partition = true;
    for (size_t q = 0; q < nq; q++) ws.plan_single_partition = ws.plan_single_partition && per_q[q].size() <= 1;
    // Segment sizing.  A segment costs a few microseconds of fixed work (pipeline refill, final
    // compaction + sort) and the dynamic schedule wastes about half a segment per CTA at the tail,
    // so with T = streaming time per CTA the best count per CTA is ~sqrt(T / 2o); o = 8 us (tune.seg_cost_x10; measured
    // 7.8 us per segment on the 1/8 stripe of C3: the threshold restarts at +inf, ~1500 appends and one compaction follow).
    const uint64_t row_bytes = (uint64_t) ix->dim_pad * 4;
    const uint64_t sm = (uint64_t) ix->sm_count;
    const double t_cta_us = (double) total_rows * (double) row_bytes / (double) sm / 44.0e3;   // ~44 GB/s per SM
    double per_cta = std::sqrt(t_cta_us / (2.0 * 0.1 * (double) ix->tune.seg_cost_x10));
    if (per_cta < 1.0) per_cta = 1.0;
    if (per_cta > 64.0) per_cta = 64.0;
    uint64_t target = (uint64_t) ((double) total_rows / ((double) sm * per_cta)) + 1;
    if (target < (uint64_t) ix->tile_rows) target = (uint64_t) ix->tile_rows;

    std::vector<RkSegment> segs;
    std::vector<uint32_t> qseg(nq + 1, 0);
    uint64_t mask_bytes = 0;
    for (size_t q = 0; q < nq; q++) {
        qseg[q] = (uint32_t) segs.size();
        mask_bytes += append_segments(ix, per_q[q], (uint32_t) q, target, segs);
    }
    qseg[nq] = (uint32_t) segs.size();
    const uint32_t nseg = (uint32_t) segs.size();

    // Execution order.  Queries that walk the same rows (same partitions, same permission mask)
    // are interleaved chunk by chunk: chunk c of every query of the group is claimed by different
    // CTAs at about the same time, so the rows one of them pulls from HBM are served to the others
    // from the 126 MB L2.  Candidate lists stay indexed by segment, so the select kernel is unaffected.
    std::vector<uint32_t> order;
    order.reserve(nseg);
    {
        std::map<std::vector<uint64_t>, std::vector<uint32_t>> groups;
        std::vector<const std::vector<uint32_t> *> group_seq;
        std::vector<uint64_t> key;
        for (size_t q = 0; q < nq; q++) {
            if (qseg[q + 1] == qseg[q]) continue;
            key.clear();
            bool any_masked = false;
            for (auto &pi : per_q[q]) {
                key.push_back(((uint64_t) pi.part << 32) | pi.flags);
                any_masked |= (pi.flags & RK_SEGF_MASKED) != 0;
            }
            key.push_back(~0ull);
            if (any_masked && uinfo[q])
                for (uint64_t w : uinfo[q]->mask) key.push_back(w);
            key.push_back((uint64_t) (qseg[q + 1] - qseg[q]));
            auto ins = groups.emplace(key, std::vector<uint32_t>());
            if (ins.second) group_seq.push_back(&ins.first->second);
            ins.first->second.push_back((uint32_t) q);
        }
        for (auto *g : group_seq) {
            const uint32_t n = qseg[(*g)[0] + 1] - qseg[(*g)[0]];
            for (uint32_t c = 0; c < n; c++)
                for (uint32_t q : *g) order.push_back(qseg[q] + c);
        }
    }

    // pinned staging: [segments][qseg][qmask][order]
    const size_t b_segs = (size_t) nseg * sizeof(RkSegment), b_qseg = (nq + 1) * 4, b_qmask = nq * (size_t) W * 8;
    const size_t b_order = (size_t) nseg * 4;
    ws.plan_idx ^= 1;
    const int pi_ = ws.plan_idx;
    if (ws.ev_plan_used[pi_]) RK_CUDA(cudaEventSynchronize(ws.ev_plan[pi_]));   // staging buffer free again?
    PinBuf &h_plan = ws.h_plan_ring[pi_];
    if (h_plan.ensure(b_segs + b_qseg + b_qmask + b_order + 128) != 0) return RK_ERR_CUDA;
    unsigned char *hp = h_plan.as<unsigned char>();
    if (nseg) memcpy(hp, segs.data(), b_segs);
    memcpy(hp + b_segs, qseg.data(), b_qseg);
    uint64_t *hm = reinterpret_cast<uint64_t *>(hp + b_segs + b_qseg);
    for (size_t q = 0; q < nq; q++)
        for (int w = 0; w < W; w++) hm[q * W + w] = uinfo[q] ? uinfo[q]->mask[w] : 0ull;
    if (ws.segs.ensure(std::max<size_t>(b_segs, 16)) != 0) return RK_ERR_CUDA;
    if (ws.qseg.ensure(b_qseg) != 0) return RK_ERR_CUDA;
    if (ws.qmask.ensure(std::max<size_t>(b_qmask, 16)) != 0) return RK_ERR_CUDA;
    if (ws.ctl.ensure(64) != 0) return RK_ERR_CUDA;
    if (nseg) RK_CUDA(cudaMemcpyAsync(ws.segs.p, hp, b_segs, cudaMemcpyHostToDevice, ws.stream));
    RK_CUDA(cudaMemcpyAsync(ws.qseg.p, hp + b_segs, b_qseg, cudaMemcpyHostToDevice, ws.stream));
    if (b_qmask) RK_CUDA(cudaMemcpyAsync(ws.qmask.p, hm, b_qmask, cudaMemcpyHostToDevice, ws.stream));
    uint32_t *hc = reinterpret_cast<uint32_t *>(hp + b_segs + b_qseg + b_qmask);
    hc = reinterpret_cast<uint32_t *>(((uintptr_t) hc + 15) & ~(uintptr_t) 15);
    hc[0] = nseg; hc[1] = 0;
    RK_CUDA(cudaMemcpyAsync(ws.ctl.p, hc, 8, cudaMemcpyHostToDevice, ws.stream));
    uint32_t *ho = hc + 4;
    if (ws.order.ensure(std::max<size_t>(b_order, 16)) != 0) return RK_ERR_CUDA;
    if (nseg) {
        memcpy(ho, order.data(), b_order);
        RK_CUDA(cudaMemcpyAsync(ws.order.p, ho, b_order, cudaMemcpyHostToDevice, ws.stream));
    }
    RK_CUDA(cudaEventRecord(ws.ev_plan[pi_], ws.stream));
    ws.ev_plan_used[pi_] = true;

    ws.segments = nseg;
    ws.seg_kinds = nseg ? 1u << RK_SEG_EXTENTS : 0u;
    ws.rows_scanned = total_rows;
    ws.algo_bytes = algo_rows * row_bytes + mask_bytes + (uint64_t) nq * 12ull * (uint64_t) o->k;
    *out_nseg = nseg;
    return 0;
}

int rk_ensure_cand(Workspace &ws, size_t nlists, int k)
{
    const size_t n = std::max<size_t>(nlists, 1) * (size_t) k;
    if (ws.cand_key.ensure(n * 4) != 0) return RK_ERR_CUDA;
    if (ws.cand_tid.ensure(n * 8) != 0) return RK_ERR_CUDA;
    if (ws.cand_row.ensure(n * 4) != 0) return RK_ERR_CUDA;
    if (ws.cand_count.ensure(std::max<size_t>(nlists, 1) * 4) != 0) return RK_ERR_CUDA;
    return 0;
}

void rk_fill_scan_params(const rk_index *ix, const Workspace &ws, RkScanParams &p, bool ivf_store, const float *dQ,
                         const rk_scan_opts *o, int k)
{
    memset(&p, 0, sizeof p);
    if (ivf_store) {
        p.store.x = ix->ivf.d_x.as<float>();
        p.store.tid = ix->ivf.d_tid.as<uint64_t>();
        p.store.rowdoc = ix->ivf.d_rowdoc.as<uint32_t>();
    } else {
        p.store.x = ix->d_x.as<float>();
        p.store.tid = ix->d_tid.as<uint64_t>();
        p.store.rowdoc = ix->d_rowdoc.as<uint32_t>();
    }
    p.store.rows = ix->nrows;
    p.doc_mask = ix->d_doc_mask.as<uint64_t>();
    p.qmask = ws.qmask.as<uint64_t>();
    p.mask_words = ix->mask_words;
    p.ext_row = ix->d_ext_row.as<uint32_t>();
    p.ext_vstart = ix->d_ext_vstart.as<uint32_t>();
    p.ext_mask = ix->d_ext_mask.as<uint64_t>();
    p.q = dQ;
    p.null_query = o->null_query;
    p.segs = ws.segs.as<RkSegment>();
    p.seg_order = ws.order.as<uint32_t>();
    p.nseg_ptr = ws.ctl.as<uint32_t>();
    p.seg_counter = ws.ctl.as<uint32_t>() + 1;
    p.cand_key = ws.cand_key.as<uint32_t>();
    p.cand_tid = ws.cand_tid.as<uint64_t>();
    p.cand_row = ws.cand_row.as<uint32_t>();
    p.cand_count = ws.cand_count.as<int32_t>();
    p.dim4 = ix->dim_pad / 4;
    p.row_bytes = ix->dim_pad * 4;
    p.tile_rows = ix->tile_rows;
    p.nstages = ix->nstages;
    p.k = k;
    p.metric = ix->metric;
    p.lb_on = ws.lb_on ? 1 : 0;
    p.lb_key = ws.lb_key;
    p.lb_tid = ws.lb_tid;
}

// ---------------------------------------------------------------------------------------------
// One query, one launch: the pgvector-shaped scan (ivfscan.c:339-389: one query per rescan).  The plan — work list and
// role mask — depends on the user's role set, the filter and the layout only, so it is built once per (role mask, filter,
// layout), kept in HBM, and a call uploads nothing but its query.  Returns 1 when the call was served, 0 when this path
// does not apply (a user with several partitions needs the dedup merge), < 0 on error.
static int search_single_fused(rk_index *ix, Workspace &ws, const float *dQ, const uint32_t *user_ids, const rk_scan_opts *o,
                               const RkSelectParams &out, int32_t *d_count, uint32_t *d_flag, uint32_t seq, const float *inline_q_host)
{
    const UserInfo *u = nullptr;
    if (user_ids) {
        auto it = ix->users.find(user_ids[0]);
        if (it != ix->users.end()) u = &it->second;
    }
    int layout = o->filter == RK_FILTER_PARTITION ? o->layout : 0;
    if (o->filter == RK_FILTER_PARTITION && layout == RK_LAYOUT_AUTO)
        layout = !ix->roleset_parts.empty() ? RK_LAYOUT_DYNAMIC : ix->want_combo_parts ? RK_LAYOUT_COMBINATION : RK_LAYOUT_ROLE;
    const uint64_t key = ((uint64_t) o->filter << 40) | ((uint64_t) layout << 36) | ((uint64_t) (o->keep_duplicates != 0) << 44) | (o->filter == RK_FILTER_NONE || !u ? 0ull : (uint64_t) u->mask_id + 1ull);
    auto pit = ws.fused_plans.find(key);
    if (pit == ws.fused_plans.end()) {
        std::vector<PlanItem> items;
        int rc = plan_user(ix, u, o, items);
        if (rc < 0) return rc;
        if (items.size() > 1 && !o->keep_duplicates) {                  // duplicates across partitions: the dedup merge
            FusedPlan na;
            na.nseg = 0xFFFFFFFFu;                                      // remembered: this role set does not take the fused path
            ws.fused_plans.emplace(key, std::move(na));
            return 0;
        }
        if (ws.fused_plans.size() >= 4096) {
            RK_CUDA(cudaStreamSynchronize(ws.stream));                  // a launch may still read a cached plan
            for (auto &kv : ws.fused_plans) kv.second.buf.release();
            ws.fused_plans.clear();
        }
        uint64_t total_rows = 0, algo_rows = 0;
        for (auto &pi : items) { total_rows += pi.rows_est; algo_rows += pi.rows_algo; }
        // two equal segments per CTA: every boundary stalls a CTA's issue for ~2 us (measured: 1-2 per CTA is best from
        // 50 MB to 5 GB; with equal segments the CTAs drain together, so there is little for a finer split to balance)
        const uint64_t row_bytes = (uint64_t) ix->dim_pad * 4, sm = (uint64_t) ix->sm_count;
        double per_cta = 2.0;
        if (ix->tune.single_segs > 0) per_cta = (double) ix->tune.single_segs;
        // a whole number of equal segments per CTA: with one partition every CTA then streams the same number of rows and
        // the grid drains together (one segment more on some CTAs is a tail as long as a segment)
        uint64_t target = (total_rows + (uint64_t) (sm * per_cta) - 1) / (uint64_t) (sm * per_cta);
        target = std::max<uint64_t>(target, (uint64_t) ix->tile_rows);
        std::vector<RkSegment> segs;
        uint64_t mask_bytes = 0;
        bool any_masked = false;
        for (auto &pi : items) any_masked |= (pi.flags & RK_SEGF_MASKED) != 0;
        // A plan that would test role masks per extent (the bitmask filter walks the table of ALL documents; impure dynamic
        // partitions) gets a PRIVATE extent table instead: only the documents this role set may see, physically adjacent
        // ones fused.  The plan is cached, so the mask test runs once per role set on the host and the kernel's producer
        // resolves a few thousand visible extents instead of testing every document of the corpus per query.
        std::vector<uint32_t> x_row, x_vstart;
        if (any_masked) {
            const int W = ix->mask_words;
            for (auto &pi : items) {
                const PartInfo &P = ix->parts[pi.part];
                const bool masked = (pi.flags & RK_SEGF_MASKED) != 0;
                PartInfo T;
                T.ext_off = (uint32_t) x_row.size();
                uint64_t v = 0, run_end = ~0ull;
                for (uint32_t d : P.docs) {
                    if (masked) {
                        uint64_t acc = 0;
                        for (int w = 0; w < W; w++) acc |= ix->doc_mask[(size_t) d * W + w] & (u ? u->mask[(size_t) w] : 0ull);
                        if (!acc) continue;
                    }
                    const uint64_t r0 = ix->doc_row_start[d], r1 = ix->doc_row_start[d + 1];
                    if (r1 == r0) continue;
                    if (r0 != run_end) { x_row.push_back((uint32_t) r0); x_vstart.push_back((uint32_t) v); T.vstart.push_back((uint32_t) v); }
                    v += r1 - r0;
                    run_end = r1;
                }
                T.vstart.push_back((uint32_t) v);
                T.n_ext = (uint32_t) T.vstart.size() - 1;
                T.rows = v;
                x_row.push_back(0u); x_vstart.push_back((uint32_t) v);        // sentinel, as in the index-wide tables
                if (v) segments_of_part(T, 0u, v, 0u, target, items.size() == 1, segs);
                mask_bytes += (uint64_t) T.n_ext * 8ull;
            }
        } else {
            mask_bytes = append_segments(ix, items, 0u, target, segs);
        }
        FusedPlan fp;
        fp.nseg = (uint32_t) segs.size();
        const int W = ix->mask_words;
        const size_t b_segs = segs.size() * sizeof(RkSegment);
        fp.off_qmask = (b_segs + 15) & ~(size_t) 15;
        fp.off_xrow = (fp.off_qmask + (size_t) W * 8 + 15) & ~(size_t) 15;
        fp.off_xvstart = fp.off_xrow + ((x_row.size() * 4 + 15) & ~(size_t) 15);
        fp.private_extents = any_masked;
        std::vector<unsigned char> host(fp.off_xvstart + x_vstart.size() * 4 + 16, 0);
        if (b_segs) memcpy(host.data(), segs.data(), b_segs);
        for (int w = 0; w < W; w++) reinterpret_cast<uint64_t *>(host.data() + fp.off_qmask)[w] = u ? u->mask[(size_t) w] : 0ull;
        if (!x_row.empty()) {
            memcpy(host.data() + fp.off_xrow, x_row.data(), x_row.size() * 4);
            memcpy(host.data() + fp.off_xvstart, x_vstart.data(), x_vstart.size() * 4);
        }
        if (fp.buf.ensure(host.size()) != 0) return RK_ERR_CUDA;
        RK_CUDA(cudaMemcpyAsync(fp.buf.p, host.data(), host.size(), cudaMemcpyHostToDevice, ws.stream));
        RK_CUDA(cudaStreamSynchronize(ws.stream));                      // `host` goes out of scope
        fp.rows_scanned = total_rows;
        fp.algo_bytes = algo_rows * row_bytes + mask_bytes;
        pit = ws.fused_plans.emplace(key, std::move(fp)).first;
    }
    FusedPlan &fp = pit->second;
    if (fp.nseg == 0xFFFFFFFFu) return 0;
    const int k = o->k;
    RkScanParams p;
    rk_fill_scan_params(ix, ws, p, false, dQ, o, k);
    const int grid = (int) std::max<uint32_t>(1u, std::min<uint32_t>((uint32_t) ix->sm_count, fp.nseg));
    // what a CTA may hand over: its k best plus ties (the scan kernel's own limit before it falls back to the exact sort)
    int ckpt = 512 / ix->tile_rows;
    ckpt = std::max(1, std::min(16, ckpt));
    const int trigger = 2048 - ckpt * ix->tile_rows;
    p.fz_cap = k + (trigger - k) / 2;
    // rows a CTA vouches for: the bound is valid once r x (CTAs reported) >= k; about three quarters of the grid should do
    p.fz_r = std::max(1, (4 * k + 3 * grid - 1) / (3 * grid));
    p.fz_seed = 0;
    if (ix->tune.single_seed > 0 && fp.rows_scanned / (uint64_t) grid >= (uint64_t) ix->tune.single_seed_min) p.fz_seed = std::max(ix->tune.single_seed, 2 * k) - 1;
    if (ws.fz_ctl.p == nullptr) {
        if (ws.fz_ctl.ensure(64) != 0) return RK_ERR_CUDA;
        RK_CUDA(cudaMemsetAsync(ws.fz_ctl.p, 0, 64, ws.stream));
    }
    const size_t flat = (size_t) ix->sm_count * (size_t) p.fz_cap;
    if (ws.fz_key.ensure(flat * 4) != 0 || ws.fz_row.ensure(flat * 4) != 0) return RK_ERR_CUDA;
    p.fused = 1;
    p.segs = fp.buf.as<RkSegment>();
    p.seg_order = nullptr;
    p.nseg_ptr = nullptr;
    p.nseg_val = fp.nseg;
    p.qmask = reinterpret_cast<const uint64_t *>(fp.buf.as<unsigned char>() + fp.off_qmask);
    if (fp.private_extents) {
        p.ext_row = reinterpret_cast<const uint32_t *>(fp.buf.as<unsigned char>() + fp.off_xrow);
        p.ext_vstart = reinterpret_cast<const uint32_t *>(fp.buf.as<unsigned char>() + fp.off_xvstart);
        p.ext_mask = nullptr;
    }
    p.seg_counter = ws.fz_ctl.as<uint32_t>() + 2;
    p.fz_ctl = ws.fz_ctl.as<uint32_t>();
    p.fz_key = ws.fz_key.as<uint32_t>();
    p.fz_row = ws.fz_row.as<uint32_t>();
    p.fz_doc_ids = ix->d_doc_ids.as<uint32_t>();
    p.fz_out_tid = out.out_tid; p.fz_out_doc = out.out_doc; p.fz_out_dist = out.out_dist; p.fz_out_key = out.out_key;
    p.fz_out_count = d_count;
    p.fz_flag = d_flag;
    p.fz_seq = seq;
    if (ix->tune.single_debug) {
        const size_t db = (size_t) (ix->sm_count + 1) * 64 + 4096;
        if (!ws.h_dbg) RK_CUDA(cudaHostAlloc((void **) &ws.h_dbg, db, cudaHostAllocDefault));
        if (ws.d_dbg.ensure(db) != 0) return RK_ERR_CUDA;
        RK_CUDA(cudaMemsetAsync(ws.d_dbg.p, 0, db, ws.stream));
        p.fz_dbg = ws.d_dbg.as<unsigned long long>();
    }
    p.cand_key = nullptr; p.cand_tid = nullptr; p.cand_row = nullptr; p.cand_count = nullptr;
    const int slot = (int) (ws.ring_n % RK_TIMING_RING);
    if (ws.timing) RK_CUDA(cudaEventRecord(ws.ring_a[slot], ws.stream));
    RK_CUDA(rk_launch_scan(p, ix->device, ws.stream, grid, inline_q_host));
    if (ws.timing) { RK_CUDA(cudaEventRecord(ws.ring_b[slot], ws.stream)); ws.timed_scan = true; ws.ring_n++; }
    ws.launches = 1;
    ws.single_fused = true;
    ws.segments = fp.nseg;
    ws.seg_kinds = fp.nseg ? 1u << RK_SEG_EXTENTS : 0u;
    ws.rows_scanned = fp.rows_scanned;
    ws.algo_bytes = fp.algo_bytes + 12ull * (uint64_t) k;
    ws.plan_single_partition = true;
    ix->total_launches += 1;
    return 1;
}

// One batch on device: dQ [nq][dim_pad].  `out` carries the destination pointers (device).
static int search_device(rk_index *ix, Workspace &ws, size_t nq, const float *dQ, const uint32_t *user_ids,
                         const rk_scan_opts *o, int probe_offset, RkSelectParams out);

static int search_device(rk_index *ix, Workspace &ws, size_t nq, const float *dQ, const uint32_t *user_ids,
                         const rk_scan_opts *o, int probe_offset, RkSelectParams out)
{
    RK_CUDA(cudaSetDevice(ix->device));
    ws.launches = 0;
    ws.seg_kinds = 0;
    ws.single_fused = false;
    ws.timed_scan = ws.timed_select = ws.timed_probe = false;
    if (&ws == &ix->ws) ix->tensor.last_plan = 0;
    if (nq == 0) return RK_OK;
    if (o->probes > 0) return rk_ivf_search(ix, ws, nq, dQ, user_ids, o, probe_offset, out);
    // RK_ENGINE_AUTO: a batch of 32 or more exact scans goes to the tensor engine (every row a group of
    // queries shares is read once and scored on the tensor cores; measured crossover on the C3 shape is
    // ~16 queries), anything smaller — and everything the tensor engine does not cover — is scanned.
    // The dynamic layout stays on the scan: its partitions need not cover a user's whole visible set.
    const bool auto_tensor = o->engine == RK_ENGINE_AUTO && nq >= 32 && &ws == &ix->ws && rk_tensor_supported(ix, o) &&
                             !(o->filter == RK_FILTER_PARTITION && (o->layout == RK_LAYOUT_DYNAMIC || (o->layout == RK_LAYOUT_AUTO && !ix->roleset_parts.empty())));
    if (o->engine == RK_ENGINE_TENSOR || auto_tensor) {
        if (!rk_tensor_supported(ix, o)) return rk_fail(RK_ERR_UNSUPPORTED, "RK_ENGINE_TENSOR needs an exact scan (probes = 0), k <= 384, no keep_duplicates / NULL query");
        if (&ws != &ix->ws) return rk_fail(RK_ERR_UNSUPPORTED, "RK_ENGINE_TENSOR is a batch engine (rk_search_batch*), not available on scan handles");
        std::vector<uint32_t> redo;
        int rc = rk_tensor_search(ix, ws, nq, dQ, user_ids, o, out, redo);
        if (rc < 0) return rc;
        ix->total_launches += ws.launches;
        if (redo.empty()) return RK_OK;
        // unproven answers go through the exact scan; the whole batch if that is most of it
        rk_scan_opts o2 = *o;
        o2.engine = RK_ENGINE_SCAN;
        const uint64_t tl = ws.launches;
        const int plan_used = ix->tensor.last_plan;
        const bool timing_was = ws.timing, ts_scan = ws.timed_scan, ts_sel = ws.timed_select;
        ws.timing = false;               // the timing ring keeps the scoring passes of this call, not the re-runs' scans
        if (redo.size() * 4 > nq) {
            rc = search_device(ix, ws, nq, dQ, user_ids, &o2, 0, out);
        } else if (redo.size() == 1) {
            const uint32_t q = redo[0];
            RkSelectParams o1 = out;
            const size_t ob = (size_t) q * o->k;
            if (o1.out_tid) o1.out_tid += ob;
            if (o1.out_doc) o1.out_doc += ob;
            if (o1.out_dist) o1.out_dist += ob;
            if (o1.out_key) o1.out_key += ob;
            if (o1.out_row) o1.out_row += ob;
            if (o1.out_count) o1.out_count += q;
            rc = search_device(ix, ws, 1, dQ + (size_t) q * ix->dim_pad, user_ids ? user_ids + q : nullptr, &o2, 0, o1);
        } else {
            // the refused queries as ONE sub-batch on the scan engine (one scan launch + one select, not a pair per query):
            // gather their vectors, search into a scratch block, scatter the rows of every output array to their places
            const size_t nr = redo.size(), k = (size_t) o->k, rb = (size_t) ix->dim_pad * 4;
            const size_t o_idx = 0, o_q = (nr * 4 + 255) & ~(size_t) 255, o_tid = o_q + ((nr * rb + 255) & ~(size_t) 255),
                         o_dist = o_tid + nr * k * 8, o_doc = o_dist + nr * k * 8, o_key = o_doc + nr * k * 4, o_row = o_key + nr * k * 4,
                         o_cnt = o_row + nr * k * 4, total = o_cnt + nr * 4 + 256;
            DevBuf &rb_ = ix->tensor.d_redo;
            if (rb_.ensure(total) != 0) return RK_ERR_CUDA;
            unsigned char *base = rb_.as<unsigned char>();
            std::vector<uint32_t> sub_users(nr);
            for (size_t i = 0; i < nr; i++) sub_users[i] = user_ids ? user_ids[redo[i]] : 0u;
            // (pageable source: the copy is staged by the runtime before the call returns)
            RK_CUDA(cudaMemcpyAsync(base + o_idx, redo.data(), nr * 4, cudaMemcpyHostToDevice, ws.stream));
            const uint32_t *d_idx = reinterpret_cast<const uint32_t *>(base + o_idx);
            RK_CUDA(rk_launch_gather_rows(dQ, reinterpret_cast<float *>(base + o_q), d_idx, nr, ix->dim_pad, ws.stream));
            RkSelectParams o1 = out;
            o1.out_tid = out.out_tid ? reinterpret_cast<uint64_t *>(base + o_tid) : nullptr;
            o1.out_dist = out.out_dist ? reinterpret_cast<double *>(base + o_dist) : nullptr;
            o1.out_doc = out.out_doc ? reinterpret_cast<uint32_t *>(base + o_doc) : nullptr;
            o1.out_key = out.out_key ? reinterpret_cast<float *>(base + o_key) : nullptr;
            o1.out_row = out.out_row ? reinterpret_cast<uint32_t *>(base + o_row) : nullptr;
            o1.out_count = out.out_count ? reinterpret_cast<int32_t *>(base + o_cnt) : nullptr;
            rc = search_device(ix, ws, nr, reinterpret_cast<const float *>(base + o_q), user_ids ? sub_users.data() : nullptr, &o2, 0, o1);
            if (rc >= 0) {
                if (out.out_tid) RK_CUDA(rk_launch_scatter_chunks(o1.out_tid, out.out_tid, d_idx, nr, (int) (2 * k), ws.stream));
                if (out.out_dist) RK_CUDA(rk_launch_scatter_chunks(o1.out_dist, out.out_dist, d_idx, nr, (int) (2 * k), ws.stream));
                if (out.out_doc) RK_CUDA(rk_launch_scatter_chunks(o1.out_doc, out.out_doc, d_idx, nr, (int) k, ws.stream));
                if (out.out_key) RK_CUDA(rk_launch_scatter_chunks(o1.out_key, out.out_key, d_idx, nr, (int) k, ws.stream));
                if (out.out_row) RK_CUDA(rk_launch_scatter_chunks(o1.out_row, out.out_row, d_idx, nr, (int) k, ws.stream));
                if (out.out_count) RK_CUDA(rk_launch_scatter_chunks(o1.out_count, out.out_count, d_idx, nr, 1, ws.stream));
                ws.launches += 7;
            }
        }
        ws.launches += tl;
        ws.timing = timing_was; ws.timed_scan = ts_scan; ws.timed_select = ts_sel; ws.timed_probe = false;
        ix->tensor.last_plan = plan_used;
        return rc;
    }

    uint32_t nseg = 0;
    int rc = build_exact_plan(ix, ws, nq, user_ids, o, &nseg);
    if (rc < 0) return rc;
    if (rk_ensure_cand(ws, nseg, o->k) != 0) return RK_ERR_CUDA;
    if (nseg) {
        RkScanParams p;
        rk_fill_scan_params(ix, ws, p, false, dQ, o, o->k);
        const int slot = (int) (ws.ring_n % RK_TIMING_RING);
        if (ws.timing) RK_CUDA(cudaEventRecord(ws.ring_a[slot], ws.stream));
        const int grid = (int) std::min<uint32_t>((uint32_t) ix->sm_count, nseg);
        RK_CUDA(rk_launch_scan(p, ix->device, ws.stream, grid));
        if (ws.timing) { RK_CUDA(cudaEventRecord(ws.ring_b[slot], ws.stream)); ws.timed_scan = true; ws.ring_n++; }
        ws.launches++;
    }
    out.cand_key = ws.cand_key.as<uint32_t>();
    out.cand_tid = ws.cand_tid.as<uint64_t>();
    out.cand_row = ws.cand_row.as<uint32_t>();
    out.cand_count = ws.cand_count.as<int32_t>();
    out.qseg = ws.qseg.as<uint32_t>();
    out.lists_per_query = 0;
    out.list_stride = o->k;
    out.k = o->k;
    out.metric = ix->metric;
    // duplicates can only come from a user's several partitions (prefilter_role.py:174-189): a plan in which every query
    // walks one partition needs no dedup pass, which lets the select cut to the k smallest keys before it sorts
    out.dedup = (o->keep_duplicates || ws.plan_single_partition) ? 0 : 1;
    out.null_query = o->null_query;
    out.key_is_float = 0;
    out.rowdoc = ix->d_rowdoc.as<uint32_t>();
    out.doc_ids = ix->d_doc_ids.as<uint32_t>();
    if (ws.timing) RK_CUDA(cudaEventRecord(ws.ev[2], ws.stream));
    RK_CUDA(rk_launch_select(out, (int) nq, ws.stream));
    if (ws.timing) { RK_CUDA(cudaEventRecord(ws.ev[3], ws.stream)); ws.timed_select = true; }
    ws.launches++;
    ix->total_launches += ws.launches;
    return RK_OK;
}

static int collect_timing(rk_index *ix, Workspace &ws)
{
    if (!ws.timing) return 0;
    if (ws.timed_scan && ws.ring_n) {
        const int slot = (int) ((ws.ring_n - 1) % RK_TIMING_RING);
        RK_CUDA(cudaEventSynchronize(ws.ring_b[slot]));
        RK_CUDA(cudaEventElapsedTime(&ix->last_scan_ms, ws.ring_a[slot], ws.ring_b[slot]));
    }
    if (ws.timed_select) { RK_CUDA(cudaEventSynchronize(ws.ev[3])); RK_CUDA(cudaEventElapsedTime(&ix->last_select_ms, ws.ev[2], ws.ev[3])); }
    if (ws.timed_probe) { RK_CUDA(cudaEventSynchronize(ws.ev[5])); RK_CUDA(cudaEventElapsedTime(&ix->last_probe_ms, ws.ev[4], ws.ev[5])); }
    return 0;
}

// Copy host queries into ws.dq, padded to dim_pad.
static int stage_queries(rk_index *ix, Workspace &ws, size_t nq, const float *Q)
{
    const size_t rb = (size_t) ix->dim_pad * 4;
    if (ws.dq.ensure(std::max<size_t>(nq * rb, 16)) != 0) return RK_ERR_CUDA;
    if (ix->dim == ix->dim_pad) {
        RK_CUDA(cudaMemcpyAsync(ws.dq.p, Q, nq * rb, cudaMemcpyHostToDevice, ws.stream));
    } else {
        RK_CUDA(cudaMemsetAsync(ws.dq.p, 0, nq * rb, ws.stream));
        RK_CUDA(cudaMemcpy2DAsync(ws.dq.p, rb, Q, (size_t) ix->dim * 4, (size_t) ix->dim * 4, nq, cudaMemcpyHostToDevice, ws.stream));
    }
    return 0;
}

// The single-query host path.  Zero-copy (default): the query and the packed outputs live in ONE mapped pinned host
// buffer that the kernel reads and writes directly; the call is a memcpy of the query, one launch and a poll of the
// completion flag the kernel stores last — no cudaMemcpy, no stream synchronisation.  With single_zero_copy = 0: pinned
// staging, one host->device copy (the query), one launch, one device->host copy (count | tid | dist | doc | key), sync.
// Returns 1 = served, 0 = not applicable, < 0 = error.
static int host_search_single(rk_index *ix, Workspace &ws, const float *Q, const uint32_t *user_ids, const rk_scan_opts *o,
                              uint64_t *out_tid, uint32_t *out_doc, double *out_dist, int32_t *out_count, float *out_key)
{
    const int k = o->k;
    const size_t rb = (size_t) ix->dim_pad * 4;
    const size_t o_tid = 16, o_dist = o_tid + (size_t) k * 8, o_doc = o_dist + (size_t) k * 8, o_key = o_doc + (size_t) k * 4;
    const size_t pack = (o_key + (size_t) k * 4 + 15) & ~(size_t) 15;
    if (!o->null_query && !rk_all_finite(Q, (size_t) ix->dim)) return rk_fail(RK_ERR_NONFINITE, "NaN or infinite value not allowed in vector");
    const bool zc = ix->tune.single_zero_copy != 0;
    // the query rides in the kernel parameters when it fits (no host->device copy at all), else one copy from pinned staging
    const bool inline_q = ix->dim_pad <= RK_INLINE_Q;
    unsigned char *hout, *dp;
    if (zc) {
        // mapped pinned buffer: [flag (64 B) | packed outputs]
        const size_t need = 64 + pack;
        if (need > ws.zc_cap) {
            RK_CUDA(cudaStreamSynchronize(ws.stream));
            if (ws.h_zc) cudaFreeHost(ws.h_zc);
            ws.h_zc = ws.d_zc = nullptr;
            const size_t want = std::max(need, ws.zc_cap * 2);
            RK_CUDA(cudaHostAlloc((void **) &ws.h_zc, want, cudaHostAllocMapped | cudaHostAllocPortable));
            RK_CUDA(cudaHostGetDevicePointer((void **) &ws.d_zc, ws.h_zc, 0));
            memset(ws.h_zc, 0, want);
            ws.zc_cap = want;
            ws.zc_seq = 0;
        }
        hout = ws.h_zc + 64;
        dp = ws.d_zc + 64;
    } else {
        if (ws.fz_pack.ensure(pack) != 0 || ws.h_fz.ensure(rb + pack) != 0) return RK_ERR_CUDA;
        hout = ws.h_fz.as<unsigned char>() + rb;
        dp = ws.fz_pack.as<unsigned char>();
    }
    float qpad[RK_INLINE_Q];
    const float *q_host = Q;
    if (inline_q) {
        if (ix->dim != ix->dim_pad) {
            memcpy(qpad, Q, (size_t) ix->dim * 4);
            for (int i = ix->dim; i < ix->dim_pad; i++) qpad[i] = 0.f;
            q_host = qpad;
        }
    } else {
        if (ws.dq.ensure(std::max<size_t>(rb, 16)) != 0 || ws.h_fz.ensure(rb + pack) != 0) return RK_ERR_CUDA;
        unsigned char *hin = ws.h_fz.as<unsigned char>();
        memcpy(hin, Q, (size_t) ix->dim * 4);
        if (ix->dim != ix->dim_pad) memset(hin + (size_t) ix->dim * 4, 0, rb - (size_t) ix->dim * 4);
        RK_CUDA(cudaMemcpyAsync(ws.dq.p, hin, rb, cudaMemcpyHostToDevice, ws.stream));
    }
    const float *dq = inline_q ? nullptr : ws.dq.as<float>();
    RkSelectParams out;
    memset(&out, 0, sizeof out);
    out.out_tid = reinterpret_cast<uint64_t *>(dp + o_tid);
    out.out_dist = reinterpret_cast<double *>(dp + o_dist);
    out.out_doc = out_doc ? reinterpret_cast<uint32_t *>(dp + o_doc) : nullptr;
    out.out_key = out_key ? reinterpret_cast<float *>(dp + o_key) : nullptr;
    ws.launches = 0;
    ws.seg_kinds = 0;
    ws.single_fused = false;
    ws.timed_scan = ws.timed_select = ws.timed_probe = false;
    if (&ws == &ix->ws) ix->tensor.last_plan = 0;
    uint32_t seq = 0;
    uint32_t *dflag = nullptr;
    if (zc) { seq = ++ws.zc_seq; if (seq == 0) seq = ++ws.zc_seq; dflag = reinterpret_cast<uint32_t *>(ws.d_zc); }
    int rc = search_single_fused(ix, ws, dq, user_ids, o, out, reinterpret_cast<int32_t *>(dp), dflag, seq, inline_q ? q_host : nullptr);
    if (rc <= 0) return rc;                                 // 0: the general path restages the query on the same stream
    if (zc) {
        // poll the flag the kernel stores after its outputs are fenced system-wide; every so often make sure the stream
        // has not died under us (a failed launch never writes the flag)
        volatile uint32_t *hflag = reinterpret_cast<volatile uint32_t *>(ws.h_zc);
        for (uint64_t spin = 1; *hflag != seq; spin++) {
            if ((spin & 0xFFFF) == 0) {
                const cudaError_t e = cudaStreamQuery(ws.stream);
                if (e != cudaSuccess && e != cudaErrorNotReady) return rk_fail(RK_ERR_CUDA, "single-query launch failed: %s", cudaGetErrorString(e));
                if (e == cudaSuccess && *hflag != seq) return rk_fail(RK_ERR_CUDA, "single-query launch finished without signalling");
            }
            __builtin_ia32_pause();
        }
        std::atomic_thread_fence(std::memory_order_acquire);
    } else {
        RK_CUDA(cudaMemcpyAsync(hout, dp, pack, cudaMemcpyDeviceToHost, ws.stream));
        RK_CUDA(cudaStreamSynchronize(ws.stream));
    }
    if (ix->tune.single_debug && ws.h_dbg) {
        // phase timeline of the launch, microseconds after the earliest CTA start: median / max over the CTAs, then the merge
        cudaMemcpyAsync(ws.h_dbg, ws.d_dbg.p, (size_t) (ix->sm_count + 1) * 64 + 4096, cudaMemcpyDeviceToHost, ws.stream);
        cudaStreamSynchronize(ws.stream);
        const int G = (int) std::max<uint32_t>(1u, std::min<uint32_t>((uint32_t) ix->sm_count, (uint32_t) ws.segments));
        unsigned long long t0 = ~0ull;
        for (int b = 0; b < G; b++) if (ws.h_dbg[b * 8]) t0 = std::min(t0, ws.h_dbg[b * 8]);
        const char *names[5] = {"start", "first tile", "scan done", "local select", "handed over"};
        for (int s_ = 0; s_ < 5; s_++) {
            std::vector<double> v;
            for (int b = 0; b < G; b++) if (ws.h_dbg[b * 8 + s_]) v.push_back((double) (ws.h_dbg[b * 8 + s_] - t0) * 1e-3);
            if (v.empty()) continue;
            std::sort(v.begin(), v.end());
            fprintf(stderr, "  cta %-13s min %7.2f med %7.2f max %7.2f us\n", names[s_], v.front(), v[v.size() / 2], v.back());
        }
        const unsigned long long *m_ = ws.h_dbg + (size_t) G * 8;
        fprintf(stderr, "  merge (cta %llu, N = %llu): start %7.2f keys %7.2f kth %7.2f gathered %7.2f written %7.2f us\n", m_[7], m_[6],
                (double) (m_[0] - t0) * 1e-3, (double) (m_[1] - t0) * 1e-3, (double) (m_[2] - t0) * 1e-3, (double) (m_[3] - t0) * 1e-3,
                (double) (m_[4] - t0) * 1e-3);
        if (ix->tune.single_debug > 1) {
            const unsigned long long *tt = ws.h_dbg + (size_t) (G + 1) * 8;
            fprintf(stderr, "  cta 0 tile arrivals:");
            for (int i = 0; i < 256 && tt[i]; i++) fprintf(stderr, " %.2f", (double) (tt[i] - t0) * 1e-3);
            fprintf(stderr, "\n  cta 0 producer (loop top / segment start / tile open / first step out):");
            for (int i = 0; i < 256 && tt[256 + i]; i++) fprintf(stderr, "%s%.2f", i % 4 ? " " : " | ", (double) (tt[256 + i] - t0) * 1e-3);
            fprintf(stderr, "\n");
        }
    }
    const int32_t cnt = *reinterpret_cast<const volatile int32_t *>(hout);
    if (cnt < 0) return 0;                                  // more ties at the k-th key than the fused merge holds: general path
    memcpy(out_tid, hout + o_tid, (size_t) k * 8);
    memcpy(out_dist, hout + o_dist, (size_t) k * 8);
    if (out_doc) memcpy(out_doc, hout + o_doc, (size_t) k * 4);
    if (out_key) memcpy(out_key, hout + o_key, (size_t) k * 4);
    *out_count = cnt;
    rc = collect_timing(ix, ws);
    return rc < 0 ? rc : 1;
}

static int host_search(rk_index *ix, Workspace &ws, size_t nq, const float *Q, const uint32_t *user_ids,
                       const rk_scan_opts *o, int probe_offset, uint64_t *out_tid, uint32_t *out_doc, double *out_dist,
                       int32_t *out_count, float *out_key = nullptr)
{
    RK_CUDA(cudaSetDevice(ix->device));
    int rc;
    if (nq == 1 && o->probes == 0 && o->engine != RK_ENGINE_TENSOR && ix->tune.single_fused) {
        rc = host_search_single(ix, ws, Q, user_ids, o, out_tid, out_doc, out_dist, out_count, out_key);
        if (rc != 0) return rc < 0 ? rc : RK_OK;          // served (or failed); 0 = this path does not apply
    }
    // the upload is started first and the host validates the same bytes while the copy engine moves them (nothing has
    // been launched on them yet when a bad element makes the call fail)
    rc = stage_queries(ix, ws, nq, Q);
    if (rc < 0) return rc;
    if (!o->null_query && !rk_all_finite(Q, nq * (size_t) ix->dim)) {
        cudaStreamSynchronize(ws.stream);
        return rk_fail(RK_ERR_NONFINITE, "NaN or infinite value not allowed in vector");
    }
    const size_t nk = nq * (size_t) o->k;
    if (ws.out_tid.ensure(std::max<size_t>(nk * 8, 16)) != 0 || ws.out_doc.ensure(std::max<size_t>(nk * 4, 16)) != 0 ||
        ws.out_dist.ensure(std::max<size_t>(nk * 8, 16)) != 0 || ws.out_count.ensure(std::max<size_t>(nq * 4, 16)) != 0)
        return RK_ERR_CUDA;
    if (out_key && ws.ex_key.ensure(std::max<size_t>(nk * 4, 16)) != 0) return RK_ERR_CUDA;
    RkSelectParams out;
    memset(&out, 0, sizeof out);
    out.out_tid = ws.out_tid.as<uint64_t>();
    out.out_doc = out_doc ? ws.out_doc.as<uint32_t>() : nullptr;
    out.out_dist = ws.out_dist.as<double>();
    out.out_count = ws.out_count.as<int32_t>();
    out.out_key = out_key ? ws.ex_key.as<float>() : nullptr;
    rc = search_device(ix, ws, nq, ws.dq.as<float>(), user_ids, o, probe_offset, out);
    if (rc < 0) return rc;
    RK_CUDA(cudaMemcpyAsync(out_tid, ws.out_tid.p, nk * 8, cudaMemcpyDeviceToHost, ws.stream));
    if (out_doc) RK_CUDA(cudaMemcpyAsync(out_doc, ws.out_doc.p, nk * 4, cudaMemcpyDeviceToHost, ws.stream));
    RK_CUDA(cudaMemcpyAsync(out_dist, ws.out_dist.p, nk * 8, cudaMemcpyDeviceToHost, ws.stream));
    RK_CUDA(cudaMemcpyAsync(out_count, ws.out_count.p, nq * 4, cudaMemcpyDeviceToHost, ws.stream));
    if (out_key) RK_CUDA(cudaMemcpyAsync(out_key, ws.ex_key.p, nk * 4, cudaMemcpyDeviceToHost, ws.stream));
    RK_CUDA(cudaStreamSynchronize(ws.stream));
    return collect_timing(ix, ws);
}

extern "C" int rk_search_batch(rk_index *ix, size_t nq, const float *Q, const uint32_t *user_ids, const rk_scan_opts *o,
                               uint64_t *out_tid, uint32_t *out_doc, double *out_dist, int32_t *out_count)
{
    if (!ix) return rk_fail(RK_ERR_INVALID, "NULL index");
    if (!ix->finalized) return rk_fail(RK_ERR_STATE, "rk_search_batch: call rk_finalize first");
    int rc = validate_opts(ix, o);
    if (rc < 0) return rc;
    if (nq == 0) return RK_OK;
    if (!Q || !out_tid || !out_dist || !out_count) return rk_fail(RK_ERR_INVALID, "rk_search_batch: NULL buffer");
    if (!user_ids && o->filter != RK_FILTER_NONE) return rk_fail(RK_ERR_INVALID, "rk_search_batch: user_ids is NULL");
    std::lock_guard<std::mutex> g(ix->batch_mu);      // the batch entry points share ix->ws
    return host_search(ix, ix->ws, nq, Q, user_ids, o, 0, out_tid, out_doc, out_dist, out_count);
}

// Device-resident queries cannot be validated by the host without a sync: a small kernel raises a flag that the next
// rk_synchronize reports (vector_in's NaN / Inf check, vector.c:102-113).
static int check_device_queries(rk_index *ix, Workspace &ws, size_t nq, const float *dQ, const rk_scan_opts *o)
{
    if (o->null_query) return 0;
    RK_CUDA(cudaSetDevice(ix->device));
    RK_CUDA(rk_launch_flag_nonfinite(dQ, nq * (size_t) ix->dim_pad, ws.qflag.as<uint32_t>(), ws.stream));
    ws.qflag_armed = true;
    ix->total_launches += 1;
    return 0;
}

extern "C" int rk_search_batch_device(rk_index *ix, size_t nq, const float *dQ, const uint32_t *user_ids,
                                      const rk_scan_opts *o, uint64_t *d_out_tid, uint32_t *d_out_doc,
                                      double *d_out_dist, int32_t *d_out_count)
{
    if (!ix) return rk_fail(RK_ERR_INVALID, "NULL index");
    if (!ix->finalized) return rk_fail(RK_ERR_STATE, "rk_search_batch_device: call rk_finalize first");
    int rc = validate_opts(ix, o);
    if (rc < 0) return rc;
    if (nq == 0) return RK_OK;
    if (!dQ) return rk_fail(RK_ERR_INVALID, "rk_search_batch_device: NULL query pointer");
    if (ix->dim != ix->dim_pad) return rk_fail(RK_ERR_UNSUPPORTED, "device queries need dim %% 4 == 0 (dim = %d)", ix->dim);
    RkSelectParams out;
    memset(&out, 0, sizeof out);
    out.out_tid = d_out_tid; out.out_doc = d_out_doc; out.out_dist = d_out_dist; out.out_count = d_out_count;
    std::lock_guard<std::mutex> g(ix->batch_mu);
    rc = check_device_queries(ix, ix->ws, nq, dQ, o);
    if (rc < 0) return rc;
    return search_device(ix, ix->ws, nq, dQ, user_ids, o, 0, out);
}

static int search_batch_shard_locked(rk_index *ix, size_t nq, const float *dQ, const uint32_t *user_ids, const rk_scan_opts *o,
                                     float *d_key, uint64_t *d_tid, uint32_t *d_doc)
{
    if (!ix->finalized) return rk_fail(RK_ERR_STATE, "rk_search_batch_shard: call rk_finalize first");
    int rc = validate_opts(ix, o);
    if (rc < 0) return rc;
    if (nq == 0) return RK_OK;
    if (!dQ || !d_key || !d_tid) return rk_fail(RK_ERR_INVALID, "rk_search_batch_shard: NULL pointer");
    if (ix->dim != ix->dim_pad) return rk_fail(RK_ERR_UNSUPPORTED, "device queries need dim %% 4 == 0 (dim = %d)", ix->dim);
    RkSelectParams out;
    memset(&out, 0, sizeof out);
    out.out_key = d_key; out.out_tid = d_tid; out.out_doc = d_doc;
    rc = check_device_queries(ix, ix->ws, nq, dQ, o);
    if (rc < 0) return rc;
    return search_device(ix, ix->ws, nq, dQ, user_ids, o, 0, out);
}

// for rk_multi.cu: host queries -> the index's device staging buffer (padded rows), asynchronous on the index stream
int rk_stage_host_queries(rk_index *ix, size_t nq, const float *Q, const float **dQ)
{
    RK_CUDA(cudaSetDevice(ix->device));
    std::lock_guard<std::mutex> g(ix->batch_mu);
    int rc = stage_queries(ix, ix->ws, nq, Q);
    if (rc < 0) return rc;
    *dQ = ix->ws.dq.as<float>();
    return RK_OK;
}

// for rk_exchange.cu (the caller holds ix->batch_mu): d_lists != NULL = the caller supplies the IVFFlat list choice
int rk_search_batch_shard_locked_for_exchange(rk_index *ix, size_t nq, const float *dQ, const uint32_t *user_ids, const rk_scan_opts *o,
                                              const uint32_t *d_lists, float *d_key, uint64_t *d_tid, uint32_t *d_doc)
{
    if (d_lists && (!o || o->probes < 1)) return rk_fail(RK_ERR_INVALID, "rk_search_batch_shard_push: a list choice needs probes >= 1");
    ix->ws.ext_probe = d_lists;
    const int rc = search_batch_shard_locked(ix, nq, dQ, user_ids, o, d_key, d_tid, d_doc);
    ix->ws.ext_probe = nullptr;
    return rc;
}

extern "C" int rk_search_batch_shard(rk_index *ix, size_t nq, const float *dQ, const uint32_t *user_ids,
                                     const rk_scan_opts *o, float *d_key, uint64_t *d_tid, uint32_t *d_doc)
{
    if (!ix) return rk_fail(RK_ERR_INVALID, "NULL index");
    std::lock_guard<std::mutex> g(ix->batch_mu);
    return search_batch_shard_locked(ix, nq, dQ, user_ids, o, d_key, d_tid, d_doc);
}

int rk_ivf_probe_only(rk_index *ix, Workspace &ws, size_t nq, const float *dQ, const rk_scan_opts *o, uint32_t *d_lists);

extern "C" int rk_ivf_probe_device(rk_index *ix, size_t nq, const float *dQ, const rk_scan_opts *o, uint32_t *d_lists)
{
    if (!ix) return rk_fail(RK_ERR_INVALID, "NULL index");
    if (!ix->finalized) return rk_fail(RK_ERR_STATE, "rk_ivf_probe_device: call rk_finalize first");
    int rc = validate_opts(ix, o);
    if (rc < 0) return rc;
    if (nq == 0) return RK_OK;
    if (!dQ || !d_lists) return rk_fail(RK_ERR_INVALID, "rk_ivf_probe_device: NULL pointer");
    if (ix->dim != ix->dim_pad) return rk_fail(RK_ERR_UNSUPPORTED, "device queries need dim %% 4 == 0 (dim = %d)", ix->dim);
    std::lock_guard<std::mutex> g(ix->batch_mu);
    return rk_ivf_probe_only(ix, ix->ws, nq, dQ, o, d_lists);
}

extern "C" int rk_search_batch_shard_probed(rk_index *ix, size_t nq, const float *dQ, const uint32_t *user_ids,
                                            const rk_scan_opts *o, const uint32_t *d_lists, float *d_key, uint64_t *d_tid,
                                            uint32_t *d_doc)
{
    if (!ix) return rk_fail(RK_ERR_INVALID, "NULL index");
    if (!d_lists) return rk_fail(RK_ERR_INVALID, "rk_search_batch_shard_probed: NULL list choice");
    if (!o || o->probes < 1) return rk_fail(RK_ERR_INVALID, "rk_search_batch_shard_probed: probes must be >= 1");
    std::lock_guard<std::mutex> g(ix->batch_mu);
    ix->ws.ext_probe = d_lists;
    const int rc = search_batch_shard_locked(ix, nq, dQ, user_ids, o, d_key, d_tid, d_doc);
    ix->ws.ext_probe = nullptr;
    return rc;
}

extern "C" int rk_merge_shards(rk_index *ix, size_t nq, int k, int n_shards, size_t shard_stride_bytes, const float *d_key,
                               const uint64_t *d_tid, const uint32_t *d_doc, uint64_t *d_out_tid, uint32_t *d_out_doc,
                               double *d_out_dist, int32_t *d_out_count)
{
    if (!ix) return rk_fail(RK_ERR_INVALID, "NULL index");
    if (k < 1 || k > RK_MAX_K || n_shards < 1) return rk_fail(RK_ERR_INVALID, "rk_merge_shards: k = %d, n_shards = %d", k, n_shards);
    if (!d_key || !d_tid) return rk_fail(RK_ERR_INVALID, "rk_merge_shards: NULL pointer");
    if (shard_stride_bytes % 8) return rk_fail(RK_ERR_INVALID, "rk_merge_shards: shard stride must be a multiple of 8 bytes");
    if (nq == 0) return RK_OK;
    std::lock_guard<std::mutex> g(ix->batch_mu);
    RK_CUDA(cudaSetDevice(ix->device));
    RkSelectParams p;
    memset(&p, 0, sizeof p);
    p.cand_key = reinterpret_cast<const uint32_t *>(d_key);
    p.cand_tid = d_tid;
    p.cand_row = d_doc ? d_doc : reinterpret_cast<const uint32_t *>(d_key);   // doc ids ride in the row slot
    p.cand_count = nullptr;
    p.lists_per_query = n_shards;
    p.list_stride = k;
    p.shard_stride_bytes = shard_stride_bytes;
    p.k = k;
    p.metric = ix->metric;
    p.dedup = 0;   // a physical row lives on exactly one shard
    p.key_is_float = 1;
    p.rowdoc = nullptr;
    p.out_tid = d_out_tid; p.out_doc = d_doc ? d_out_doc : nullptr; p.out_dist = d_out_dist; p.out_count = d_out_count;
    Workspace &ws = ix->ws;
    if (ws.timing) RK_CUDA(cudaEventRecord(ws.ev[2], ws.stream));
    RK_CUDA(rk_launch_select(p, (int) nq, ws.stream));
    if (ws.timing) { RK_CUDA(cudaEventRecord(ws.ev[3], ws.stream)); ws.timed_select = true; }
    ws.launches = 1;
    ix->total_launches += 1;
    return RK_OK;
}

// ---------------------------------------------------------------------------------------------
// the index-scan iterator (pgvector/src/ivfscan.c:241-406)

extern "C" rk_scan *rk_beginscan(rk_index *ix)
{
    if (!ix) { rk_fail(RK_ERR_INVALID, "NULL index"); return nullptr; }
    if (!ix->finalized) { rk_fail(RK_ERR_STATE, "rk_beginscan: call rk_finalize first"); return nullptr; }
    rk_scan *sc = new rk_scan();
    sc->ix = ix;
    if (sc->ws.init(ix->device) != 0) { delete sc; return nullptr; }
    return sc;
}

// Fetches the next k tuples of the scan.  resume = false: the next batch of `probes` lists (or, for an exact scan, the
// first k tuples); resume = true: the k tuples of the CURRENT batch that follow the last pair handed back — the scan is
// re-run with a lower bound (key, tid) > last, so a batch of any size streams out k tuples at a time, in order, exactly
// as ivfflatgettuple drains its tuplesort (ivfscan.c:375-388) before it moves on to the next `probes` lists.
static int scan_fetch_batch(rk_scan *sc, bool resume)
{
    rk_index *ix = sc->ix;
    const int k = sc->opts.k;
    sc->r_tid.resize(k); sc->r_doc.resize(k); sc->r_dist.resize(k); sc->r_key.resize(k);
    int32_t cnt = 0;
    rk_scan_opts o = sc->opts;
    if (!resume) {
        sc->batch_offset = sc->probes_done;
        sc->batch_probes = sc->probes_eff;
        if (sc->batch_probes > 0) sc->batch_probes = std::min(sc->batch_probes, sc->max_probes_eff - sc->probes_done);   // ivfscan.c:128 listIndex < maxProbes
        sc->probes_done += sc->probes_eff;
        sc->ws.lb_on = false;
    } else {
        sc->ws.lb_on = true;
        sc->ws.lb_key = rk_float_to_key(sc->r_key[(size_t) sc->r_count - 1]);
        sc->ws.lb_tid = sc->r_tid[(size_t) sc->r_count - 1];
    }
    o.probes = sc->batch_probes;
    int rc = host_search(ix, sc->ws, 1, sc->q.data(), &sc->user, &o, sc->batch_offset, sc->r_tid.data(), sc->r_doc.data(),
                         sc->r_dist.data(), &cnt, sc->r_key.data());
    sc->ws.lb_on = false;
    if (rc < 0) return rc;
    sc->r_count = cnt;
    sc->r_pos = 0;
    // a full answer from an index scan: the probed lists may hold more tuples beyond the k-th.  (An exact scan stands for
    // `ORDER BY distance LIMIT k`, prefilter_role.py:162-172: k IS the statement's limit, so it ends after k tuples.)
    sc->batch_more = cnt == k && sc->batch_probes > 0;
    return 0;
}

extern "C" int rk_rescan(rk_scan *sc, const float *q, int32_t q_dim, uint32_t user_id, const rk_scan_opts *o)
{
    if (!sc) return rk_fail(RK_ERR_INVALID, "NULL scan");
    rk_index *ix = sc->ix;
    int rc = validate_opts(ix, o);
    if (rc < 0) return rc;
    if (!q && !o->null_query) return rk_fail(RK_ERR_NO_ORDER, "cannot scan without order");   // ivfscan.c:360-361
    if (q && q_dim != ix->dim) return rk_fail(RK_ERR_DIM, "different vector dimensions %d and %d", ix->dim, q_dim);   // vector.c:61-67
    sc->q.assign((size_t) ix->dim, 0.0f);
    if (q) memcpy(sc->q.data(), q, (size_t) ix->dim * 4);
    sc->user = user_id;
    sc->opts = *o;
    sc->probes_done = 0;
    if (o->probes > 0) {
        // ivfscan.c:257-266: maxProbes = max(max_probes, probes) when iterating, both clamped to lists
        int lists = ix->ivf.lists;
        sc->probes_eff = std::min(o->probes, lists);
        sc->max_probes_eff = std::min(std::max(o->max_probes, o->probes), lists);
    } else {
        sc->probes_eff = 0;
        sc->max_probes_eff = 0;
    }
    sc->bound = true;
    return scan_fetch_batch(sc, false);
}

extern "C" int rk_gettuple(rk_scan *sc, uint64_t *tid, uint32_t *doc, double *distance)
{
    if (!sc) return rk_fail(RK_ERR_INVALID, "NULL scan");
    if (!sc->bound) return rk_fail(RK_ERR_NO_ORDER, "cannot scan without order");
    while (sc->r_pos >= sc->r_count) {
        if (sc->batch_more && sc->r_count > 0) {
            // the current batch was cut at k tuples: continue it after the last pair returned
            int rc = scan_fetch_batch(sc, true);
            if (rc < 0) return rc;
            continue;
        }
        // ivfscan.c:375-381: refill from the next `probes` lists until max_probes is reached
        if (sc->probes_eff == 0 || sc->probes_done >= sc->max_probes_eff) return 0;
        int rc = scan_fetch_batch(sc, false);
        if (rc < 0) return rc;
    }
    if (tid) *tid = sc->r_tid[sc->r_pos];
    if (doc) *doc = sc->r_doc[sc->r_pos];
    if (distance) *distance = sc->r_dist[sc->r_pos];
    sc->r_pos++;
    return 1;
}

extern "C" int rk_gettuple_batch(rk_scan *sc, int max_n, uint64_t *tid, uint32_t *doc, double *distance)
{
    if (!sc) return rk_fail(RK_ERR_INVALID, "NULL scan");
    if (max_n < 0) return rk_fail(RK_ERR_INVALID, "rk_gettuple_batch: max_n = %d", max_n);
    int n = 0;
    while (n < max_n) {
        if (sc->bound && sc->r_pos < sc->r_count) {
            // straight from the ring
            const int take = std::min(max_n - n, sc->r_count - sc->r_pos);
            if (tid) memcpy(tid + n, sc->r_tid.data() + sc->r_pos, (size_t) take * 8);
            if (doc) memcpy(doc + n, sc->r_doc.data() + sc->r_pos, (size_t) take * 4);
            if (distance) memcpy(distance + n, sc->r_dist.data() + sc->r_pos, (size_t) take * 8);
            sc->r_pos += take;
            n += take;
            continue;
        }
        const int rc = rk_gettuple(sc, tid ? tid + n : nullptr, doc ? doc + n : nullptr, distance ? distance + n : nullptr);   // refills
        if (rc < 0) return rc;
        if (rc == 0) break;
        n++;
    }
    return n;
}

extern "C" void rk_endscan(rk_scan *sc)
{
    if (!sc) return;
    cudaSetDevice(sc->ix->device);
    if (sc->ws.stream) cudaStreamSynchronize(sc->ws.stream);
    sc->ws.destroy();
    delete sc;
}

// ---------------------------------------------------------------------------------------------
// plumbing

extern "C" int rk_set_stream(rk_index *ix, void *cuda_stream)
{
    if (!ix) return rk_fail(RK_ERR_INVALID, "NULL index");
    std::lock_guard<std::mutex> g(ix->batch_mu);
    ix->ws.stream = cuda_stream ? (cudaStream_t) cuda_stream : ix->ws.own_stream;
    return RK_OK;
}

extern "C" int rk_synchronize(rk_index *ix)
{
    if (!ix) return rk_fail(RK_ERR_INVALID, "NULL index");
    RK_CUDA(cudaSetDevice(ix->device));
    std::lock_guard<std::mutex> g(ix->batch_mu);
    Workspace &ws = ix->ws;
    const bool armed = ws.qflag_armed;
    if (armed) {
        RK_CUDA(cudaMemcpyAsync(ws.h_qflag, ws.qflag.p, 4, cudaMemcpyDeviceToHost, ws.stream));
        RK_CUDA(cudaMemsetAsync(ws.qflag.p, 0, 4, ws.stream));
        ws.qflag_armed = false;
    }
    RK_CUDA(cudaStreamSynchronize(ws.stream));
    int rc = collect_timing(ix, ws);
    if (rc < 0) return rc;
    if (armed && *ws.h_qflag) {
        *ws.h_qflag = 0;
        return rk_fail(RK_ERR_NONFINITE, "NaN or infinite value not allowed in vector (device-resident queries of a call since the last rk_synchronize)");
    }
    return RK_OK;
}

extern "C" int rk_enable_timing(rk_index *ix, int on)
{
    if (!ix) return rk_fail(RK_ERR_INVALID, "NULL index");
    ix->ws.timing = on != 0;
    return RK_OK;
}

extern "C" int rk_get_scan_times(rk_index *ix, float *out_ms, int max_n)
{
    if (!ix || (!out_ms && max_n > 0)) return rk_fail(RK_ERR_INVALID, "NULL argument");
    RK_CUDA(cudaSetDevice(ix->device));
    RK_CUDA(cudaStreamSynchronize(ix->ws.stream));
    Workspace &ws = ix->ws;
    int n = (int) std::min<uint64_t>(ws.ring_n, RK_TIMING_RING);
    if (n > max_n) n = max_n;
    for (int i = 0; i < n; i++) {
        const int slot = (int) ((ws.ring_n - (uint64_t) n + (uint64_t) i) % RK_TIMING_RING);
        RK_CUDA(cudaEventElapsedTime(&out_ms[i], ws.ring_a[slot], ws.ring_b[slot]));
    }
    return n;
}

extern "C" int rk_get_stats(rk_index *ix, rk_stats *out)
{
    if (!ix || !out) return rk_fail(RK_ERR_INVALID, "NULL argument");
    memset(out, 0, sizeof *out);
    out->rows = ix->nrows;
    out->docs = ix->doc_ids.size();
    out->partitions = ix->parts.size();
    out->device_bytes = ix->device_bytes;
    out->launches = ix->ws.launches;
    out->segments = ix->ws.segments;
    out->rows_scanned = ix->ws.rows_scanned;
    out->algo_bytes = ix->ws.algo_bytes;
    out->scan_ms = ix->last_scan_ms;
    out->select_ms = ix->last_select_ms;
    out->probe_ms = ix->last_probe_ms;
    out->total_launches = ix->total_launches;
    out->tensor_redo = ix->tensor.last_redo;
    out->tensor_exact = ix->tensor.last_exact ? 1 : 0;
    out->tensor_plan = (uint64_t) ix->tensor.last_plan;
    out->tensor_flop = (uint64_t) ix->tensor.last_flop;
    out->tensor_plan_bytes = ix->tensor.last_plan_bytes;
    out->segment_kinds = ix->ws.seg_kinds;
    out->single_fused = ix->ws.single_fused ? 1 : 0;
    out->ivf_assign_tc_rows = ix->ivf.assign_tc_rows;
    out->ivf_assign_fallback_rows = ix->ivf.assign_fallback_rows;
    return RK_OK;
}
