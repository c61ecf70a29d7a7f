// rk_tensor.cu — host side of the batched tensor-core engine (rk_bscan.cu).
//
// Replaces the per-query loop of the reference harness (basic_benchmark/common_function.py:1366-1382)
// and its ground-truth computation (:671-759, :859-1139) for a whole batch of queries at once: the
// batch is cut into tiles of 128 queries, the row store into ranges of 128-row tiles, and every
// (query tile, row range) unit nominates candidates on the tensor cores; a select kernel merges the
// nominations per query and rk_rescore_kernel produces the exact answer or flags the query for the
// exact scan path.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <map>

#include "rk_host.h"

#define RK_CUDA(call)                                                                                    \
    do {                                                                                                 \
        cudaError_t e__ = (call);                                                                        \
        if (e__ != cudaSuccess)                                                                          \
            return rk_fail(RK_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), __FILE__, __LINE__); \
    } while (0)

// One-time per index: squared row norms, per-row role masks, TF32-exactness of the corpus.
static int tensor_prepare(rk_index *ix)
{
    TensorState &ts = ix->tensor;
    if (ts.ready) return 0;
    const size_t n = ix->nrows;
    const int W = ix->mask_words;
    cudaStream_t st = 0;
    if (ts.d_xnorm.ensure(std::max<size_t>(n * 4, 16)) != 0 || ts.d_rowmask.ensure(std::max<size_t>(n * (size_t) W * 8, 16)) != 0 ||
        ts.d_probe.ensure(64) != 0)
        return RK_ERR_CUDA;
    RK_CUDA(rk_launch_row_norms(ix->d_x.as<float>(), n, ix->dim_pad, ts.d_xnorm.as<float>(), st));
    RK_CUDA(rk_launch_rowmask(ix->d_rowdoc.as<uint32_t>(), ix->d_doc_mask.as<uint64_t>(), n, W, ts.d_rowmask.as<uint64_t>(), st));
    RK_CUDA(cudaMemsetAsync(ts.d_probe.p, 0, 64, st));
    RK_CUDA(rk_launch_tf32_probe(ix->d_x.as<float>(), n * (size_t) ix->dim_pad, ts.d_probe.as<unsigned long long>(), st));
    RK_CUDA(rk_launch_max_f32(ts.d_xnorm.as<float>(), n, ts.d_probe.as<unsigned>() + 8, st));
    unsigned long long h[5] = {0, 0, 0, 0, 0};
    RK_CUDA(cudaMemcpy(h, ts.d_probe.p, sizeof h, cudaMemcpyDeviceToHost));
    ts.x_integer = h[0] == 0;
    uint32_t b = (uint32_t) h[1];
    memcpy(&ts.x_absmax, &b, 4);
    b = (uint32_t) (h[4] & 0xFFFFFFFFull);
    memcpy(&ts.xnorm_max, &b, 4);
    ix->device_bytes += ts.d_xnorm.cap + ts.d_rowmask.cap;
    ts.ready = true;
    return 0;
}

int rk_tensor_supported(const rk_index *ix, const rk_scan_opts *o)
{
    if (o->probes != 0 || o->keep_duplicates || o->null_query) return 0;
    if (o->k > 256) return 0;
    (void) ix;
    return 1;
}

// Tiles (first rows of 128-row boxes) that cover every row visible under `mask`, in physical order.
// Documents are contiguous extents, so a visible document starts a tile unless the previous tile
// already reaches into it; rows a tile drags in beyond the visible ones are rejected by the
// epilogue's mask test.  Cached per role mask (there are as many as distinct role sets).
static int tiles_for_mask(rk_index *ix, const std::vector<uint64_t> &mask, uint32_t *off, uint32_t *cnt, bool *grew)
{
    TensorState &ts = ix->tensor;
    auto it = ts.tile_cache.find(mask);
    if (it != ts.tile_cache.end()) { *off = it->second.first; *cnt = it->second.second; return 0; }
    const int W = ix->mask_words;
    const size_t D = ix->doc_ids.size();
    const uint32_t o = (uint32_t) ts.h_tiles.size();
    uint64_t covered = 0;
    for (size_t d = 0; d < D; d++) {
        uint64_t acc = 0;
        for (int w = 0; w < W; w++) acc |= ix->doc_mask[d * W + w] & mask[w];
        if (!acc) continue;
        const uint64_t r0 = ix->doc_row_start[d], r1 = ix->doc_row_start[d + 1];
        if (covered >= r1) continue;
        uint64_t start = std::max(r0, covered);
        while (start < r1) { ts.h_tiles.push_back((uint32_t) start); start += 128; }
        covered = start;
    }
    *off = o;
    *cnt = (uint32_t) ts.h_tiles.size() - o;
    ts.tile_cache[mask] = std::make_pair(*off, *cnt);
    *grew = true;
    return 0;
}

struct TGroup {
    std::vector<uint32_t> queries;
    const UserInfo *user = nullptr;
    uint32_t tl_off = 0, tl_n = 0;
};

// Runs the batch on the tensor engine.  `out` carries the final device destinations.  On return
// `redo` lists the queries whose answer could not be proven (to be re-run on the exact scan path).
int rk_tensor_search(rk_index *ix, Workspace &ws, size_t nq, const float *dQ, const uint32_t *user_ids, const rk_scan_opts *o,
                     const RkSelectParams &out, std::vector<uint32_t> &redo)
{
    redo.clear();
    int rc = tensor_prepare(ix);
    if (rc < 0) return rc;
    TensorState &ts = ix->tensor;
    const int W = ix->mask_words;
    const int k = o->k;
    int kp = std::max(2 * k, k + 128);
    if (kp > 512) kp = 512;
    const size_t n = ix->nrows;
    const uint32_t T = (uint32_t) ((n + 127) / 128);
    const int sm = ix->sm_count;
    const bool use_mask = o->filter != RK_FILTER_NONE;
    cudaStream_t st = ws.stream;

    // ---- groups: queries that see the same rows (same role mask) share one pass over those rows
    std::vector<TGroup> groups;
    std::vector<uint32_t> empty_queries;           // users without roles: nothing visible
    {
        std::map<std::vector<uint64_t>, uint32_t> by_mask;
        std::vector<uint64_t> key((size_t) W);
        for (size_t q = 0; q < nq; q++) {
            const UserInfo *u = nullptr;
            if (use_mask) {
                if (user_ids) {
                    auto it = ix->users.find(user_ids[q]);
                    if (it != ix->users.end()) u = &it->second;
                }
                if (!u || u->allowed_rows == 0) { empty_queries.push_back((uint32_t) q); continue; }
                key = u->mask;
            } else {
                std::fill(key.begin(), key.end(), ~0ull);
            }
            auto ins = by_mask.emplace(key, (uint32_t) groups.size());
            if (ins.second) { groups.emplace_back(); groups.back().user = u; }
            groups[ins.first->second].queries.push_back((uint32_t) q);
        }
    }
    // grouped (visit only tiles with visible rows, once per group) or dense (every tile, all queries
    // together)?  Compare the bytes each plan stages into shared memory: (128 + padded queries) per tile.
    bool grouped = use_mask;
    bool grew = false;
    if (grouped) {
        double cost_g = 0.0;
        for (auto &g : groups) {
            rc = tiles_for_mask(ix, g.user->mask, &g.tl_off, &g.tl_n, &grew);
            if (rc < 0) return rc;
            const size_t qn = g.queries.size();
            cost_g += (double) g.tl_n * (double) ((qn + 127) / 128) * (128.0 + (double) std::min<size_t>(128, (qn + 15) & ~(size_t) 15));
        }
        const double cost_d = (double) T * (double) ((nq + 127) / 128) * 256.0;
        if (cost_g > cost_d) grouped = false;
    }
    if (!grouped && groups.size() != 1) {
        // dense: one group holding every query that can see something; the mask test is per (row, query)
        TGroup all;
        for (auto &g : groups) all.queries.insert(all.queries.end(), g.queries.begin(), g.queries.end());
        std::sort(all.queries.begin(), all.queries.end());
        groups.clear();
        groups.push_back(std::move(all));
    }
    if (!grouped)
        for (auto &g : groups) { g.tl_off = 0; g.tl_n = T; }
    if (grouped && grew) {
        if (ts.d_tiles.ensure(std::max<size_t>(ts.h_tiles.size() * 4, 16)) != 0) return RK_ERR_CUDA;
        RK_CUDA(cudaMemcpyAsync(ts.d_tiles.p, ts.h_tiles.data(), ts.h_tiles.size() * 4, cudaMemcpyHostToDevice, st));
        RK_CUDA(cudaStreamSynchronize(st));      // h_tiles may reallocate on the next call
    }

    // ---- slots (group-ordered queries), units, lists
    size_t nslots = 0;
    for (auto &g : groups) nslots += g.queries.size();
    double total_work = 0.0;
    for (auto &g : groups) total_work += (double) g.tl_n * (double) ((g.queries.size() + 127) / 128);
    // tiles per unit: ~8 units per CTA, but a unit ends with one compaction + sort per query, so it
    // must be long enough to amortise them; with few units, size them to fill whole rounds of CTAs
    size_t max_gq = 0;
    for (auto &g : groups) max_gq = std::max(max_gq, std::min<size_t>(128, g.queries.size()));
    uint32_t tpu = (uint32_t) std::max(32.0 + 2.0 * (double) max_gq, total_work / ((double) sm * 8.0));
    uint32_t single_R = 0;
    {
        const double est_units = total_work / (double) tpu;
        if (est_units < 4.0 * sm) {
            const double m = std::max(1.0, std::floor(est_units / sm));
            tpu = (uint32_t) std::ceil(total_work / (m * sm));
            if (groups.size() == 1) {
                const uint32_t qt = (uint32_t) ((groups[0].queries.size() + 127) / 128);
                single_R = std::max<uint32_t>(1, (uint32_t) (m * sm) / std::max<uint32_t>(qt, 1));
                single_R = std::min<uint32_t>(single_R, std::max<uint32_t>(1, groups[0].tl_n));
            }
        }
    }
    std::vector<RkBUnit> units;
    std::vector<uint32_t> slot_q(nslots), lid0(nslots), lstep(nslots), nlists(nslots);
    std::vector<double> unit_work;
    uint32_t slot = 0, list_total = 0;
    for (auto &g : groups) {
        const size_t gq = g.queries.size();
        for (size_t b = 0; b < gq; b += 128) {
            const uint32_t qn = (uint32_t) std::min<size_t>(128, gq - b);
            const uint32_t R = single_R ? single_R : std::max<uint32_t>(1, (g.tl_n + tpu - 1) / tpu);
            for (uint32_t i = 0; i < qn; i++) {
                slot_q[slot + i] = g.queries[b + i];
                lid0[slot + i] = list_total + i;
                lstep[slot + i] = qn;
                nlists[slot + i] = R;
            }
            for (uint32_t r = 0; r < R; r++) {
                RkBUnit u;
                memset(&u, 0, sizeof u);
                u.q0 = slot; u.nq = qn;
                u.tile_begin = (uint32_t) ((uint64_t) g.tl_n * r / R);
                u.tile_end = (uint32_t) ((uint64_t) g.tl_n * (r + 1) / R);
                u.tl_off = g.tl_off;
                u.list_base = list_total;
                list_total += qn;
                units.push_back(u);
                unit_work.push_back((double) (u.tile_end - u.tile_begin) * (128.0 + (double) ((qn + 15) & ~15u)) + 2000.0 * qn);
            }
            slot += qn;
        }
    }
    const uint32_t nunits = (uint32_t) units.size();
    const int grid = (int) std::min<uint32_t>((uint32_t) sm, std::max<uint32_t>(nunits, 1));
    // balanced static schedule: longest unit first onto the least loaded CTA
    std::vector<uint32_t> order(nunits), cta_off((size_t) grid + 1, 0), cta_units(nunits);
    {
        for (uint32_t i = 0; i < nunits; i++) order[i] = i;
        std::sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) { return unit_work[a] > unit_work[b]; });
        std::vector<std::vector<uint32_t>> per((size_t) grid);
        std::vector<std::pair<double, int>> heap;            // (load, cta) min-heap
        for (int c = 0; c < grid; c++) heap.push_back({0.0, c});
        auto cmp = [](const std::pair<double, int> &a, const std::pair<double, int> &b) { return a.first > b.first; };
        std::make_heap(heap.begin(), heap.end(), cmp);
        for (uint32_t u : order) {
            std::pop_heap(heap.begin(), heap.end(), cmp);
            auto &top = heap.back();
            per[(size_t) top.second].push_back(u);
            top.first += unit_work[u];
            std::push_heap(heap.begin(), heap.end(), cmp);
        }
        uint32_t pos = 0;
        for (int c = 0; c < grid; c++) {
            cta_off[(size_t) c] = pos;
            for (uint32_t u : per[(size_t) c]) cta_units[pos++] = u;
        }
        cta_off[(size_t) grid] = pos;
    }

    // ---- staging: [units][cta_off][cta_units][slot_q][lid0][lstep][nlists][qmask per slot]
    const size_t b_units = (size_t) nunits * sizeof(RkBUnit), b_coff = ((size_t) grid + 1) * 4, b_cu = (size_t) nunits * 4;
    const size_t b_slot = nslots * 4, b_qmask = nslots * (size_t) W * 8;
    const size_t total_stage = b_units + b_coff + b_cu + 4 * b_slot + b_qmask + 256;
    ws.plan_idx ^= 1;
    const int pi_ = ws.plan_idx;
    if (ws.ev_plan_used[pi_]) RK_CUDA(cudaEventSynchronize(ws.ev_plan[pi_]));
    PinBuf &hp_ = ws.h_plan_ring[pi_];
    if (hp_.ensure(total_stage) != 0) return RK_ERR_CUDA;
    if (ts.d_plan.ensure(total_stage) != 0) return RK_ERR_CUDA;
    unsigned char *hp = hp_.as<unsigned char>();
    size_t off = 0;
    auto put = [&](const void *src, size_t bytes) -> size_t {
        const size_t at = off;
        if (bytes) memcpy(hp + at, src, bytes);
        off = (off + bytes + 15) & ~(size_t) 15;
        return at;
    };
    const size_t o_units = put(units.data(), b_units), o_coff = put(cta_off.data(), b_coff), o_cu = put(cta_units.data(), b_cu);
    const size_t o_slotq = put(slot_q.data(), b_slot), o_lid0 = put(lid0.data(), b_slot), o_lstep = put(lstep.data(), b_slot);
    const size_t o_nl = put(nlists.data(), b_slot);
    const size_t o_qmask = off;
    {
        uint64_t *hm = reinterpret_cast<uint64_t *>(hp + o_qmask);
        uint32_t s = 0;
        for (auto &g : groups) {
            // grouped: the group's mask; dense: each query's own mask
            for (uint32_t q : g.queries) {
                const UserInfo *u = g.user;
                if (!grouped && use_mask && user_ids) {
                    auto it = ix->users.find(user_ids[q]);
                    u = it == ix->users.end() ? nullptr : &it->second;
                }
                for (int w = 0; w < W; w++) hm[(size_t) s * W + w] = u ? u->mask[w] : 0ull;
                s++;
            }
        }
        off += b_qmask;
    }
    const unsigned char *dplan = ts.d_plan.as<unsigned char>();
    if (ts.d_qp.ensure(std::max<size_t>(nslots * (size_t) ix->dim_pad * 4, 16)) != 0 ||
        ts.d_qnorm.ensure(std::max<size_t>(nslots * 4, 16)) != 0 || ts.d_scratch.ensure(rk_bscan_scratch_bytes(grid)) != 0 ||
        ts.d_cand_key.ensure(std::max<size_t>((size_t) list_total * kp * 4, 16)) != 0 ||
        ts.d_cand_row.ensure(std::max<size_t>((size_t) list_total * kp * 4, 16)) != 0 ||
        ts.d_cand_count.ensure(std::max<size_t>((size_t) list_total * 4, 16)) != 0 ||
        ts.d_sel_key.ensure(std::max<size_t>(nslots * (size_t) kp * 4, 16)) != 0 ||
        ts.d_sel_row.ensure(std::max<size_t>(nslots * (size_t) kp * 4, 16)) != 0 || ts.d_sel_count.ensure(std::max<size_t>(nslots * 4, 16)) != 0 ||
        ts.d_flags.ensure(std::max<size_t>((nq + nslots) * 4, 16)) != 0 || ts.h_flags.ensure(std::max<size_t>(nq * 4 + 128, 128)) != 0)
        return RK_ERR_CUDA;
    RK_CUDA(cudaMemcpyAsync(ts.d_plan.p, hp, off, cudaMemcpyHostToDevice, st));
    RK_CUDA(cudaEventRecord(ws.ev_plan[pi_], st));
    ws.ev_plan_used[pi_] = true;
    RK_CUDA(cudaMemsetAsync(ts.d_flags.p, 0, (nq + nslots) * 4, st));      // [uncertain per query][trunc per slot]
    RK_CUDA(cudaMemsetAsync(ts.d_probe.p, 0, 64, st));

    // queries nobody can answer from (no visible rows): empty results straight away
    if (!empty_queries.empty() || nslots == 0) {
        RkSelectParams ep = out;
        ep.cand_key = nullptr; ep.cand_tid = nullptr; ep.cand_row = nullptr; ep.cand_count = nullptr;
        if (ws.qseg.ensure((nq + 1) * 4) != 0) return RK_ERR_CUDA;
        RK_CUDA(cudaMemsetAsync(ws.qseg.p, 0, (nq + 1) * 4, st));           // every query owns zero lists
        ep.qseg = ws.qseg.as<uint32_t>();
        ep.lists_per_query = 0; ep.lid0 = nullptr; ep.list_stride = k; ep.k = k; ep.metric = ix->metric; ep.dedup = 0;
        ep.key_is_float = 0; ep.rowdoc = nullptr; ep.doc_ids = nullptr;
        RK_CUDA(rk_launch_select(ep, (int) nq, st));                         // fills padding + count 0 for all; real answers overwrite below
        ws.launches++;
    }
    if (nslots == 0) {
        ts.last_redo = 0; ts.last_exact = true; ts.last_flop = 0.0;
        return RK_OK;
    }

    // group-ordered query matrix, its norms and TF32-exactness
    const uint32_t *d_slotq = reinterpret_cast<const uint32_t *>(dplan + o_slotq);
    RK_CUDA(rk_launch_gather_rows(dQ, ts.d_qp.as<float>(), d_slotq, nslots, ix->dim_pad, st));
    const float *dQp = ts.d_qp.as<float>();
    RK_CUDA(rk_launch_row_norms(dQp, nslots, ix->dim_pad, ts.d_qnorm.as<float>(), st));
    RK_CUDA(rk_launch_tf32_probe(dQp, nslots * (size_t) ix->dim_pad, ts.d_probe.as<unsigned long long>(), st));
    RK_CUDA(rk_launch_max_f32(ts.d_qnorm.as<float>(), nslots, ts.d_probe.as<unsigned>() + 8, st));
    unsigned long long *hprobe = reinterpret_cast<unsigned long long *>(ts.h_flags.as<unsigned char>() + ((nq * 4 + 15) & ~(size_t) 15));
    RK_CUDA(cudaMemcpyAsync(hprobe, ts.d_probe.p, 40, cudaMemcpyDeviceToHost, st));
    RK_CUDA(cudaStreamSynchronize(st));
    float qnorm_max;
    { uint32_t b = (uint32_t) (hprobe[4] & 0xFFFFFFFFull); memcpy(&qnorm_max, &b, 4); }
    // exact iff operands are small integers (exact in TF32, products exact in fp32) and every partial
    // sum stays below 2^24: |x|^2 + |q|^2 + 2 |x.q| <= (|x| + |q|)^2 by Cauchy-Schwarz
    const double nx = std::sqrt((double) ts.xnorm_max), nqm = std::sqrt((double) qnorm_max);
    const bool exact = ts.x_integer && hprobe[0] == 0 && (nx + nqm) * (nx + nqm) < 16777216.0;
    // TF32 keeps 10 mantissa bits and the MMA truncates: each operand is off by < 2^-10 relative, a
    // product by < 2^-9; |x.q| error <= 2^-9 * |x||q|; the L2 form doubles it
    const float eps_scale = exact ? 0.f : (ix->metric == RK_METRIC_L2 ? 2.f : 1.f) * (1.0f / 512.0f);

    RkBScanParams bp;
    memset(&bp, 0, sizeof bp);
    bp.units = reinterpret_cast<const RkBUnit *>(dplan + o_units);
    bp.nunits = nunits;
    bp.cta_off = reinterpret_cast<const uint32_t *>(dplan + o_coff);
    bp.cta_units = reinterpret_cast<const uint32_t *>(dplan + o_cu);
    bp.tile_rows = grouped ? ts.d_tiles.as<uint32_t>() : nullptr;
    bp.nrows = n;
    bp.dim_pad = ix->dim_pad;
    bp.kchunks = (ix->dim_pad + 31) / 32;
    bp.metric = ix->metric;
    bp.kp = kp;
    bp.mask_words = W;
    bp.use_mask = use_mask ? 1 : 0;
    bp.xnorm = ts.d_xnorm.as<float>();
    bp.rowmask = ts.d_rowmask.as<uint64_t>();
    bp.qnorm = ts.d_qnorm.as<float>();
    bp.qmask = reinterpret_cast<const uint64_t *>(dplan + o_qmask);
    bp.scratch = ts.d_scratch.as<uint2>();
    bp.cand_key = ts.d_cand_key.as<uint32_t>();
    bp.cand_row = ts.d_cand_row.as<uint32_t>();
    bp.cand_count = ts.d_cand_count.as<int32_t>();
    bp.trunc = ts.d_flags.as<int32_t>() + nq;
    const int slot_ev = (int) (ws.ring_n % RK_TIMING_RING);
    if (ws.timing) RK_CUDA(cudaEventRecord(ws.ring_a[slot_ev], st));
    RK_CUDA(rk_launch_bscan(bp, ix->d_x.as<float>(), dQp, nslots, grid, st));
    if (ws.timing) { RK_CUDA(cudaEventRecord(ws.ring_b[slot_ev], st)); ws.timed_scan = true; ws.ring_n++; }

    // merge the nominations of each slot: top-kp by (approximate key, row)
    RkSelectParams sp;
    memset(&sp, 0, sizeof sp);
    sp.cand_key = ts.d_cand_key.as<uint32_t>();
    sp.cand_tid = nullptr;
    sp.cand_row = ts.d_cand_row.as<uint32_t>();
    sp.cand_count = ts.d_cand_count.as<int32_t>();
    sp.lid0 = reinterpret_cast<const uint32_t *>(dplan + o_lid0);
    sp.lstep = reinterpret_cast<const uint32_t *>(dplan + o_lstep);
    sp.nlists = reinterpret_cast<const uint32_t *>(dplan + o_nl);
    sp.list_stride = kp;
    sp.k = kp;
    sp.metric = ix->metric;
    sp.dedup = 0;
    sp.out_key = ts.d_sel_key.as<float>();
    sp.out_row = ts.d_sel_row.as<uint32_t>();
    sp.out_count = ts.d_sel_count.as<int32_t>();
    if (ws.timing) RK_CUDA(cudaEventRecord(ws.ev[2], st));
    RK_CUDA(rk_launch_select(sp, (int) nslots, st));

    RkRescoreParams rp;
    memset(&rp, 0, sizeof rp);
    rp.x = ix->d_x.as<float>();
    rp.tid = ix->d_tid.as<uint64_t>();
    rp.rowdoc = ix->d_rowdoc.as<uint32_t>();
    rp.doc_ids = ix->d_doc_ids.as<uint32_t>();
    rp.q = dQp;
    rp.qnorm = ts.d_qnorm.as<float>();
    rp.slot_q = d_slotq;
    rp.dim_pad = ix->dim_pad;
    rp.k = k;
    rp.kp = kp;
    rp.metric = ix->metric;
    rp.null_query = 0;
    rp.sel_key = ts.d_sel_key.as<float>();
    rp.sel_row = ts.d_sel_row.as<uint32_t>();
    rp.sel_count = ts.d_sel_count.as<int32_t>();
    rp.trunc = ts.d_flags.as<int32_t>() + nq;
    rp.eps_scale = eps_scale;
    rp.xnorm_max = ts.xnorm_max;
    rp.out_tid = out.out_tid;
    rp.out_doc = out.out_doc;
    rp.out_dist = out.out_dist;
    rp.out_key = out.out_key;
    rp.out_count = out.out_count;
    rp.uncertain = ts.d_flags.as<int32_t>();
    RK_CUDA(rk_launch_rescore(rp, (int) nslots, st));
    if (ws.timing) { RK_CUDA(cudaEventRecord(ws.ev[3], st)); ws.timed_select = true; }
    int32_t *hflags = ts.h_flags.as<int32_t>();
    RK_CUDA(cudaMemcpyAsync(hflags, ts.d_flags.p, nq * 4, cudaMemcpyDeviceToHost, st));
    RK_CUDA(cudaStreamSynchronize(st));
    for (size_t q = 0; q < nq; q++)
        if (hflags[q]) redo.push_back((uint32_t) q);

    ws.launches += 7;
    ws.segments = nunits;
    uint64_t tile_visits = 0, unique_tiles = 0;
    for (auto &g : groups) { tile_visits += (uint64_t) g.tl_n * ((g.queries.size() + 127) / 128); unique_tiles += g.tl_n; }
    ws.rows_scanned = tile_visits * 128;
    // bytes this plan must move: every visited tile once per query tile, plus the queries
    ws.algo_bytes = (uint64_t) ix->dim_pad * 4ull * (tile_visits * 128ull + nslots);
    ts.last_flop = 2.0 * (double) ix->dim * 128.0 * (double) tile_visits * 128.0;
    ts.last_redo = (uint32_t) redo.size();
    ts.last_exact = exact;
    ts.last_grouped = grouped;
    (void) unique_tiles;
    return RK_OK;
}
