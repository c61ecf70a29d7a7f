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
#include <cstdlib>
#include <cstring>
#include <chrono>
#include <cstdio>
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
    // norms and the 16-byte row masks are padded by one tile: the kernel stages them with bulk copies of
    // whole tiles, and a tile may reach past the last row
    const size_t n_pad = n + 128;
    if (ts.d_xnorm.ensure(n_pad * 4) != 0 || ts.d_rowmask2.ensure(n_pad * 16) != 0 || ts.d_probe.ensure(64) != 0) return RK_ERR_CUDA;
    if (W > 2 && ts.d_rowmask.ensure(std::max<size_t>(n * (size_t) W * 8, 16)) != 0) return RK_ERR_CUDA;
    RK_CUDA(cudaMemsetAsync(ts.d_xnorm.p, 0, n_pad * 4, st));
    RK_CUDA(rk_launch_row_norms(ix->d_x.as<float>(), n, ix->dim_pad, ts.d_xnorm.as<float>(), st));
    if (W > 2) RK_CUDA(rk_launch_rowmask(ix->d_rowdoc.as<uint32_t>(), ix->d_doc_mask.as<uint64_t>(), n, W, ts.d_rowmask.as<uint64_t>(), st));
    RK_CUDA(rk_launch_rowmask2(ix->d_rowdoc.as<uint32_t>(), ix->d_doc_mask.as<uint64_t>(), n, n_pad, W, ts.d_rowmask2.as<uint64_t>(), st));
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
    ix->device_bytes += ts.d_xnorm.cap + ts.d_rowmask.cap + ts.d_rowmask2.cap;

    // Atoms: documents with the same role mask are visible to exactly the same users.  When there are
    // few of them (role hierarchies: one per role) a batch is scored atom by atom — every atom's rows
    // once, against all the queries that may see it — instead of once per distinct user.
    {
        const size_t D = ix->doc_ids.size();
        std::map<std::vector<uint64_t>, uint32_t> by_mask;
        std::vector<uint32_t> doc_atom(D, 0xFFFFFFFFu);
        std::vector<uint64_t> key((size_t) W);
        bool ok = true;
        for (size_t d = 0; d < D && ok; d++) {
            bool any = false;
            for (int w = 0; w < W; w++) { key[(size_t) w] = ix->doc_mask[d * W + w]; any |= key[(size_t) w] != 0; }
            if (!any) continue;                                      // nobody can see it
            auto ins = by_mask.emplace(key, (uint32_t) ts.atoms.size());
            if (ins.second) {
                if (ts.atoms.size() >= 4096) { ok = false; break; }
                ts.atoms.emplace_back();
                ts.atoms.back().mask = key;
            }
            doc_atom[d] = ins.first->second;
        }
        if (!ok) ts.atoms.clear();
        if (!ts.atoms.empty()) {
            // tile lists: per atom the 128-row tiles covering its documents, in physical order
            std::vector<std::vector<uint32_t>> tl(ts.atoms.size());
            std::vector<uint64_t> covered(ts.atoms.size(), 0);
            for (size_t d = 0; d < D; d++) {
                const uint32_t a = doc_atom[d];
                if (a == 0xFFFFFFFFu) continue;
                const uint64_t r0 = ix->doc_row_start[d], r1 = ix->doc_row_start[d + 1];
                ts.atoms[a].rows += r1 - r0;
                if (covered[a] >= r1) continue;
                uint64_t start = std::max(r0, covered[a]) & ~(uint64_t) 3;
                while (start < r1) { tl[a].push_back((uint32_t) start); start += 128; }
                covered[a] = start;
            }
            for (size_t a = 0; a < ts.atoms.size(); a++) {
                ts.atoms[a].tl_off = (uint32_t) ts.h_tiles.size();
                ts.atoms[a].tl_n = (uint32_t) tl[a].size();
                ts.h_tiles.insert(ts.h_tiles.end(), tl[a].begin(), tl[a].end());
            }
            ts.tiles_dirty = true;
        }
    }
    ts.ready = true;
    return 0;
}

int rk_tensor_supported(const rk_index *ix, const rk_scan_opts *o)
{
    if (o->probes != 0 || o->keep_duplicates || o->null_query) return 0;
    // nomination lists hold at most 512 entries and the sufficiency proof wants kp >= k + 128: k <= 384 (round 2: was 256, when
    // kp = 2k was required of every k)
    if (o->k > 384) return 0;
    (void) ix;
    return 1;
}

// Tiles (first rows of 128-row boxes) that cover every row visible under `mask`, in physical order.
// Documents are contiguous extents, so a visible document starts a tile unless the previous tile
// already reaches into it; rows a tile drags in beyond the visible ones are rejected by the
// epilogue's mask test.  Cached per role mask (there are as many as distinct role sets).
static int tiles_for_mask(rk_index *ix, const std::vector<uint64_t> &mask, uint32_t *off, uint32_t *cnt, bool *grew);

static int tiles_for_user(rk_index *ix, const UserInfo &u, uint32_t *off, uint32_t *cnt, bool *grew)
{
    TensorState &ts = ix->tensor;
    if (ts.tile_by_mask_id.size() != ix->n_mask_ids) ts.tile_by_mask_id.assign(ix->n_mask_ids, std::make_pair(0u, ~0u));
    auto &e = ts.tile_by_mask_id[u.mask_id];
    if (e.second == ~0u) {
        int rc = tiles_for_mask(ix, u.mask, &e.first, &e.second, grew);
        if (rc < 0) return rc;
    }
    *off = e.first; *cnt = e.second;
    return 0;
}

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
        // tile starts are multiples of 4 rows: the tile's norms are fetched with a 16-byte aligned bulk copy
        uint64_t start = std::max(r0, covered) & ~(uint64_t) 3;
        while (start < r1) { ts.h_tiles.push_back((uint32_t) start); start += 128; }
        covered = start;
    }
    *off = o;
    *cnt = (uint32_t) ts.h_tiles.size() - o;
    ts.tile_cache[mask] = std::make_pair(*off, *cnt);
    *grew = true;
    ts.tiles_dirty = true;      // uploaded by the next call whose plan walks tile lists
    return 0;
}

struct TGroup {
    std::vector<uint32_t> queries;
    const UserInfo *user = nullptr;
    uint32_t tl_off = 0, tl_n = 0;
    uint64_t min_visible = 0;      // fewest rows any query of the group can see
    const std::vector<uint64_t> *own_mask = nullptr;   // atoms plan: the document class the group's units may nominate
};

// One launch of the scoring kernel: units, their balanced assignment to CTAs, and which candidate
// lists each query slot owns afterwards.
struct TPass {
    std::vector<RkBUnit> units;
    std::vector<uint32_t> cta_off, cta_units, lid0, lstep, nlists;
    uint32_t list_total = 0;
    int grid = 1;
    double tile_visits = 0.0;
    double pairs = 0.0;            // (query, row) pairs scored: sum over units of queries x tiles x 128
    size_t o_units = 0, o_coff = 0, o_cu = 0, o_lid0 = 0, o_lstep = 0, o_nl = 0;     // offsets in the staged plan
};

// Cuts every (group, tile of 128 queries) pair into units over the group's tiles visited with stride
// step[g] (0 = the group takes no part in this pass) and balances the units onto the CTAs.
// Paired mode (rk_bscan2_kernel): query tiles are `qt` = 256 wide, every range of tiles becomes a PAIR of units
// (2p walks the even visits of the range, 2p + 1 the odd ones) and the schedule places pairs onto clusters.
static void build_pass(const std::vector<TGroup> &groups, const std::vector<uint32_t> &step, int sm, uint32_t LPQ, size_t nslots,
                       TPass &ps, bool paired, int force_tpu)
{
    const size_t qt = paired ? 256 : 128;
    const uint32_t upr = paired ? 2u : 1u;           // units per range of tiles
    double total_work = 0.0;
    size_t max_gq = 0;
    for (size_t gi = 0; gi < groups.size(); gi++) {
        if (!step[gi]) continue;
        const TGroup &g = groups[gi];
        const double visits = (double) ((g.tl_n + step[gi] - 1) / step[gi]);
        total_work += visits * (double) ((g.queries.size() + qt - 1) / qt);
        ps.pairs += visits * 128.0 * (double) g.queries.size();
        max_gq = std::max(max_gq, std::min<size_t>(qt, g.queries.size()));
    }
    ps.tile_visits = total_work;
    // tiles per unit: ~8 units per CTA, but a unit ends with one compaction + sort per query, so it
    // must be long enough to amortise them; with few units, size them to fill whole rounds of CTAs
    uint32_t tpu = (uint32_t) std::max(32.0 + 2.0 * (double) max_gq, total_work / ((double) sm * 8.0));
    {
        const double est_units = total_work / (double) tpu;
        if (est_units < 4.0 * sm) {
            // few units: they must fill whole rounds of CTAs — one unit too many doubles the makespan.
            // Grow the unit length until the units of all (group, query tile) pairs fit m rounds.
            const double m = std::max(1.0, std::floor(est_units / sm));
            tpu = (uint32_t) std::max(1.0, std::ceil(total_work / (m * sm)));
            // (with many (group, query tile) pairs the units differ in length anyway and the
            // longest-first schedule below balances them: more, shorter units are better there)
            size_t npairs = 0;
            for (size_t gi = 0; gi < groups.size(); gi++)
                if (step[gi]) npairs += (groups[gi].queries.size() + qt - 1) / qt;
            for (int it = 0; it < 64 && npairs * 2 <= (size_t) sm; it++) {
                uint64_t units = 0;
                for (size_t gi = 0; gi < groups.size(); gi++) {
                    if (!step[gi]) continue;
                    const uint64_t visits = (groups[gi].tl_n + step[gi] - 1) / step[gi];
                    units += upr * std::max<uint64_t>(1, (visits + upr * tpu - 1) / (upr * tpu)) * ((groups[gi].queries.size() + qt - 1) / qt);
                }
                if ((double) units <= m * sm) break;
                tpu += std::max<uint32_t>(1, tpu / 64);
            }
        }
    }
    if (force_tpu > 0) tpu = (uint32_t) force_tpu;      // debugging aid (rk_set_option "tensor_tpu"): tiles per unit
    ps.lid0.assign(nslots, 0); ps.lstep.assign(nslots, 0); ps.nlists.assign(nslots, 0);
    std::vector<double> unit_work;
    uint32_t slot = 0;
    for (size_t gi = 0; gi < groups.size(); gi++) {
        const TGroup &g = groups[gi];
        const size_t gq = g.queries.size();
        if (!step[gi]) { slot += (uint32_t) gq; continue; }
        const uint32_t st = step[gi];
        const uint32_t visits = (g.tl_n + st - 1) / st;
        for (size_t b = 0; b < gq; b += qt) {
            const uint32_t qn = (uint32_t) std::min<size_t>(qt, gq - b);
            const uint32_t R = std::max<uint32_t>(1, (visits + upr * tpu - 1) / (upr * tpu));
            for (uint32_t i = 0; i < qn; i++) {
                ps.lid0[slot + i] = ps.list_total + i;
                ps.lstep[slot + i] = qn;
                ps.nlists[slot + i] = R * upr * LPQ;    // every unit writes one list per epilogue set
            }
            for (uint32_t r = 0; r < R; r++) {
                const uint32_t lo = (uint32_t) ((uint64_t) visits * r / R), hi = (uint32_t) ((uint64_t) visits * (r + 1) / R);
                for (uint32_t h = 0; h < upr; h++) {
                    RkBUnit u;
                    memset(&u, 0, sizeof u);
                    u.q0 = slot; u.nq = qn;
                    u.tile_begin = (lo + h) * st;
                    u.tile_end = hi == visits ? g.tl_n : hi * st;
                    u.tile_step = st * upr;
                    u.tl_off = g.tl_off;
                    u.mask_id = g.own_mask ? (uint32_t) gi + 1u : 0u;
                    u.list_base = ps.list_total;
                    ps.list_total += qn * LPQ;
                    ps.units.push_back(u);
                }
                // per CTA: tiles walked x bytes staged per tile, plus the unit-end sorts
                unit_work.push_back((double) ((hi - lo + upr - 1) / upr) * (128.0 + (double) (((qn + 15) & ~15u) / upr)) + 2000.0 * qn / (paired ? 2.0 : 1.0));
            }
            slot += qn;
        }
    }
    // schedulable items: units, or unit pairs onto clusters of two CTAs
    const uint32_t nunits = (uint32_t) unit_work.size();
    const int bins = (int) std::min<uint32_t>((uint32_t) (paired ? sm / 2 : sm), std::max<uint32_t>(nunits, 1));
    ps.grid = bins;
    // balanced static schedule: longest unit first onto the least loaded CTA
    std::vector<uint32_t> order(nunits);
    ps.cta_off.assign((size_t) ps.grid + 1, 0);
    ps.cta_units.assign(nunits, 0);
    for (uint32_t i = 0; i < nunits; i++) order[i] = i;
    std::sort(order.begin(), order.end(), [&](uint32_t x, uint32_t y) { return unit_work[x] > unit_work[y]; });
    std::vector<std::vector<uint32_t>> per((size_t) ps.grid);
    std::vector<std::pair<double, int>> heap;            // (load, cta) min-heap
    for (int c = 0; c < ps.grid; c++) heap.push_back({0.0, c});
    auto cmp = [](const std::pair<double, int> &x, const std::pair<double, int> &y) { return x.first > y.first; };
    std::make_heap(heap.begin(), heap.end(), cmp);
    for (uint32_t u : order) {
        std::pop_heap(heap.begin(), heap.end(), cmp);
        auto &top = heap.back();
        per[(size_t) top.second].push_back(u);
        top.first += unit_work[u];
        std::push_heap(heap.begin(), heap.end(), cmp);
    }
    uint32_t pos = 0;
    for (int c = 0; c < ps.grid; c++) {
        ps.cta_off[(size_t) c] = pos;
        for (uint32_t u : per[(size_t) c]) ps.cta_units[pos++] = u;
    }
    ps.cta_off[(size_t) ps.grid] = pos;
    if (paired) ps.grid *= 2;        // cta_off stays per cluster
}

// Runs the batch on the tensor engine.  `out` carries the final device destinations.  On return
// `redo` lists the queries whose answer could not be proven (to be re-run on the exact scan path).
int rk_tensor_search(rk_index *ix, Workspace &ws, size_t nq, const float *dQ, const uint32_t *user_ids, const rk_scan_opts *o,
                     const RkSelectParams &out, std::vector<uint32_t> &redo)
{
    redo.clear();
    // option tensor_host_timing: host microseconds per phase of this call on stderr (tuning aid)
    const RkTuning &tune = ix->tune;
    const bool host_timing = tune.tensor_host_timing != 0;
    auto t_prev = std::chrono::steady_clock::now();
    double t_phase[6] = {0, 0, 0, 0, 0, 0};
    auto tick = [&](int i) {
        if (!host_timing) return;
        const auto now = std::chrono::steady_clock::now();
        t_phase[i] += std::chrono::duration<double, std::micro>(now - t_prev).count();
        t_prev = now;
    };
    int rc = tensor_prepare(ix);
    if (rc < 0) return rc;
    TensorState &ts = ix->tensor;
    const int W = ix->mask_words;
    const int k = o->k;
    // nominations per query: the proof needs the kp-th approximate value to clear the exact k-th by the error bound; the
    // coarser bf16 nomination takes a longer list (option tensor_kp_extra adds to either)
    const bool bf16_likely = (ix->tune.tensor_pair >= 0 ? ix->tune.tensor_pair != 0 : ix->dim_pad >= 256) &&
                             (ix->tune.tensor_bf16 == 1 || (ix->tune.tensor_bf16 < 0 && !ix->tensor.bf16_off));
    int kp = std::max(2 * k, k + 128) + ix->tune.tensor_kp_extra + (bf16_likely ? ix->tune.tensor_kp_bf16 : 0);
    // Exact nomination needs no error margin: when the rows are small integers (known since tensor_prepare) and the previous
    // batch's queries were too, the approximate order IS the exact order and k + a few nominations (ties at the k-th value)
    // are enough — fewer appends, fewer compactions, a shorter merge and re-score (C3 batched kernel 2.06 -> 1.93 ms).  A batch
    // that turns out inexact after all is still answered correctly: the proof then works with its usual error bound on the
    // shorter list and refuses more queries (re-run on the scan engine), and the next batch is back to the full list.
    if (ix->tune.tensor_kp_exact >= 0 && ts.x_integer && ts.last_exact)
        kp = std::min(kp, k + std::max(ix->tune.tensor_kp_exact, 4) + ix->tune.tensor_kp_extra);
    if (kp > 512) kp = 512;
    const size_t n = ix->nrows;
    const uint32_t T = (uint32_t) ((n + 127) / 128);
    const int sm = ix->sm_count;
    const bool use_mask = o->filter != RK_FILTER_NONE;
    cudaStream_t st = ws.stream;

    // ---- groups: queries that see the same rows (same role mask) share one pass over those rows
    std::vector<TGroup> groups;
    {
        // users with the same role mask carry the same dense mask id (rk_finalize): grouping is a table lookup
        std::vector<int32_t> gid((size_t) ix->n_mask_ids + 1, -1);
        for (size_t q = 0; q < nq; q++) {
            const UserInfo *u = nullptr;
            size_t slot = ix->n_mask_ids;                            // unfiltered: one group
            if (use_mask) {
                if (user_ids) {
                    auto it = ix->users.find(user_ids[q]);
                    if (it != ix->users.end()) u = &it->second;
                }
                if (!u || u->allowed_rows == 0) continue;          // nothing visible: the query owns no list
                slot = u->mask_id;
            }
            if (gid[slot] < 0) {
                gid[slot] = (int32_t) groups.size();
                groups.emplace_back();
                groups.back().user = u;
                groups.back().min_visible = u ? u->allowed_rows : n;
            }
            groups[(size_t) gid[slot]].queries.push_back((uint32_t) q);
        }
    }
    // Three ways to walk the rows; the plan that stages the fewest bytes into shared memory wins
    // ((128 rows + padded queries) per tile visit):
    //   dense    every tile against all queries together, mask test per (row, query)
    //   grouped  per distinct user mask: only the tiles holding rows that user sees
    //   atoms    per document class (same role mask): its tiles once, against every query that sees it
    // Which scoring kernel: long rows make the one-CTA kernel L2-feed-bound (32 flop per staged byte), so from 256
    // dimensions on the CTA-pair kernel (64 flop per staged byte, query tiles of 256) takes the batch; short rows
    // are bound by the epilogue's select work, which the pair does not change.  Option tensor_pair = 0|1 forces it (tests).
    bool paired = ix->dim_pad >= 256;
    if (tune.tensor_pair >= 0) paired = tune.tensor_pair != 0;
    const size_t qt = paired ? 256 : 128;
    auto visit_cost = [&](double tiles, size_t qn) {
        const double qcols = (double) std::min<size_t>(qt, (qn + 15) & ~(size_t) 15);
        return tiles * (double) ((qn + qt - 1) / qt) * (128.0 + (paired ? 0.5 * qcols : qcols));
    };
    enum { PLAN_DENSE, PLAN_GROUPED, PLAN_ATOMS } plan = PLAN_DENSE;
    bool grew = false;
    std::vector<std::vector<uint32_t>> atom_groups;          // atoms plan: user groups that see each atom
    if (use_mask && !groups.empty()) {
        double cost_g = 0.0, cost_a = 0.0;
        size_t nq_live = 0;
        for (auto &g : groups) {
            rc = tiles_for_user(ix, *g.user, &g.tl_off, &g.tl_n, &grew);
            if (rc < 0) return rc;
            cost_g += visit_cost((double) g.tl_n, g.queries.size());
            nq_live += g.queries.size();
        }
        const double cost_d = visit_cost((double) T, nq_live);
        double best = cost_d;
        if (cost_g < best) { best = cost_g; plan = PLAN_GROUPED; }
        if (!ts.atoms.empty()) {
            atom_groups.resize(ts.atoms.size());
            for (size_t a = 0; a < ts.atoms.size(); a++) {
                size_t qn = 0;
                for (size_t gi = 0; gi < groups.size(); gi++) {
                    uint64_t acc = 0;
                    for (int w = 0; w < W; w++) acc |= ts.atoms[a].mask[(size_t) w] & groups[gi].user->mask[(size_t) w];
                    if (acc) { atom_groups[a].push_back((uint32_t) gi); qn += groups[gi].queries.size(); }
                }
                if (qn) cost_a += visit_cost((double) ts.atoms[a].tl_n, qn);
            }
            if (cost_a < 0.9 * best) { best = cost_a; plan = PLAN_ATOMS; }
        }
        // test aid: option tensor_plan = 1 dense | 2 grouped | 3 atoms forces a plan (all three give the same answers)
        if (tune.tensor_plan == 1) plan = PLAN_DENSE;
        else if (tune.tensor_plan == 2) plan = PLAN_GROUPED;
        else if (tune.tensor_plan == 3 && !ts.atoms.empty()) plan = PLAN_ATOMS;
    }
    const bool per_slot_mask = plan != PLAN_GROUPED;       // which mask a slot carries: its query's own, or the group's
    if (plan == PLAN_DENSE && groups.size() > 1) {
        // one group holding every query that can see something
        TGroup all;
        all.min_visible = n;
        for (auto &g : groups) {
            all.queries.insert(all.queries.end(), g.queries.begin(), g.queries.end());
            all.min_visible = std::min(all.min_visible, g.min_visible);
        }
        std::sort(all.queries.begin(), all.queries.end());
        groups.clear();
        groups.push_back(std::move(all));
    }
    if (plan == PLAN_DENSE)
        for (auto &g : groups) { g.tl_off = 0; g.tl_n = T; }
    if (plan == PLAN_ATOMS) {
        // the units' groups become the atoms; a query appears once per atom it can see
        std::vector<TGroup> ag;
        for (size_t a = 0; a < ts.atoms.size(); a++) {
            if (atom_groups[a].empty()) continue;
            TGroup g;
            g.tl_off = ts.atoms[a].tl_off; g.tl_n = ts.atoms[a].tl_n;
            g.min_visible = ts.atoms[a].rows;
            g.own_mask = &ts.atoms[a].mask;
            for (uint32_t gi : atom_groups[a]) g.queries.insert(g.queries.end(), groups[gi].queries.begin(), groups[gi].queries.end());
            ag.push_back(std::move(g));
        }
        groups.swap(ag);
    }
    if (plan != PLAN_DENSE && (grew || ts.tiles_dirty)) {
        if (ts.d_tiles.ensure(std::max<size_t>(ts.h_tiles.size() * 4, 16)) != 0) return RK_ERR_CUDA;
        RK_CUDA(cudaMemcpyAsync(ts.d_tiles.p, ts.h_tiles.data(), ts.h_tiles.size() * 4, cudaMemcpyHostToDevice, st));
        RK_CUDA(cudaStreamSynchronize(st));      // h_tiles may reallocate on the next call
        ts.tiles_dirty = false;
    }

    tick(0);      // groups + plan choice
    // ---- slots (group-ordered queries; a query owns one slot per group it appears in) and the two passes
    size_t nslots = 0;
    for (auto &g : groups) nslots += g.queries.size();
    const uint32_t LPQ = (uint32_t) rk_bscan_lists_per_query();
    std::vector<uint32_t> slot_q(nslots);
    {
        uint32_t s = 0;
        for (auto &g : groups)
            for (uint32_t q : g.queries) slot_q[s++] = q;
    }
    // Sampling pass: the kp-th smallest value over every S-th tile bounds the kp-th smallest over all
    // tiles from above, so the main pass can start every unit at that threshold instead of +inf: far
    // fewer (query, row) pairs survive the threshold test, and units may be cut as small as load
    // balance wants.
    std::vector<uint32_t> step_main(groups.size(), 1), step_sample(groups.size(), 0);
    bool sampled = false;
    // test / tuning aids: option tensor_sample = 0 disables the sampling pass, tensor_sdiv and tensor_smin move
    // the point where it starts to pay (parity tests use them to exercise it on small corpora)
    const int env_sample = tune.tensor_sample;
    const int env_sdiv = tune.tensor_sdiv > 0 ? tune.tensor_sdiv : 400;
    const uint32_t env_smin = (uint32_t) tune.tensor_smin;
    for (size_t gi = 0; gi < groups.size() && env_sample; gi++) {
        const TGroup &g = groups[gi];
        if (g.tl_n < env_smin) continue;
        // about 1 in 400 of the (visible row, query) pairs should survive the seeded threshold: kp S / visible
        uint64_t S = std::min<uint64_t>(64, g.min_visible / ((uint64_t) env_sdiv * (uint64_t) kp));
        if (S < 4 && tune.tensor_sdiv == 0) {
            // Too few visible rows for that, but a coarser seed still pays: a unit that starts at +inf collects about
            // kp ln(rows / kp) entries per query and stops the whole CTA for a compaction every few tiles.  Seeded
            // from every S-th tile a unit collects about kp S / (ranges the tiles are cut into): keep that near
            // 300, well under one buffer, at a pass that costs 1 / S of the main one.
            const double qblocks = std::max(1.0, std::ceil((double) nslots / (double) qt));
            const double ranges = std::max(1.0, (paired ? sm / 2 : sm) / qblocks);
            S = (uint64_t) std::min(64.0, 300.0 * ranges / (double) kp);
            S = std::min<uint64_t>(S, g.min_visible / (4ull * (uint64_t) kp));     // the sample itself must hold >= 4 kp visible rows
            if (S < 8) S = 0;
        }
        if (S >= 4) { step_sample[gi] = (uint32_t) S; sampled = true; }
    }
    TPass pm_, ps_;
    build_pass(groups, step_main, sm, LPQ, nslots, pm_, paired, tune.tensor_tpu);
    if (sampled) build_pass(groups, step_sample, sm, LPQ, nslots, ps_, paired, tune.tensor_tpu);
    // final merge: every query owns the main-pass lists of all its slots
    std::vector<uint32_t> qoff(nq + 1, 0), qlist;
    {
        for (size_t s_ = 0; s_ < nslots; s_++) qoff[slot_q[s_] + 1] += pm_.nlists[s_];
        for (size_t q = 0; q < nq; q++) qoff[q + 1] += qoff[q];
        qlist.resize(qoff[nq]);
        std::vector<uint32_t> fill(qoff.begin(), qoff.end() - 1);
        for (size_t s_ = 0; s_ < nslots; s_++)
            for (uint32_t j = 0; j < pm_.nlists[s_]; j++) qlist[fill[slot_q[s_]]++] = pm_.lid0[s_] + j * pm_.lstep[s_];
    }

    tick(1);      // passes built
    // ---- staging: per pass [units][cta_off][cta_units][lid0][lstep][nlists], then [slot_q][qoff][qlist][qmask per slot]
    const size_t b_slot = nslots * 4, b_qmask = nslots * (size_t) W * 8;
    auto pass_bytes = [&](const TPass &t) { return t.units.size() * sizeof(RkBUnit) + (t.cta_off.size() + t.cta_units.size()) * 4 + 3 * b_slot + 128; };
    const size_t total_stage = pass_bytes(pm_) + (sampled ? pass_bytes(ps_) : 0) + b_slot + (qoff.size() + qlist.size()) * 4 + b_qmask +
                               groups.size() * (size_t) W * 8 + 256;
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
    auto stage_pass = [&](TPass &t) {
        t.o_units = put(t.units.data(), t.units.size() * sizeof(RkBUnit));
        t.o_coff = put(t.cta_off.data(), t.cta_off.size() * 4);
        t.o_cu = put(t.cta_units.data(), t.cta_units.size() * 4);
        t.o_lid0 = put(t.lid0.data(), b_slot);
        t.o_lstep = put(t.lstep.data(), b_slot);
        t.o_nl = put(t.nlists.data(), b_slot);
    };
    stage_pass(pm_);
    if (sampled) stage_pass(ps_);
    size_t o_umask = 0;
    if (plan == PLAN_ATOMS) {
        std::vector<uint64_t> um(groups.size() * (size_t) W);
        for (size_t gi = 0; gi < groups.size(); gi++)
            for (int w = 0; w < W; w++) um[gi * W + w] = (*groups[gi].own_mask)[(size_t) w];
        o_umask = put(um.data(), um.size() * 8);
    }
    const size_t o_slotq = put(slot_q.data(), b_slot);
    const size_t o_qoff = put(qoff.data(), qoff.size() * 4);
    const size_t o_qlist = put(qlist.data(), qlist.size() * 4);
    const size_t o_qmask = off;
    {
        uint64_t *hm = reinterpret_cast<uint64_t *>(hp + o_qmask);
        uint32_t s_ = 0;
        for (auto &g : groups) {
            for (uint32_t q : g.queries) {
                const UserInfo *u = g.user;
                if (per_slot_mask && use_mask && user_ids) {
                    auto it = ix->users.find(user_ids[q]);
                    u = it == ix->users.end() ? nullptr : &it->second;
                }
                for (int w = 0; w < W; w++) hm[(size_t) s_ * W + w] = u ? u->mask[(size_t) w] : 0ull;
                s_++;
            }
        }
        off += b_qmask;
    }
    const unsigned char *dplan = ts.d_plan.as<unsigned char>();
    const uint32_t list_max = std::max(pm_.list_total, ps_.list_total);
    const int grid_max = std::max(pm_.grid, ps_.grid);
    if (ts.d_qp.ensure(std::max<size_t>(nslots * (size_t) ix->dim_pad * 4, 16)) != 0 ||
        ts.d_qnorm.ensure(std::max<size_t>(nq * 4, 16)) != 0 || ts.d_scratch.ensure(paired ? rk_bscan2_scratch_bytes(grid_max) : rk_bscan_scratch_bytes(grid_max)) != 0 ||
        ts.d_cand_key.ensure(std::max<size_t>((size_t) list_max * kp * 4, 16)) != 0 ||
        ts.d_cand_row.ensure(std::max<size_t>((size_t) list_max * kp * 4, 16)) != 0 ||
        ts.d_cand_count.ensure(std::max<size_t>((size_t) list_max * 4, 16)) != 0 ||
        ts.d_sel_key.ensure(std::max<size_t>(nq * (size_t) kp * 4, 16)) != 0 ||
        ts.d_sel_row.ensure(std::max<size_t>(nq * (size_t) kp * 4, 16)) != 0 || ts.d_sel_count.ensure(std::max<size_t>(nq * 4, 16)) != 0 ||
        ts.d_tau_key.ensure(std::max<size_t>(nslots * (size_t) kp * 4, 16)) != 0 || ts.d_tau_count.ensure(std::max<size_t>(nslots * 4, 16)) != 0 ||
        ts.d_flags.ensure(std::max<size_t>(4 * nq * 4, 16)) != 0 || ts.h_flags.ensure(std::max<size_t>(nq * 4 + 128, 128)) != 0)
        return RK_ERR_CUDA;
    RK_CUDA(cudaMemcpyAsync(ts.d_plan.p, hp, off, cudaMemcpyHostToDevice, st));
    RK_CUDA(cudaEventRecord(ws.ev_plan[pi_], st));
    ws.ev_plan_used[pi_] = true;
    RK_CUDA(cudaMemsetAsync(ts.d_flags.p, 0, 3 * nq * 4, st));      // [uncertain][trunc][sample-pass trunc], each per query
    RK_CUDA(cudaMemsetAsync(ts.d_flags.as<int32_t>() + 3 * nq, 0xFF, nq * 4, st));      // [shared thresholds of the main pass] = RK_KEY_SENTINEL
    RK_CUDA(cudaMemsetAsync(ts.d_probe.p, 0, 64, st));
    uint64_t launches = 0;

    // norms and TF32-exactness of the queries; the group-ordered copy of the query matrix the TMA reads
    const uint32_t *d_slotq = reinterpret_cast<const uint32_t *>(dplan + o_slotq);
    RK_CUDA(rk_launch_row_norms(dQ, nq, ix->dim_pad, ts.d_qnorm.as<float>(), st));
    RK_CUDA(rk_launch_tf32_probe(dQ, nq * (size_t) ix->dim_pad, ts.d_probe.as<unsigned long long>(), st));
    RK_CUDA(rk_launch_max_f32(ts.d_qnorm.as<float>(), nq, ts.d_probe.as<unsigned>() + 8, st));
    launches += 3;
    unsigned long long *hprobe = reinterpret_cast<unsigned long long *>(ts.h_flags.as<unsigned char>() + ((nq * 4 + 15) & ~(size_t) 15));
    RK_CUDA(cudaMemcpyAsync(hprobe, ts.d_probe.p, 40, cudaMemcpyDeviceToHost, st));
    // (the probe result is only needed for the rescore launch: the scoring passes are queued first)
    const float *dQp = ts.d_qp.as<float>();
    if (nslots) { RK_CUDA(rk_launch_gather_rows(dQ, ts.d_qp.as<float>(), d_slotq, nslots, ix->dim_pad, st)); launches++; }
    // bf16 nomination (CTA-pair kernel only; option tensor_bf16: 1 = on, 0 / -1 = TF32): half the operand bytes through the
    // shared-memory port that bounds the TF32 main pass and twice the MMA rate.  The exact fp32 re-score and the sufficiency
    // proof make the ANSWER independent of the nomination precision; what changes is the proof's margin (eps below), i.e.
    // how often a query has to be re-run.  The shadow copy of the row store is built on first use.
    // -1 (default): on for the pair kernel until a batch has had more than 2 % of its queries refused by the proof — the
    // margin depends on the data — after which this index goes back to TF32; 1 = always, 0 = never
    const bool bf16 = paired && (tune.tensor_bf16 == 1 || (tune.tensor_bf16 < 0 && !ts.bf16_off));
    if (bf16) {
        const int dim_b = (ix->dim_pad + 7) & ~7;
        if (ts.dim_b != dim_b || ts.d_xb.p == nullptr) {
            if (ts.d_xb.ensure((n + 256) * (size_t) dim_b * 2) != 0) return RK_ERR_CUDA;
            RK_CUDA(rk_launch_rows_to_bf16(ix->d_x.as<float>(), n, ix->dim_pad, dim_b, ts.d_xb.p, st));
            ts.dim_b = dim_b;
            ix->device_bytes += ts.d_xb.cap;
            launches++;
        }
        if (ts.d_qpb.ensure(std::max<size_t>(nslots * (size_t) dim_b * 2, 16)) != 0) return RK_ERR_CUDA;
        if (nslots) { RK_CUDA(rk_launch_rows_to_bf16(dQp, nslots, ix->dim_pad, dim_b, ts.d_qpb.p, st)); launches++; }
    }
    const void *x_nom = bf16 ? ts.d_xb.p : ix->d_x.p;
    const void *q_nom = bf16 ? ts.d_qpb.p : (const void *) dQp;

    tick(2);      // staged, uploads and query preparation enqueued
    RkBScanParams bp;
    memset(&bp, 0, sizeof bp);
    bp.tile_rows = plan != PLAN_DENSE ? ts.d_tiles.as<uint32_t>() : nullptr;
    bp.nrows = n;
    bp.dim_pad = ix->dim_pad;
    bp.kchunks = bf16 ? (ts.dim_b + 63) / 64 : (ix->dim_pad + 31) / 32;
    bp.bf16 = bf16 ? 1 : 0;
    bp.loose_extra = (uint32_t) ix->tune.tensor_loose;
    bp.round_mark = (uint32_t) std::max(ix->tune.tensor_round_mark, kp + 64);      // never compact a buffer that holds little more than kp
    bp.metric = ix->metric;
    bp.kp = kp;
    bp.mask_words = W;
    bp.use_mask = use_mask ? 1 : 0;
    bp.xnorm = ts.d_xnorm.as<float>();
    bp.rowmask = ts.d_rowmask.as<uint64_t>();
    bp.rowmask2 = ts.d_rowmask2.as<uint64_t>();
    bp.qnorm = nullptr;
    bp.qmask = reinterpret_cast<const uint64_t *>(dplan + o_qmask);
    bp.slot_q = d_slotq;
    bp.unit_mask = plan == PLAN_ATOMS ? reinterpret_cast<const uint64_t *>(dplan + o_umask) : nullptr;
    bp.scratch = ts.d_scratch.as<uint2>();
    bp.cand_key = ts.d_cand_key.as<uint32_t>();
    bp.cand_row = ts.d_cand_row.as<uint32_t>();
    bp.cand_count = ts.d_cand_count.as<int32_t>();
    auto bind_pass = [&](const TPass &t) {
        bp.units = reinterpret_cast<const RkBUnit *>(dplan + t.o_units);
        bp.nunits = (uint32_t) t.units.size();
        bp.cta_off = reinterpret_cast<const uint32_t *>(dplan + t.o_coff);
        bp.cta_units = reinterpret_cast<const uint32_t *>(dplan + t.o_cu);
    };
    // merge of nominations: top-kp by (approximate key, row)
    RkSelectParams sp;
    memset(&sp, 0, sizeof sp);
    sp.cand_key = ts.d_cand_key.as<uint32_t>();
    sp.cand_tid = nullptr;
    sp.cand_row = ts.d_cand_row.as<uint32_t>();
    sp.cand_count = ts.d_cand_count.as<int32_t>();
    sp.list_stride = kp;
    sp.k = kp;
    sp.metric = ix->metric;
    sp.dedup = 0;
    sp.unsorted = 1;       // unit lists come unsorted, largest entry last (bs_warp_emit_unsorted)
    sp.out_maxlast = 1;    // both consumers (the threshold seed, the exact re-score) need the kp best as a set + its largest value

    const int slot_ev = (int) (ws.ring_n % RK_TIMING_RING);
    if (ws.timing) RK_CUDA(cudaEventRecord(ws.ring_a[slot_ev], st));
    if (sampled) {
        bind_pass(ps_);
        bp.trunc = ts.d_flags.as<int32_t>() + 2 * nq;      // ties dropped here only loosen the seed
        bp.tau_key = nullptr; bp.tau_count = nullptr;
        bp.qtau = nullptr;
        if (paired) RK_CUDA(rk_launch_bscan2(bp, x_nom, q_nom, nslots, ps_.grid, st, ts.dim_b));
        else RK_CUDA(rk_launch_bscan(bp, ix->d_x.as<float>(), dQp, nslots, ps_.grid, st));
        // per slot: the sample's kp smallest values
        sp.lid0 = reinterpret_cast<const uint32_t *>(dplan + ps_.o_lid0);
        sp.lstep = reinterpret_cast<const uint32_t *>(dplan + ps_.o_lstep);
        sp.nlists = reinterpret_cast<const uint32_t *>(dplan + ps_.o_nl);
        sp.out_key = ts.d_tau_key.as<float>();
        sp.out_row = nullptr;
        sp.out_count = ts.d_tau_count.as<int32_t>();
        RK_CUDA(rk_launch_select(sp, (int) nslots, st));
        launches += 2;
    }
    if (nslots) {
        bind_pass(pm_);
        bp.trunc = ts.d_flags.as<int32_t>() + nq;
        bp.qtau = ix->tune.tensor_qtau ? reinterpret_cast<uint32_t *>(ts.d_flags.as<int32_t>() + 3 * nq) : nullptr;
        bp.tau_key = sampled ? ts.d_tau_key.as<float>() : nullptr;
        bp.tau_count = sampled ? ts.d_tau_count.as<int32_t>() : nullptr;
        if (paired) RK_CUDA(rk_launch_bscan2(bp, x_nom, q_nom, nslots, pm_.grid, st, ts.dim_b));
        else RK_CUDA(rk_launch_bscan(bp, ix->d_x.as<float>(), dQp, nslots, pm_.grid, st));
        launches++;
    }
    if (ws.timing) { RK_CUDA(cudaEventRecord(ws.ring_b[slot_ev], st)); ws.timed_scan = true; ws.ring_n++; }

    // per query: the kp best nominations over all its slots' lists
    sp.lid0 = nullptr; sp.lstep = nullptr; sp.nlists = nullptr;
    sp.qseg = reinterpret_cast<const uint32_t *>(dplan + o_qoff);
    sp.list_ids = reinterpret_cast<const uint32_t *>(dplan + o_qlist);
    sp.out_key = ts.d_sel_key.as<float>();
    sp.out_row = ts.d_sel_row.as<uint32_t>();
    sp.out_count = ts.d_sel_count.as<int32_t>();
    if (ws.timing) RK_CUDA(cudaEventRecord(ws.ev[2], st));
    RK_CUDA(rk_launch_select(sp, (int) nq, st));
    launches++;

    // TF32 keeps 10 mantissa bits and the MMA truncates: each operand is off by < 2^-10 relative, a
    // product by < 2^-9; |x.q| error <= 2^-9 * |x||q|; the L2 form doubles it.  Whether this batch is exact instead
    // (small-integer corpus and queries) is decided by the re-score kernel from the probe results on the device.
    // bf16 (round to nearest, 8 significant bits): each operand off by <= 2^-8 relative (half a unit in the last place), a
    // product by < 2^-7.
    const float eps_scale = (ix->metric == RK_METRIC_L2 ? 2.f : 1.f) * (bf16 ? 1.0f / 128.0f : 1.0f / 512.0f);

    // exact re-score of every query's nominations in the caller's order (a query without lists gets an
    // empty result)
    RkRescoreParams rp;
    memset(&rp, 0, sizeof rp);
    rp.x = ix->d_x.as<float>();
    rp.tid = ix->d_tid.as<uint64_t>();
    rp.rowdoc = ix->d_rowdoc.as<uint32_t>();
    rp.doc_ids = ix->d_doc_ids.as<uint32_t>();
    rp.q = dQ;
    rp.qnorm = ts.d_qnorm.as<float>();
    rp.slot_q = nullptr;
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
    rp.probe = ts.d_probe.as<unsigned long long>();
    rp.x_integer = ts.x_integer ? 1 : 0;
    rp.exact_absmax = bf16 ? 256.f : 2047.f;
    rp.x_absmax = ts.x_absmax;
    rp.wide_rows = ix->tile_rows < 64 ? 1 : 0;
    rp.xnorm_max = ts.xnorm_max;
    rp.out_tid = out.out_tid;
    rp.out_doc = out.out_doc;
    rp.out_dist = out.out_dist;
    rp.out_key = out.out_key;
    rp.out_count = out.out_count;
    rp.uncertain = ts.d_flags.as<int32_t>();
    RK_CUDA(rk_launch_rescore(rp, (int) nq, st));
    launches++;
    if (ws.timing) { RK_CUDA(cudaEventRecord(ws.ev[3], st)); ws.timed_select = true; }
    int32_t *hflags = ts.h_flags.as<int32_t>();
    RK_CUDA(cudaMemcpyAsync(hflags, ts.d_flags.p, nq * 4, cudaMemcpyDeviceToHost, st));
    tick(3);      // everything enqueued
    RK_CUDA(cudaStreamSynchronize(st));
    tick(4);      // device done
    for (size_t q = 0; q < nq; q++)
        if (hflags[q]) redo.push_back((uint32_t) q);
    if (bf16 && tune.tensor_bf16 < 0 && nq >= 50 && redo.size() * 50 > nq) ts.bf16_off = true;
    float qnorm_max;
    { uint32_t b = (uint32_t) (hprobe[4] & 0xFFFFFFFFull); memcpy(&qnorm_max, &b, 4); }
    const double nx = std::sqrt((double) ts.xnorm_max), nqm = std::sqrt((double) qnorm_max);
    float q_absmax;
    { uint32_t b = (uint32_t) (hprobe[1] & 0xFFFFFFFFull); memcpy(&q_absmax, &b, 4); }
    const bool exact = ts.x_integer && hprobe[0] == 0 && (nx + nqm) * (nx + nqm) < 16777216.0 &&
                       std::max(ts.x_absmax, q_absmax) <= (bf16 ? 256.f : 2047.f);      // (statistics only; the re-score kernel decides)

    ws.launches += launches;
    ws.segments = pm_.units.size() + ps_.units.size();
    const uint64_t tile_visits = (uint64_t) (pm_.tile_visits + 0.5);
    ws.rows_scanned = tile_visits * 128;
    // SURVEY.md §8(d), batched exact: bytes 4d (R_g + nq_g) over the groups the plan forms — every row a
    // group sees once (min_visible is exact for every plan: all rows / the user's rows / the atom's rows)
    uint64_t group_rows = 0;
    for (auto &g : groups) group_rows += plan == PLAN_DENSE ? n : g.min_visible;
    ws.algo_bytes = (uint64_t) ix->dim_pad * 4ull * (group_rows + nslots);
    // what the plan actually stages: every visited 128-row tile once per 128-query tile, plus the queries
    ts.last_plan_bytes = (uint64_t) ix->dim_pad * 4ull * (tile_visits * 128ull + nslots);
    // SURVEY.md §8(d), batched exact: F = 2 nq_g R_g d over the groups the plan forms (main pass)
    ts.last_flop = 2.0 * (double) ix->dim * pm_.pairs;
    ts.last_redo = (uint32_t) redo.size();
    ts.last_exact = exact;
    if (host_timing)
        fprintf(stderr, "rk_tensor_search host us: groups %.0f passes %.0f stage %.0f launch %.0f wait %.0f\n", t_phase[0], t_phase[1],
                t_phase[2], t_phase[3], t_phase[4]);
    ts.last_plan = 1 + (int) plan + (sampled ? 16 : 0) + (paired ? 32 : 0) + (bf16 ? 64 : 0);
    return RK_OK;
}
