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
    const uint32_t QT = (uint32_t) ((nq + 127) / 128);
    const int sm = ix->sm_count;
    // row ranges: the smallest R (longest units, fewest per-unit costs) whose static schedule keeps
    // >= 90 % of the SMs busy; otherwise the best one found
    uint32_t Rmax = std::max<uint32_t>(1, T / 16);
    uint32_t bestR = 1;
    double bestEff = 0.0;
    for (uint32_t R = 1; R <= Rmax && R <= 4096; R++) {
        const uint64_t units = (uint64_t) R * QT;
        const uint64_t rounds = (units + sm - 1) / sm;
        const double eff = (double) units / (double) (rounds * sm);
        if (eff > bestEff + 1e-9) { bestEff = eff; bestR = R; }
        if (eff >= 0.9) break;
    }
    const uint32_t R = bestR;
    const uint32_t nunits = R * QT;
    const int grid = (int) std::min<uint32_t>((uint32_t) sm, nunits);

    // ---- staging: units, user masks
    const size_t b_units = (size_t) nunits * sizeof(RkBUnit), b_qmask = nq * (size_t) W * 8;
    ws.plan_idx ^= 1;
    const int pi_ = ws.plan_idx;
    if (ws.ev_plan_used[pi_]) RK_CUDA(cudaEventSynchronize(ws.ev_plan[pi_]));
    PinBuf &hp_ = ws.h_plan_ring[pi_];
    if (hp_.ensure(b_units + b_qmask + 64) != 0) return RK_ERR_CUDA;
    RkBUnit *hu = hp_.as<RkBUnit>();
    for (uint32_t r = 0; r < R; r++)
        for (uint32_t qt = 0; qt < QT; qt++) {
            RkBUnit &u = hu[(size_t) r * QT + qt];                 // range-major: lists of one query are R blocks apart
            u.q0 = qt * 128;
            u.nq = (uint32_t) std::min<size_t>(128, nq - (size_t) qt * 128);
            u.tile_begin = (uint32_t) ((uint64_t) T * r / R);
            u.tile_end = (uint32_t) ((uint64_t) T * (r + 1) / R);
        }
    uint64_t *hm = reinterpret_cast<uint64_t *>(reinterpret_cast<unsigned char *>(hu) + ((b_units + 15) & ~(size_t) 15));
    const bool use_mask = o->filter != RK_FILTER_NONE;
    for (size_t q = 0; q < nq; q++) {
        const UserInfo *u = nullptr;
        if (user_ids && use_mask) {
            auto it = ix->users.find(user_ids[q]);
            if (it != ix->users.end()) u = &it->second;
        }
        for (int w = 0; w < W; w++) hm[q * W + w] = u ? u->mask[w] : 0ull;
    }
    const size_t nlists = (size_t) nunits * 128;
    if (ts.d_units.ensure(std::max<size_t>(b_units, 16)) != 0 || ws.qmask.ensure(std::max<size_t>(b_qmask, 16)) != 0 ||
        ts.d_qnorm.ensure(std::max<size_t>(nq * 4, 16)) != 0 || ts.d_scratch.ensure(rk_bscan_scratch_bytes(grid)) != 0 ||
        ts.d_cand_key.ensure(nlists * (size_t) kp * 4) != 0 || ts.d_cand_row.ensure(nlists * (size_t) kp * 4) != 0 ||
        ts.d_cand_count.ensure(nlists * 4) != 0 || ts.d_sel_key.ensure(std::max<size_t>(nq * (size_t) kp * 4, 16)) != 0 ||
        ts.d_sel_row.ensure(std::max<size_t>(nq * (size_t) kp * 4, 16)) != 0 || ts.d_sel_count.ensure(std::max<size_t>(nq * 4, 16)) != 0 ||
        ts.d_flags.ensure(std::max<size_t>(nq * 8, 16)) != 0 || ts.h_flags.ensure(std::max<size_t>(nq * 4 + 128, 128)) != 0)
        return RK_ERR_CUDA;
    cudaStream_t st = ws.stream;
    RK_CUDA(cudaMemcpyAsync(ts.d_units.p, hu, b_units, cudaMemcpyHostToDevice, st));
    if (b_qmask) RK_CUDA(cudaMemcpyAsync(ws.qmask.p, hm, b_qmask, cudaMemcpyHostToDevice, st));
    RK_CUDA(cudaEventRecord(ws.ev_plan[pi_], st));
    ws.ev_plan_used[pi_] = true;
    RK_CUDA(cudaMemsetAsync(ts.d_cand_count.p, 0, nlists * 4, st));
    RK_CUDA(cudaMemsetAsync(ts.d_flags.p, 0, nq * 8, st));           // [uncertain nq][trunc nq]
    RK_CUDA(cudaMemsetAsync(ts.d_probe.p, 0, 64, st));
    // query norms and the TF32-exactness of the queries
    RK_CUDA(rk_launch_row_norms(dQ, nq, ix->dim_pad, ts.d_qnorm.as<float>(), st));
    RK_CUDA(rk_launch_tf32_probe(dQ, nq * (size_t) ix->dim_pad, ts.d_probe.as<unsigned long long>(), st));
    RK_CUDA(rk_launch_max_f32(ts.d_qnorm.as<float>(), nq, ts.d_probe.as<unsigned>() + 8, st));
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
    bp.units = ts.d_units.as<RkBUnit>();
    bp.nunits = nunits;
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
    bp.qmask = ws.qmask.as<uint64_t>();
    bp.scratch = ts.d_scratch.as<uint2>();
    bp.cand_key = ts.d_cand_key.as<uint32_t>();
    bp.cand_row = ts.d_cand_row.as<uint32_t>();
    bp.cand_count = ts.d_cand_count.as<int32_t>();
    bp.trunc = ts.d_flags.as<int32_t>() + nq;
    const int slot = (int) (ws.ring_n % RK_TIMING_RING);
    if (ws.timing) RK_CUDA(cudaEventRecord(ws.ring_a[slot], st));
    RK_CUDA(rk_launch_bscan(bp, ix->d_x.as<float>(), dQ, nq, grid, st));
    if (ws.timing) { RK_CUDA(cudaEventRecord(ws.ring_b[slot], st)); ws.timed_scan = true; ws.ring_n++; }

    // merge the R nominations of each query: top-kp by (approximate key, row)
    RkSelectParams sp;
    memset(&sp, 0, sizeof sp);
    sp.cand_key = ts.d_cand_key.as<uint32_t>();
    sp.cand_tid = nullptr;
    sp.cand_row = ts.d_cand_row.as<uint32_t>();
    sp.cand_count = ts.d_cand_count.as<int32_t>();
    sp.lists_per_query = (int) R;
    sp.list_stride = kp;
    sp.count_stride = (int) (QT * 128);
    sp.shard_stride_bytes = (size_t) QT * 128 * (size_t) kp * 4;
    sp.k = kp;
    sp.metric = ix->metric;
    sp.dedup = 0;
    sp.out_key = ts.d_sel_key.as<float>();
    sp.out_row = ts.d_sel_row.as<uint32_t>();
    sp.out_count = ts.d_sel_count.as<int32_t>();
    if (ws.timing) RK_CUDA(cudaEventRecord(ws.ev[2], st));
    RK_CUDA(rk_launch_select(sp, (int) nq, st));

    RkRescoreParams rp;
    memset(&rp, 0, sizeof rp);
    rp.x = ix->d_x.as<float>();
    rp.tid = ix->d_tid.as<uint64_t>();
    rp.rowdoc = ix->d_rowdoc.as<uint32_t>();
    rp.doc_ids = ix->d_doc_ids.as<uint32_t>();
    rp.q = dQ;
    rp.qnorm = ts.d_qnorm.as<float>();
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
    RK_CUDA(rk_launch_rescore(rp, (int) nq, st));
    if (ws.timing) { RK_CUDA(cudaEventRecord(ws.ev[3], st)); ws.timed_select = true; }
    int32_t *hflags = ts.h_flags.as<int32_t>();
    RK_CUDA(cudaMemcpyAsync(hflags, ts.d_flags.p, nq * 4, cudaMemcpyDeviceToHost, st));
    RK_CUDA(cudaStreamSynchronize(st));
    for (size_t q = 0; q < nq; q++)
        if (hflags[q]) redo.push_back((uint32_t) q);

    ws.launches += 6;
    ws.segments = nunits;
    ws.rows_scanned = (uint64_t) n * nq;
    // dense contraction: 2 * nq * rows * dim flop; bytes = 4d * (rows * QT + nq)
    ws.algo_bytes = (uint64_t) ix->dim_pad * 4ull * ((uint64_t) n * QT + nq);
    ts.last_flop = 2.0 * (double) nq * (double) n * (double) ix->dim;
    ts.last_redo = (uint32_t) redo.size();
    ts.last_exact = exact;
    return RK_OK;
}
