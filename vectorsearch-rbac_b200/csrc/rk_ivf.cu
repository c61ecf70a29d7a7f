// rk_ivf.cu — IVFFlat host side: CREATE INDEX (sample, k-means, assign, pack lists) and the list-probe
// search, behind rk_build_ivf / rk_build_ivf_from_centers / rk_ivf_info and the probes > 0 branch
// of every search entry point.
//
// Reference (paths relative to the reference checkout):
//   build ...... pgvector/src/ivfbuild.c:404-441 (sample max(10000, 50*lists) rows), ivfkmeans.c (k-means;
//                spherical for vector_ip_ops: sql/vector.sql:298-303 FUNCTION 3/4), ivfbuild.c:141-198
//                (assign every tuple to its nearest centre by FUNCTION 1), :251-310 (pack lists)
//   scan ....... pgvector/src/ivfscan.c:36-107 (choose lists), :112-176 (score every tuple of the
//                probed lists), :257-266 (probe clamps), :375-381 (iterative refill)
// The executor's WHERE / RLS post-filter on what the index returns becomes the
// role-mask test inside the list scan: same rows, same order.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <map>
#include <random>
#include <unordered_set>

#include "rk_host.h"

#define RK_CUDA(call)                                                                                    \
    do {                                                                                                 \
        cudaError_t e__ = (call);                                                                        \
        if (e__ != cudaSuccess)                                                                          \
            return rk_fail(RK_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), __FILE__, __LINE__); \
    } while (0)

#define RK_IVF_MAX_LISTS_SMEM 16384   // the probe kernel sorts (key, list) pairs of one query in shared memory
#define RK_IVF_MAX_SEGS (1u << 18)    // segments per sub-batch (bounds the candidate buffers)

// Nearest centre of n rows (ivfbuild.c:141-198, ivfkmeans.c:258-494): on the tensor cores when the shape allows
// (rk_assign_tc.cu: TF32 nomination + exact re-score + proof, refused rows through the fp32 kernel — identical results), else
// the fp32 kernel.
static int assign_rows(rk_index *ix, const float *X, size_t n, const float *C, int lists, int metric, uint32_t *out, cudaStream_t st)
{
    if (ix->tune.ivf_assign_tc && rk_assign_tc_supported(n, lists, ix->dim_pad)) {
        const size_t chunk = std::min<size_t>(n, (size_t) 1 << 21);
        DevBuf scratch;
        if (scratch.ensure(rk_assign_tc_scratch_bytes(chunk, lists)) != 0) return RK_ERR_CUDA;
        size_t nf = 0;
        cudaError_t e = rk_assign_tc(X, n, C, lists, ix->dim_pad, metric, out, nullptr, scratch.p, chunk, ix->sm_count, st, &nf);
        scratch.release();
        if (e != cudaSuccess) return rk_fail(RK_ERR_CUDA, "tensor-core assignment failed: %s", cudaGetErrorString(e));
        ix->ivf.assign_tc_rows += n;
        ix->ivf.assign_fallback_rows += nf;
        return RK_OK;
    }
    RK_CUDA(rk_launch_assign(X, n, C, lists, ix->dim_pad, metric, out, nullptr, st));
    return RK_OK;
}

// Upload padded centres, assign every row (or take the caller's assignment, `given` by load order), pack the list-sorted
// copy of the store.
static int build_from_device_centers(rk_index *ix, int lists, const float *d_centers, const uint32_t *given = nullptr)
{
    IvfIndex &iv = ix->ivf;
    const size_t n = ix->nrows;
    const size_t rb = (size_t) ix->dim_pad * 4;
    cudaStream_t st = 0;
    iv.row_list.assign(n, 0);
    if (given) {
        for (size_t i = 0; i < n; i++) {
            if (given[i] >= (uint32_t) lists) return rk_fail(RK_ERR_INVALID, "rk_build_ivf_from_lists: row %zu is in list %u of %d", i, given[i], lists);
            iv.row_list[ix->load_to_row[i]] = given[i];
        }
    } else {
        DevBuf d_row_list;
        if (d_row_list.ensure(std::max<size_t>(n * 4, 16)) != 0) return RK_ERR_CUDA;
        { const int rc_ = assign_rows(ix, ix->d_x.as<float>(), n, d_centers, lists, ix->metric, d_row_list.as<uint32_t>(), st); if (rc_ < 0) return rc_; }
        if (n) RK_CUDA(cudaMemcpy(iv.row_list.data(), d_row_list.p, n * 4, cudaMemcpyDeviceToHost));
        d_row_list.release();
    }
    // Document classes: documents with the same role mask are visible to exactly the same users.  With at most 128 of them
    // (role hierarchies: one per role) every list keeps its rows ordered by class, so the rows a user may see in a list are
    // one extent per visible class and the list scan needs no per-row permission test.  Class `ncls` holds the documents
    // nobody can see.  (Row order inside a list is not observable: results are ordered by (distance, tuple id).)
    const int W = ix->mask_words;
    const size_t D = ix->doc_ids.size();
    std::vector<uint32_t> doc_cls(D, 0);
    iv.ncls = 0;
    iv.cls_mask.clear();
    {
        std::map<std::vector<uint64_t>, uint32_t> by_mask;
        std::vector<uint64_t> key((size_t) W);
        bool ok = true;
        for (size_t dd = 0; dd < D && ok; dd++) {
            bool any = false;
            for (int w = 0; w < W; w++) { key[(size_t) w] = ix->doc_mask[dd * W + w]; any |= key[(size_t) w] != 0; }
            if (!any) { doc_cls[dd] = 0xFFFFFFFFu; continue; }
            auto ins = by_mask.emplace(key, (uint32_t) by_mask.size());
            if (ins.second) {
                if (by_mask.size() > 128) { ok = false; break; }
                iv.cls_mask.push_back(key);
            }
            doc_cls[dd] = ins.first->second;
        }
        if (ok && !by_mask.empty()) iv.ncls = (int) by_mask.size();
        else iv.cls_mask.clear();
    }
    const uint32_t NB = (uint32_t) iv.ncls + 1;            // buckets per list (ncls == 0: one bucket = the list itself)
    std::vector<uint32_t> row_cls;
    if (iv.ncls) {
        row_cls.resize(n);
        for (size_t dd = 0; dd < D; dd++) {
            const uint32_t c = doc_cls[dd] == 0xFFFFFFFFu ? (uint32_t) iv.ncls : doc_cls[dd];
            for (uint64_t r = ix->doc_row_start[dd]; r < ix->doc_row_start[dd + 1]; r++) row_cls[r] = c;
        }
    }
    // counting sort by (list, class), stable in finalized row order (ivfbuild.c:251-310 packs tuples list by list)
    std::vector<uint32_t> bucket_start((size_t) lists * NB + 1, 0);
    for (size_t r = 0; r < n; r++) bucket_start[(size_t) iv.row_list[r] * NB + (iv.ncls ? row_cls[r] : 0u) + 1]++;
    for (size_t b = 0; b < (size_t) lists * NB; b++) bucket_start[b + 1] += bucket_start[b];
    iv.list_start.assign((size_t) lists + 1, 0);
    iv.max_list_len = 0;
    for (int l = 0; l <= lists; l++) iv.list_start[l] = bucket_start[(size_t) l * NB];
    for (int l = 0; l < lists; l++) iv.max_list_len = std::max(iv.max_list_len, iv.list_start[l + 1] - iv.list_start[l]);
    std::vector<uint32_t> perm(n), cursor(bucket_start.begin(), bucket_start.end() - 1);
    for (size_t r = 0; r < n; r++) perm[cursor[(size_t) iv.row_list[r] * NB + (iv.ncls ? row_cls[r] : 0u)]++] = (uint32_t) r;
    if (iv.ncls) {
        if (iv.d_cls_off.ensure(bucket_start.size() * 4) != 0) return RK_ERR_CUDA;
        RK_CUDA(cudaMemcpy(iv.d_cls_off.p, bucket_start.data(), bucket_start.size() * 4, cudaMemcpyHostToDevice));
        // the 128-bit set of classes each distinct user role mask may see, computed HERE — one thread, n_mask_ids known —
        // so concurrent scans (any number of rk_scan handles, include/rbacknn.h:17-19) only ever read the table
        iv.qcls_by_mask_id.assign((size_t) ix->n_mask_ids * 2, 0ull);
        std::vector<char> done(ix->n_mask_ids, 0);
        for (auto &kv : ix->users) {
            const UserInfo &u = kv.second;
            if (done[u.mask_id]) continue;
            done[u.mask_id] = 1;
            uint64_t a0 = 0, a1 = 0;
            for (int c = 0; c < iv.ncls; c++) {
                uint64_t acc = 0;
                for (int w = 0; w < W; w++) acc |= iv.cls_mask[(size_t) c][(size_t) w] & u.mask[(size_t) w];
                if (acc) { if (c < 64) a0 |= 1ull << c; else a1 |= 1ull << (c - 64); }
            }
            iv.qcls_by_mask_id[(size_t) u.mask_id * 2] = a0;
            iv.qcls_by_mask_id[(size_t) u.mask_id * 2 + 1] = a1;
        }
    }
    DevBuf d_perm;
    if (d_perm.ensure(std::max<size_t>(n * 4, 16)) != 0) return RK_ERR_CUDA;
    if (n) RK_CUDA(cudaMemcpy(d_perm.p, perm.data(), n * 4, cudaMemcpyHostToDevice));
    if (iv.d_x.ensure(std::max<size_t>(n * rb, 16)) != 0 || iv.d_tid.ensure(std::max<size_t>(n * 8, 16)) != 0 ||
        iv.d_rowdoc.ensure(std::max<size_t>(n * 4, 16)) != 0 || iv.d_list_start.ensure(((size_t) lists + 1) * 4) != 0)
        return RK_ERR_CUDA;
    RK_CUDA(rk_launch_gather_rows(ix->d_x.as<float>(), iv.d_x.as<float>(), d_perm.as<uint32_t>(), n, ix->dim_pad, st));
    RK_CUDA(rk_launch_gather_u64(ix->d_tid.as<uint64_t>(), iv.d_tid.as<uint64_t>(), d_perm.as<uint32_t>(), n, st));
    RK_CUDA(rk_launch_gather_u32(ix->d_rowdoc.as<uint32_t>(), iv.d_rowdoc.as<uint32_t>(), d_perm.as<uint32_t>(), n, st));
    RK_CUDA(cudaMemcpy(iv.d_list_start.p, iv.list_start.data(), ((size_t) lists + 1) * 4, cudaMemcpyHostToDevice));
    RK_CUDA(cudaDeviceSynchronize());
    d_perm.release();
    if (iv.d_centers.p != d_centers) {
        if (iv.d_centers.ensure((size_t) lists * rb) != 0) return RK_ERR_CUDA;
        RK_CUDA(cudaMemcpy(iv.d_centers.p, d_centers, (size_t) lists * rb, cudaMemcpyDeviceToDevice));
    }
    // keep an unpadded host copy for rk_ivf_info
    std::vector<float> padded((size_t) lists * ix->dim_pad);
    RK_CUDA(cudaMemcpy(padded.data(), iv.d_centers.p, (size_t) lists * rb, cudaMemcpyDeviceToHost));
    iv.centers.assign((size_t) lists * ix->dim, 0.f);
    for (int l = 0; l < lists; l++)
        memcpy(&iv.centers[(size_t) l * ix->dim], &padded[(size_t) l * ix->dim_pad], (size_t) ix->dim * 4);
    iv.lists = lists;
    iv.built = true;
    ix->device_bytes += iv.d_x.cap + iv.d_tid.cap + iv.d_rowdoc.cap + iv.d_centers.cap + iv.d_list_start.cap;
    return RK_OK;
}

static int check_build_args(rk_index *ix, int lists, const char *who)
{
    if (!ix) return rk_fail(RK_ERR_INVALID, "%s: NULL index", who);
    if (!ix->finalized) return rk_fail(RK_ERR_STATE, "%s: call rk_finalize first", who);
    if (ix->ivf.built) return rk_fail(RK_ERR_STATE, "%s: an IVFFlat index already exists", who);
    if (lists < 1 || lists > RK_MAX_LISTS) return rk_fail(RK_ERR_INVALID, "value %d out of bounds for option \"lists\" (1 .. %d)", lists, RK_MAX_LISTS);   // ivfflat.c:36-38
    if (lists > RK_IVF_MAX_LISTS_SMEM) return rk_fail(RK_ERR_UNSUPPORTED, "%s: lists > %d not supported by the probe kernel", who, RK_IVF_MAX_LISTS_SMEM);
    if (ix->dim > 2000) return rk_fail(RK_ERR_UNSUPPORTED, "column cannot have more than 2000 dimensions for ivfflat index");   // ivfflat.h:24
    return 0;
}

static int build_from_host_centers(rk_index *ix, int lists, const float *centers, const uint32_t *given, const char *who)
{
    int rc = check_build_args(ix, lists, who);
    if (rc < 0) return rc;
    if (!centers) return rk_fail(RK_ERR_INVALID, "%s: NULL centers", who);
    for (size_t i = 0; i < (size_t) lists * ix->dim; i++)
        if (!std::isfinite(centers[i])) return rk_fail(RK_ERR_NONFINITE, "NaN or infinite value not allowed in vector");
    RK_CUDA(cudaSetDevice(ix->device));
    const size_t rb = (size_t) ix->dim_pad * 4;
    if (ix->ivf.d_centers.ensure((size_t) lists * rb) != 0) return RK_ERR_CUDA;
    RK_CUDA(cudaMemset(ix->ivf.d_centers.p, 0, (size_t) lists * rb));
    RK_CUDA(cudaMemcpy2D(ix->ivf.d_centers.p, rb, centers, (size_t) ix->dim * 4, (size_t) ix->dim * 4, lists, cudaMemcpyHostToDevice));
    return build_from_device_centers(ix, lists, ix->ivf.d_centers.as<float>(), given);
}

extern "C" int rk_build_ivf_from_centers(rk_index *ix, int lists, const float *centers)
{
    return build_from_host_centers(ix, lists, centers, nullptr, "rk_build_ivf_from_centers");
}

// An index that already exists on disk (rk_ivfpages_read): the lists are taken as stored — ivfscan.c:128-168 scans a list's
// pages whatever centre its tuples are nearest to — so nothing is re-assigned.
extern "C" int rk_build_ivf_from_lists(rk_index *ix, int lists, const float *centers, const uint32_t *row_list)
{
    if (!row_list) return rk_fail(RK_ERR_INVALID, "rk_build_ivf_from_lists: NULL row_list");
    return build_from_host_centers(ix, lists, centers, row_list, "rk_build_ivf_from_lists");
}

extern "C" int rk_build_ivf(rk_index *ix, int lists, uint64_t seed)
{
    int rc = check_build_args(ix, lists, "rk_build_ivf");
    if (rc < 0) return rc;
    RK_CUDA(cudaSetDevice(ix->device));
    const size_t n = ix->nrows;
    if (n == 0) return rk_fail(RK_ERR_STATE, "rk_build_ivf: the index holds no rows");
    const size_t rb = (size_t) ix->dim_pad * 4;
    cudaStream_t st = 0;
    // ivfbuild.c:412-414: sample 50 rows per list, at least 10000
    size_t ns = std::max<size_t>(10000, (size_t) 50 * lists);
    if (ns > n) ns = n;
    std::vector<uint32_t> sample(ns);
    std::mt19937_64 rng(seed);
    if (ix->tune.ivf_sample_identity) {
        for (size_t i = 0; i < ns; i++) sample[i] = (uint32_t) i;
    } else if (ns == n) {
        for (size_t i = 0; i < n; i++) sample[i] = (uint32_t) i;
        std::shuffle(sample.begin(), sample.end(), rng);
    } else {
        std::unordered_set<uint32_t> seen;
        seen.reserve(ns * 2);
        size_t i = 0;
        while (i < ns) {
            uint32_t r = (uint32_t) (rng() % n);
            if (seen.insert(r).second) sample[i++] = r;
        }
    }
    DevBuf d_sidx, d_s, d_label, d_prev, d_sums, d_counts, d_changed;
    if (d_sidx.ensure(ns * 4) != 0 || d_s.ensure(ns * rb) != 0 || d_label.ensure(ns * 4) != 0 || d_prev.ensure(ns * 4) != 0 ||
        d_sums.ensure((size_t) lists * rb) != 0 || d_counts.ensure((size_t) lists * 4) != 0 || d_changed.ensure(16) != 0 ||
        ix->ivf.d_centers.ensure((size_t) lists * rb) != 0)
        return RK_ERR_CUDA;
    RK_CUDA(cudaMemcpy(d_sidx.p, sample.data(), ns * 4, cudaMemcpyHostToDevice));
    RK_CUDA(rk_launch_gather_rows(ix->d_x.as<float>(), d_s.as<float>(), d_sidx.as<uint32_t>(), ns, ix->dim_pad, st));
    const int spherical = ix->metric == RK_METRIC_IP;      // vector_ip_ops: FUNCTION 3 spherical distance, FUNCTION 4 norm
    if (spherical) RK_CUDA(rk_launch_km_normalize_rows(d_s.as<float>(), ns, ix->dim_pad, st));
    // initial centres: k-means++ over the sample; with the option off, or no more samples than lists (pgvector warns "index created
    // with little data"), the first `lists` sampled rows (the sample is a random permutation), wrapping around
    float *C = ix->ivf.d_centers.as<float>();
    if (ix->tune.ivf_kmeanspp && ns > (size_t) lists) {
        // k-means++ (ivfkmeans.c:21-89): each next centre is a sample drawn with probability proportional to its squared
        // distance from the centres chosen so far
        DevBuf d_w, d_bs, d_ch;
        if (d_w.ensure(ns * 4) != 0 || d_bs.ensure((size_t) rk_kpp_blocks((uint32_t) ns) * 8) != 0 || d_ch.ensure((size_t) lists * 4) != 0) return RK_ERR_CUDA;
        RK_CUDA(rk_launch_kmeanspp(d_s.as<float>(), (uint32_t) ns, ix->dim_pad, lists, spherical, seed, d_w.as<float>(), d_bs.as<double>(),
                                   d_ch.as<uint32_t>(), st));
        RK_CUDA(rk_launch_gather_rows(d_s.as<float>(), C, d_ch.as<uint32_t>(), (size_t) lists, ix->dim_pad, st));
        RK_CUDA(cudaStreamSynchronize(st));
        d_w.release(); d_bs.release(); d_ch.release();
    } else {
        for (size_t l = 0; l < (size_t) lists; l += ns) {
            const size_t cnt = std::min<size_t>(ns, (size_t) lists - l);
            RK_CUDA(cudaMemcpyAsync((unsigned char *) C + l * rb, d_s.p, cnt * rb, cudaMemcpyDeviceToDevice, st));
        }
    }
    RK_CUDA(cudaMemsetAsync(d_prev.p, 0xFF, ns * 4, st));
    const int max_iter = ix->tune.ivf_kmeans_iters;
    for (int it = 0; it < max_iter; it++) {
        // on normalised data the nearest centre by L2 is the one with the largest dot product
        { const int rc_ = assign_rows(ix, d_s.as<float>(), ns, C, lists, spherical ? 1 : 0, d_label.as<uint32_t>(), st); if (rc_ < 0) return rc_; }
        RK_CUDA(cudaMemsetAsync(d_changed.p, 0, 8, st));
        RK_CUDA(rk_launch_count_changed(d_label.as<uint32_t>(), d_prev.as<uint32_t>(), ns, d_changed.as<unsigned long long>(), st));
        unsigned long long changed = 0;
        RK_CUDA(cudaMemcpyAsync(&changed, d_changed.p, 8, cudaMemcpyDeviceToHost, st));
        RK_CUDA(cudaStreamSynchronize(st));
        if (changed * 1000 < ns && it > 0) break;          // < 0.1 % of the sample moved
        RK_CUDA(cudaMemsetAsync(d_sums.p, 0, (size_t) lists * rb, st));
        RK_CUDA(cudaMemsetAsync(d_counts.p, 0, (size_t) lists * 4, st));
        RK_CUDA(rk_launch_km_accumulate(d_s.as<float>(), ns, ix->dim_pad, d_label.as<uint32_t>(), d_sums.as<float>(), d_counts.as<uint32_t>(), st));
        RK_CUDA(rk_launch_km_finalize(C, d_sums.as<float>(), d_counts.as<uint32_t>(), lists, ix->dim_pad, d_s.as<float>(), ns,
                                      seed * 1315423911ull + (uint64_t) it, spherical, st));
        RK_CUDA(cudaMemcpyAsync(d_prev.p, d_label.p, ns * 4, cudaMemcpyDeviceToDevice, st));
    }
    RK_CUDA(cudaStreamSynchronize(st));
    DevBuf *tmp[] = {&d_sidx, &d_s, &d_label, &d_prev, &d_sums, &d_counts, &d_changed};
    for (auto *b : tmp) b->release();
    return build_from_device_centers(ix, lists, C);
}

extern "C" int rk_ivf_info(const rk_index *ix, int *lists, float *centers, uint32_t *row_list)
{
    if (!ix) return rk_fail(RK_ERR_INVALID, "NULL index");
    if (!ix->ivf.built) return rk_fail(RK_ERR_STATE, "rk_ivf_info: no IVFFlat index was built");
    if (lists) *lists = ix->ivf.lists;
    if (centers) memcpy(centers, ix->ivf.centers.data(), ix->ivf.centers.size() * 4);
    if (row_list)
        for (size_t i = 0; i < ix->nrows; i++) row_list[i] = ix->ivf.row_list[ix->load_to_row[i]];
    return RK_OK;
}

// ---------------------------------------------------------------------------------------------
// list choice alone (the sharded index splits it by query across ranks)

int rk_ivf_probe_only(rk_index *ix, Workspace &ws, size_t nq, const float *dQ, const rk_scan_opts *o, uint32_t *d_lists)
{
    IvfIndex &iv = ix->ivf;
    if (!iv.built) return rk_fail(RK_ERR_STATE, "rk_ivf_probe_device: no IVFFlat index was built");
    const int probes = std::min(o->probes, iv.lists);
    if (probes < 1) return rk_fail(RK_ERR_INVALID, "rk_ivf_probe_device: probes = %d", o->probes);
    RK_CUDA(cudaSetDevice(ix->device));
    // key scratch: nq x lists x 4 B, in slices of at most 256 MB
    const size_t qb = std::max<size_t>(1, std::min<size_t>(nq, ((size_t) 256 << 20) / ((size_t) iv.lists * 4)));
    if (ws.probe_key.ensure(std::max<size_t>(rk_ivf_probe_keys_bytes((int) qb, iv.lists), 16)) != 0) return RK_ERR_CUDA;
    for (size_t q0 = 0; q0 < nq; q0 += qb) {
        const size_t n = std::min(qb, nq - q0);
        RK_CUDA(rk_launch_ivf_probe(dQ + q0 * (size_t) ix->dim_pad, (int) n, iv.d_centers.as<float>(), iv.lists, ix->dim_pad, ix->metric,
                                    o->null_query, 0, probes, d_lists + q0 * (size_t) probes,
                                    ix->tune.ivf_probe_single ? nullptr : ws.probe_key.as<uint32_t>(), ws.stream));
        ix->total_launches += 2;
    }
    return RK_OK;
}

// ---------------------------------------------------------------------------------------------
// search

void rk_fill_scan_params(const rk_index *ix, const Workspace &ws, RkScanParams &p, bool ivf_store, const float *dQ,
                         const rk_scan_opts *o, int k);
int rk_ensure_cand(Workspace &ws, size_t nlists, int k);

int rk_ivf_search(rk_index *ix, Workspace &ws, size_t nq_all, const float *dQ_all, const uint32_t *user_ids,
                  const rk_scan_opts *o, int probe_offset, const RkSelectParams &out_all)
{
    IvfIndex &iv = ix->ivf;
    const int lists = iv.lists;
    const int W = ix->mask_words;
    const int k = o->k;
    int probes = std::min(o->probes, lists - probe_offset);        // ivfscan.c:262-266
    if (probes < 0) probes = 0;
    const uint32_t kind = o->filter == RK_FILTER_NONE ? RK_SEG_ROWS : RK_SEG_ROWMASK;
    const uint64_t row_bytes = (uint64_t) ix->dim_pad * 4;
    // chunking of long lists so that a small batch still spreads over all SMs
    const double avg_len = lists ? (double) ix->nrows / lists : 0.0;
    const double total_est = (double) nq_all * std::max(probes, 1) * avg_len;
    uint64_t rows_per_seg = std::max<uint64_t>((256 * 1024) / row_bytes, (uint64_t) (total_est / (ix->sm_count * 4.0)));
    rows_per_seg = std::max<uint64_t>(rows_per_seg, (uint64_t) ix->tile_rows);
    int chunks = (int) ((iv.max_list_len + rows_per_seg - 1) / rows_per_seg);
    chunks = std::max(1, std::min(chunks, 64));
    const uint32_t chunk_rows = std::max<uint32_t>(1u, (iv.max_list_len + chunks - 1) / chunks);
    // Class-ordered lists (at most 128 document classes): segments are (query, a few probed lists) and the producer walks
    // one extent per visible class; they do not depend on the probe result, so the host writes them.
    const bool cls_mode = iv.ncls > 0 && o->filter != RK_FILTER_NONE && probes > 0;
    int cls_group = 1, cls_spq = 0;
    if (cls_mode) {
        // enough segments to fill the SMs four times over, at most one per probed list
        int spq = (int) std::min<size_t>((size_t) probes, std::max<size_t>(1, ((size_t) ix->sm_count * ix->tune.ivf_seg_factor + nq_all - 1) / std::max<size_t>(nq_all, 1)));
        cls_group = (probes + spq - 1) / spq;
        cls_spq = (probes + cls_group - 1) / cls_group;
    }
    const size_t per_q = cls_mode ? (size_t) cls_spq : (size_t) std::max(probes, 0) * chunks;
    size_t qb = per_q ? std::max<size_t>(1, RK_IVF_MAX_SEGS / per_q) : nq_all;
    if (qb > nq_all) qb = nq_all;

    ws.segments = 0;
    ws.seg_kinds = probes > 0 ? 1u << (cls_mode ? RK_SEG_IVFQ : kind) : 0u;
    ws.rows_scanned = (uint64_t) total_est;
    ws.algo_bytes = (uint64_t) nq_all * ((uint64_t) lists * row_bytes + 12ull * k) + (uint64_t) (total_est * row_bytes);
    for (size_t q0 = 0; q0 < nq_all; q0 += qb) {
        const size_t nq = std::min(qb, nq_all - q0);
        const float *dQ = dQ_all + q0 * (size_t) ix->dim_pad;
        const uint32_t nseg = (uint32_t) (nq * per_q);
        // staging: [qseg][qmask][ctl] (+ [segments][visible classes per query] for class-ordered lists)
        const size_t b_qseg = (nq + 1) * 4, b_qmask = nq * (size_t) W * 8;
        const size_t b_csegs = cls_mode ? (size_t) nseg * sizeof(RkSegment) : 0, b_qcls = cls_mode ? nq * 16 : 0;
        ws.plan_idx ^= 1;
        const int pi_ = ws.plan_idx;
        if (ws.ev_plan_used[pi_]) RK_CUDA(cudaEventSynchronize(ws.ev_plan[pi_]));
        PinBuf &hp_ = ws.h_plan_ring[pi_];
        if (hp_.ensure(b_qseg + b_qmask + b_csegs + b_qcls + 256) != 0) return RK_ERR_CUDA;
        unsigned char *hp = hp_.as<unsigned char>();
        uint32_t *hq = reinterpret_cast<uint32_t *>(hp);
        for (size_t q = 0; q <= nq; q++) hq[q] = (uint32_t) (q * per_q);
        uint64_t *hm = reinterpret_cast<uint64_t *>(hp + ((b_qseg + 7) & ~(size_t) 7));
        for (size_t q = 0; q < nq; q++) {
            const UserInfo *u = nullptr;
            if (user_ids) {
                auto it = ix->users.find(user_ids[q0 + q]);
                if (it != ix->users.end()) u = &it->second;
            }
            for (int w = 0; w < W; w++) hm[q * W + w] = u ? u->mask[w] : 0ull;
        }
        uint32_t *hc = reinterpret_cast<uint32_t *>(reinterpret_cast<unsigned char *>(hm) + ((b_qmask + 15) & ~(size_t) 15));
        hc[0] = nseg; hc[1] = 0;
        RkSegment *hs = reinterpret_cast<RkSegment *>((reinterpret_cast<uintptr_t>(hc + 4) + 15) & ~(uintptr_t) 15);
        uint64_t *hcl = reinterpret_cast<uint64_t *>(reinterpret_cast<unsigned char *>(hs) + b_csegs);
        if (cls_mode) {
            for (size_t q = 0; q < nq; q++) {
                for (int s_ = 0; s_ < cls_spq; s_++) {
                    RkSegment &sg = hs[q * (size_t) cls_spq + s_];
                    memset(&sg, 0, sizeof sg);
                    sg.query = (uint32_t) q; sg.kind = RK_SEG_IVFQ;
                    sg.begin = (uint64_t) s_ * cls_group;
                    sg.end = (uint64_t) std::min(probes, (s_ + 1) * cls_group);
                }
                // classes this user may see (precomputed per distinct role mask by the build)
                uint64_t c0 = 0, c1 = 0;
                const UserInfo *u = nullptr;
                if (user_ids) {
                    auto it = ix->users.find(user_ids[q0 + q]);
                    if (it != ix->users.end()) u = &it->second;
                }
                if (u) { c0 = iv.qcls_by_mask_id[(size_t) u->mask_id * 2]; c1 = iv.qcls_by_mask_id[(size_t) u->mask_id * 2 + 1]; }
                hcl[q * 2] = c0; hcl[q * 2 + 1] = c1;
            }
        }
        if (ws.qseg.ensure(b_qseg) != 0 || ws.qmask.ensure(std::max<size_t>(b_qmask, 16)) != 0 || ws.ctl.ensure(64) != 0 ||
            ws.segs.ensure(std::max<size_t>((size_t) nseg * sizeof(RkSegment), 16)) != 0 ||
            ws.probe_list.ensure(std::max<size_t>(nq * (size_t) std::max(probes, 1) * 4, 16)) != 0 ||
            ws.probe_key.ensure(std::max<size_t>(rk_ivf_probe_keys_bytes((int) nq, lists), 16)) != 0)
            return RK_ERR_CUDA;
        RK_CUDA(cudaMemcpyAsync(ws.qseg.p, hq, b_qseg, cudaMemcpyHostToDevice, ws.stream));
        if (b_qmask) RK_CUDA(cudaMemcpyAsync(ws.qmask.p, hm, b_qmask, cudaMemcpyHostToDevice, ws.stream));
        RK_CUDA(cudaMemcpyAsync(ws.ctl.p, hc, 8, cudaMemcpyHostToDevice, ws.stream));
        if (cls_mode) {
            if (ws.order.ensure(std::max<size_t>(b_qcls, 16)) != 0) return RK_ERR_CUDA;
            RK_CUDA(cudaMemcpyAsync(ws.segs.p, hs, b_csegs, cudaMemcpyHostToDevice, ws.stream));
            RK_CUDA(cudaMemcpyAsync(ws.order.p, hcl, b_qcls, cudaMemcpyHostToDevice, ws.stream));
        }
        RK_CUDA(cudaEventRecord(ws.ev_plan[pi_], ws.stream));
        ws.ev_plan_used[pi_] = true;
        if (rk_ensure_cand(ws, nseg, k) != 0) return RK_ERR_CUDA;

        if (nseg) {
            if (ws.timing && q0 == 0) RK_CUDA(cudaEventRecord(ws.ev[4], ws.stream));
            const uint32_t *plist = ws.ext_probe ? ws.ext_probe + q0 * (size_t) probes : ws.probe_list.as<uint32_t>();
            if (!ws.ext_probe)
                RK_CUDA(rk_launch_ivf_probe(dQ, (int) nq, iv.d_centers.as<float>(), lists, ix->dim_pad, ix->metric, o->null_query,
                                            probe_offset, probes, ws.probe_list.as<uint32_t>(),
                                            ix->tune.ivf_probe_single ? nullptr : ws.probe_key.as<uint32_t>(), ws.stream));
            if (!cls_mode)
                RK_CUDA(rk_launch_ivf_plan(plist, iv.d_list_start.as<uint32_t>(), (int) nq, probes, chunks,
                                           chunk_rows, kind, ws.segs.as<RkSegment>(), ws.stream));
            if (ws.timing && q0 == 0) { RK_CUDA(cudaEventRecord(ws.ev[5], ws.stream)); ws.timed_probe = true; }
            RkScanParams p;
            rk_fill_scan_params(ix, ws, p, true, dQ, o, k);
            p.seg_order = nullptr;
            if (cls_mode) {
                p.ivf_probe_list = plist;
                p.ivf_probes = probes;
                p.ivf_cls_off = iv.d_cls_off.as<uint32_t>();
                p.ivf_ncls = iv.ncls;
                p.ivf_qcls = ws.order.as<uint64_t>();
            }
            const int slot = (int) (ws.ring_n % RK_TIMING_RING);
            if (ws.timing) RK_CUDA(cudaEventRecord(ws.ring_a[slot], ws.stream));
            const int grid = (int) std::min<uint32_t>((uint32_t) ix->sm_count, nseg);
            RK_CUDA(rk_launch_scan(p, ix->device, ws.stream, grid));
            if (ws.timing) { RK_CUDA(cudaEventRecord(ws.ring_b[slot], ws.stream)); ws.timed_scan = true; ws.ring_n++; }
            // list choice (keys + ordering kernels for a batch, one kernel for a single query), work list, list scan
            ws.launches += (ws.ext_probe ? 0 : (nq >= 4 ? 2 : 1)) + (cls_mode ? 0 : 1) + 1;
        }
        RkSelectParams out = out_all;
        const size_t ob = q0 * (size_t) k;
        if (out.out_tid) out.out_tid += ob;
        if (out.out_doc) out.out_doc += ob;
        if (out.out_dist) out.out_dist += ob;
        if (out.out_key) out.out_key += ob;
        if (out.out_row) out.out_row += ob;
        if (out.out_count) out.out_count += q0;
        out.cand_key = ws.cand_key.as<uint32_t>();
        out.cand_tid = ws.cand_tid.as<uint64_t>();
        out.cand_row = ws.cand_row.as<uint32_t>();
        out.cand_count = ws.cand_count.as<int32_t>();
        out.qseg = ws.qseg.as<uint32_t>();
        out.lists_per_query = 0;
        out.list_stride = k;
        out.shard_stride_bytes = 0;
        out.k = k;
        out.metric = ix->metric;
        out.dedup = 0;                         // a tuple sits in exactly one list
        out.null_query = o->null_query;
        out.key_is_float = 0;
        out.rowdoc = iv.d_rowdoc.as<uint32_t>();
        out.doc_ids = ix->d_doc_ids.as<uint32_t>();
        if (ws.timing && q0 == 0) RK_CUDA(cudaEventRecord(ws.ev[2], ws.stream));
        RK_CUDA(rk_launch_select(out, (int) nq, ws.stream));
        if (ws.timing && q0 == 0) { RK_CUDA(cudaEventRecord(ws.ev[3], ws.stream)); ws.timed_select = true; }
        ws.launches++;
        ws.segments += nseg;
    }
    ix->total_launches += ws.launches;
    return RK_OK;
}
