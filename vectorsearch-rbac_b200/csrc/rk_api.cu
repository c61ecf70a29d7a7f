// rk_api.cu — host side of librbacknn behind the C-ABI of include/rbacknn.h.
//
// Holds what PostgreSQL's catalog + heap hold for the reference (rows, PermissionAssignment,
// UserRoles, the role / combination / dynamic partition layouts), turns a batch of
// (query vector, user) pairs into a work list for the persistent scan kernel, and mirrors the
// pgvector index-scan contract (begin / rescan / gettuple / endscan).  No CPU scoring path exists
// here: every distance is computed by rk_scan.cu on the GPU.
#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <map>
#include <mutex>
#include <string>
#include <unordered_map>
#include <vector>

#include "../../include/rbacknn.h"
#include "rk_host.h"

// ---------------------------------------------------------------------------------------------
// errors

static thread_local std::string g_err;

int rk_fail(int code, const char *fmt, ...)
{
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_err = buf;
    return code;
}

#define RK_CUDA(call)                                                                                    \
    do {                                                                                                 \
        cudaError_t e__ = (call);                                                                        \
        if (e__ != cudaSuccess)                                                                          \
            return rk_fail(RK_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), __FILE__, __LINE__); \
    } while (0)

extern "C" const char *rk_last_error(void) { return g_err.c_str(); }
extern "C" int rk_abi_version(void) { return RK_ABI_VERSION; }

// ---------------------------------------------------------------------------------------------
// buffers

int DevBuf::ensure(size_t bytes)
{
    if (bytes <= cap) return 0;
    size_t want = std::max(bytes, cap + cap / 2);
    want = (want + 255) & ~(size_t) 255;
    void *np = nullptr;
    RK_CUDA(cudaMalloc(&np, want));
    if (p) cudaFree(p);
    p = np;
    cap = want;
    return 0;
}
void DevBuf::release()
{
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
}
int PinBuf::ensure(size_t bytes)
{
    if (bytes <= cap) return 0;
    size_t want = std::max(bytes, cap * 2);
    void *np = nullptr;
    RK_CUDA(cudaHostAlloc(&np, want, cudaHostAllocDefault));
    if (p) cudaFreeHost(p);
    p = np;
    cap = want;
    return 0;
}
void PinBuf::release()
{
    if (p) cudaFreeHost(p);
    p = nullptr;
    cap = 0;
}

int Workspace::init(int device_)
{
    device = device_;
    RK_CUDA(cudaSetDevice(device));
    RK_CUDA(cudaStreamCreateWithFlags(&own_stream, cudaStreamNonBlocking));
    stream = own_stream;
    for (auto &e : ev) RK_CUDA(cudaEventCreate(&e));
    for (auto &e : ring_a) RK_CUDA(cudaEventCreate(&e));
    for (auto &e : ring_b) RK_CUDA(cudaEventCreate(&e));
    for (auto &e : ev_plan) RK_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    if (qflag.ensure(16) != 0) return RK_ERR_CUDA;
    RK_CUDA(cudaMemset(qflag.p, 0, 16));
    RK_CUDA(cudaHostAlloc((void **) &h_qflag, 16, cudaHostAllocDefault));
    *h_qflag = 0;
    return 0;
}
void Workspace::destroy()
{
    if (h_qflag) cudaFreeHost(h_qflag);
    h_qflag = nullptr;
    for (auto &kv : fused_plans) kv.second.buf.release();
    fused_plans.clear();
    fz_ctl.release(); fz_key.release(); fz_row.release(); fz_pack.release();
    h_fz.release();
    if (h_dbg) cudaFreeHost(h_dbg);
    h_dbg = nullptr;
    d_dbg.release();
    if (h_zc) cudaFreeHost(h_zc);
    h_zc = d_zc = nullptr;
    zc_cap = 0;
    DevBuf *bufs[] = {&qflag, &segs, &order, &ctl, &cand_key, &cand_tid, &cand_row, &cand_count, &qseg, &qmask, &dq, &out_tid,
                      &out_doc, &out_dist, &out_count, &probe_key, &probe_list, &ex_key, &ex_tid, &ex_doc};
    for (auto *b : bufs) b->release();
    h_plan_ring[0].release();
    h_plan_ring[1].release();
    h_out.release();
    for (auto &e : ev_plan)
        if (e) cudaEventDestroy(e);
    for (auto &e : ev)
        if (e) cudaEventDestroy(e);
    for (auto &e : ring_a)
        if (e) cudaEventDestroy(e);
    for (auto &e : ring_b)
        if (e) cudaEventDestroy(e);
    if (own_stream) cudaStreamDestroy(own_stream);
    own_stream = nullptr;
}

// ---------------------------------------------------------------------------------------------
// lifecycle

extern "C" rk_index *rk_open(const rk_config *cfg)
{
    if (!cfg) { rk_fail(RK_ERR_INVALID, "rk_open: cfg is NULL"); return nullptr; }
    if (cfg->abi_version != RK_ABI_VERSION) { rk_fail(RK_ERR_INVALID, "rk_open: ABI version %d, library is %d", cfg->abi_version, RK_ABI_VERSION); return nullptr; }
    if (cfg->dim < 1 || cfg->dim > RK_MAX_DIM) { rk_fail(RK_ERR_UNSUPPORTED, "rk_open: dim %d outside 1..%d", cfg->dim, RK_MAX_DIM); return nullptr; }
    if (cfg->metric != RK_METRIC_L2 && cfg->metric != RK_METRIC_IP) { rk_fail(RK_ERR_INVALID, "rk_open: unknown metric %d", cfg->metric); return nullptr; }
    if (cfg->k_max < 1 || cfg->k_max > RK_MAX_K) { rk_fail(RK_ERR_INVALID, "rk_open: k_max %d outside 1..%d", cfg->k_max, RK_MAX_K); return nullptr; }
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev <= 0) {
        rk_fail(RK_ERR_NO_DEVICE, "rk_open: no CUDA device (%s); this engine has no CPU path", e == cudaSuccess ? "0 devices" : cudaGetErrorString(e));
        return nullptr;
    }
    if (cfg->device < 0 || cfg->device >= ndev) { rk_fail(RK_ERR_INVALID, "rk_open: device %d of %d", cfg->device, ndev); return nullptr; }
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, cfg->device) != cudaSuccess || prop.major != 10) {
        rk_fail(RK_ERR_NO_DEVICE, "rk_open: device %d is sm_%d%d; kernels are built for sm_100a only", cfg->device, prop.major, prop.minor);
        return nullptr;
    }
    rk_index *ix = new rk_index();
    ix->cfg = *cfg;
    ix->dim = cfg->dim;
    ix->dim_pad = (cfg->dim + 3) & ~3;
    ix->metric = cfg->metric;
    ix->device = cfg->device;
    ix->k_max = cfg->k_max;
    ix->sm_count = prop.multiProcessorCount;
    rk_scan_geometry(ix->dim_pad, &ix->tile_rows, &ix->nstages, &ix->scan_smem);
    // tuning / test switches: the environment is read here, once, never on the per-call path
    {
        RkTuning &t = ix->tune;
        auto env_int = [](const char *name, int dflt) { const char *e = getenv(name); return e && *e ? atoi(e) : dflt; };
        t.tensor_tpu = std::max(0, env_int("RK_TENSOR_TPU", 0));
        t.tensor_pair = env_int("RK_TENSOR_PAIR", -1);
        if (const char *e = getenv("RK_TENSOR_PLAN"))
            t.tensor_plan = !strcmp(e, "dense") ? 1 : !strcmp(e, "grouped") ? 2 : !strcmp(e, "atoms") ? 3 : 0;
        t.tensor_sample = env_int("RK_TENSOR_SAMPLE", 1);
        t.tensor_sdiv = std::max(0, env_int("RK_TENSOR_SDIV", 0));
        t.tensor_smin = std::max(1, env_int("RK_TENSOR_SMIN", 1024));
        t.tensor_host_timing = getenv("RK_TENSOR_HOST_TIMING") != nullptr;
        t.tensor_bf16 = env_int("RK_TENSOR_BF16", -1);
        t.tensor_kp_extra = std::max(0, env_int("RK_TENSOR_KP_EXTRA", 0));
        t.tensor_kp_exact = env_int("RK_TENSOR_KP_EXACT", 28);
        t.class_order = env_int("RK_CLASS_ORDER", 1);
        t.tensor_loose = std::max(0, env_int("RK_TENSOR_LOOSE", 128));
        t.tensor_round_mark = std::min(768, std::max(1, env_int("RK_TENSOR_ROUND_MARK", 512)));
        t.tensor_kp_bf16 = std::max(0, env_int("RK_TENSOR_KP_BF16", 64));
        t.tensor_qtau = env_int("RK_TENSOR_QTAU", 1);
        t.ivf_seg_factor = std::max(1, env_int("RK_IVF_SEG_FACTOR", 4));
        t.ivf_probe_single = getenv("RK_IVF_PROBE_SINGLE") != nullptr;
        t.ivf_assign_tc = env_int("RK_IVF_ASSIGN_TC", 1);
        t.ivf_kmeanspp = env_int("RK_IVF_KMEANSPP", 1);
        t.ivf_kmeans_iters = std::max(0, env_int("RK_IVF_KMEANS_ITERS", 25));
        t.ivf_sample_identity = env_int("RK_IVF_SAMPLE_IDENTITY", 0);
        t.single_fused = env_int("RK_SINGLE_FUSED", 1);
        t.seg_cost_x10 = std::max(1, env_int("RK_SEG_COST_X10", 80));
        t.single_zero_copy = env_int("RK_SINGLE_ZERO_COPY", 1);
        t.single_segs = env_int("RK_SINGLE_SEGS", 0);
        t.single_seed = env_int("RK_SINGLE_SEED", 0);
        t.single_seed_min = env_int("RK_SINGLE_SEED_MIN", 2048);
        t.single_debug = env_int("RK_SINGLE_DEBUG", 0);
    }
    if (ix->ws.init(cfg->device) != 0) { delete ix; return nullptr; }
    return ix;
}

extern "C" int rk_set_option(rk_index *ix, const char *name, int64_t value)
{
    if (!ix || !name) return rk_fail(RK_ERR_INVALID, "rk_set_option: NULL argument");
    RkTuning &t = ix->tune;
    const int v = (int) value;
    std::lock_guard<std::mutex> g(ix->batch_mu);
    if (!strcmp(name, "tensor_tpu")) t.tensor_tpu = std::max(0, v);
    else if (!strcmp(name, "tensor_pair")) t.tensor_pair = v;
    else if (!strcmp(name, "tensor_plan")) t.tensor_plan = (v >= 0 && v <= 3) ? v : 0;
    else if (!strcmp(name, "tensor_sample")) t.tensor_sample = v;
    else if (!strcmp(name, "tensor_sdiv")) t.tensor_sdiv = std::max(0, v);
    else if (!strcmp(name, "tensor_smin")) t.tensor_smin = std::max(1, v);
    else if (!strcmp(name, "tensor_host_timing")) t.tensor_host_timing = v != 0;
    else if (!strcmp(name, "tensor_bf16")) t.tensor_bf16 = v;
    else if (!strcmp(name, "tensor_kp_extra")) t.tensor_kp_extra = std::max(0, v);
    else if (!strcmp(name, "tensor_kp_exact")) t.tensor_kp_exact = v;
    else if (!strcmp(name, "class_order")) t.class_order = v;
    else if (!strcmp(name, "tensor_loose")) t.tensor_loose = std::max(0, v);
    else if (!strcmp(name, "tensor_round_mark")) t.tensor_round_mark = std::min(768, std::max(1, v));
    else if (!strcmp(name, "tensor_kp_bf16")) t.tensor_kp_bf16 = std::max(0, v);
    else if (!strcmp(name, "tensor_qtau")) t.tensor_qtau = v;
    else if (!strcmp(name, "ivf_seg_factor")) t.ivf_seg_factor = std::max(1, v);
    else if (!strcmp(name, "ivf_probe_single")) t.ivf_probe_single = v != 0;
    else if (!strcmp(name, "ivf_assign_tc")) t.ivf_assign_tc = v;
    else if (!strcmp(name, "ivf_kmeanspp")) t.ivf_kmeanspp = v;
    else if (!strcmp(name, "ivf_kmeans_iters")) t.ivf_kmeans_iters = std::max(0, v);
    else if (!strcmp(name, "ivf_sample_identity")) t.ivf_sample_identity = v;
    else if (!strcmp(name, "single_fused")) t.single_fused = v;
    else if (!strcmp(name, "seg_cost_x10")) t.seg_cost_x10 = std::max(1, v);
    else if (!strcmp(name, "single_zero_copy")) t.single_zero_copy = v;
    else if (!strcmp(name, "single_segs")) t.single_segs = v;
    else if (!strcmp(name, "single_seed")) t.single_seed = v;
    else if (!strcmp(name, "single_seed_min")) t.single_seed_min = v;
    else if (!strcmp(name, "single_debug")) t.single_debug = v;
    else return rk_fail(RK_ERR_INVALID, "rk_set_option: unknown option \"%s\"", name);
    return RK_OK;
}

extern "C" void rk_close(rk_index *ix)
{
    if (!ix) return;
    cudaSetDevice(ix->device);
    cudaDeviceSynchronize();
    ix->ws.destroy();
    DevBuf *bufs[] = {&ix->d_x, &ix->d_tid, &ix->d_rowdoc, &ix->d_doc_ids, &ix->d_doc_mask, &ix->d_ext_row,
                      &ix->d_ext_vstart, &ix->d_ext_mask, &ix->d_part_ext_off, &ix->ivf.d_x, &ix->ivf.d_tid,
                      &ix->ivf.d_rowdoc, &ix->ivf.d_centers, &ix->ivf.d_center_tid, &ix->ivf.d_list_start};
    for (auto *b : bufs) b->release();
    TensorState &ts = ix->tensor;
    DevBuf *tb[] = {&ts.d_xnorm, &ts.d_rowmask, &ts.d_probe, &ts.d_units, &ts.d_qnorm, &ts.d_scratch, &ts.d_cand_key, &ts.d_cand_row,
                    &ts.d_cand_count, &ts.d_sel_key, &ts.d_sel_row, &ts.d_sel_count, &ts.d_flags, &ts.d_plan, &ts.d_qp, &ts.d_tiles,
                    &ts.d_rowmask2, &ts.d_tau_key, &ts.d_tau_count, &ts.d_xb, &ts.d_qpb, &ts.d_redo};
    for (auto *b : tb) b->release();
    ts.h_flags.release();
    delete ix;
}

// ---------------------------------------------------------------------------------------------
// loading

static int check_finite(const float *x, size_t n)
{
    return rk_all_finite(x, n) ? 0 : -1;
}

extern "C" int rk_load_rows(rk_index *ix, const float *x, const uint64_t *tid, const uint32_t *doc, size_t n)
{
    if (!ix || (!x && n) || (!tid && n) || (!doc && n)) return rk_fail(RK_ERR_INVALID, "rk_load_rows: NULL argument");
    if (ix->finalized) return rk_fail(RK_ERR_STATE, "rk_load_rows: index already finalized");
    if (n == 0) return RK_OK;
    if (ix->nrows + n >= 0xFFFFFFF0ull) return rk_fail(RK_ERR_UNSUPPORTED, "rk_load_rows: more than 2^32 rows");
    // vector_in rejects NaN and infinity (pgvector/src/vector.c:102-113)
    if (check_finite(x, n * (size_t) ix->dim) != 0) return rk_fail(RK_ERR_NONFINITE, "rk_load_rows: NaN or infinite value not allowed in vector");
    RK_CUDA(cudaSetDevice(ix->device));
    const size_t rb = (size_t) ix->dim_pad * 4;
    const size_t need = (ix->nrows + n) * rb;
    if (need > ix->d_x.cap) {
        // grow, keeping what is already loaded
        DevBuf nb;
        size_t want = std::max(need, ix->d_x.cap * 2);
        if (nb.ensure(want) != 0) return RK_ERR_CUDA;
        if (ix->nrows) RK_CUDA(cudaMemcpy(nb.p, ix->d_x.p, ix->nrows * rb, cudaMemcpyDeviceToDevice));
        ix->d_x.release();
        ix->d_x = nb;
    }
    unsigned char *dst = (unsigned char *) ix->d_x.p + ix->nrows * rb;
    if (ix->dim == ix->dim_pad) {
        RK_CUDA(cudaMemcpy(dst, x, n * rb, cudaMemcpyHostToDevice));
    } else {
        RK_CUDA(cudaMemset(dst, 0, n * rb));
        RK_CUDA(cudaMemcpy2D(dst, rb, x, (size_t) ix->dim * 4, (size_t) ix->dim * 4, n, cudaMemcpyHostToDevice));
    }
    ix->h_tid.insert(ix->h_tid.end(), tid, tid + n);
    ix->h_doc.insert(ix->h_doc.end(), doc, doc + n);
    ix->nrows += n;
    return RK_OK;
}

extern "C" int rk_load_rows_device(rk_index *ix, const float *d_x, const uint64_t *tid, const uint32_t *doc, size_t n)
{
    if (!ix || (!d_x && n) || (!tid && n) || (!doc && n)) return rk_fail(RK_ERR_INVALID, "rk_load_rows_device: NULL argument");
    if (ix->finalized) return rk_fail(RK_ERR_STATE, "rk_load_rows_device: index already finalized");
    if (n == 0) return RK_OK;
    if (ix->nrows + n >= 0xFFFFFFF0ull) return rk_fail(RK_ERR_UNSUPPORTED, "rk_load_rows_device: more than 2^32 rows");
    RK_CUDA(cudaSetDevice(ix->device));
    unsigned long long bad = 0;
    RK_CUDA(rk_count_nonfinite(d_x, n * (size_t) ix->dim, &bad, 0));
    if (bad) return rk_fail(RK_ERR_NONFINITE, "rk_load_rows_device: NaN or infinite value not allowed in vector");
    const size_t rb = (size_t) ix->dim_pad * 4;
    const size_t need = (ix->nrows + n) * rb;
    if (need > ix->d_x.cap) {
        DevBuf nb;
        if (nb.ensure(std::max(need, ix->nrows ? ix->d_x.cap * 2 : need)) != 0) return RK_ERR_CUDA;
        if (ix->nrows) RK_CUDA(cudaMemcpy(nb.p, ix->d_x.p, ix->nrows * rb, cudaMemcpyDeviceToDevice));
        ix->d_x.release();
        ix->d_x = nb;
    }
    unsigned char *dst = (unsigned char *) ix->d_x.p + ix->nrows * rb;
    if (ix->dim == ix->dim_pad) {
        RK_CUDA(cudaMemcpy(dst, d_x, n * rb, cudaMemcpyDeviceToDevice));
    } else {
        RK_CUDA(rk_launch_pad_rows(d_x, reinterpret_cast<float *>(dst), n, ix->dim, ix->dim_pad, 0));
        RK_CUDA(cudaDeviceSynchronize());
    }
    ix->h_tid.insert(ix->h_tid.end(), tid, tid + n);
    ix->h_doc.insert(ix->h_doc.end(), doc, doc + n);
    ix->nrows += n;
    return RK_OK;
}

extern "C" int rk_reserve_rows(rk_index *ix, size_t n)
{
    if (!ix) return rk_fail(RK_ERR_INVALID, "rk_reserve_rows: NULL index");
    if (ix->finalized) return rk_fail(RK_ERR_STATE, "rk_reserve_rows: index already finalized");
    RK_CUDA(cudaSetDevice(ix->device));
    const size_t rb = (size_t) ix->dim_pad * 4;
    if (n * rb > ix->d_x.cap) {
        DevBuf nb;
        if (nb.ensure(n * rb) != 0) return RK_ERR_CUDA;
        if (ix->nrows) RK_CUDA(cudaMemcpy(nb.p, ix->d_x.p, ix->nrows * rb, cudaMemcpyDeviceToDevice));
        ix->d_x.release();
        ix->d_x = nb;
    }
    ix->h_tid.reserve(n);
    ix->h_doc.reserve(n);
    return RK_OK;
}

extern "C" int rk_set_doc_roles(rk_index *ix, const uint32_t *doc, const uint32_t *role, size_t n)
{
    if (!ix || ((!doc || !role) && n)) return rk_fail(RK_ERR_INVALID, "rk_set_doc_roles: NULL argument");
    if (ix->finalized) return rk_fail(RK_ERR_STATE, "rk_set_doc_roles: index already finalized");
    for (size_t i = 0; i < n; i++) ix->doc_roles.push_back({doc[i], role[i]});
    return RK_OK;
}

extern "C" int rk_set_user_roles(rk_index *ix, const uint32_t *user, const uint32_t *role, size_t n)
{
    if (!ix || ((!user || !role) && n)) return rk_fail(RK_ERR_INVALID, "rk_set_user_roles: NULL argument");
    if (ix->finalized) return rk_fail(RK_ERR_STATE, "rk_set_user_roles: index already finalized");
    for (size_t i = 0; i < n; i++) ix->user_roles.push_back({user[i], role[i]});
    return RK_OK;
}

extern "C" int rk_build_role_partitions(rk_index *ix)
{
    if (!ix) return rk_fail(RK_ERR_INVALID, "NULL index");
    if (ix->finalized) return rk_fail(RK_ERR_STATE, "rk_build_role_partitions: index already finalized");
    ix->want_role_parts = true;
    return RK_OK;
}

extern "C" int rk_build_combination_partitions(rk_index *ix)
{
    if (!ix) return rk_fail(RK_ERR_INVALID, "NULL index");
    if (ix->finalized) return rk_fail(RK_ERR_STATE, "rk_build_combination_partitions: index already finalized");
    ix->want_combo_parts = true;
    return RK_OK;
}

extern "C" int rk_define_partition(rk_index *ix, uint32_t part_id, const uint32_t *docs, size_t n)
{
    if (!ix || (!docs && n)) return rk_fail(RK_ERR_INVALID, "rk_define_partition: NULL argument");
    if (ix->finalized) return rk_fail(RK_ERR_STATE, "rk_define_partition: index already finalized");
    auto &v = ix->part_defs[part_id];
    v.insert(v.end(), docs, docs + n);
    return RK_OK;
}

extern "C" int rk_map_roleset_to_partitions(rk_index *ix, const uint32_t *roles, size_t nroles, const uint32_t *parts, size_t nparts)
{
    if (!ix || (!roles && nroles) || (!parts && nparts)) return rk_fail(RK_ERR_INVALID, "rk_map_roleset_to_partitions: NULL argument");
    if (ix->finalized) return rk_fail(RK_ERR_STATE, "rk_map_roleset_to_partitions: index already finalized");
    std::vector<uint32_t> key(roles, roles + nroles);
    std::sort(key.begin(), key.end());
    key.erase(std::unique(key.begin(), key.end()), key.end());
    auto &v = ix->roleset_parts[key];
    v.insert(v.end(), parts, parts + nparts);
    return RK_OK;
}

// ---------------------------------------------------------------------------------------------
// finalize: freeze the layout and upload it

template <typename T>
static int upload(DevBuf &b, const std::vector<T> &v, uint64_t *acc)
{
    size_t bytes = std::max<size_t>(v.size() * sizeof(T), 16);
    if (b.ensure(bytes) != 0) return RK_ERR_CUDA;
    if (!v.empty()) RK_CUDA(cudaMemcpy(b.p, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice));
    if (acc) *acc += b.cap;
    return 0;
}

static void role_mask_of(const rk_index *ix, const std::vector<uint32_t> &roles, uint64_t *mask)
{
    for (int w = 0; w < ix->mask_words; w++) mask[w] = 0;
    for (uint32_t r : roles) {
        auto it = std::lower_bound(ix->role_ids.begin(), ix->role_ids.end(), r);
        if (it == ix->role_ids.end() || *it != r) continue;
        size_t b = (size_t) (it - ix->role_ids.begin());
        mask[b >> 6] |= 1ull << (b & 63);
    }
}

static inline bool mask_hit(const uint64_t *a, const uint64_t *b, int w)
{
    uint64_t acc = 0;
    for (int i = 0; i < w; i++) acc |= a[i] & b[i];
    return acc != 0;
}

extern "C" int rk_finalize(rk_index *ix)
{
    if (!ix) return rk_fail(RK_ERR_INVALID, "NULL index");
    if (ix->finalized) return rk_fail(RK_ERR_STATE, "rk_finalize: already finalized");
    RK_CUDA(cudaSetDevice(ix->device));
    const size_t n = ix->nrows;
    const size_t rb = (size_t) ix->dim_pad * 4;

    // roles: dense bit positions in ascending raw id
    for (auto &pr : ix->doc_roles) ix->role_ids.push_back(pr.second);
    for (auto &pr : ix->user_roles) ix->role_ids.push_back(pr.second);
    std::sort(ix->role_ids.begin(), ix->role_ids.end());
    ix->role_ids.erase(std::unique(ix->role_ids.begin(), ix->role_ids.end()), ix->role_ids.end());
    ix->mask_words = std::max<int>(1, (int) ((ix->role_ids.size() + 63) / 64));
    if (ix->mask_words > 16) return rk_fail(RK_ERR_UNSUPPORTED, "rk_finalize: %zu roles, at most 1024 supported", ix->role_ids.size());
    const int W = ix->mask_words;

    // documents.  The dense document index — which is also the physical order of the rows — goes by (role mask, raw id):
    // documents with the same role mask (a document class: visible to exactly the same users) are neighbours, so what a user
    // may see is a few long runs of rows instead of documents scattered over the whole store.  The scan's producer fuses
    // adjacent extents into one bulk copy, a single query's private extent table shrinks to a handful of runs, and a unit of
    // the tensor engine's per-class plan covers whole 128-row tiles of its own class (before: a document of 100 rows dragged
    // 28 rows of some other class into every tile).  Results do not depend on the order (ties break by tuple id).
    std::vector<uint32_t> raw_sorted = ix->h_doc;
    std::sort(raw_sorted.begin(), raw_sorted.end());
    raw_sorted.erase(std::unique(raw_sorted.begin(), raw_sorted.end()), raw_sorted.end());
    const size_t D = raw_sorted.size();
    auto raw_index = [&](uint32_t raw) -> size_t {       // position in raw_sorted, or D
        auto it = std::lower_bound(raw_sorted.begin(), raw_sorted.end(), raw);
        return it != raw_sorted.end() && *it == raw ? (size_t) (it - raw_sorted.begin()) : D;
    };
    std::vector<uint64_t> mask0(std::max<size_t>(D, 1) * W, 0);
    for (auto &pr : ix->doc_roles) {
        const size_t d = raw_index(pr.first);
        if (d == D) continue;   // permission for a document with no rows
        size_t b = (size_t) (std::lower_bound(ix->role_ids.begin(), ix->role_ids.end(), pr.second) - ix->role_ids.begin());
        mask0[d * W + (b >> 6)] |= 1ull << (b & 63);
    }
    std::vector<uint32_t> doc_order(D), new_of_raw(D);
    for (size_t d = 0; d < D; d++) doc_order[d] = (uint32_t) d;
    bool class_order = ix->tune.class_order != 0;
    if (class_order) {
        // (re-ordering the rows takes a second copy of the row store for a moment: a store that fills more than half of the
        // free HBM keeps the order by id)
        size_t fr = 0, tot = 0;
        if (cudaMemGetInfo(&fr, &tot) == cudaSuccess && fr < n * rb + ((size_t) 256 << 20)) class_order = false;
    }
    if (class_order)
        std::stable_sort(doc_order.begin(), doc_order.end(), [&](uint32_t a, uint32_t b) {
            return std::lexicographical_compare(&mask0[(size_t) a * W], &mask0[(size_t) a * W] + W, &mask0[(size_t) b * W], &mask0[(size_t) b * W] + W);
        });
    if (class_order && ix->tune.class_order == 1) {
        // worth it only when classes are groups: with (nearly) one mask per document — random RBAC — there are no runs to gain,
        // and the tensor engine's sampling pass (every S-th tile) would no longer see a uniform sample of each user's rows
        // (C2 measured 5 % slower).  class_order = 2 forces the order (tests).
        size_t classes = D ? 1 : 0;
        for (size_t i = 1; i < D; i++)
            classes += !std::equal(&mask0[(size_t) doc_order[i] * W], &mask0[(size_t) doc_order[i] * W] + W, &mask0[(size_t) doc_order[i - 1] * W]);
        if (classes * 4 > D)
            for (size_t d = 0; d < D; d++) doc_order[d] = (uint32_t) d;
    }
    ix->doc_ids.resize(D);
    ix->doc_mask.assign(std::max<size_t>(D, 1) * W, 0);
    for (size_t i = 0; i < D; i++) {
        new_of_raw[doc_order[i]] = (uint32_t) i;
        ix->doc_ids[i] = raw_sorted[doc_order[i]];
        for (int w = 0; w < W; w++) ix->doc_mask[i * W + w] = mask0[(size_t) doc_order[i] * W + w];
    }
    std::vector<uint32_t> dense(n);
    bool sorted = true;
    for (size_t i = 0; i < n; i++) {
        dense[i] = new_of_raw[raw_index(ix->h_doc[i])];
        if (i && dense[i] < dense[i - 1]) sorted = false;
    }
    // rows ordered by document (stable), so a document is one contiguous extent
    ix->load_to_row.resize(n);
    if (sorted) {
        for (size_t i = 0; i < n; i++) ix->load_to_row[i] = (uint32_t) i;
        ix->h_rowdoc.swap(dense);
    } else {
        std::vector<uint32_t> perm(n);   // finalized row -> load index
        for (size_t i = 0; i < n; i++) perm[i] = (uint32_t) i;
        std::stable_sort(perm.begin(), perm.end(), [&](uint32_t a, uint32_t b) { return dense[a] < dense[b]; });
        std::vector<uint64_t> t2(n);
        ix->h_rowdoc.resize(n);
        for (size_t i = 0; i < n; i++) {
            t2[i] = ix->h_tid[perm[i]];
            ix->h_rowdoc[i] = dense[perm[i]];
            ix->load_to_row[perm[i]] = (uint32_t) i;
        }
        ix->h_tid.swap(t2);
        DevBuf d_perm, nx;
        if (upload(d_perm, perm, nullptr) != 0) return RK_ERR_CUDA;
        if (nx.ensure(std::max<size_t>(n * rb, 16)) != 0) return RK_ERR_CUDA;
        RK_CUDA(rk_launch_gather_rows(ix->d_x.as<float>(), nx.as<float>(), d_perm.as<uint32_t>(), n, ix->dim_pad, 0));
        RK_CUDA(cudaDeviceSynchronize());
        ix->d_x.release();
        ix->d_x = nx;
        d_perm.release();
    }
    if (ix->d_x.p == nullptr && ix->d_x.ensure(16) != 0) return RK_ERR_CUDA;
    ix->doc_row_start.assign(D + 1, 0);
    for (size_t i = 0; i < n; i++) ix->doc_row_start[ix->h_rowdoc[i] + 1]++;
    for (size_t d = 0; d < D; d++) ix->doc_row_start[d + 1] += ix->doc_row_start[d];

    // users
    for (auto &pr : ix->user_roles) ix->users[pr.first].roles.push_back(pr.second);
    std::map<std::vector<uint32_t>, std::pair<uint64_t, uint64_t>> set_rows;   // role set -> (rows, docs)
    std::map<std::vector<uint64_t>, uint32_t> mask_ids;
    for (auto &kv : ix->users) {
        UserInfo &u = kv.second;
        std::sort(u.roles.begin(), u.roles.end());
        u.roles.erase(std::unique(u.roles.begin(), u.roles.end()), u.roles.end());
        u.mask.assign(W, 0);
        role_mask_of(ix, u.roles, u.mask.data());
        auto it = set_rows.find(u.roles);
        if (it == set_rows.end()) {
            uint64_t rows = 0, docs = 0;
            for (size_t d = 0; d < D; d++)
                if (mask_hit(&ix->doc_mask[d * W], u.mask.data(), W)) {
                    rows += ix->doc_row_start[d + 1] - ix->doc_row_start[d];
                    docs++;
                }
            it = set_rows.emplace(u.roles, std::make_pair(rows, docs)).first;
        }
        u.allowed_rows = it->second.first;
        u.allowed_docs = it->second.second;
        u.mask_id = mask_ids.emplace(u.mask, (uint32_t) mask_ids.size()).first->second;
    }
    ix->n_mask_ids = (uint32_t) mask_ids.size();

    // partitions; [0] = every document (bitmask / unfiltered scans walk it)
    ix->parts.clear();
    ix->parts.emplace_back();
    ix->parts[0].docs.resize(D);
    for (size_t d = 0; d < D; d++) ix->parts[0].docs[d] = (uint32_t) d;
    auto docs_matching = [&](const uint64_t *m, std::vector<uint32_t> &out) {
        out.clear();
        for (size_t d = 0; d < D; d++)
            if (mask_hit(&ix->doc_mask[d * W], m, W)) out.push_back((uint32_t) d);
    };
    if (ix->want_role_parts) {
        // initialize_partitions.py:302-311 — documentblocks_role_<r>
        std::vector<uint64_t> m(W);
        for (uint32_t r : ix->role_ids) {
            role_mask_of(ix, std::vector<uint32_t>{r}, m.data());
            PartInfo pi;
            docs_matching(m.data(), pi.docs);
            ix->role_part[r] = (uint32_t) ix->parts.size();
            ix->parts.push_back(std::move(pi));
        }
    }
    if (ix->want_combo_parts) {
        // initialize_partitions.py:549-555, 682-693 — one table per distinct role set, DISTINCT union
        std::vector<uint64_t> m(W);
        for (auto &kv : ix->users) {
            const auto &rs = kv.second.roles;
            if (rs.empty() || ix->combo_part.count(rs)) continue;
            role_mask_of(ix, rs, m.data());
            PartInfo pi;
            docs_matching(m.data(), pi.docs);
            ix->combo_part[rs] = (uint32_t) ix->parts.size();
            ix->parts.push_back(std::move(pi));
        }
    }
    for (auto &kv : ix->part_defs) {
        // load_result_to_database.py:207-250 — documentblocks_partition_<p>
        PartInfo pi;
        for (uint32_t raw : kv.second) {
            const size_t d = raw_index(raw);
            if (d != D) pi.docs.push_back(new_of_raw[d]);
        }
        std::sort(pi.docs.begin(), pi.docs.end());
        pi.docs.erase(std::unique(pi.docs.begin(), pi.docs.end()), pi.docs.end());
        ix->dyn_part[kv.first] = (uint32_t) ix->parts.size();
        ix->parts.push_back(std::move(pi));
    }
    for (auto &kv : ix->roleset_parts)
        for (uint32_t pid : kv.second)
            if (!ix->dyn_part.count(pid)) return rk_fail(RK_ERR_INVALID, "rk_finalize: role set maps to undefined partition %u", pid);

    // extent tables: one extent per document, plus a sentinel per partition
    std::vector<uint32_t> ext_row, ext_vstart, part_off;
    std::vector<uint64_t> ext_mask;
    for (auto &pi : ix->parts) {
        pi.ext_off = (uint32_t) ext_row.size();
        part_off.push_back(pi.ext_off);
        uint64_t v = 0;
        pi.vstart.reserve(pi.docs.size() + 1);
        for (uint32_t d : pi.docs) {
            uint32_t r0 = ix->doc_row_start[d], r1 = ix->doc_row_start[d + 1];
            ext_row.push_back(r0);
            ext_vstart.push_back((uint32_t) v);
            pi.vstart.push_back((uint32_t) v);
            for (int w = 0; w < W; w++) ext_mask.push_back(ix->doc_mask[(size_t) d * W + w]);
            v += r1 - r0;
        }
        pi.vstart.push_back((uint32_t) v);
        pi.n_ext = (uint32_t) pi.docs.size();
        pi.rows = v;
        if (v >= 0xFFFFFFFFull) return rk_fail(RK_ERR_UNSUPPORTED, "rk_finalize: partition with 2^32 rows");
        ext_row.push_back(0);
        ext_vstart.push_back((uint32_t) v);
        for (int w = 0; w < W; w++) ext_mask.push_back(0);
    }
    part_off.push_back((uint32_t) ext_row.size());

    ix->device_bytes = ix->d_x.cap;
    if (upload(ix->d_tid, ix->h_tid, &ix->device_bytes) != 0) return RK_ERR_CUDA;
    if (upload(ix->d_rowdoc, ix->h_rowdoc, &ix->device_bytes) != 0) return RK_ERR_CUDA;
    if (upload(ix->d_doc_ids, ix->doc_ids, &ix->device_bytes) != 0) return RK_ERR_CUDA;
    if (upload(ix->d_doc_mask, ix->doc_mask, &ix->device_bytes) != 0) return RK_ERR_CUDA;
    if (upload(ix->d_ext_row, ext_row, &ix->device_bytes) != 0) return RK_ERR_CUDA;
    if (upload(ix->d_ext_vstart, ext_vstart, &ix->device_bytes) != 0) return RK_ERR_CUDA;
    if (upload(ix->d_ext_mask, ext_mask, &ix->device_bytes) != 0) return RK_ERR_CUDA;
    if (upload(ix->d_part_ext_off, part_off, &ix->device_bytes) != 0) return RK_ERR_CUDA;
    std::vector<uint32_t>().swap(ix->h_doc);
    ix->finalized = true;
    return RK_OK;
}
