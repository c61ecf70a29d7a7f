// rk_multi.cu — several GPUs behind ONE handle in one process: what `rk_config{…, n_gpus, …}` was to be.
//
// The reference's nearest analogue is FAISS's IndexShards — host threads over sub-indexes with a CPU merge
// (logical_partition_benchmark/benchmark/faiss/faiss/IndexShards.cpp:197-256).  Here a C caller (the PostgreSQL AM shim, the
// C++ benches of INTEGRATION.md §3) gets the document-striped layout of sharded.py without torch or NCCL: rk_multi owns one
// rk_index per device (document d lives on shard d mod G, so every role / combination / dynamic partition splits evenly),
// one rk_exchange ring per shard wired to the others by peer pointers, and one host thread per device for a search:
//   phase 1  stage the queries, search the shard, push the packed result into every peer's ring (rk_search_batch_shard_push)
//   phase 2  G-way merge out of the ring (rk_merge_shards_pull) — every shard merges, which acknowledges its ring slots;
//            shard 0 copies the answer to the caller's host buffers.
// The phases are separated by a host-side join: the call is synchronous anyway, and on a box where one physical GPU is
// listed twice (tests) no merge kernel may spin on a flag whose producer still waits for its SMs.
#include <thread>

#include "rk_host.h"
#include "rk_kernels.h"

#define RK_CUDA(call)                                                                                    \
    do {                                                                                                 \
        cudaError_t e__ = (call);                                                                        \
        if (e__ != cudaSuccess)                                                                          \
            return rk_fail(RK_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), __FILE__, __LINE__); \
    } while (0)

struct rk_exchange;
extern "C" {
int rk_exchange_create(rk_index *ix, size_t nq_max, int k_max, int world, int rank, int depth, rk_exchange **out);
int rk_exchange_destroy(rk_exchange *ex);
void *rk_exchange_base(rk_exchange *ex);
int rk_exchange_connect_ptr(rk_exchange *ex, int peer, void *peer_base, int peer_device);
int rk_search_batch_shard_push(rk_index *ix, rk_exchange *ex, size_t nq, const float *dQ, const uint32_t *user_ids,
                               const rk_scan_opts *o, const uint32_t *d_lists);
int rk_merge_shards_pull(rk_index *ix, rk_exchange *ex, size_t nq, int k, void *cuda_stream, uint64_t *d_out_tid, uint32_t *d_out_doc,
                         double *d_out_dist, int32_t *d_out_count);
int rk_exchange_status(rk_exchange *ex, void *cuda_stream);
}
int rk_stage_host_queries(rk_index *ix, size_t nq, const float *Q, const float **dQ);

struct rk_multi {
    int G = 0;
    size_t nq_max = 0;
    int k_max = 0;
    std::vector<rk_index *> ix;
    std::vector<rk_exchange *> ex;
    std::vector<DevBuf> o_tid, o_doc, o_dist, o_cnt;
    bool finalized = false;
    std::mutex mu;
};

// runs f(g) for every shard on its own host thread; returns the first error
template <typename F>
static int for_each_shard(rk_multi *m, F f)
{
    std::vector<int> rc((size_t) m->G, RK_OK);
    std::vector<std::string> msg((size_t) m->G);
    std::vector<std::thread> th;
    for (int g = 1; g < m->G; g++)
        th.emplace_back([&, g] { rc[g] = f(g); if (rc[g] < 0) msg[g] = rk_last_error(); });
    rc[0] = f(0);
    if (rc[0] < 0) msg[0] = rk_last_error();
    for (auto &t : th) t.join();
    for (int g = 0; g < m->G; g++)
        if (rc[g] < 0) return rk_fail(rc[g], "shard %d: %s", g, msg[g].c_str());      // (the error text is thread-local: carry it over)
    return RK_OK;
}

extern "C" rk_multi *rk_multi_open(const rk_config *cfg, const int32_t *devices, int32_t n_gpus, size_t nq_max)
{
    if (!cfg || !devices || n_gpus < 1 || n_gpus > RK_MAX_PEERS || nq_max < 1) { rk_fail(RK_ERR_INVALID, "rk_multi_open: bad argument"); return nullptr; }
    rk_multi *m = new rk_multi();
    m->G = n_gpus; m->nq_max = nq_max; m->k_max = cfg->k_max;
    for (int g = 0; g < n_gpus; g++) {
        rk_config c = *cfg;
        c.device = devices[g];
        rk_index *ix = rk_open(&c);
        if (!ix) { for (auto *p : m->ix) rk_close(p); delete m; return nullptr; }
        m->ix.push_back(ix);
    }
    m->ex.assign((size_t) n_gpus, nullptr);
    m->o_tid.resize((size_t) n_gpus); m->o_doc.resize((size_t) n_gpus); m->o_dist.resize((size_t) n_gpus); m->o_cnt.resize((size_t) n_gpus);
    return m;
}

extern "C" void rk_multi_close(rk_multi *m)
{
    if (!m) return;
    for (int g = 0; g < m->G; g++) {
        cudaSetDevice(m->ix[g]->device);
        cudaDeviceSynchronize();
    }
    for (int g = 0; g < m->G; g++) {
        if (m->ex[g]) rk_exchange_destroy(m->ex[g]);
        cudaSetDevice(m->ix[g]->device);
        m->o_tid[g].release(); m->o_doc[g].release(); m->o_dist[g].release(); m->o_cnt[g].release();
    }
    for (auto *p : m->ix) rk_close(p);
    delete m;
}

extern "C" int rk_multi_n_gpus(const rk_multi *m) { return m ? m->G : 0; }
extern "C" rk_index *rk_multi_shard(rk_multi *m, int g) { return (m && g >= 0 && g < m->G) ? m->ix[g] : nullptr; }

extern "C" int rk_multi_load_rows(rk_multi *m, const float *x, const uint64_t *tid, const uint32_t *doc, size_t n)
{
    if (!m || (!x && n) || (!tid && n) || (!doc && n)) return rk_fail(RK_ERR_INVALID, "rk_multi_load_rows: NULL argument");
    if (m->finalized) return rk_fail(RK_ERR_STATE, "rk_multi_load_rows: index already finalized");
    const size_t dim = (size_t) m->ix[0]->dim;
    // a document's rows stay together on shard (document_id mod G)
    return for_each_shard(m, [&](int g) -> int {
        std::vector<float> xs;
        std::vector<uint64_t> ts;
        std::vector<uint32_t> ds;
        for (size_t i = 0; i < n; i++) {
            if ((int) (doc[i] % (uint32_t) m->G) != g) continue;
            xs.insert(xs.end(), x + i * dim, x + (i + 1) * dim);
            ts.push_back(tid[i]);
            ds.push_back(doc[i]);
        }
        return ts.empty() ? RK_OK : rk_load_rows(m->ix[g], xs.data(), ts.data(), ds.data(), ts.size());
    });
}

extern "C" int rk_multi_set_doc_roles(rk_multi *m, const uint32_t *doc, const uint32_t *role, size_t n)
{
    if (!m) return rk_fail(RK_ERR_INVALID, "NULL index");
    // every shard registers ALL pairs (as each rank of sharded.py does): role numbering and partition ids stay global, a
    // document without rows on a shard is an empty extent there
    return for_each_shard(m, [&](int g) -> int { return rk_set_doc_roles(m->ix[g], doc, role, n); });
}

extern "C" int rk_multi_set_user_roles(rk_multi *m, const uint32_t *user, const uint32_t *role, size_t n)
{
    if (!m) return rk_fail(RK_ERR_INVALID, "NULL index");
    return for_each_shard(m, [&](int g) -> int { return rk_set_user_roles(m->ix[g], user, role, n); });
}

extern "C" int rk_multi_build_role_partitions(rk_multi *m)
{
    if (!m) return rk_fail(RK_ERR_INVALID, "NULL index");
    return for_each_shard(m, [&](int g) -> int { return rk_build_role_partitions(m->ix[g]); });
}

extern "C" int rk_multi_build_combination_partitions(rk_multi *m)
{
    if (!m) return rk_fail(RK_ERR_INVALID, "NULL index");
    return for_each_shard(m, [&](int g) -> int { return rk_build_combination_partitions(m->ix[g]); });
}

extern "C" int rk_multi_define_partition(rk_multi *m, uint32_t part_id, const uint32_t *docs, size_t n)
{
    if (!m) return rk_fail(RK_ERR_INVALID, "NULL index");
    return for_each_shard(m, [&](int g) -> int { return rk_define_partition(m->ix[g], part_id, docs, n); });
}

extern "C" int rk_multi_map_roleset_to_partitions(rk_multi *m, const uint32_t *roles, size_t nroles, const uint32_t *parts, size_t nparts)
{
    if (!m) return rk_fail(RK_ERR_INVALID, "NULL index");
    return for_each_shard(m, [&](int g) -> int { return rk_map_roleset_to_partitions(m->ix[g], roles, nroles, parts, nparts); });
}

extern "C" int rk_multi_finalize(rk_multi *m)
{
    if (!m) return rk_fail(RK_ERR_INVALID, "NULL index");
    if (m->finalized) return RK_OK;
    int rc = for_each_shard(m, [&](int g) -> int { return rk_finalize(m->ix[g]); });
    if (rc < 0) return rc;
    // the rings, wired all-to-all by pointer (one process: no IPC handles)
    const size_t nk = m->nq_max * (size_t) m->k_max;
    for (int g = 0; g < m->G; g++) {
        rc = rk_exchange_create(m->ix[g], m->nq_max, m->k_max, m->G, g, 4, &m->ex[g]);
        if (rc < 0) return rc;
        RK_CUDA(cudaSetDevice(m->ix[g]->device));
        if (m->o_tid[g].ensure(nk * 8) != 0 || m->o_doc[g].ensure(nk * 4) != 0 || m->o_dist[g].ensure(nk * 8) != 0 ||
            m->o_cnt[g].ensure(m->nq_max * 4) != 0)
            return RK_ERR_CUDA;
    }
    for (int a = 0; a < m->G; a++)
        for (int b = 0; b < m->G; b++) {
            if (a == b) continue;
            rc = rk_exchange_connect_ptr(m->ex[a], b, rk_exchange_base(m->ex[b]), m->ix[b]->device);
            if (rc < 0) return rc;
        }
    m->finalized = true;
    return RK_OK;
}

// IVFFlat over the stripes: shard 0 trains the centres on its stripe — a uniform 1/G sample of the table
// pgvector samples from (ivfbuild.c:404-441) — and every other shard assigns its rows to them (ivfbuild.c:141-198); all shards
// then choose the same lists for a query.
extern "C" int rk_multi_build_ivf(rk_multi *m, int lists, uint64_t seed)
{
    if (!m || !m->finalized) return rk_fail(RK_ERR_STATE, "rk_multi_build_ivf: call rk_multi_finalize first");
    int rc = rk_build_ivf(m->ix[0], lists, seed);
    if (rc < 0 || m->G == 1) return rc;
    int got = 0;
    rc = rk_ivf_info(m->ix[0], &got, nullptr, nullptr);
    if (rc < 0) return rc;
    std::vector<float> centres((size_t) got * (size_t) m->ix[0]->dim);
    rc = rk_ivf_info(m->ix[0], &got, centres.data(), nullptr);
    if (rc < 0) return rc;
    return for_each_shard(m, [&](int g) -> int { return g == 0 ? RK_OK : rk_build_ivf_from_centers(m->ix[g], got, centres.data()); });
}

extern "C" int rk_multi_search_batch(rk_multi *m, size_t nq, const float *Q, const uint32_t *user_ids, const rk_scan_opts *o,
                                     uint64_t *out_tid, uint32_t *out_doc, double *out_dist, int32_t *out_count)
{
    if (!m) return rk_fail(RK_ERR_INVALID, "NULL index");
    if (!m->finalized) return rk_fail(RK_ERR_STATE, "rk_multi_search_batch: call rk_multi_finalize first");
    if (!o || !Q || !out_tid || !out_dist || !out_count) return rk_fail(RK_ERR_INVALID, "rk_multi_search_batch: NULL argument");
    if (nq == 0) return RK_OK;
    if (nq > m->nq_max || o->k > m->k_max) return rk_fail(RK_ERR_INVALID, "rk_multi_search_batch: nq = %zu / k = %d exceed what rk_multi_open reserved (%zu / %d)", nq, o->k, m->nq_max, m->k_max);
    if (!o->null_query && !rk_all_finite(Q, nq * (size_t) m->ix[0]->dim)) return rk_fail(RK_ERR_NONFINITE, "NaN or infinite value not allowed in vector");
    std::lock_guard<std::mutex> lk(m->mu);
    const size_t nk = nq * (size_t) o->k;
    // phase 1: every shard searches and pushes
    int rc = for_each_shard(m, [&](int g) -> int {
        rk_index *ix = m->ix[g];
        const float *dQ = nullptr;
        int r = rk_stage_host_queries(ix, nq, Q, &dQ);
        if (r < 0) return r;
        r = rk_search_batch_shard_push(ix, m->ex[g], nq, dQ, user_ids, o, nullptr);
        if (r < 0) return r;
        RK_CUDA(cudaStreamSynchronize(ix->ws.stream));
        return RK_OK;
    });
    if (rc < 0) return rc;
    // phase 2: every shard merges (and so acknowledges its ring slot); shard 0 hands the answer back
    return for_each_shard(m, [&](int g) -> int {
        rk_index *ix = m->ix[g];
        int r = rk_merge_shards_pull(ix, m->ex[g], nq, o->k, nullptr, m->o_tid[g].as<uint64_t>(), out_doc ? m->o_doc[g].as<uint32_t>() : nullptr,
                                     m->o_dist[g].as<double>(), m->o_cnt[g].as<int32_t>());
        if (r < 0) return r;
        RK_CUDA(cudaSetDevice(ix->device));
        if (g == 0) {
            RK_CUDA(cudaMemcpyAsync(out_tid, m->o_tid[0].p, nk * 8, cudaMemcpyDeviceToHost, ix->ws.stream));
            if (out_doc) RK_CUDA(cudaMemcpyAsync(out_doc, m->o_doc[0].p, nk * 4, cudaMemcpyDeviceToHost, ix->ws.stream));
            RK_CUDA(cudaMemcpyAsync(out_dist, m->o_dist[0].p, nk * 8, cudaMemcpyDeviceToHost, ix->ws.stream));
            RK_CUDA(cudaMemcpyAsync(out_count, m->o_cnt[0].p, nq * 4, cudaMemcpyDeviceToHost, ix->ws.stream));
        }
        return rk_exchange_status(m->ex[g], nullptr);      // synchronises the stream; reports a peer wait that timed out
    });
}
