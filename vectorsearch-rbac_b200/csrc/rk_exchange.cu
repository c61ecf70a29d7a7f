// rk_exchange.cu — the multi-GPU candidate exchange over peer memory (NVLink / NVSwitch), without a collective library on
// the data path.
//
// The reference has no multi-GPU path (nearest analogue: FAISS's host-thread shards with a CPU merge,
// logical_partition_benchmark/benchmark/faiss/faiss/IndexShards.cpp:197-256).  Here every rank owns a RING of receive
// slots in its own HBM, [depth][world] blocks of one packed shard result {key f32 | doc u32 | tid u64}[nq * k] each, plus an
// arrival flag and an acknowledgement flag per (slot, rank).  A step with sequence number s uses slot s mod depth:
//
//   local search   writes this rank's result straight into its own block of the slot (rk_search_batch_shard_push),
//   push kernel    stores that block into the same block of every PEER's slot over NVLink (16-byte stores, a few CTAs per
//                  peer), fences system-wide and raises the peer's arrival flag [slot][rank] = s — after waiting, if need
//                  be, for the peer's acknowledgement of step s - depth (the slot's previous use),
//   merge kernel   (rk_select_kernel in shard mode, rk_merge_shards_pull) waits for the arrival flags of all peers, merges
//                  the `world` lists per query, and its last block acknowledges [slot][rank] = s on every peer.
//
// Nothing but flags synchronises the ranks, so a rank that is ahead keeps scanning (up to `depth` steps) instead of
// waiting in a collective: the merge of step s can run on a side stream under the scan of step s + 1.  Memory comes from
// cudaMalloc and travels between processes as a CUDA IPC handle (one process per GPU), or as a plain pointer with peer
// access enabled (several GPUs in one process).  All waits are bounded (rk_wait_flag_sys).
#include "rk_host.h"
#include "rk_kernels.h"

#define RK_CUDA(call)                                                                                    \
    do {                                                                                                 \
        cudaError_t e__ = (call);                                                                        \
        if (e__ != cudaSuccess)                                                                          \
            return rk_fail(RK_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), __FILE__, __LINE__); \
    } while (0)

struct rk_exchange {
    rk_index *ix = nullptr;
    int world = 0, rank = 0, depth = 0;
    size_t nq_max = 0;
    int k_max = 0;
    size_t block_bytes = 0;          // one rank's packed result: 16 * nq_max * k_max
    size_t flags_off = 0, acks_off = 0, ctl_off = 0, total_bytes = 0;
    unsigned char *base = nullptr;   // this rank's allocation
    unsigned char *peer[RK_MAX_PEERS] = {nullptr};
    bool ipc_opened[RK_MAX_PEERS] = {false};
    uint32_t seq_push = 0, seq_pull = 0;
    uint32_t *h_err = nullptr;       // pinned copy target of the error word

    unsigned char *slot_block(unsigned char *b, uint32_t seq, int src) const
    {
        return b + ((size_t) (seq % (uint32_t) depth) * (size_t) world + (size_t) src) * block_bytes;
    }
    uint32_t *flag(unsigned char *b, uint32_t seq, int src) const
    {
        return reinterpret_cast<uint32_t *>(b + flags_off) + (size_t) (seq % (uint32_t) depth) * world + src;
    }
    uint32_t *ack(unsigned char *b, uint32_t seq, int by) const
    {
        return reinterpret_cast<uint32_t *>(b + acks_off) + (size_t) (seq % (uint32_t) depth) * world + by;
    }
    uint32_t *ctl(int i) const { return reinterpret_cast<uint32_t *>(base + ctl_off) + i; }     // [0] error, [1] merge blocks done, [2 + g] push CTAs done
};

struct RkPushParams {
    const uint4 *src;                // this rank's block (own slot)
    uint4 *dst[RK_MAX_PEERS];        // the same block of the slot on peer g
    uint32_t *flag[RK_MAX_PEERS];    // arrival flag [slot][rank] on peer g
    const uint32_t *ack[RK_MAX_PEERS];   // LOCAL acknowledgement [slot][g]: peer g has merged the slot's previous step
    int peer_rank[RK_MAX_PEERS];
    int n_peers, ctas_per_peer;
    size_t n16;                      // 16-byte words to copy
    uint32_t seq, ack_need;          // ack_need = seq - depth (0 = slot never used)
    uint32_t *done;                  // [n_peers] CTAs finished (returns to 0)
    uint32_t *xerr;
};

__global__ void __launch_bounds__(256) rk_push_kernel(const RkPushParams p)
{
    const int g = blockIdx.x / p.ctas_per_peer, c = blockIdx.x % p.ctas_per_peer;
    __shared__ int s_ok;
    if (threadIdx.x == 0) {
        s_ok = 1;
        if (p.ack_need != 0 && !rk_wait_flag_sys(p.ack[g], p.ack_need)) { *p.xerr = 2u; s_ok = 0; }
    }
    __syncthreads();
    if (s_ok) {
        const size_t per = (p.n16 + p.ctas_per_peer - 1) / p.ctas_per_peer;
        const size_t a = per * c, b = a + per < p.n16 ? a + per : p.n16;
        uint4 *d = p.dst[g];
        for (size_t i0 = a + threadIdx.x; i0 < b; i0 += 256 * 4) {
            uint4 v[4];
#pragma unroll
            for (int j = 0; j < 4; j++) if (i0 + (size_t) j * 256 < b) v[j] = __ldcg(p.src + i0 + (size_t) j * 256);
#pragma unroll
            for (int j = 0; j < 4; j++) if (i0 + (size_t) j * 256 < b) d[i0 + (size_t) j * 256] = v[j];
        }
    }
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        if (atomicAdd(p.done + g, 1u) == (uint32_t) p.ctas_per_peer - 1) {
            p.done[g] = 0;
            __threadfence_system();
            if (s_ok) rk_st_release_sys(p.flag[g], p.seq);
        }
    }
}

extern "C" int rk_exchange_create(rk_index *ix, size_t nq_max, int k_max, int world, int rank, int depth, rk_exchange **out)
{
    if (!ix || !out) return rk_fail(RK_ERR_INVALID, "rk_exchange_create: NULL argument");
    if (world < 1 || world > RK_MAX_PEERS || rank < 0 || rank >= world) return rk_fail(RK_ERR_INVALID, "rk_exchange_create: world = %d, rank = %d", world, rank);
    if (depth < 2 || depth > 16) return rk_fail(RK_ERR_INVALID, "rk_exchange_create: depth must be 2..16");
    if (nq_max < 1 || k_max < 1 || k_max > RK_MAX_K) return rk_fail(RK_ERR_INVALID, "rk_exchange_create: nq_max / k_max");
    RK_CUDA(cudaSetDevice(ix->device));
    rk_exchange *ex = new rk_exchange();
    ex->ix = ix; ex->world = world; ex->rank = rank; ex->depth = depth; ex->nq_max = nq_max; ex->k_max = k_max;
    ex->block_bytes = ((size_t) 16 * nq_max * (size_t) k_max + 255) & ~(size_t) 255;
    ex->flags_off = ex->block_bytes * (size_t) world * (size_t) depth;
    ex->acks_off = ex->flags_off + (((size_t) depth * world * 4 + 255) & ~(size_t) 255);
    ex->ctl_off = ex->acks_off + (((size_t) depth * world * 4 + 255) & ~(size_t) 255);
    ex->total_bytes = ex->ctl_off + 256;
    if (cudaMalloc((void **) &ex->base, ex->total_bytes) != cudaSuccess) { delete ex; return rk_fail(RK_ERR_CUDA, "rk_exchange_create: cudaMalloc of %zu bytes failed", ex->total_bytes); }
    cudaMemset(ex->base + ex->flags_off, 0, ex->total_bytes - ex->flags_off);
    cudaHostAlloc((void **) &ex->h_err, 16, cudaHostAllocDefault);
    if (ex->h_err) *ex->h_err = 0;
    cudaDeviceSynchronize();
    ex->peer[rank] = ex->base;
    *out = ex;
    return RK_OK;
}

extern "C" int rk_exchange_destroy(rk_exchange *ex)
{
    if (!ex) return RK_OK;
    cudaSetDevice(ex->ix->device);
    cudaDeviceSynchronize();
    for (int g = 0; g < ex->world; g++)
        if (ex->ipc_opened[g] && ex->peer[g]) cudaIpcCloseMemHandle(ex->peer[g]);
    if (ex->base) cudaFree(ex->base);
    if (ex->h_err) cudaFreeHost(ex->h_err);
    delete ex;
    return RK_OK;
}

extern "C" int rk_exchange_get_handle(rk_exchange *ex, void *handle64)
{
    if (!ex || !handle64) return rk_fail(RK_ERR_INVALID, "rk_exchange_get_handle: NULL argument");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "CUDA IPC handle size");
    cudaIpcMemHandle_t h;
    RK_CUDA(cudaSetDevice(ex->ix->device));
    RK_CUDA(cudaIpcGetMemHandle(&h, ex->base));
    memcpy(handle64, &h, 64);
    return RK_OK;
}

extern "C" void *rk_exchange_base(rk_exchange *ex) { return ex ? ex->base : nullptr; }
extern "C" size_t rk_exchange_block_bytes(rk_exchange *ex) { return ex ? ex->block_bytes : 0; }

extern "C" int rk_exchange_connect_ipc(rk_exchange *ex, int peer, const void *handle64)
{
    if (!ex || !handle64 || peer < 0 || peer >= ex->world) return rk_fail(RK_ERR_INVALID, "rk_exchange_connect_ipc: bad argument");
    if (peer == ex->rank) return RK_OK;
    cudaIpcMemHandle_t h;
    memcpy(&h, handle64, 64);
    RK_CUDA(cudaSetDevice(ex->ix->device));
    void *p = nullptr;
    cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) { cudaGetLastError(); return rk_fail(RK_ERR_CUDA, "rk_exchange_connect_ipc: cudaIpcOpenMemHandle(rank %d): %s", peer, cudaGetErrorString(e)); }
    ex->peer[peer] = (unsigned char *) p;
    ex->ipc_opened[peer] = true;
    return RK_OK;
}

extern "C" int rk_exchange_connect_ptr(rk_exchange *ex, int peer, void *peer_base, int peer_device)
{
    if (!ex || !peer_base || peer < 0 || peer >= ex->world) return rk_fail(RK_ERR_INVALID, "rk_exchange_connect_ptr: bad argument");
    if (peer == ex->rank) return RK_OK;
    RK_CUDA(cudaSetDevice(ex->ix->device));
    if (peer_device >= 0 && peer_device != ex->ix->device) {
        cudaError_t e = cudaDeviceEnablePeerAccess(peer_device, 0);
        if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); return rk_fail(RK_ERR_CUDA, "rk_exchange_connect_ptr: no peer access %d -> %d: %s", ex->ix->device, peer_device, cudaGetErrorString(e)); }
        cudaGetLastError();
    }
    ex->peer[peer] = (unsigned char *) peer_base;
    return RK_OK;
}

int rk_search_batch_shard_locked_for_exchange(rk_index *ix, size_t nq, const float *dQ, const uint32_t *user_ids, const rk_scan_opts *o,
                                              const uint32_t *d_lists, float *d_key, uint64_t *d_tid, uint32_t *d_doc);

extern "C" int rk_search_batch_shard_push(rk_index *ix, rk_exchange *ex, size_t nq, const float *dQ, const uint32_t *user_ids,
                                          const rk_scan_opts *o, const uint32_t *d_lists)
{
    if (!ix || !ex || ex->ix != ix) return rk_fail(RK_ERR_INVALID, "rk_search_batch_shard_push: bad handle");
    if (!o || nq < 1 || nq > ex->nq_max || o->k > ex->k_max) return rk_fail(RK_ERR_INVALID, "rk_search_batch_shard_push: nq / k exceed the exchange's capacity");
    for (int g = 0; g < ex->world; g++)
        if (!ex->peer[g]) return rk_fail(RK_ERR_STATE, "rk_search_batch_shard_push: rank %d is not connected", g);
    const uint32_t seq = ++ex->seq_push;
    const size_t n = nq * (size_t) o->k;
    unsigned char *mine = ex->slot_block(ex->base, seq, ex->rank);
    std::lock_guard<std::mutex> lk(ix->batch_mu);
    int rc = rk_search_batch_shard_locked_for_exchange(ix, nq, dQ, user_ids, o, d_lists, reinterpret_cast<float *>(mine),
                                                       reinterpret_cast<uint64_t *>(mine + 8 * n), reinterpret_cast<uint32_t *>(mine + 4 * n));
    if (rc < 0) { --ex->seq_push; return rc; }
    if (ex->world == 1) return RK_OK;
    RkPushParams p;
    memset(&p, 0, sizeof p);
    p.src = reinterpret_cast<const uint4 *>(mine);
    p.n16 = n;                                   // 16 bytes per (query, rank) entry
    p.seq = seq;
    p.ack_need = seq > (uint32_t) ex->depth ? seq - (uint32_t) ex->depth : 0u;
    p.ctas_per_peer = 4;
    p.done = ex->ctl(2);
    p.xerr = ex->ctl(0);
    int np = 0;
    for (int g = 0; g < ex->world; g++) {
        if (g == ex->rank) continue;
        p.dst[np] = reinterpret_cast<uint4 *>(ex->slot_block(ex->peer[g], seq, ex->rank));
        p.flag[np] = ex->flag(ex->peer[g], seq, ex->rank);
        p.ack[np] = ex->ack(ex->base, seq, g);
        p.peer_rank[np] = g;
        np++;
    }
    p.n_peers = np;
    rk_push_kernel<<<np * p.ctas_per_peer, 256, 0, ix->ws.stream>>>(p);
    RK_CUDA(cudaGetLastError());
    ix->total_launches += 1;
    ix->ws.launches += 1;
    return RK_OK;
}

extern "C" int rk_merge_shards_pull(rk_index *ix, rk_exchange *ex, size_t nq, int k, void *cuda_stream, uint64_t *d_out_tid,
                                    uint32_t *d_out_doc, double *d_out_dist, int32_t *d_out_count)
{
    if (!ix || !ex || ex->ix != ix) return rk_fail(RK_ERR_INVALID, "rk_merge_shards_pull: bad handle");
    if (nq < 1 || nq > ex->nq_max || k < 1 || k > ex->k_max) return rk_fail(RK_ERR_INVALID, "rk_merge_shards_pull: nq / k exceed the exchange's capacity");
    if (ex->seq_pull >= ex->seq_push) return rk_fail(RK_ERR_STATE, "rk_merge_shards_pull: no pushed step is waiting");
    const uint32_t seq = ++ex->seq_pull;
    const size_t n = nq * (size_t) k;
    RK_CUDA(cudaSetDevice(ix->device));
    unsigned char *b0 = ex->slot_block(ex->base, seq, 0);
    RkSelectParams p;
    memset(&p, 0, sizeof p);
    p.cand_key = reinterpret_cast<const uint32_t *>(b0);
    p.cand_tid = reinterpret_cast<const uint64_t *>(b0 + 8 * n);
    p.cand_row = reinterpret_cast<const uint32_t *>(b0 + 4 * n);
    p.lists_per_query = ex->world;
    p.list_stride = k;
    p.shard_stride_bytes = ex->block_bytes;
    p.k = k;
    p.metric = ix->metric;
    p.key_is_float = 1;
    p.out_tid = d_out_tid; p.out_doc = d_out_doc; p.out_dist = d_out_dist; p.out_count = d_out_count;
    if (ex->world > 1) {
        p.wait_flag = ex->flag(ex->base, seq, 0);
        p.wait_n = ex->world;
        p.wait_self = ex->rank;
        p.wait_seq = seq;
        int np = 0;
        for (int g = 0; g < ex->world; g++)
            if (g != ex->rank) p.ack_flag[np++] = ex->ack(ex->peer[g], seq, ex->rank);
        p.ack_n = np;
        p.ack_seq = seq;
        p.ack_done = ex->ctl(1);
        p.xerr = ex->ctl(0);
    }
    cudaStream_t st = cuda_stream ? (cudaStream_t) cuda_stream : ix->ws.stream;
    RK_CUDA(rk_launch_select(p, (int) nq, st));
    ix->total_launches += 1;
    return RK_OK;
}

// 0 = no peer wait has timed out since the exchange was created; synchronises `cuda_stream` (NULL = the index stream).
extern "C" int rk_exchange_status(rk_exchange *ex, void *cuda_stream)
{
    if (!ex) return rk_fail(RK_ERR_INVALID, "NULL exchange");
    RK_CUDA(cudaSetDevice(ex->ix->device));
    cudaStream_t st = cuda_stream ? (cudaStream_t) cuda_stream : ex->ix->ws.stream;
    RK_CUDA(cudaMemcpyAsync(ex->h_err, ex->ctl(0), 4, cudaMemcpyDeviceToHost, st));
    RK_CUDA(cudaStreamSynchronize(st));
    if (*ex->h_err) return rk_fail(RK_ERR_CUDA, "peer exchange: a wait for rank data timed out (code %u: 1 = arrival flag in the merge, 2 = acknowledgement in the push)", *ex->h_err);
    return RK_OK;
}
