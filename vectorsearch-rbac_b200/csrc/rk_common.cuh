// rk_common.cuh — shared device/host definitions of librbacknn (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define RK_SEG_EXTENTS 0u   // virtual row range [begin,end) of an extent table (partition)
#define RK_SEG_ROWS 2u      // physical row range [begin,end), no predicate
#define RK_SEG_ROWMASK 3u   // physical row range [begin,end), role mask tested per row through rowdoc[]
#define RK_SEG_IVFQ 4u      // probed lists [begin,end) of the query's probe order; every list's rows are ordered by document class
                            // (distinct role mask), so the rows a user may see are a few extents per list: ivf_cls_off[]

#define RK_SEGF_MASKED 1u   // test the extent's role mask against the user's before loading it
#define RK_SEGF_MERGE_OK 2u // the query walks ONE partition: no tuple can reach it twice, so a CTA that runs two of its
                            // segments back to back keeps one candidate buffer (and one threshold) across them

#define RK_KEY_SENTINEL 0xFFFFFFFFu

// One unit of scan work: rows of ONE query's visible set.  Segments of a query are contiguous in
// the work list so that the select kernel finds them as [qseg[q], qseg[q+1]).
struct RkSegment {
    uint32_t query;   // index into the batch
    uint16_t kind;    // RK_SEG_*
    uint16_t flags;   // RK_SEGF_*
    uint32_t ext_lo;  // RK_SEG_EXTENTS: global index of the first extent touching [begin, end) ...
    uint32_t ext_hi;  // ... and one past the last (resolved on the host when the plan is built)
    uint64_t begin;   // virtual row range inside the partition (EXTENTS) or physical rows (ROWS / ROWMASK)
    uint64_t end;
};

// Device view of one row store (the main doc-sorted store or the IVF list-sorted copy).
struct RkStoreView {
    const float *x;            // [rows][dim_pad]
    const uint64_t *tid;       // [rows]
    const uint32_t *rowdoc;    // [rows] dense document index
    uint64_t rows;
};

struct RkScanParams {
    RkStoreView store;
    // permission data (main store is sorted by document)
    const uint64_t *doc_mask;       // [docs][mask_words] (row-mask segments go through store.rowdoc)
    const uint64_t *qmask;          // [nq][mask_words] role mask of each query's user
    int mask_words;
    // extent tables
    const uint32_t *ext_row;        // physical first row of each extent (+ sentinel per partition)
    const uint32_t *ext_vstart;     // virtual row offset of each extent inside its partition (+ sentinel = partition rows)
    const uint64_t *ext_mask;       // [extents][mask_words] role mask of the extent's document
    // IVFFlat lists ordered by document class (RK_SEG_IVFQ)
    const uint32_t *ivf_probe_list; // [nq][ivf_probes] lists of each query, nearest first
    const uint32_t *ivf_cls_off;    // [lists][ivf_ncls + 1] (+ 1 sentinel) first row of each (list, class) in the list-sorted store
    const uint64_t *ivf_qcls;       // [nq][2] classes the query's user may see (bit c of 128)
    int ivf_probes, ivf_ncls;
    // queries
    const float *q;                 // [nq][dim_pad]
    int null_query;                 // ORDER BY v <-> NULL: all distances 0
    // work list
    const RkSegment *segs;
    const uint32_t *seg_order;      // execution order: position -> segment index (NULL = identity)
    const uint32_t *nseg_ptr;       // number of segments (device resident so plans can be built on device)
    uint32_t *seg_counter;          // dynamic scheduler
    // per-segment output (k_pad entries each)
    uint32_t *cand_key;
    uint64_t *cand_tid;
    uint32_t *cand_row;
    int32_t *cand_count;
    // geometry
    int dim4;        // dim_pad / 4
    int row_bytes;   // dim_pad * 4
    int tile_rows;   // rows per pipeline stage
    int nstages;
    int k;
    int metric;      // 0 = L2, 1 = IP
    // streaming scans (rk_gettuple past the k-th tuple, ivfscan.c:375-388): keep only tuples whose (key, tid) is
    // strictly greater than the last pair already handed back
    int lb_on;
    uint32_t lb_key;
    uint64_t lb_tid;
    // Fused single-query mode: ONE launch per query.  Every segment belongs to the same query and no tuple can reach it
    // twice, so a CTA keeps one candidate buffer through all its segments, hands its k best (unsorted) to a flat global
    // list when the work list is drained, and the LAST CTA to finish selects the answer from that list and writes the
    // outputs — the scan + 148-list select of a pgvector-shaped scan (one query per ivfflatrescan, ivfscan.c:339-389)
    // without a second launch.  nseg comes by value and the control words are reset by the last CTA, so a call uploads
    // nothing but the query.
    int fused;
    uint32_t nseg_val;              // number of segments when nseg_ptr is NULL
    int fz_cap;                     // entries one CTA may hand over
    int fz_r;                       // rows a CTA vouches for in the cooperative threshold (about 3k / grid)
    int fz_seed;                    // > 0: the first compaction of a CTA runs as soon as it holds this many candidates
    uint32_t *fz_ctl;               // [0] entries handed over, [1] CTAs done, [2] segment counter, [4] rows vouched for (low
                                    // half) | CTAs reported (high half), [6] largest r-th key (all 0 between launches)
    uint32_t *fz_key;               // [grid * fz_cap]
    uint32_t *fz_row;
    const uint32_t *fz_doc_ids;     // dense document -> raw id
    uint64_t *fz_out_tid;           // [k] outputs (any may be NULL except status / count)
    uint32_t *fz_out_doc;
    double *fz_out_dist;
    float *fz_out_key;
    int32_t *fz_out_count;          // [0] tuples returned; -1 = the merge could not finish (too many ties): rerun unfused
    // Zero-copy hand-shake with the host (query and outputs live in mapped pinned host memory): the query is read straight
    // from the host buffer by the consumer warps, and once the outputs are written and fenced system-wide the last thread
    // stores fz_seq to *fz_flag, which the host polls — no cudaMemcpy and no stream synchronisation on the call path.
    uint32_t *fz_flag;
    uint32_t fz_seq;
    unsigned long long *fz_dbg;     // tuning aid (option single_debug): [grid + 1][8] globaltimer stamps of the launch's phases
};

#define RK_MAX_PEERS 16

struct RkSelectParams {
    const uint32_t *cand_key;
    const uint64_t *cand_tid;
    const uint32_t *cand_row;
    const int32_t *cand_count;
    const uint32_t *qseg;       // [nq+1] segment range per query
    int lists_per_query;        // if > 0: query q owns lists [q*L, (q+1)*L) (shard merge / fixed plans) and qseg is ignored
    int list_stride;            // entries between consecutive lists
    // list-id mode (tensor engine): query q owns lists lid0[q] + j * lstep[q], j < nlists[q]; a list's
    // entries start at id * list_stride and its count is cand_count[id]
    const uint32_t *lid0, *lstep, *nlists;
    // explicit mode: query q owns lists list_ids[qseg[q] .. qseg[q+1])
    const uint32_t *list_ids;
    int count_stride;           // fixed-list mode: cand_count[j * count_stride + q] (0 = number of queries launched)
    size_t shard_stride_bytes;  // fixed-list mode: bytes between list j and j+1 of one array (0 = dense nq*k elements)
    int k;
    int metric;
    int dedup;
    int null_query;
    int key_is_float;           // cand_key holds fp32 ordering values (exchange form) instead of uint keys
    int unsorted;               // lists are not sorted; only their LAST entry is their largest (tensor-engine nominations)
    int out_maxlast;            // the consumer only needs the k best as a SET with the largest entry last (no dedup): skip the sort
    // row -> document id (may be NULL when cand_row already holds document ids)
    const uint32_t *rowdoc;
    const uint32_t *doc_ids;
    // outputs (any may be NULL)
    uint64_t *out_tid;
    uint32_t *out_doc;
    double *out_dist;
    int32_t *out_count;
    float *out_key;             // exchange form: fp32 ordering value
    uint32_t *out_row;
    // Peer exchange (rk_exchange.cu): the lists were written into this GPU's memory by the other ranks over NVLink.
    // Before a block reads them it waits until wait_flag[j] has reached wait_seq for every j < wait_n except wait_self
    // (system-scope acquire; bounded — on timeout *xerr is set and the block goes on); when the last block of the launch
    // is done, ack_seq is stored to ack_flag[g] on every peer g < ack_n (system-scope release): the ring slot may be reused.
    const uint32_t *wait_flag;
    int wait_n, wait_self;
    uint32_t wait_seq;
    uint32_t *ack_flag[RK_MAX_PEERS];
    int ack_n;
    uint32_t ack_seq;
    uint32_t *ack_done;         // local counter of finished blocks (returns to 0)
    uint32_t *xerr;
};

// ---- batched tensor-core engine (rk_bscan.cu) ------------------------------------------------
// One unit = one tile of up to 128 queries x a contiguous range of 128-row tiles of the row store.
struct RkBUnit {
    uint32_t q0, nq;             // query slots [q0, q0 + nq) of the (group-ordered) query matrix
    uint32_t tile_begin, tile_end;   // tiles [begin, end) of the unit's tile list (or of the dense tiling)
    uint32_t tl_off;             // offset of the group's tile list in tile_rows[]
    uint32_t list_base;          // candidate list of slot q0 + i is list_base + i
    uint32_t tile_step;          // the unit visits tiles tile_begin, tile_begin + tile_step, ... (> 1: sampling pass)
    uint32_t mask_id;            // > 0: nominate only rows whose role mask equals unit_mask[mask_id - 1] (atoms plan)
};

struct RkBScanParams {
    const RkBUnit *units;
    uint32_t nunits;
    const uint32_t *cta_off;     // [grid + 1] balanced static schedule: CTA b runs units cta_units[cta_off[b] .. cta_off[b+1])
    const uint32_t *cta_units;
    const uint32_t *tile_rows;   // first row of every listed tile (NULL = dense tiling, tile t starts at row 128 t)
    uint64_t nrows;
    int dim_pad, kchunks, metric, kp, mask_words, use_mask;
    const float *xnorm;          // [rows + 128] squared norms (L2 only), zero padded
    const uint64_t *rowmask;     // [rows][mask_words] role mask of each row's document (words >= 2 read from here)
    const uint64_t *rowmask2;    // [rows + 128][2] first two mask words of each row, zero padded
    const float *qnorm;          // [nq]
    const uint64_t *qmask;       // [nq][mask_words]
    uint2 *scratch;              // [grid][128][1024] (approximate key, row) candidate buffers
    uint32_t *cand_key;          // [lists][kp] sorted by (key, row)
    uint32_t *cand_row;
    int32_t *cand_count;         // [lists]
    int32_t *trunc;              // [queries] set when ties at a threshold had to be dropped
    // optional seed thresholds (sampling pass): slot s starts at tau_key[s * kp + kp - 1] when tau_count[s] >= kp
    const float *tau_key;
    const int32_t *tau_count;
    const uint64_t *unit_mask;   // [masks][mask_words] exact row masks of the atoms plan's units (see RkBUnit::mask_id)
    const uint32_t *slot_q;      // slot -> query of the caller's batch (trunc[] is indexed by query); NULL = identity
    // Shared threshold of a query over ALL its units (a query of the atoms plan owns one unit per document class it sees, on
    // different CTAs): qtau[query] = smallest proven upper bound of the query's kp-th best approximate value.  Every
    // compaction lowers it (atomicMin), units start from it and re-read it every few tiles — so the units of a query nominate
    // kp rows between them, not kp each.  RK_KEY_SENTINEL before the launch; NULL = every unit on its own.
    uint32_t *qtau;
    // bf16 nomination (CTA-pair kernel): the tensor maps describe bf16 shadow copies of the rows and the queries, a k-chunk is
    // 64 elements (the same 128 bytes per row) and the MMA is kind::f16 with bf16 operands, fp32 accumulation
    int bf16;
    // compaction tuning: a mid-unit compaction keeps up to kp + loose_extra entries; once a compaction round has been called
    // (some buffer above the trigger) every buffer above round_mark is compacted in it
    uint32_t loose_extra, round_mark;
};

struct RkRescoreParams {
    const float *x;              // row store
    const uint64_t *tid;
    const uint32_t *rowdoc, *doc_ids;
    const float *q;              // [slots][dim_pad] (group-ordered)
    const float *qnorm;
    const uint64_t *unit_mask;   // [masks][mask_words] exact row masks of the atoms plan's units (see RkBUnit::mask_id)
    const uint32_t *slot_q;      // slot -> query index of the caller's batch (outputs are written there); NULL = identity
    int dim_pad, k, kp, metric, null_query;
    const float *sel_key;        // [nq][kp] approximate ordering values, ascending
    const uint32_t *sel_row;
    const int32_t *sel_count;
    const int32_t *trunc;
    float eps_scale, xnorm_max;  // error bound of the tensor-core value: eps_scale * sqrt(xnorm_max * qnorm) (0 = exact)
    // when set, the kernel decides itself whether TF32 was exact for this batch (no host round trip between the passes):
    // probe[0] = query elements that are not small integers, low word of probe[4] = bits of the largest squared query norm
    const unsigned long long *probe;
    int x_integer;               // the corpus holds only small integers
    float exact_absmax;          // largest |value| the nomination format holds exactly for integers (TF32: 2047, bf16: 256)
    float x_absmax;              // largest |value| of the corpus
    int wide_rows;               // the scan scores a row with a whole warp (tile_rows < 64): use its summation order
    uint64_t *out_tid;
    uint32_t *out_doc;
    double *out_dist;
    float *out_key;
    int32_t *out_count;
    int32_t *uncertain;          // [nq] 1 = the candidate set is not provably sufficient: re-run on the exact path
};

// ---------------------------------------------------------------------------------------------
// order-preserving float <-> uint key

__host__ __device__ __forceinline__ uint32_t rk_float_to_key(float v)
{
#ifdef __CUDA_ARCH__
    uint32_t u = __float_as_uint(v);
#else
    union { float f; uint32_t u; } c; c.f = v; uint32_t u = c.u;
#endif
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

__host__ __device__ __forceinline__ float rk_key_to_float(uint32_t k)
{
    uint32_t u = (k & 0x80000000u) ? (k & 0x7FFFFFFFu) : ~k;
#ifdef __CUDA_ARCH__
    return __uint_as_float(u);
#else
    union { float f; uint32_t u; } c; c.u = u; return c.f;
#endif
}

#ifdef __CUDACC__
__device__ __forceinline__ unsigned long long rk_globaltimer()
{
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// ---------------------------------------------------------------------------------------------
// PTX wrappers: mbarrier + 1-D bulk async copy (TMA unit, SASS UBLKCP)

__device__ __forceinline__ uint32_t rk_ld_volatile_u32(const uint32_t *p)
{
    uint32_t v;
    asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// ---- system-scope flags between GPUs (peer exchange, rk_exchange.cu) ----
__device__ __forceinline__ uint32_t rk_ld_acquire_sys(const uint32_t *p)
{
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

__device__ __forceinline__ void rk_st_release_sys(uint32_t *p, uint32_t v)
{
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// Waits until *flag has reached seq (sequence numbers only grow; compared modulo 2^32).  Bounded: false after
// RK_FLAG_TIMEOUT_NS, so a lost peer shows up as an error of the call, never as a hung GPU.
#define RK_FLAG_TIMEOUT_NS 4000000000ull
__device__ __forceinline__ bool rk_wait_flag_sys(const uint32_t *flag, uint32_t seq)
{
    if ((int32_t) (rk_ld_acquire_sys(flag) - seq) >= 0) return true;
    const unsigned long long t_end = rk_globaltimer() + RK_FLAG_TIMEOUT_NS;
    for (;;) {
        if ((int32_t) (rk_ld_acquire_sys(flag) - seq) >= 0) return true;
        if (rk_globaltimer() > t_end) return false;
        __nanosleep(200);
    }
}

__device__ __forceinline__ uint32_t rk_smem_u32(const void *p)
{
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

__device__ __forceinline__ void rk_mbar_init(uint64_t *bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(rk_smem_u32(bar)), "r"(count) : "memory");
}

__device__ __forceinline__ void rk_fence_barrier_init()
{
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}

__device__ __forceinline__ void rk_fence_proxy_async()
{
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

__device__ __forceinline__ void rk_mbar_arrive(uint64_t *bar)
{
    asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.shared::cta.b64 st, [%0];\n\t}" ::"r"(rk_smem_u32(bar)) : "memory");
}

__device__ __forceinline__ void rk_mbar_arrive_expect_tx(uint64_t *bar, uint32_t bytes)
{
    asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1;\n\t}" ::"r"(rk_smem_u32(bar)),
                 "r"(bytes)
                 : "memory");
}

__device__ __forceinline__ void rk_mbar_wait(uint64_t *bar, uint32_t parity)
{
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "RK_WAIT_LOOP:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra RK_WAIT_DONE;\n\t"
        "bra RK_WAIT_LOOP;\n\t"
        "RK_WAIT_DONE:\n\t"
        "}" ::"r"(rk_smem_u32(bar)),
        "r"(parity)
        : "memory");
}

// global -> shared bulk copy completing on an mbarrier.  dst, src and bytes must be multiples of 16.
__device__ __forceinline__ void rk_bulk_g2s(void *smem_dst, const void *gmem_src, uint32_t bytes, uint64_t *bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     rk_smem_u32(smem_dst)),
                 "l"(gmem_src), "r"(bytes), "r"(rk_smem_u32(bar))
                 : "memory");
}

// L2 eviction policies: row data streams through once (evict_first); the small extent / mask
// tables are re-read by every segment (evict_last) and must survive the stream.
__device__ __forceinline__ uint64_t rk_policy_evict_first()
{
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}

__device__ __forceinline__ uint64_t rk_policy_evict_last()
{
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}

__device__ __forceinline__ void rk_bulk_g2s_hint(void *smem_dst, const void *gmem_src, uint32_t bytes, uint64_t *bar, uint64_t pol)
{
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(
            rk_smem_u32(smem_dst)),
        "l"(gmem_src), "r"(bytes), "r"(rk_smem_u32(bar)), "l"(pol)
        : "memory");
}

__device__ __forceinline__ uint32_t rk_ld_u32_hint(const uint32_t *p, uint64_t pol)
{
    uint32_t v;
    asm volatile("ld.global.nc.L2::cache_hint.u32 %0, [%1], %2;" : "=r"(v) : "l"(p), "l"(pol));
    return v;
}

__device__ __forceinline__ uint64_t rk_ld_u64_hint(const uint64_t *p, uint64_t pol)
{
    uint64_t v;
    asm volatile("ld.global.nc.L2::cache_hint.u64 %0, [%1], %2;" : "=l"(v) : "l"(p), "l"(pol));
    return v;
}

__device__ __forceinline__ void rk_named_bar_sync(int id, int nthreads)
{
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// Exact fp32 sum of one row against one query as the scan kernel forms it: 8 lanes per row, lane l8
// takes float4 columns l8, l8+8, ...; partial sums (x+y)+(z+w), then xor-shuffles 4, 2, 1.  Every
// kernel that reports a distance uses this order so that all engines agree bit for bit.
template <int METRIC>
__device__ __forceinline__ void rk_acc4(float4 &acc, const float4 &x, const float4 &q)
{
    if (METRIC == 0) {
        float d0 = x.x - q.x, d1 = x.y - q.y, d2 = x.z - q.z, d3 = x.w - q.w;
        acc.x = fmaf(d0, d0, acc.x); acc.y = fmaf(d1, d1, acc.y);
        acc.z = fmaf(d2, d2, acc.z); acc.w = fmaf(d3, d3, acc.w);
    } else {
        acc.x = fmaf(x.x, q.x, acc.x); acc.y = fmaf(x.y, q.y, acc.y);
        acc.z = fmaf(x.z, q.z, acc.z); acc.w = fmaf(x.w, q.w, acc.w);
    }
}

template <int METRIC>
__device__ __forceinline__ float rk_row_sum8(const float4 *__restrict__ row, const float4 *__restrict__ q, int dim4, int l8)
{
    float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int c = l8; c < dim4; c += 8) rk_acc4<METRIC>(a, row[c], q[c]);
    float s = (a.x + a.y) + (a.z + a.w);
    s += __shfl_xor_sync(0xffffffffu, s, 4);
    s += __shfl_xor_sync(0xffffffffu, s, 2);
    s += __shfl_xor_sync(0xffffffffu, s, 1);
    return s;
}

// whole warp per row: the scan kernel's order for wide rows (tiles of fewer than 64 rows)
template <int METRIC>
__device__ __forceinline__ float rk_row_sum32(const float4 *__restrict__ row, const float4 *__restrict__ q, int dim4, int lane)
{
    float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int c = lane; c < dim4; c += 32) rk_acc4<METRIC>(a, row[c], q[c]);
    float s = (a.x + a.y) + (a.z + a.w);
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    return s;
}

// Named barrier that also ORs a predicate over all participating threads; every thread receives
// the same result.
__device__ __forceinline__ bool rk_named_bar_red_or(int id, int nthreads, bool pred)
{
    uint32_t r;
    asm volatile(
        "{\n\t"
        ".reg .pred p, q;\n\t"
        "setp.ne.u32 q, %3, 0;\n\t"
        "barrier.cta.red.or.pred p, %1, %2, q;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t"
        "}"
        : "=r"(r)
        : "r"(id), "r"(nthreads), "r"((uint32_t) pred)
        : "memory");
    return r != 0;
}
#endif  // __CUDACC__
