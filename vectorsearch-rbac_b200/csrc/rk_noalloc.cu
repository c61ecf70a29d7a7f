// rk_noalloc.cu — rk_search_rows_noalloc: an exact unfiltered top-k scan that allocates NOTHING.
//
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid allocation).
//
// The ordinary path (rk_search_batch_device) keeps the scan kernel's work list and its per-segment candidate lists in
// cudaMalloc'ed workspace buffers that grow on demand and uploads a host-built plan through pinned staging.  Here the same
// two kernels (rk_scan_kernel over RK_SEG_ROWS segments, rk_select_kernel) run with that scratch in ONE file-scope
// `static __device__` array of fixed size, and the work list is written by a small kernel on the device, so the call makes
// no cudaMalloc / cudaFree / cudaHostAlloc and no host-to-device copy: three launches on the index stream.
//
// Limits (RK_NOALLOC_MAX_NQ, RK_NOALLOC_MAX_K in include/rbacknn.h): nq <= 16 queries, k <= 128; the array is
// about 4.9 MB and every GPU that loads the library holds it until the process exits.  All calls share the array and it
// keeps the previous call's contents: the plan kernel rewrites every word a call reads (segments, query offsets, the
// scheduler's counter, the list counts), so calls in stream order are independent — but two calls that OVERLAP on the GPU
// (different indexes / streams, or several host threads) race on it.  Serialise them.
#include "rk_host.h"
#include "rk_kernels.h"

#define RK_CUDA(call)                                                                                    \
    do {                                                                                                 \
        cudaError_t e__ = (call);                                                                        \
        if (e__ != cudaSuccess)                                                                          \
            return rk_fail(RK_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), __FILE__, __LINE__); \
    } while (0)

#define NA_MAX_NQ RK_NOALLOC_MAX_NQ
#define NA_MAX_K RK_NOALLOC_MAX_K
#define NA_SEGS_PER_Q 148                       // at most one segment per SM and query
#define NA_MAX_SEGS (NA_MAX_NQ * NA_SEGS_PER_Q)

struct NaScratch {
    uint64_t cand_tid[(size_t) NA_MAX_SEGS * NA_MAX_K];
    uint32_t cand_key[(size_t) NA_MAX_SEGS * NA_MAX_K];
    uint32_t cand_row[(size_t) NA_MAX_SEGS * NA_MAX_K];
    RkSegment segs[NA_MAX_SEGS];
    int32_t cand_count[NA_MAX_SEGS];
    uint32_t qseg[NA_MAX_NQ + 1];
    uint32_t ctl[8];                            // [0] the scan kernel's segment counter
};

// In host code this name is only a stand-in for the symbol, NOT a device address: host code goes through
// cudaGetSymbolAddress (na_scratch_ptr below); device code names it directly (na_plan_kernel).
static __device__ __align__(256) NaScratch g_na_scratch;

// Work list of one call: every query walks physical rows [0, nrows) in `per_q` equal segments.  Also resets everything the
// scan / select kernels read before they write it.
__global__ void na_plan_kernel(uint32_t nq, uint32_t per_q, uint64_t nrows)
{
    NaScratch &s = g_na_scratch;
    const uint32_t nseg = nq * per_q;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < nseg; i += gridDim.x * blockDim.x) {
        const uint32_t q = i / per_q, c = i - q * per_q;
        RkSegment g;
        g.query = q;
        g.kind = (uint16_t) RK_SEG_ROWS;
        g.flags = 0;
        g.ext_lo = 0; g.ext_hi = 0;
        g.begin = nrows * c / per_q;
        g.end = nrows * (c + 1) / per_q;
        s.segs[i] = g;
        s.cand_count[i] = 0;
    }
    if (blockIdx.x == 0) {
        for (uint32_t q = threadIdx.x; q <= nq; q += blockDim.x) s.qseg[q] = q * per_q;
        if (threadIdx.x < 8) s.ctl[threadIdx.x] = 0;
    }
}

// Device address of the array on `device` (which must be current): looked up once per device.
static int na_scratch_ptr(int device, NaScratch **out)
{
    static NaScratch *ptr[64];
    static std::mutex mu;
    if (device < 0 || device >= 64) return rk_fail(RK_ERR_INVALID, "rk_search_rows_noalloc: device ordinal %d", device);
    std::lock_guard<std::mutex> g(mu);
    if (!ptr[device]) {
        NaScratch *p = nullptr;
        RK_CUDA(cudaGetSymbolAddress((void **) &p, g_na_scratch));
        ptr[device] = p;
    }
    *out = ptr[device];
    return RK_OK;
}

extern "C" int rk_search_rows_noalloc(rk_index *ix, size_t nq, const float *dQ, int32_t k, uint64_t *d_out_tid, uint32_t *d_out_doc,
                                      double *d_out_dist, int32_t *d_out_count)
{
    if (!ix) return rk_fail(RK_ERR_INVALID, "NULL index");
    if (!ix->finalized) return rk_fail(RK_ERR_STATE, "rk_search_rows_noalloc: call rk_finalize first");
    if (nq == 0) return RK_OK;
    if (nq > (size_t) NA_MAX_NQ) return rk_fail(RK_ERR_INVALID, "rk_search_rows_noalloc: nq = %zu exceeds RK_NOALLOC_MAX_NQ = %d (the static scratch is sized for that; use rk_search_batch_device)", nq, NA_MAX_NQ);
    if (k < 1 || k > NA_MAX_K || k > ix->k_max) return rk_fail(RK_ERR_INVALID, "rk_search_rows_noalloc: k = %d exceeds RK_NOALLOC_MAX_K = %d or the index's k_max = %d", k, NA_MAX_K, ix->k_max);
    if (!dQ || !d_out_tid || !d_out_dist || !d_out_count) return rk_fail(RK_ERR_INVALID, "rk_search_rows_noalloc: NULL pointer");
    if (ix->dim != ix->dim_pad) return rk_fail(RK_ERR_UNSUPPORTED, "device queries need dim %% 4 == 0 (dim = %d)", ix->dim);
    RK_CUDA(cudaSetDevice(ix->device));
    NaScratch *sc = nullptr;
    int rc = na_scratch_ptr(ix->device, &sc);
    if (rc < 0) return rc;
    cudaStream_t st = ix->ws.stream;
    const uint64_t nrows = ix->nrows;
    if (nrows == 0) {
        RK_CUDA(cudaMemsetAsync(d_out_count, 0, nq * 4, st));
        return RK_OK;
    }
    // equal segments, at most one per SM and query, none shorter than a tile
    uint32_t per_q = (uint32_t) std::min<uint64_t>((uint64_t) std::min(ix->sm_count, NA_SEGS_PER_Q), std::max<uint64_t>(1, nrows / (uint64_t) ix->tile_rows));
    const uint32_t nseg = (uint32_t) nq * per_q;
    na_plan_kernel<<<(nseg + 255) / 256, 256, 0, st>>>((uint32_t) nq, per_q, nrows);
    RK_CUDA(cudaGetLastError());

    RkScanParams p;
    memset(&p, 0, sizeof p);
    p.store.x = ix->d_x.as<float>();
    p.store.tid = ix->d_tid.as<uint64_t>();
    p.store.rowdoc = ix->d_rowdoc.as<uint32_t>();
    p.store.rows = nrows;
    p.mask_words = ix->mask_words;
    p.q = dQ;
    p.segs = sc->segs;
    p.seg_order = nullptr;
    p.nseg_ptr = nullptr;
    p.nseg_val = nseg;
    p.seg_counter = sc->ctl;
    p.cand_key = sc->cand_key;
    p.cand_tid = sc->cand_tid;
    p.cand_row = sc->cand_row;
    p.cand_count = sc->cand_count;
    p.dim4 = ix->dim_pad / 4;
    p.row_bytes = ix->dim_pad * 4;
    p.tile_rows = ix->tile_rows;
    p.nstages = ix->nstages;
    p.k = k;
    p.metric = ix->metric;
    RK_CUDA(rk_launch_scan(p, ix->device, st, (int) std::min<uint32_t>((uint32_t) ix->sm_count, nseg)));

    RkSelectParams sp;
    memset(&sp, 0, sizeof sp);
    sp.cand_key = sc->cand_key;
    sp.cand_tid = sc->cand_tid;
    sp.cand_row = sc->cand_row;
    sp.cand_count = sc->cand_count;
    sp.qseg = sc->qseg;
    sp.list_stride = k;
    sp.k = k;
    sp.metric = ix->metric;
    sp.dedup = 0;                               // a physical row is visited once
    sp.rowdoc = ix->d_rowdoc.as<uint32_t>();
    sp.doc_ids = ix->d_doc_ids.as<uint32_t>();
    sp.out_tid = d_out_tid;
    sp.out_doc = d_out_doc;
    sp.out_dist = d_out_dist;
    sp.out_count = d_out_count;
    RK_CUDA(rk_launch_select(sp, (int) nq, st));
    ix->total_launches += 3;
    return RK_OK;
}
