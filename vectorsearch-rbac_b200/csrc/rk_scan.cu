// rk_scan.cu — the filtered distance scan: permission prefilter ⊕ L2 / inner-product scoring ⊕ top-k.
//
// Replaces, for one batch of queries, what the reference runs as PostgreSQL SeqScan + l2_distance()
// per heap row + top-N sort (pgvector/src/vector.c:549-578, 596-636, driven by
// controller/baseline/prefilter/prefilter_role.py:162-172) and as pgvector's GetScanItems list scan
// (pgvector/src/ivfscan.c:112-176).  The row-level-security predicate
// (controller/baseline/pg_row_security/row_level_security.py:55-65) is evaluated by the producer
// warp BEFORE a row's bytes are requested, so masked-out rows cost no HBM traffic.
//
// Shape: persistent kernel, one CTA per SM.  Warp 8 is the producer: it pops segments from the
// work list, resolves them to runs of physical rows (extent table, document mask or row mask) and
// streams those rows into a shared-memory ring with 1-D bulk async copies (TMA unit, UBLKCP)
// completing on mbarriers.  Warps 0-7 are consumers: 8 lanes per row, fp32 FMA, shuffle reduce,
// threshold test against the CTA's running k-th best, append to a shared candidate buffer that is
// compacted by a bitonic sort when it fills and at each segment end.
#include "rk_common.cuh"
#include "rk_select.cuh"
#include "rk_kernels.h"

#define RK_CONS_WARPS 8
#define RK_CONS_THREADS (RK_CONS_WARPS * 32)
#define RK_SCAN_THREADS (RK_CONS_THREADS + 32)

#define RK_FLAG_FIRST 1u
#define RK_FLAG_LAST 2u
#define RK_FLAG_END 4u

struct RkStageMeta {
    uint32_t nrows, seg, query, flags;
};

// Shared-memory carve-up, identical on host and device.
struct RkScanSmem {
    uint32_t tile_bytes, off_tiles, off_qslot, off_qcur, off_rowid, off_ckey, off_crow, off_ctid, off_meta, off_bar,
        off_misc, total;
};

__host__ __device__ inline RkScanSmem rk_scan_smem_layout(int row_bytes, int tile_rows, int nstages)
{
    RkScanSmem s;
    uint32_t o = 0;
    s.tile_bytes = (uint32_t) ((tile_rows * row_bytes + 127) & ~127);
    s.off_tiles = o; o += s.tile_bytes * nstages;
    s.off_qslot = o; o += (uint32_t) row_bytes * nstages;
    s.off_qcur = o; o += (uint32_t) row_bytes;
    o = (o + 127) & ~127u;
    s.off_ctid = o; o += RK_CAND_CAP * 8;
    s.off_ckey = o; o += RK_CAND_CAP * 4;
    s.off_crow = o; o += RK_CAND_CAP * 4;
    s.off_rowid = o; o += (uint32_t) (tile_rows * 4 * nstages);
    o = (o + 15) & ~15u;
    s.off_meta = o; o += 16u * nstages;
    s.off_bar = o; o += 16u * nstages;
    s.off_misc = o; o += 64;
    s.total = o;
    return s;
}

size_t rk_scan_smem_bytes(int row_bytes, int tile_rows, int nstages)
{
    return rk_scan_smem_layout(row_bytes, tile_rows, nstages).total;
}

// ---------------------------------------------------------------------------------------------

struct RkProducer {
    const RkScanParams &p;
    const RkScanSmem &L;
    unsigned char *smem;
    uint64_t *full, *empty;
    RkStageMeta *meta;
    int lane;
    int stage;
    uint32_t phase;
    uint32_t cur_rows, tx;
    uint32_t seg_id, query;
    bool first;

    __device__ __forceinline__ void open_tile()
    {
        rk_mbar_wait(&empty[stage], phase ^ 1u);
        cur_rows = 0;
        tx = 0;
    }

    __device__ __forceinline__ void close_tile(uint32_t flags)
    {
        if (first) {
            flags |= RK_FLAG_FIRST;
            if (lane == 0) {
                rk_bulk_g2s(smem + L.off_qslot + (size_t) stage * p.row_bytes,
                            reinterpret_cast<const unsigned char *>(p.q) + (size_t) query * p.row_bytes,
                            (uint32_t) p.row_bytes, &full[stage]);
            }
            tx += (uint32_t) p.row_bytes;
            first = false;
        }
        if (lane == 0) {
            RkStageMeta m;
            m.nrows = cur_rows; m.seg = seg_id; m.query = query; m.flags = flags;
            meta[stage] = m;
        }
        __syncwarp();
        if (lane == 0) {
            if (tx) rk_mbar_arrive_expect_tx(&full[stage], tx);
            else rk_mbar_arrive(&full[stage]);
        }
        if (++stage == p.nstages) { stage = 0; phase ^= 1u; }
    }

    // Stream `len` physically contiguous rows starting at `phys` into the ring.
    __device__ __forceinline__ void emit_run(uint64_t phys, uint64_t len)
    {
        while (len > 0) {
            uint32_t space = (uint32_t) p.tile_rows - cur_rows;
            uint32_t n = len < space ? (uint32_t) len : space;
            if (lane == 0) {
                rk_bulk_g2s(smem + L.off_tiles + (size_t) stage * L.tile_bytes + (size_t) cur_rows * p.row_bytes,
                            reinterpret_cast<const unsigned char *>(p.store.x) + phys * (uint64_t) p.row_bytes,
                            n * (uint32_t) p.row_bytes, &full[stage]);
            }
            uint32_t *rid = reinterpret_cast<uint32_t *>(smem + L.off_rowid) + (size_t) stage * p.tile_rows + cur_rows;
            for (uint32_t i = lane; i < n; i += 32) rid[i] = (uint32_t) (phys + i);
            tx += n * (uint32_t) p.row_bytes;
            cur_rows += n;
            phys += n;
            len -= n;
            if (cur_rows == (uint32_t) p.tile_rows) {
                close_tile(0);
                open_tile();
            }
        }
    }

    __device__ __forceinline__ bool mask_hit(const uint64_t *m, const uint64_t *um) const
    {
        uint64_t acc = 0;
        for (int w = 0; w < p.mask_words; w++) acc |= m[w] & um[w];
        return acc != 0;
    }

    __device__ void run_segment(const RkSegment &s)
    {
        query = s.query;
        first = true;
        open_tile();
        if (s.kind == RK_SEG_ROWS) {
            emit_run(s.begin, s.end - s.begin);
        } else if (s.kind == RK_SEG_EXTENTS) {
            // Partition = table of per-document extents.  32 extents are resolved per step, one per
            // lane (coalesced loads of the extent table and, when the segment is masked, of the
            // extent's 128-bit role mask); allowed extents are then streamed in order.
            const uint32_t e0 = p.part_ext_off[s.src];
            const uint32_t e1 = p.part_ext_off[s.src + 1] - 1;  // sentinel index
            const bool masked = (s.flags & RK_SEGF_MASKED) != 0;
            const uint64_t *um = p.qmask + (size_t) query * p.mask_words;
            uint32_t lo = e0, hi = e1;                       // last extent with vstart <= begin
            while (hi - lo > 1) {
                const uint32_t mid = (lo + hi) >> 1;
                if ((uint64_t) p.ext_vstart[mid] <= s.begin) lo = mid; else hi = mid;
            }
            uint64_t run_start = 0, run_len = 0;
            for (uint32_t eb = lo; eb < e1; eb += 32) {
                const uint32_t e = eb + lane;
                bool ok = false, past = true;
                uint64_t v0 = 0, v1 = 0, prow = 0;
                if (e < e1) {
                    const uint64_t vs = p.ext_vstart[e], ve = p.ext_vstart[e + 1];
                    past = vs >= s.end;
                    v0 = vs > s.begin ? vs : s.begin;
                    v1 = ve < s.end ? ve : s.end;
                    if (v1 > v0) {
                        ok = masked ? mask_hit(p.ext_mask + (size_t) e * p.mask_words, um) : true;
                        prow = (uint64_t) p.ext_row[e] + (v0 - vs);
                    }
                }
                unsigned bal = __ballot_sync(0xffffffffu, ok);
                while (bal) {
                    const int b = __ffs(bal) - 1;
                    bal &= bal - 1;
                    const uint64_t r0 = __shfl_sync(0xffffffffu, prow, b);
                    const uint64_t n = __shfl_sync(0xffffffffu, v1 - v0, b);
                    if (run_len && run_start + run_len == r0) {
                        run_len += n;                        // physically adjacent: one longer copy
                    } else {
                        if (run_len) emit_run(run_start, run_len);
                        run_start = r0;
                        run_len = n;
                    }
                }
                if (__any_sync(0xffffffffu, past)) break;
            }
            if (run_len) emit_run(run_start, run_len);
        } else {  // RK_SEG_ROWMASK
            const uint64_t *um = p.qmask + (size_t) query * p.mask_words;
            uint64_t run_start = 0, run_len = 0;
            for (uint64_t r0 = s.begin; r0 < s.end; r0 += 32) {
                uint64_t r = r0 + lane;
                bool ok = false;
                if (r < s.end) ok = mask_hit(p.doc_mask + (size_t) p.store.rowdoc[r] * p.mask_words, um);
                unsigned bal = __ballot_sync(0xffffffffu, ok);
                while (bal) {
                    int b = __ffs(bal) - 1;
                    unsigned rest = bal >> b;              // run of consecutive set bits starting at b
                    int n = (~rest == 0u) ? 32 - b : __ffs(~rest) - 1;
                    if (run_len && run_start + run_len == r0 + b) {
                        run_len += n;
                    } else {
                        if (run_len) emit_run(run_start, run_len);
                        run_start = r0 + b;
                        run_len = n;
                    }
                    bal = (b + n >= 32) ? 0u : (bal & ~((1u << (b + n)) - 1u));
                }
            }
            if (run_len) emit_run(run_start, run_len);
        }
        close_tile(RK_FLAG_LAST);
    }
};

// ---------------------------------------------------------------------------------------------

template <int METRIC>
__device__ __forceinline__ void rk_acc(float4 &acc, const float4 &x, const float4 &q)
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
__global__ void __launch_bounds__(RK_SCAN_THREADS, 1) rk_scan_kernel(const RkScanParams p)
{
    extern __shared__ __align__(128) unsigned char smem[];
    const RkScanSmem L = rk_scan_smem_layout(p.row_bytes, p.tile_rows, p.nstages);
    uint64_t *full = reinterpret_cast<uint64_t *>(smem + L.off_bar);
    uint64_t *empty = full + p.nstages;
    RkStageMeta *meta = reinterpret_cast<RkStageMeta *>(smem + L.off_meta);
    volatile uint32_t *s_count = reinterpret_cast<volatile uint32_t *>(smem + L.off_misc);
    volatile uint32_t *s_tau = s_count + 1;

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    if (threadIdx.x == 0) {
        for (int s = 0; s < p.nstages; s++) {
            rk_mbar_init(&full[s], 1);
            rk_mbar_init(&empty[s], RK_CONS_WARPS);
        }
        *s_count = 0;
        *s_tau = RK_KEY_SENTINEL;
        rk_fence_barrier_init();
    }
    __syncthreads();

    const uint32_t nseg = *p.nseg_ptr;

    if (warp == RK_CONS_WARPS) {
        // ===================== producer warp =====================
        RkProducer pr{p, L, smem, full, empty, meta, lane, 0, 0u, 0u, 0u, 0u, 0u, true};
        for (;;) {
            uint32_t sid = 0;
            if (lane == 0) sid = atomicAdd(p.seg_counter, 1u);
            sid = __shfl_sync(0xffffffffu, sid, 0);
            if (sid >= nseg) break;
            pr.seg_id = sid;
            RkSegment s = p.segs[sid];
            pr.run_segment(s);
        }
        pr.open_tile();
        pr.first = false;
        pr.close_tile(RK_FLAG_END);
        return;
    }

    // ===================== consumer warps =====================
    RkCandBuf cb;
    cb.key = reinterpret_cast<uint32_t *>(smem + L.off_ckey);
    cb.row = reinterpret_cast<uint32_t *>(smem + L.off_crow);
    cb.tid = reinterpret_cast<uint64_t *>(smem + L.off_ctid);
    const float4 *qcur = reinterpret_cast<const float4 *>(smem + L.off_qcur);
    const int t = threadIdx.x;
    const int g = lane >> 3, l8 = lane & 7;
    auto csync = [] { rk_named_bar_sync(1, RK_CONS_THREADS); };

    int stage = 0;
    uint32_t phase = 0;
    for (;;) {
        rk_mbar_wait(&full[stage], phase);
        const RkStageMeta m = meta[stage];
        if (m.flags & RK_FLAG_END) break;
        if (m.flags & RK_FLAG_FIRST) {
            const float4 *qs = reinterpret_cast<const float4 *>(smem + L.off_qslot + (size_t) stage * p.row_bytes);
            float4 *qd = reinterpret_cast<float4 *>(smem + L.off_qcur);
            for (int i = t; i < p.dim4; i += RK_CONS_THREADS) qd[i] = qs[i];
            csync();
        }
        const float4 *tile = reinterpret_cast<const float4 *>(smem + L.off_tiles + (size_t) stage * L.tile_bytes);
        const uint32_t *rid = reinterpret_cast<const uint32_t *>(smem + L.off_rowid) + (size_t) stage * p.tile_rows;
        const int nrows = (int) m.nrows;
        for (int r0 = warp * 4; r0 < nrows; r0 += RK_CONS_WARPS * 8) {
            const int ra = r0 + g, rb = ra + RK_CONS_WARPS * 4;
            const bool va = ra < nrows, vb = rb < nrows;
            const float4 *pa = tile + (size_t) (va ? ra : 0) * p.dim4;
            const float4 *pb = tile + (size_t) (vb ? rb : 0) * p.dim4;
            float4 aa = make_float4(0.f, 0.f, 0.f, 0.f), ab = aa;
            if (!p.null_query) {
#pragma unroll 4
                for (int c = l8; c < p.dim4; c += 8) {
                    const float4 q = qcur[c];
                    const float4 xa = pa[c];
                    const float4 xb = pb[c];
                    rk_acc<METRIC>(aa, xa, q);
                    rk_acc<METRIC>(ab, xb, q);
                }
            }
            float sa = (aa.x + aa.y) + (aa.z + aa.w);
            float sb = (ab.x + ab.y) + (ab.z + ab.w);
            sa += __shfl_xor_sync(0xffffffffu, sa, 4); sb += __shfl_xor_sync(0xffffffffu, sb, 4);
            sa += __shfl_xor_sync(0xffffffffu, sa, 2); sb += __shfl_xor_sync(0xffffffffu, sb, 2);
            sa += __shfl_xor_sync(0xffffffffu, sa, 1); sb += __shfl_xor_sync(0xffffffffu, sb, 1);
            if (l8 == 0) {
                const uint32_t tau = *s_tau;
                if (va) {
                    const uint32_t key = rk_float_to_key(METRIC == 0 ? sa : -sa);
                    if (key <= tau) {
                        const uint32_t slot = atomicAdd((uint32_t *) s_count, 1u);
                        const uint32_t row = rid[ra];
                        cb.key[slot] = key; cb.row[slot] = row; cb.tid[slot] = p.store.tid[row];
                    }
                }
                if (vb) {
                    const uint32_t key = rk_float_to_key(METRIC == 0 ? sb : -sb);
                    if (key <= tau) {
                        const uint32_t slot = atomicAdd((uint32_t *) s_count, 1u);
                        const uint32_t row = rid[rb];
                        cb.key[slot] = key; cb.row[slot] = row; cb.tid[slot] = p.store.tid[row];
                    }
                }
            }
        }
        __syncwarp();
        if (lane == 0) rk_mbar_arrive(&empty[stage]);
        csync();
        const int n = (int) *s_count;
        const bool last = (m.flags & RK_FLAG_LAST) != 0;
        if (last || n > RK_CAND_CAP - p.tile_rows) {
            rk_bitonic_sort(cb, n, t, RK_CONS_THREADS, csync);
            const int keep = n < p.k ? n : p.k;
            if (last) {
                const size_t base = (size_t) m.seg * p.k;
                for (int i = t; i < keep; i += RK_CONS_THREADS) {
                    p.cand_key[base + i] = cb.key[i];
                    p.cand_tid[base + i] = cb.tid[i];
                    p.cand_row[base + i] = cb.row[i];
                }
                if (t == 0) {
                    p.cand_count[m.seg] = keep;
                    *s_count = 0;
                    *s_tau = RK_KEY_SENTINEL;
                }
            } else if (t == 0) {
                *s_count = (uint32_t) keep;
                *s_tau = (n >= p.k) ? cb.key[p.k - 1] : RK_KEY_SENTINEL;
            }
            csync();
        }
        if (++stage == p.nstages) { stage = 0; phase ^= 1u; }
    }
}

// ---------------------------------------------------------------------------------------------
// host launcher

static int g_sm_count[64];

int rk_device_sm_count(int device)
{
    if (device < 0 || device >= 64) return 148;
    if (g_sm_count[device] == 0) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, device) != cudaSuccess || n <= 0) n = 148;
        g_sm_count[device] = n;
    }
    return g_sm_count[device];
}

void rk_scan_geometry(int dim_pad, int *tile_rows, int *nstages, size_t *smem)
{
    const int row_bytes = dim_pad * 4;
    int tr = 32768 / row_bytes;
    if (tr < 1) tr = 1;
    if (tr > 256) tr = 256;
    int ns = 8;
    const size_t budget = 220 * 1024;
    while (ns > 2 && rk_scan_smem_bytes(row_bytes, tr, ns) > budget) ns--;
    *tile_rows = tr;
    *nstages = ns;
    *smem = rk_scan_smem_bytes(row_bytes, tr, ns);
}

cudaError_t rk_launch_scan(const RkScanParams &p, int device, cudaStream_t stream, int grid_override)
{
    const size_t smem = rk_scan_smem_bytes(p.row_bytes, p.tile_rows, p.nstages);
    static bool attr_set[64][2];
    auto kern = p.metric == 0 ? rk_scan_kernel<0> : rk_scan_kernel<1>;
    if (!attr_set[device & 63][p.metric & 1]) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
        if (e != cudaSuccess) return e;
        attr_set[device & 63][p.metric & 1] = true;
    }
    int grid = grid_override > 0 ? grid_override : rk_device_sm_count(device);
    kern<<<grid, RK_SCAN_THREADS, smem, stream>>>(p);
    return cudaGetLastError();
}
