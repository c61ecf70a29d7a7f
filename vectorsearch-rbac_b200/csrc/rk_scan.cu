// rk_scan.cu — the filtered distance scan: permission prefilter ⊕ L2 / inner-product scoring ⊕ top-k.
//
// Replaces, for one batch of queries, what the reference runs as PostgreSQL SeqScan + l2_distance()
// per heap row + top-N sort (pgvector/src/vector.c:549-578, 596-636, driven by
// controller/baseline/prefilter/prefilter_role.py:162-172) and as pgvector's GetScanItems list scan
// (pgvector/src/ivfscan.c:112-176).  The row-level-security predicate
// (controller/baseline/pg_row_security/row_level_security.py:55-65) is evaluated by the producer
// warp BEFORE a row's bytes are requested, so masked-out rows cost no HBM traffic.
//
// Shape: persistent kernel, one CTA per SM, dynamic segment scheduler.
//   warp 8 (producer)  pops segments (claimed two ahead so the counter round trip and the segment
//       descriptor load are off the critical path), resolves them to runs of physical rows (extent
//       table + role-mask test, 128 extents per step) and streams those rows into a shared-memory
//       ring with 1-D bulk async copies (TMA unit, SASS UBLKCP) completing on mbarriers.
//   warps 0-7 (consumers)  run decoupled: each waits on a stage's full barrier, scores its own rows
//       of the tile (8 lanes per row, fp32 FMA, shuffle reduce), tests them against the CTA's running
//       k-th best and appends survivors (key, row) to a shared candidate buffer with one
//       warp-aggregated atomic, then releases the stage.  They meet only at a checkpoint every
//       RK_CKPT_ROWS rows and at segment ends, where one reducing barrier decides uniformly whether
//       the buffer is compacted.  Compaction is an exact 4-pass radix select of the k-th smallest key
//       (not a sort); tuple ids are gathered and the survivors sorted only when a segment ends.
#include "rk_common.cuh"
#include "rk_select.cuh"
#include "rk_kernels.h"

#include <cstring>

#define RK_CONS_WARPS 8
#define RK_CONS_THREADS (RK_CONS_WARPS * 32)
#define RK_SCAN_THREADS (RK_CONS_THREADS + 32)
#define RK_CKPT_ROWS 512     // consumer warps meet at least every this many rows of a segment
#define RK_EPL 4             // extents resolved per producer lane per step

#ifndef RK_SCAN_XPRE
#define RK_SCAN_XPRE 0
#endif
#ifndef RK_SCAN_SEL8
#define RK_SCAN_SEL8 0
#endif

#define RK_FLAG_FIRST 1u
#define RK_FLAG_LAST 2u
#define RK_FLAG_END 4u
#define RK_FLAG_CONT 8u     // with LAST: the CTA's next segment continues the same query's single partition — keep the buffer

struct RkStageMeta {
    uint32_t nrows, seg, query, flags;
};

// Shared-memory carve-up, identical on host and device.
struct RkScanSmem {
    uint32_t tile_bytes, off_tiles, off_qslot, off_qcur, off_rowid, off_ckey, off_crow, off_ctid, off_meta, off_bar,
        off_hist, off_misc, total;
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
    s.off_hist = o; o += 256 * 4;
    s.off_misc = o; o += 128;
    s.total = o;
    return s;
}

size_t rk_scan_smem_bytes(int row_bytes, int tile_rows, int nstages)
{
    return rk_scan_smem_layout(row_bytes, tile_rows, nstages).total;
}

// ---------------------------------------------------------------------------------------------
// producer

template <bool FUSED>
struct RkProducer {
    // cross-segment prefetch of the next segment's first extents: on in the fused single-query kernel, whose segments are few
    // and whose start-up latency is what matters; the batch kernel has no registers left for it (RK_SCAN_XPRE to try)
    static constexpr bool XPRE = FUSED || RK_SCAN_XPRE;
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
    bool cont_next;     // the segment after this one may share its candidate buffer (set by the claim loop)
    bool first;
    uint64_t pol_stream, pol_keep;

    __device__ __forceinline__ void open_tile()
    {
        rk_mbar_wait(&empty[stage], phase ^ 1u);
        cur_rows = 0;
        tx = 0;
    }

    __device__ __forceinline__ void close_tile(uint32_t flags)
    {
        if (first && !FUSED) {
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
                rk_bulk_g2s_hint(smem + L.off_tiles + (size_t) stage * L.tile_bytes + (size_t) cur_rows * p.row_bytes,
                                 reinterpret_cast<const unsigned char *>(p.store.x) + phys * (uint64_t) p.row_bytes,
                                 n * (uint32_t) p.row_bytes, &full[stage], pol_stream);
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
        for (int w = 0; w < p.mask_words; w++) acc |= __ldg(m + w) & um[w];
        return acc != 0;
    }

    // row / virtual offsets of one resolve step (RK_EPL extents per lane)
    struct ExtStep { uint32_t vs[RK_EPL], ve[RK_EPL], er[RK_EPL]; };
    int dbg_n;
    __device__ __forceinline__ void stamp()
    {
        if (FUSED && p.fz_dbg && blockIdx.x == 0 && lane == 0 && dbg_n < 256) p.fz_dbg[(size_t) (gridDim.x + 1) * 8 + 256 + dbg_n++] = rk_globaltimer();
    }
    ExtStep pre;            // first step of the next segment, requested while the previous segment's last rows were issued
    bool pre_valid;

    __device__ __forceinline__ void ext_load(const RkSegment &s, uint32_t eb, ExtStep &x) const
    {
        // unconditional, independent loads (indices clamped): one memory round trip per step, not one per extent
#pragma unroll
        for (int j = 0; j < RK_EPL; j++) {
            const uint32_t e = eb + j * 32 + lane;
            const uint32_t ec = e < s.ext_hi ? e : s.ext_hi - 1;
            x.vs[j] = rk_ld_u32_hint(p.ext_vstart + ec, pol_keep);
            x.ve[j] = rk_ld_u32_hint(p.ext_vstart + ec + 1, pol_keep);
            x.er[j] = rk_ld_u32_hint(p.ext_row + ec, pol_keep);
        }
    }

    // `next`: the segment this CTA runs after `s` (nullptr = none)
    __device__ __forceinline__ void run_segment(const RkSegment &s, const RkSegment *next)
    {
        query = s.query;
        first = true;
        stamp();
        open_tile();
        stamp();
        if (s.kind == RK_SEG_ROWS) {
            emit_run(s.begin, s.end - s.begin);
        } else if (s.kind == RK_SEG_IVFQ) {
            // probed lists [begin, end) of this query; in every list the rows of one document class are contiguous, so
            // what the user may see is one extent per visible class: no per-row test, masked-out rows cost nothing
            // The class offsets of the NEXT list are fetched (probe entry -> offsets: two round trips) while the extents
            // of the current one stream, so the producer does not stall between lists.
            const uint64_t c0 = p.ivf_qcls[2 * (size_t) s.query], c1 = p.ivf_qcls[2 * (size_t) s.query + 1];
            const int nc = p.ivf_ncls;
            const uint32_t *pl = p.ivf_probe_list + (size_t) s.query * p.ivf_probes;
            bool vis[4];
#pragma unroll
            for (int b4 = 0; b4 < 4; b4++) {
                const int c = b4 * 32 + lane;
                vis[b4] = c < nc && (((c < 64 ? c0 >> (c & 63) : c1 >> (c & 63)) & 1ull) != 0);
            }
            uint32_t lo[4], hi[4], nlo[4], nhi[4];
            uint32_t run_begin = 0, run_end = 0;
            auto fetch = [&](uint32_t pi, uint32_t (&a_)[4], uint32_t (&b_)[4]) {
                const uint32_t *co = p.ivf_cls_off + (size_t) pl[pi] * (nc + 1);
#pragma unroll
                for (int b4 = 0; b4 < 4; b4++) {
                    a_[b4] = 0; b_[b4] = 0;
                    if (vis[b4]) { a_[b4] = co[b4 * 32 + lane]; b_[b4] = co[b4 * 32 + lane + 1]; }
                }
            };
            if ((uint32_t) s.begin < (uint32_t) s.end) fetch((uint32_t) s.begin, nlo, nhi);
            for (uint32_t pi = (uint32_t) s.begin; pi < (uint32_t) s.end; pi++) {
#pragma unroll
                for (int b4 = 0; b4 < 4; b4++) { lo[b4] = nlo[b4]; hi[b4] = nhi[b4]; }
                if (pi + 1 < (uint32_t) s.end) fetch(pi + 1, nlo, nhi);
#pragma unroll
                for (int b4 = 0; b4 < 4; b4++) {
                    uint32_t m = __ballot_sync(0xffffffffu, hi[b4] > lo[b4]);
                    while (m) {
                        const int j = __ffs((int) m) - 1;
                        m &= m - 1;
                        const uint32_t a = __shfl_sync(0xffffffffu, lo[b4], j), b = __shfl_sync(0xffffffffu, hi[b4], j);
                        if (a == run_end) { run_end = b; continue; }        // neighbouring visible classes: one copy
                        if (run_end > run_begin) emit_run(run_begin, run_end - run_begin);
                        run_begin = a; run_end = b;
                    }
                }
            }
            if (run_end > run_begin) emit_run(run_begin, run_end - run_begin);
        } else if (s.kind == RK_SEG_EXTENTS) {
            // Partition = table of per-document extents; the host resolved which extents touch
            // [begin, end).  RK_EPL * 32 extents are resolved per step: every lane fetches the row / virtual offsets of its
            // RK_EPL extents (all loads independent) and tests their role masks; allowed extents are streamed in order and
            // physically adjacent ones are fused into one longer copy.  The offsets of the NEXT step — the first step of the
            // next segment when this one ends — are requested before the current step's rows are issued, so that the table
            // round trip is hidden behind the copies instead of draining the ring at every step and segment boundary.
            const bool masked = (s.flags & RK_SEGF_MASKED) != 0;
            const uint64_t *um = p.qmask + (size_t) query * p.mask_words;
            uint64_t run_start = 0, run_len = 0;
            ExtStep x;
            if (XPRE && pre_valid) x = pre; else if (s.ext_lo < s.ext_hi) ext_load(s, s.ext_lo, x);
            if (XPRE) pre_valid = false;
            for (uint32_t eb = s.ext_lo; eb < s.ext_hi; eb += 32 * RK_EPL) {
                ExtStep nx;
                if (eb + 32 * RK_EPL < s.ext_hi) {
                    ext_load(s, eb + 32 * RK_EPL, nx);
                } else if (XPRE && next && next->kind == RK_SEG_EXTENTS && next->ext_lo < next->ext_hi) {
                    ext_load(*next, next->ext_lo, nx);
                    pre_valid = true;
                }
                uint32_t ec[RK_EPL];
                bool ok[RK_EPL];
#pragma unroll
                for (int j = 0; j < RK_EPL; j++) {
                    const uint32_t e = eb + j * 32 + lane;
                    ok[j] = e < s.ext_hi;
                    ec[j] = ok[j] ? e : s.ext_hi - 1;
                }
                if (masked) {
                    uint64_t hit[RK_EPL];
#pragma unroll
                    for (int j = 0; j < RK_EPL; j++) hit[j] = 0;
                    for (int w = 0; w < p.mask_words; w++) {
                        const uint64_t u = __ldg(um + w);
#pragma unroll
                        for (int j = 0; j < RK_EPL; j++)
                            hit[j] |= rk_ld_u64_hint(p.ext_mask + (size_t) ec[j] * p.mask_words + w, pol_keep) & u;
                    }
#pragma unroll
                    for (int j = 0; j < RK_EPL; j++) ok[j] = ok[j] && hit[j] != 0;
                }
                uint64_t prow[RK_EPL];
                uint32_t plen[RK_EPL];
#pragma unroll
                for (int j = 0; j < RK_EPL; j++) {
                    const uint64_t vs = x.vs[j], ve = x.ve[j];
                    const uint64_t v0 = vs > s.begin ? vs : s.begin;
                    const uint64_t v1 = ve < s.end ? ve : s.end;
                    const bool take = ok[j] && v1 > v0;
                    prow[j] = (uint64_t) x.er[j] + (v0 - vs);
                    plen[j] = take ? (uint32_t) (v1 - v0) : 0u;
                }
#pragma unroll
                for (int j = 0; j < RK_EPL; j++) {
                    unsigned bal = __ballot_sync(0xffffffffu, plen[j] != 0);
                    while (bal) {
                        const int b = __ffs(bal) - 1;
                        bal &= bal - 1;
                        const uint64_t r0 = __shfl_sync(0xffffffffu, prow[j], b);
                        const uint64_t n = __shfl_sync(0xffffffffu, plen[j], b);
                        if (run_len && run_start + run_len == r0) {
                            run_len += n;                        // physically adjacent: one longer copy
                        } else {
                            if (run_len) emit_run(run_start, run_len);
                            run_start = r0;
                            run_len = n;
                        }
                    }
                }
                // keep the pipeline fed: do not hold more than a few tiles of rows back while resolving
                if (run_len >= 4u * (uint32_t) p.tile_rows) {
                    const uint64_t n = run_len - (run_len % (uint32_t) p.tile_rows);
                    emit_run(run_start, n);
                    run_start += n;
                    run_len -= n;
                }
                if (eb == s.ext_lo) stamp();
                x = nx;
            }
            if (XPRE && pre_valid) pre = x;
            if (run_len) emit_run(run_start, run_len);
        } else {  // RK_SEG_ROWMASK
            // Rows of a store that is NOT sorted by document (IVF lists): the role mask is reached
            // through rowdoc[].  RK_EPL * 32 rows per step, two rounds of independent loads.
            const uint64_t *um = p.qmask + (size_t) query * p.mask_words;
            uint64_t run_start = 0, run_len = 0;
            for (uint64_t r0 = s.begin; r0 < s.end; r0 += 32 * RK_EPL) {
                uint32_t dd[RK_EPL];
                bool ok[RK_EPL];
#pragma unroll
                for (int j = 0; j < RK_EPL; j++) {
                    const uint64_t r = r0 + (uint64_t) j * 32 + lane;
                    ok[j] = r < s.end;
                    dd[j] = __ldg(p.store.rowdoc + (ok[j] ? r : s.end - 1));
                }
                uint64_t hit[RK_EPL];
#pragma unroll
                for (int j = 0; j < RK_EPL; j++) hit[j] = 0;
                for (int w = 0; w < p.mask_words; w++) {
                    const uint64_t u = __ldg(um + w);
#pragma unroll
                    for (int j = 0; j < RK_EPL; j++)
                        hit[j] |= rk_ld_u64_hint(p.doc_mask + (size_t) dd[j] * p.mask_words + w, pol_keep) & u;
                }
#pragma unroll
                for (int j = 0; j < RK_EPL; j++) {
                    unsigned bal = __ballot_sync(0xffffffffu, ok[j] && hit[j] != 0);
                    const uint64_t rb = r0 + (uint64_t) j * 32;
                    while (bal) {
                        int b = __ffs(bal) - 1;
                        unsigned rest = bal >> b;              // run of consecutive set bits starting at b
                        int n = (~rest == 0u) ? 32 - b : __ffs(~rest) - 1;
                        if (run_len && run_start + run_len == rb + b) {
                            run_len += n;
                        } else {
                            if (run_len) emit_run(run_start, run_len);
                            run_start = rb + b;
                            run_len = n;
                        }
                        bal = (b + n >= 32) ? 0u : (bal & ~((1u << (b + n)) - 1u));
                    }
                }
            }
            if (run_len) emit_run(run_start, run_len);
        }
        close_tile(RK_FLAG_LAST | (cont_next ? RK_FLAG_CONT : 0u));
    }
};

// ---------------------------------------------------------------------------------------------
// consumer helpers

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

// Inclusive scan of one value per consumer thread (RK_CONS_THREADS threads); wsum = 8 words of smem.
// Returns the inclusive prefix; *total receives the sum over all threads.  Two consumer barriers.
template <typename Sync>
__device__ __forceinline__ uint32_t rk_block_scan(uint32_t v, int t, volatile uint32_t *wsum, uint32_t *total, Sync sync)
{
    uint32_t inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t u = __shfl_up_sync(0xffffffffu, inc, o);
        if ((t & 31) >= o) inc += u;
    }
    sync();                                   // previous users of wsum are done
    if ((t & 31) == 31) wsum[t >> 5] = inc;
    sync();
    uint32_t off = 0, tot = 0;
#pragma unroll
    for (int w = 0; w < RK_CONS_WARPS; w++) {
        const uint32_t x = wsum[w];
        if (w < (t >> 5)) off += x;
        tot += x;
    }
    *total = tot;
    return off + inc;
}

// Exact k-th smallest (1-based) of key[0..n), n >= k, by a 4-pass 8-bit radix select over the
// order-preserving uint keys.  All RK_CONS_THREADS consumer threads participate.
template <typename Sync>
__device__ __forceinline__ uint32_t rk_radix_kth(const uint32_t *key, int n, int k, uint32_t *hist, volatile uint32_t *wsum,
                                                 volatile uint32_t *sel, int t, Sync sync)
{
    uint32_t prefix = 0, mask = 0, kk = (uint32_t) k;
#pragma unroll 1
    for (int shift = 24; shift >= 0; shift -= 8) {
        hist[t] = 0;
        sync();
        for (int i = t; i < n; i += RK_CONS_THREADS) {
            const uint32_t kv = key[i];
            if ((kv & mask) == prefix) atomicAdd(&hist[(kv >> shift) & 255u], 1u);
        }
        sync();
        const uint32_t c = hist[t];
        uint32_t tot;
        const uint32_t incl = rk_block_scan(c, t, wsum, &tot, sync);
        const uint32_t excl = incl - c;
        if (excl < kk && kk <= incl) { sel[0] = (uint32_t) t; sel[1] = kk - excl; }
        sync();
        prefix |= sel[0] << shift;
        mask |= 0xFFu << shift;
        kk = sel[1];
    }
    return prefix;
}

// Smallest and largest of key[0..n) over the consumer threads (n >= 1).  wsum: 16 words of scratch.  Two barriers.
template <typename Sync>
__device__ __forceinline__ void rk_block_minmax(const uint32_t *key, int n, volatile uint32_t *mm, int t, uint32_t *lo_out, uint32_t *hi_out, Sync sync)
{
    uint32_t lo = RK_KEY_SENTINEL, hi = 0;
    for (int i = t; i < n; i += RK_CONS_THREADS) { const uint32_t kv = key[i]; lo = min(lo, kv); hi = max(hi, kv); }
    lo = __reduce_min_sync(0xffffffffu, lo);
    hi = __reduce_max_sync(0xffffffffu, hi);
    sync();
    if ((t & 31) == 0) { mm[t >> 5] = lo; mm[8 + (t >> 5)] = hi; }
    sync();
    lo = mm[0]; hi = mm[8];
#pragma unroll
    for (int w = 1; w < RK_CONS_WARPS; w++) { lo = min(lo, mm[w]); hi = max(hi, mm[8 + w]); }
    *lo_out = lo; *hi_out = hi;
}

// k-th smallest (1-based) of key[0..n), n >= k, for the once-per-launch epilogue: digits of up to 12 bits over
// `key - min`, starting at the top bit the spread of the keys occupies — one or two passes for a candidate set instead of
// four 8-bit passes over the full word.  hist: 4096 words (the idle tile ring).
// loose = true stops as soon as the bin holding the k-th key has at most 64 entries and returns that bin's upper end: at
// least k keys are <= the result and at most 63 of them exceed the exact k-th — enough for a threshold whose survivors are
// ordered exactly afterwards.
template <typename Sync>
__device__ __forceinline__ uint32_t rk_radix_kth_wide(const uint32_t *key, int n, int k, uint32_t *hist, volatile uint32_t *wsum,
                                                      volatile uint32_t *sel, volatile uint32_t *mm, int t, Sync sync, bool loose = false)
{
    uint32_t lo, hi;
    rk_block_minmax(key, n, mm, t, &lo, &hi, sync);
    const uint32_t span = hi - lo;
    if (span == 0) return lo;
    const int bits = 32 - __clz(span);
    uint32_t prefix = 0, kk = (uint32_t) k;
    int prev = 32;                                   // bits [prev, 32) of key - min are fixed to `prefix` (32 = none yet)
    int shift = bits > 12 ? bits - 12 : 0;
#pragma unroll 1
    for (;;) {
        const int width = (prev == 32 ? bits : prev) - shift;          // <= 12
        for (int i = t; i < 4096; i += RK_CONS_THREADS) hist[i] = 0;
        sync();
        for (int i = t; i < n; i += RK_CONS_THREADS) {
            const uint32_t kv = key[i] - lo;
            if (prev == 32 || (kv >> prev) == (prefix >> prev)) atomicAdd(&hist[(kv >> shift) & ((1u << width) - 1u)], 1u);
        }
        sync();
        uint32_t h[16], c = 0;
#pragma unroll
        for (int b_ = 0; b_ < 16; b_++) { h[b_] = hist[t * 16 + b_]; c += h[b_]; }
        uint32_t tot;
        const uint32_t incl = rk_block_scan(c, t, wsum, &tot, sync);
        uint32_t run = incl - c;
        if (run < kk && kk <= incl) {
#pragma unroll
            for (int b_ = 0; b_ < 16; b_++) {
                if (run < kk && kk <= run + h[b_]) { sel[0] = (uint32_t) (t * 16 + b_); sel[1] = kk - run; mm[0] = h[b_]; }
                run += h[b_];
            }
        }
        sync();
        prefix |= sel[0] << shift;
        kk = sel[1];
        const uint32_t in_bin = mm[0];
        sync();
        if (shift == 0) break;
        if (loose && in_bin <= 64u) {
            const uint32_t up = prefix | ((1u << shift) - 1u);      // the bin's largest value
            return up > span ? hi : up + lo;
        }
        prev = shift;
        shift = shift > 12 ? shift - 12 : 0;
    }
    return prefix + lo;
}

// Keep the entries of (key,row)[0..n) with key <= tau, packed to the front (order not preserved
// across threads' blocks but deterministic).  Returns the number kept.
template <typename Sync>
__device__ __forceinline__ int rk_compact_le(RkCandBuf cb, int n, uint32_t tau, int t, volatile uint32_t *wsum, Sync sync)
{
    constexpr int PER = RK_CAND_CAP / RK_CONS_THREADS;   // 8
    uint32_t mk[PER], mr[PER];
    int mc = 0;
    const int b0 = t * PER;
#pragma unroll
    for (int i = 0; i < PER; i++) {
        const int idx = b0 + i;
        if (idx < n) {
            const uint32_t kv = cb.key[idx];
            if (kv <= tau) { mk[mc] = kv; mr[mc] = cb.row[idx]; mc++; }
        }
    }
    uint32_t tot;
    const uint32_t incl = rk_block_scan((uint32_t) mc, t, wsum, &tot, sync);   // its barriers also order the reads above
    int pos = (int) (incl - (uint32_t) mc);
#pragma unroll
    for (int i = 0; i < PER; i++)
        if (i < mc) { cb.key[pos + i] = mk[i]; cb.row[pos + i] = mr[i]; }
    sync();
    return (int) tot;
}

// ---------------------------------------------------------------------------------------------

// The query of a fused single-query launch rides in the kernel parameters (constant bank): it arrives with the launch
// command itself, so a call needs no host->device copy at all.  Rows of up to RK_INLINE_Q floats.
struct RkInlineQuery {
    float v[RK_INLINE_Q];
};

// FUSED / LB are compile-time so that the batch kernel (neither) carries none of the single-query epilogue's or the
// streaming lower bound's code and registers: at 288 threads the register file allows 168 per thread, and the batch
// kernel's inner loop sits right at that limit.
template <int METRIC, bool WIDE, bool FUSED, bool LB>
__device__ __forceinline__ void rk_scan_body(const RkScanParams &p, const float *inline_q)
{
    extern __shared__ __align__(128) unsigned char smem[];
    const RkScanSmem L = rk_scan_smem_layout(p.row_bytes, p.tile_rows, p.nstages);
    uint64_t *full = reinterpret_cast<uint64_t *>(smem + L.off_bar);
    uint64_t *empty = full + p.nstages;
    RkStageMeta *meta = reinterpret_cast<RkStageMeta *>(smem + L.off_meta);
    volatile uint32_t *s_count = reinterpret_cast<volatile uint32_t *>(smem + L.off_misc);
    volatile uint32_t *s_tau = s_count + 1;
    volatile uint32_t *s_sel = s_count + 2;     // [2] radix-select hand-off
    volatile uint32_t *s_wsum = s_count + 8;    // [8] block-scan scratch
    uint32_t *s_hist = reinterpret_cast<uint32_t *>(smem + L.off_hist);

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

    const uint32_t nseg = p.nseg_ptr ? *p.nseg_ptr : p.nseg_val;

    if (warp == RK_CONS_WARPS) {
        // ===================== producer warp =====================
        RkProducer<FUSED> pr{p, L, smem, full, empty, meta, lane, 0, 0u, 0u, 0u, 0u, 0u, false, true, rk_policy_evict_first(),
                      rk_policy_evict_last()};
        auto seg_at = [&](uint32_t pos) -> uint32_t { return p.seg_order ? p.seg_order[pos] : pos; };
        pr.pre_valid = false;
        pr.dbg_n = 0;
        // Claims run three segments ahead: while segment i streams, the descriptor of i + 1 is already in registers (its
        // query decides whether i may hand its candidate buffer on, and its first extents are requested before i ends),
        // the descriptor of i + 2 is in flight and the position of i + 3 is being claimed — no load or atomic round trip
        // sits between two segments.  The first three positions of every CTA are static (b, b + grid, b + 2 grid), so no
        // atomic sits before the first row is requested either; the shared counter hands out positions from 3 * grid on.
        const uint32_t sid_dyn0 = 3u * gridDim.x;
        uint32_t sid0 = blockIdx.x, sid1 = sid0 + gridDim.x, sid2 = sid1 + gridDim.x;
        RkSegment cur, nxt;
        uint32_t cur_id = 0, nxt_id = 0;
        if (sid0 < nseg) { cur_id = seg_at(sid0); cur = p.segs[cur_id]; }
        if (sid1 < nseg) { nxt_id = seg_at(sid1); nxt = p.segs[nxt_id]; }
        while (sid0 < nseg) {
            pr.stamp();
            RkSegment nn;
            uint32_t nn_id = 0, claim = 0;
            const bool more = sid2 < nseg;
            if (more) {
                nn_id = seg_at(sid2);
                nn = p.segs[nn_id];
                if (lane == 0) claim = atomicAdd(p.seg_counter, 1u);       // (the result is not touched before the segment is out)
            }
            pr.seg_id = cur_id;
            pr.cont_next = sid1 < nseg && nxt.query == cur.query && (cur.flags & nxt.flags & RK_SEGF_MERGE_OK) != 0;
            pr.run_segment(cur, sid1 < nseg ? &nxt : nullptr);
            const uint32_t sid3 = more ? sid_dyn0 + __shfl_sync(0xffffffffu, claim, 0) : sid2;
            sid0 = sid1; sid1 = sid2; sid2 = sid3;
            cur = nxt; cur_id = nxt_id;
            nxt = nn; nxt_id = nn_id;
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
    // checkpoint every `ckpt` tiles: at most ckpt * tile_rows appends can land between two
    // checkpoints, so the buffer cannot overflow if it is compacted above `trigger`
    int ckpt = RK_CKPT_ROWS / p.tile_rows;
    if (ckpt < 1) ckpt = 1;
    if (ckpt > 16) ckpt = 16;
    const uint32_t trigger = (uint32_t) (RK_CAND_CAP - ckpt * p.tile_rows);
    // after a radix-select compaction more than this many survivors means "too many ties at the
    // k-th key": fall back to the exact (key, tid) sort that keeps precisely k
    const int tie_limit = p.k + ((int) trigger - p.k) / 2;

    unsigned long long *dbg = FUSED && p.fz_dbg ? p.fz_dbg + (size_t) blockIdx.x * 8 : nullptr;
    if (dbg && t == 0) dbg[0] = rk_globaltimer();                   // CTA start
    bool dbg_first = dbg != nullptr;
    int dbg_tiles = 0;
    if (FUSED) {
        // one query for the whole launch: the consumers fetch it themselves — from the kernel parameters when it rides
        // there, else from HBM — while the producer resolves its first extents
        float4 *qd = reinterpret_cast<float4 *>(smem + L.off_qcur);
        if (inline_q) {
            const float4 *q4 = reinterpret_cast<const float4 *>(inline_q);
            for (int i = t; i < p.dim4; i += RK_CONS_THREADS) qd[i] = q4[i];
        } else {
            const float4 *q4 = reinterpret_cast<const float4 *>(p.q);
            for (int i = t; i < p.dim4; i += RK_CONS_THREADS) qd[i] = __ldcv(&q4[i]);
        }
        csync();
    }
    volatile uint32_t *s_mm = s_count + 16;                                  // [16] min / max scratch
    int stage = 0;
    uint32_t phase = 0;
    int since_ckpt = 0;
    // a long single-query scan sets its threshold early: until the first compaction every row is a candidate
    const uint32_t trigger0 = FUSED && p.fz_seed > 0 && (uint32_t) p.fz_seed < trigger ? (uint32_t) p.fz_seed : trigger;
    bool seeded = false;
    for (;;) {
        rk_mbar_wait(&full[stage], phase);
        const RkStageMeta m = meta[stage];
        if (dbg_first && t == 0) { dbg[1] = rk_globaltimer(); dbg_first = false; }       // first tile landed
        if (dbg && blockIdx.x == 0 && t == 0 && dbg_tiles < 256) p.fz_dbg[(size_t) (gridDim.x + 1) * 8 + dbg_tiles++] = rk_globaltimer();
        if (m.flags & RK_FLAG_END) break;
        if (m.flags & RK_FLAG_FIRST) {
            // Every warp copies the segment's query for itself (identical values from all warps).
            // All warps have left the previous segment through its final barriers, so nobody still
            // reads the old contents.
            const float4 *qs = reinterpret_cast<const float4 *>(smem + L.off_qslot + (size_t) stage * p.row_bytes);
            float4 *qd = reinterpret_cast<float4 *>(smem + L.off_qcur);
            for (int i = lane; i < p.dim4; i += 32) qd[i] = qs[i];
            __syncwarp();
        }
        const float4 *tile = reinterpret_cast<const float4 *>(smem + L.off_tiles + (size_t) stage * L.tile_bytes);
        const uint32_t *rid = reinterpret_cast<const uint32_t *>(smem + L.off_rowid) + (size_t) stage * p.tile_rows;
        const int nrows = (int) m.nrows;
        // Rows per warp step: narrow rows (tiles of 64 rows or more) take 8 lanes each, 2 x 4 rows per warp; wide rows
        // (few rows per tile) take the whole warp, 2 rows per warp, so that every warp has work in every tile.
        constexpr bool lanes8 = !WIDE;                    // WIDE <=> p.tile_rows < 64 (rk_launch_scan)
        const int r_first = lanes8 ? warp * 4 : warp, r_step = lanes8 ? RK_CONS_WARPS * 8 : RK_CONS_WARPS * 2;
        const int r_pair = lanes8 ? RK_CONS_WARPS * 4 : RK_CONS_WARPS;
        for (int r0 = r_first; r0 < nrows; r0 += r_step) {
            const int ra = lanes8 ? r0 + g : r0, rb = ra + r_pair;
            const bool va = ra < nrows, vb = rb < nrows;
            const float4 *pa = tile + (size_t) (va ? ra : 0) * p.dim4;
            const float4 *pb = tile + (size_t) (vb ? rb : 0) * p.dim4;
            float4 aa = make_float4(0.f, 0.f, 0.f, 0.f), ab = aa;
            float sa, sb;
            if (lanes8) {
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
                sa = (aa.x + aa.y) + (aa.z + aa.w);
                sb = (ab.x + ab.y) + (ab.z + ab.w);
            } else {
                if (!p.null_query) {
#pragma unroll 2
                    for (int c = lane; c < p.dim4; c += 32) {
                        const float4 q = qcur[c];
                        const float4 xa = pa[c];
                        const float4 xb = pb[c];
                        rk_acc<METRIC>(aa, xa, q);
                        rk_acc<METRIC>(ab, xb, q);
                    }
                }
                sa = (aa.x + aa.y) + (aa.z + aa.w);
                sb = (ab.x + ab.y) + (ab.z + ab.w);
                sa += __shfl_xor_sync(0xffffffffu, sa, 16); sb += __shfl_xor_sync(0xffffffffu, sb, 16);
                sa += __shfl_xor_sync(0xffffffffu, sa, 8); sb += __shfl_xor_sync(0xffffffffu, sb, 8);
            }
            sa += __shfl_xor_sync(0xffffffffu, sa, 4); sb += __shfl_xor_sync(0xffffffffu, sb, 4);
            sa += __shfl_xor_sync(0xffffffffu, sa, 2); sb += __shfl_xor_sync(0xffffffffu, sb, 2);
            sa += __shfl_xor_sync(0xffffffffu, sa, 1); sb += __shfl_xor_sync(0xffffffffu, sb, 1);
            const bool lead = lanes8 ? l8 == 0 : lane == 0;
            // survivors of this warp step: one aggregated atomic for the whole warp
            const uint32_t tau = *s_tau;
            const uint32_t ka = rk_float_to_key(METRIC == 0 ? sa : -sa);
            const uint32_t kb = rk_float_to_key(METRIC == 0 ? sb : -sb);
            bool ca = lead && va && ka <= tau;
            bool cb2 = lead && vb && kb <= tau;
            if (LB) {
                // a streaming scan resumes after (lb_key, lb_tid): everything up to that pair was already returned
                if (ca && (ka < p.lb_key || (ka == p.lb_key && p.store.tid[rid[ra]] <= p.lb_tid))) ca = false;
                if (cb2 && (kb < p.lb_key || (kb == p.lb_key && p.store.tid[rid[rb]] <= p.lb_tid))) cb2 = false;
            }
            const unsigned ma = __ballot_sync(0xffffffffu, ca), mb = __ballot_sync(0xffffffffu, cb2);
            if (ma | mb) {
                const int na = __popc(ma);
                uint32_t base = 0;
                if (lane == 0) base = atomicAdd((uint32_t *) s_count, (uint32_t) (na + __popc(mb)));
                base = __shfl_sync(0xffffffffu, base, 0);
                const unsigned below = (1u << lane) - 1u;
                if (ca) { const uint32_t slot = base + __popc(ma & below); cb.key[slot] = ka; cb.row[slot] = rid[ra]; }
                if (cb2) { const uint32_t slot = base + na + __popc(mb & below); cb.key[slot] = kb; cb.row[slot] = rid[rb]; }
            }
        }
        __syncwarp();
        if (lane == 0) rk_mbar_arrive(&empty[stage]);
        const bool last = (m.flags & RK_FLAG_LAST) != 0;
        // a segment whose successor on this CTA continues the same single-partition query ends without a final
        // select: the buffer and the threshold carry over, and this segment's own list stays empty
        // (fused single-query mode: every segment is the same query's, the buffer lives until the work list is drained)
        const bool cont = last && ((m.flags & RK_FLAG_CONT) != 0 || FUSED);
        const bool fin = last && !cont;
        if (last || ++since_ckpt == ckpt) {
            since_ckpt = 0;
            // one reducing barrier: every warp gets the same verdict whatever it read itself
            const bool full_buf = rk_named_bar_red_or(1, RK_CONS_THREADS, *s_count > (seeded ? trigger : trigger0));
            if (fin || full_buf) {
                int n = (int) *s_count;              // stable: nobody appends before the next barrier
                uint32_t tau = RK_KEY_SENTINEL;
                if (n > p.k) {
                    // 12-bit digits over the spread of the keys; the 4096-bin histogram lies over the tuple-id slots, which
                    // are only filled for the exact sort below.  A compaction in mid-segment needs a threshold only (loose:
                    // up to 63 keys above the k-th stay), the end of a segment the exact k-th key.
                    if (!FUSED && RK_SCAN_SEL8) tau = rk_radix_kth(cb.key, n, p.k, s_hist, s_wsum, s_sel, t, csync);
                    else tau = rk_radix_kth_wide(cb.key, n, p.k, reinterpret_cast<uint32_t *>(cb.tid), s_wsum, s_sel, s_mm, t, csync, !fin);
                    n = rk_compact_le(cb, n, tau, t, s_wsum, csync);
                }
                if (fin || n > tie_limit) {
                    // exact order: (key, tuple id) ascending over the survivors
                    for (int i = t; i < n; i += RK_CONS_THREADS) cb.tid[i] = p.store.tid[cb.row[i]];
                    rk_bitonic_sort(cb, n, t, RK_CONS_THREADS, csync);
                    if (n > p.k) n = p.k;
                }
                seeded = true;
                if (fin) {
                    const size_t base = (size_t) m.seg * p.k;
                    for (int i = t; i < n; i += RK_CONS_THREADS) {
                        p.cand_key[base + i] = cb.key[i];
                        p.cand_tid[base + i] = cb.tid[i];
                        p.cand_row[base + i] = cb.row[i];
                    }
                    if (t == 0) {
                        p.cand_count[m.seg] = n;
                        *s_count = 0;
                        *s_tau = RK_KEY_SENTINEL;
                    }
                } else if (t == 0) {
                    *s_count = (uint32_t) n;
                    // (after the exact sort n == k entries remain and the k-th key is still the threshold)
                    *s_tau = tau;
                }
                csync();
            }
            if (cont && t == 0 && !FUSED) p.cand_count[m.seg] = 0;
        }
        if (++stage == p.nstages) { stage = 0; phase ^= 1u; }
    }
    if (!FUSED) return;

    // ===================== fused single-query epilogue =====================
    // (1) Cooperative threshold.  Every CTA publishes the r-th smallest key it found (r ~ 3k / grid): B = the largest such
    // key over the CTAs that have reported and S = how many rows those CTAs vouch for (r each).  As soon as S >= k, at least k
    // rows of the query have a key <= B, so nothing above B can be in the answer and each CTA hands over only its
    // candidates <= B — a few dozen entries in all instead of k per CTA, which is what makes the last CTA's merge short.
    // Writers update B before S and the reader reads S before B, so any S >= k it sees is covered by the B it reads.  The
    // wait is bounded (other kernels may hold SMs this grid is waiting for): on timeout, or when every CTA has reported
    // without reaching k rows, the CTA falls back to handing over its own k best.
    csync();                                  // every warp has left the loop: no append is in flight
    if (dbg && t == 0) dbg[2] = rk_globaltimer();                   // scan done
    uint32_t *whist = reinterpret_cast<uint32_t *>(smem + L.off_tiles);     // 4096-bin histogram: the tile ring is idle from here on
    int n = (int) *s_count;
    if (gridDim.x > 1) {
        const int rc_ = n < p.fz_r ? n : p.fz_r;
        uint32_t vc = 0;
        if (rc_ == 1) { uint32_t hi_; rk_block_minmax(cb.key, n, s_mm, t, &vc, &hi_, csync); }       // r = 1: the CTA's best key
        else if (rc_ > 1) vc = rk_radix_kth_wide(cb.key, n, rc_, whist, s_wsum, s_sel, s_mm, t, csync);
        if (t == 0) {
            // fz_ctl[4] = rows vouched for (low 16 bits) | CTAs reported (high 16 bits): one atomic publishes both and its
            // return value is already a snapshot
            if (rc_ > 0) {
                atomicMax(&p.fz_ctl[6], vc);
                __threadfence();
            }
            uint32_t sr = atomicAdd(&p.fz_ctl[4], (uint32_t) rc_ | 0x10000u) + ((uint32_t) rc_ | 0x10000u);
            const unsigned long long t_end = rk_globaltimer() + 20000ull;      // 20 us
            while ((sr & 0xFFFFu) < (uint32_t) p.k && (sr >> 16) < gridDim.x && rk_globaltimer() <= t_end) sr = rk_ld_volatile_u32(&p.fz_ctl[4]);
            __threadfence();
            s_sel[0] = (sr & 0xFFFFu) >= (uint32_t) p.k ? rk_ld_volatile_u32(&p.fz_ctl[6]) : RK_KEY_SENTINEL;
        }
        csync();
        const uint32_t bound = s_sel[0];
        csync();
        if (bound != RK_KEY_SENTINEL && n > 0) n = rk_compact_le(cb, n, bound, t, s_wsum, csync);
    }
    // (2) at most its k best (threshold only, no order) go to the flat global list
    if (n > p.k) {
        const uint32_t tau = rk_radix_kth_wide(cb.key, n, p.k, whist, s_wsum, s_sel, s_mm, t, csync, true);
        n = rk_compact_le(cb, n, tau, t, s_wsum, csync);
        if (n > p.fz_cap) {
            // more ties at the k-th key than a CTA may hand over: exact (key, tuple id) order keeps precisely k
            for (int i = t; i < n; i += RK_CONS_THREADS) cb.tid[i] = p.store.tid[cb.row[i]];
            rk_bitonic_sort(cb, n, t, RK_CONS_THREADS, csync);
            n = p.k;
        }
    }
    if (dbg && t == 0) dbg[3] = rk_globaltimer();                   // local select done
    if (t == 0) s_sel[0] = n ? atomicAdd(&p.fz_ctl[0], (uint32_t) n) : 0u;
    csync();
    {
        const uint32_t base = s_sel[0];
        for (int i = t; i < n; i += RK_CONS_THREADS) {
            __stcg(&p.fz_key[base + i], cb.key[i]);
            __stcg(&p.fz_row[base + i], cb.row[i]);
        }
    }
    __threadfence();
    csync();
    if (t == 0) s_sel[1] = atomicAdd(&p.fz_ctl[1], 1u);
    csync();
    if (dbg && t == 0) dbg[4] = rk_globaltimer();                   // handed over
    if (s_sel[1] != gridDim.x - 1) return;
    dbg = p.fz_dbg ? p.fz_dbg + (size_t) gridDim.x * 8 : nullptr;
    if (dbg && t == 0) { dbg[0] = rk_globaltimer(); dbg[7] = blockIdx.x; }

    // (3) the last CTA to finish selects the answer from everything handed over, orders it by (key, tuple id) and writes it.
    __threadfence();
    const int N = (int) __ldcg(&p.fz_ctl[0]);
    int m;
    bool overflow = false;
    if (N <= RK_CAND_CAP) {
        // the usual case (the cooperative threshold left a few hundred entries): straight into the candidate buffer, the
        // CTA's own threshold select over it
        for (int i = t; i < N; i += RK_CONS_THREADS) { cb.key[i] = __ldcg(&p.fz_key[i]); cb.row[i] = __ldcg(&p.fz_row[i]); }
        csync();
        if (dbg && t == 0) { dbg[1] = rk_globaltimer(); dbg[6] = (unsigned long long) N; }
        m = N;
        if (m > p.k) {
            // (a loose threshold will do: the survivors are ordered exactly below)
            const uint32_t tau = rk_radix_kth_wide(cb.key, m, p.k, whist, s_wsum, s_sel, s_mm, t, csync, true);
            m = rk_compact_le(cb, m, tau, t, s_wsum, csync);
        }
        if (dbg && t == 0) dbg[2] = rk_globaltimer();
    } else {
    uint32_t kth = RK_KEY_SENTINEL;
    // The keys are staged ONCE in shared memory — the tile ring is idle by now (every tile was consumed before END) and holds
    // up to 48 K keys — so the radix passes run from shared memory in compact loops (no instruction-cache-cold unrolled code
    // on this once-per-launch path).  The select runs over `key - min` and only over the bytes the spread of the keys occupies
    // (the k best of neighbouring CTAs are close: two or three passes instead of four).  Hand-overs beyond the ring's capacity
    // (large k with many ties) re-read the keys from L2 on every pass.
    uint32_t *skeys = reinterpret_cast<uint32_t *>(smem + L.off_tiles);
    const bool staged = (size_t) N * 4 <= (size_t) L.tile_bytes * (size_t) p.nstages;
    if (staged) {
#pragma unroll 8
        for (int i = t; i < N; i += RK_CONS_THREADS) skeys[i] = __ldcg(&p.fz_key[i]);
        csync();
    }
    if (dbg && t == 0) dbg[1] = rk_globaltimer();                   // keys staged
    if (N > p.k) {
        uint32_t kmin = 0, span = 0xFFFFFFFFu;
        if (staged) {
            uint32_t lo = RK_KEY_SENTINEL, hi = 0;
            for (int i = t; i < N; i += RK_CONS_THREADS) { const uint32_t kv = skeys[i]; lo = min(lo, kv); hi = max(hi, kv); }
            lo = __reduce_min_sync(0xffffffffu, lo);
            hi = __reduce_max_sync(0xffffffffu, hi);
            if (lane == 0) { s_hist[warp] = lo; s_hist[8 + warp] = hi; }
            csync();
            lo = s_hist[0]; hi = s_hist[8];
#pragma unroll
            for (int w = 1; w < RK_CONS_WARPS; w++) { lo = min(lo, s_hist[w]); hi = max(hi, s_hist[8 + w]); }
            csync();
            kmin = lo;
            span = hi - lo;
        }
        const int npass = span == 0 ? 0 : (39 - __clz(span)) / 8;          // bytes the spread occupies
        uint32_t prefix = 0, mask = 0, kk = (uint32_t) p.k;
#pragma unroll 1
        for (int shift = 8 * (npass - 1); shift >= 0; shift -= 8) {
            s_hist[t] = 0;
            csync();
            if (staged) {
#pragma unroll 4
                for (int i = t; i < N; i += RK_CONS_THREADS) {
                    const uint32_t kv = skeys[i] - kmin;
                    if ((kv & mask) == prefix) atomicAdd(&s_hist[(kv >> shift) & 255u], 1u);
                }
            } else {
                for (int i0 = t; i0 < N; i0 += RK_CONS_THREADS * 8) {
                    uint32_t kv[8];
#pragma unroll
                    for (int j = 0; j < 8; j++) { const int i = i0 + j * RK_CONS_THREADS; kv[j] = i < N ? __ldcg(&p.fz_key[i]) : 0u; }
#pragma unroll
                    for (int j = 0; j < 8; j++)
                        if (i0 + j * RK_CONS_THREADS < N && (kv[j] & mask) == prefix) atomicAdd(&s_hist[(kv[j] >> shift) & 255u], 1u);
                }
            }
            csync();
            const uint32_t c = s_hist[t];
            uint32_t tot;
            const uint32_t incl = rk_block_scan(c, t, s_wsum, &tot, csync);
            const uint32_t excl = incl - c;
            if (excl < kk && kk <= incl) { s_sel[0] = (uint32_t) t; s_sel[1] = kk - excl; }
            csync();
            prefix |= s_sel[0] << shift;
            mask |= 0xFFu << shift;
            kk = s_sel[1];
        }
        kth = prefix + kmin;
    }
    if (t == 0) *s_count = 0;
    csync();
    if (dbg && t == 0) { dbg[2] = rk_globaltimer(); dbg[6] = (unsigned long long) N; }      // k-th key known
    if (staged) {
        for (int i = t; i < N; i += RK_CONS_THREADS) {
            const uint32_t kv = skeys[i];
            if (kv <= kth) {
                const uint32_t slot = atomicAdd((uint32_t *) s_count, 1u);
                if (slot < RK_CAND_CAP) { cb.key[slot] = kv; cb.row[slot] = __ldcg(&p.fz_row[i]); }
            }
        }
    } else {
        for (int i0 = t; i0 < N; i0 += RK_CONS_THREADS * 8) {
            uint32_t kv[8];
#pragma unroll
            for (int j = 0; j < 8; j++) { const int i = i0 + j * RK_CONS_THREADS; kv[j] = i < N ? __ldcg(&p.fz_key[i]) : RK_KEY_SENTINEL; }
#pragma unroll
            for (int j = 0; j < 8; j++) {
                const int i = i0 + j * RK_CONS_THREADS;
                if (i < N && kv[j] <= kth) {
                    const uint32_t slot = atomicAdd((uint32_t *) s_count, 1u);
                    if (slot < RK_CAND_CAP) { cb.key[slot] = kv[j]; cb.row[slot] = __ldcg(&p.fz_row[i]); }
                }
            }
        }
    }
    csync();
    m = (int) *s_count;
    overflow = m > RK_CAND_CAP;                // more than RK_CAND_CAP entries tie at the k-th key: the host reruns unfused
    if (overflow) m = 0;
    }
    // tuple ids and (when asked for) document ids of the survivors: fetched together, before the ordering needs them
    uint32_t *sdoc = whist;                    // [RK_CAND_CAP] document ids, in the idle tile ring
    for (int i = t; i < m; i += RK_CONS_THREADS) {
        const uint32_t r = cb.row[i];
        const uint64_t tid = p.store.tid[r];
        if (p.fz_out_doc) sdoc[i] = p.fz_doc_ids[p.store.rowdoc[r]];
        cb.tid[i] = tid;
    }
    csync();
    if (dbg && t == 0) dbg[3] = rk_globaltimer();                   // gathered, tuple ids fetched
    const int cnt = overflow ? 0 : (m < p.k ? m : p.k);
    auto emit = [&](int r, uint32_t key, uint32_t doc, uint64_t tid) {
        const float ord = rk_key_to_float(key);
        if (p.fz_out_tid) p.fz_out_tid[r] = tid;
        if (p.fz_out_key) p.fz_out_key[r] = ord;
        if (p.fz_out_doc) p.fz_out_doc[r] = doc;
        if (p.fz_out_dist) p.fz_out_dist[r] = p.null_query ? 0.0 : (METRIC == 0 ? sqrt((double) ord) : (double) ord);   // vector.c:577, :635
    };
    if (m <= 512) {
        // rank sort: an entry's position is the number of entries ordered before it — no barriers, m^2 / 256 compares per
        // thread
        for (int i = t; i < m; i += RK_CONS_THREADS) {
            const uint32_t ki = cb.key[i];
            const uint64_t ti = cb.tid[i];
            int r = 0;
            // (identical (key, tuple id) pairs — the ground-truth merge keeps duplicates — rank by position)
            for (int j = 0; j < m; j++) {
                const uint32_t kj = cb.key[j];
                const uint64_t tj = cb.tid[j];
                r += (rk_cand_less(kj, tj, ki, ti) || (kj == ki && tj == ti && j < i)) ? 1 : 0;
            }
            if (r < cnt) emit(r, ki, p.fz_out_doc ? sdoc[i] : 0u, ti);
        }
    } else {
        // (the sort moves key / row / tuple id; the document ids are looked up again afterwards)
        rk_bitonic_sort(cb, m, t, RK_CONS_THREADS, csync);
        for (int i = t; i < cnt; i += RK_CONS_THREADS)
            emit(i, cb.key[i], p.fz_out_doc ? p.fz_doc_ids[p.store.rowdoc[cb.row[i]]] : 0u, cb.tid[i]);
    }
    for (int i = cnt + t; i < p.k; i += RK_CONS_THREADS) {
        if (p.fz_out_tid) p.fz_out_tid[i] = 0xFFFFFFFFFFFFFFFFull;
        if (p.fz_out_key) p.fz_out_key[i] = __int_as_float(0x7f800000);
        if (p.fz_out_doc) p.fz_out_doc[i] = 0xFFFFFFFFu;
        if (p.fz_out_dist) p.fz_out_dist[i] = __longlong_as_double(0x7ff0000000000000ll);
    }
    if (t == 0) {
        *p.fz_out_count = overflow ? -1 : cnt;
        // ready for the next launch: nothing is uploaded per call but the query
        p.fz_ctl[0] = 0;
        p.fz_ctl[1] = 0;
        p.fz_ctl[4] = 0;
        p.fz_ctl[6] = 0;
        *p.seg_counter = 0;
    }
    if (dbg && t == 0) dbg[4] = rk_globaltimer();                   // outputs written
    if (p.fz_flag) {
        __threadfence_system();                 // every thread's outputs are visible to the host ...
        csync();
        if (t == 0) *reinterpret_cast<volatile uint32_t *>(p.fz_flag) = p.fz_seq;      // ... before it sees the flag
    }
}

template <int METRIC, bool WIDE, bool FUSED, bool LB>
__global__ void __launch_bounds__(RK_SCAN_THREADS, 1) rk_scan_kernel(const RkScanParams p)
{
    rk_scan_body<METRIC, WIDE, FUSED, LB>(p, nullptr);
}

template <int METRIC, bool WIDE, bool LB>
__global__ void __launch_bounds__(RK_SCAN_THREADS, 1) rk_scan_single_kernel(const __grid_constant__ RkScanParams p,
                                                                            const __grid_constant__ RkInlineQuery iq)
{
    rk_scan_body<METRIC, WIDE, true, LB>(p, iq.v);
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
    if (tr > 256) tr = 256;
    if (tr < 64) {
        // wide rows: a warp scores two rows at a time, so tiles of up to 48 KB holding a multiple of 8 rows where possible
        tr = 49152 / row_bytes;
        if (tr >= 8) tr &= ~7;
    }
    if (tr < 1) tr = 1;
    int ns = 8;
    const size_t budget = 220 * 1024;
    while (ns > 2 && rk_scan_smem_bytes(row_bytes, tr, ns) > budget) ns--;
    *tile_rows = tr;
    *nstages = ns;
    *smem = rk_scan_smem_bytes(row_bytes, tr, ns);
}

template <typename K>
static cudaError_t rk_scan_attr(K kern, bool *done)
{
    if (*done) return cudaSuccess;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e == cudaSuccess) *done = true;
    return e;
}

// variant index: metric | wide << 1 | lower bound << 2 | fused << 3
template <int V>
static cudaError_t rk_launch_scan_variant(const RkScanParams &p, int grid, size_t smem, cudaStream_t stream, bool *attr_done,
                                          const float *inline_q_host)
{
    constexpr int M = V & 1;
    constexpr bool W = (V & 2) != 0, LB = (V & 4) != 0, F = (V & 8) != 0;
    if (F && inline_q_host) {
        // fused single-query launch with the query in the parameters
        cudaError_t e = rk_scan_attr(rk_scan_single_kernel<M, W, LB>, attr_done + 16);
        if (e != cudaSuccess) return e;
        RkInlineQuery iq;
        memcpy(iq.v, inline_q_host, (size_t) p.row_bytes);
        rk_scan_single_kernel<M, W, LB><<<grid, RK_SCAN_THREADS, smem, stream>>>(p, iq);
        return cudaGetLastError();
    }
    cudaError_t e = rk_scan_attr(rk_scan_kernel<M, W, F, LB>, attr_done);
    if (e != cudaSuccess) return e;
    rk_scan_kernel<M, W, F, LB><<<grid, RK_SCAN_THREADS, smem, stream>>>(p);
    return cudaGetLastError();
}

cudaError_t rk_launch_scan(const RkScanParams &p, int device, cudaStream_t stream, int grid_override, const float *inline_q_host)
{
    const size_t smem = rk_scan_smem_bytes(p.row_bytes, p.tile_rows, p.nstages);
    static bool attr_set[64][32][2];
    const bool wide = p.tile_rows < 64;
    const int grid = grid_override > 0 ? grid_override : rk_device_sm_count(device);
    const int v = (p.metric & 1) | (wide ? 2 : 0) | (p.lb_on ? 4 : 0) | (p.fused ? 8 : 0);
    bool *done = &attr_set[device & 63][v][0];
    // (attr_done + 16 in the variant = the single-kernel's flag: lay the two flags of a variant 16 bools apart)
    static bool attr_single[64][32];
    bool flags[17] = {done[0]};
    flags[16] = attr_single[device & 63][v];
    cudaError_t e;
    switch (v) {
#define RK_SCAN_CASE(V) case V: e = rk_launch_scan_variant<V>(p, grid, smem, stream, flags, inline_q_host); break;
        RK_SCAN_CASE(0) RK_SCAN_CASE(1) RK_SCAN_CASE(2) RK_SCAN_CASE(3) RK_SCAN_CASE(4) RK_SCAN_CASE(5) RK_SCAN_CASE(6) RK_SCAN_CASE(7)
        RK_SCAN_CASE(8) RK_SCAN_CASE(9) RK_SCAN_CASE(10) RK_SCAN_CASE(11) RK_SCAN_CASE(12) RK_SCAN_CASE(13) RK_SCAN_CASE(14) RK_SCAN_CASE(15)
#undef RK_SCAN_CASE
        default: return cudaErrorInvalidValue;
    }
    done[0] = flags[0];
    attr_single[device & 63][v] = flags[16];
    return e;
}
