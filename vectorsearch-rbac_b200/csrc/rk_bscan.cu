// rk_bscan.cu — batched scoring on the 5th-generation tensor cores (tcgen05 + TMEM + TMA).
//
// For a batch of queries the distance computation is a dense contraction: with D = X · Qᵀ
//   L2:  ‖x − q‖² = ‖x‖² − 2·D + ‖q‖²        (pgvector/src/vector.c:549-563)
//   IP:  −x·q     = −D                         (pgvector/src/vector.c:596-606, :626-636)
// so every 128-row tile of the row store is multiplied against a tile of up to 128 queries with
// `tcgen05.mma.cta_group::1.kind::tf32` (fp32 operands straight from HBM, TF32 inside the MMA, fp32
// accumulation in TMEM).  The permission predicate (row_level_security.py:55-65) is a per-(row, query)
// role-mask test in the epilogue.  Tensor-core values are APPROXIMATE (TF32), so this kernel only
// nominates candidates: per (unit, query) the `kp` rows with the smallest approximate ordering value
// (‖x‖² − 2·D for L2 — the query's own norm is a constant of the ordering and is added back by
// rk_rescore_kernel — and −D for IP).
// rk_rescore_kernel then recomputes those rows' distances in exact fp32 with the scan kernel's
// summation order, takes the final top-k by (distance, tuple id) and proves — or refuses to — that
// no row outside the candidate set could belong to the answer (refused queries are re-run on the
// exact scan path by the host).  On integer-valued data (SIFT shape) TF32 is exact and the proof
// needs no margin.
//
// Shape: persistent, one 320-thread CTA per SM, static unit schedule.
//   warp 0   TMA producer: per k-chunk of 32 floats one 128x32 box of X and one of Q
//            (cp.async.bulk.tensor.2d, 128-byte swizzle) into a ring of stages.
//   warp 1   MMA issuer: 4 x (M=128, N<=128, K=8) tcgen05.mma per stage, accumulators double-buffered
//            in TMEM (2 x 128 columns); tcgen05.commit releases stages and publishes accumulators.
//   warps 2-9  epilogue, two per TMEM lane quarter (a warp reads lanes 32 (warp % 4) .. + 31 only), each scoring every
//            other 16-column chunk — with one warp per scheduler the epilogue ran at 0.1 IPC, every dependent instruction's
//            latency exposed: tcgen05.ld (one TMEM lane = one row), distance, mask test, threshold test
//            against the running per-query tau, append to per-query candidate buffers (global
//            scratch, L2 resident); warp-parallel radix-select compaction when a buffer fills.
#include <cuda_bf16.h>

#include "rk_bscan_dev.cuh"
#include "rk_kernels.h"

#define BS_M 128
#define BS_N 128
#define BS_KC 32
#ifndef BS_STAGES
#define BS_STAGES 5
#endif
#define BS_EW 8                         // epilogue warps: two per TMEM lane quarter, splitting the 16-column chunks
#define BS_THREADS (64 + BS_EW * 32)
#define BS_STAGE_BYTES 32768u          // X box 16 KB + Q box 16 KB

// ---------------------------------------------------------------------------------------------

struct BsSmem {
    uint32_t off_stage, off_bar, off_tmem, off_tau, off_tauf, off_cnt, off_trunc, off_qa, off_um, off_hist, off_sel, off_sort, total;
};

__host__ __device__ inline BsSmem bs_smem_layout(int mask_words)
{
    BsSmem s;
    uint32_t o = 0;
    s.off_stage = o; o += BS_STAGES * BS_STAGE_BYTES;
    s.off_bar = o; o += 8u * (2 * BS_STAGES + 4);
    s.off_tmem = o; o += 16;
    s.off_tau = o; o += BS_N * 4;
    s.off_tauf = o; o += BS_N * 4;     // the same thresholds as floats (16-byte aligned: read four at a time)
    s.off_cnt = o; o += BS_N * 4;
    o = (o + 15) & ~15u;
    s.off_trunc = o; o += BS_N * 8;
    s.off_qa = o; o += BS_N * 16;       // first two mask words of each query as one 16-byte load
    s.off_um = o; o += BS_N * 8u * (uint32_t) mask_words;
    s.off_hist = o; o += BS_EW * 256 * 4;
    s.off_sel = o; o += BS_EW * 16;
    o = (o + 15) & ~15u;
    s.off_sort = o; o += BS_EW * BS_SORT_MAX * 8;
    s.total = o + 1024;     // slack for the manual 1024-byte alignment of the stage ring
    return s;
}


#ifdef BS_TRACE
// experiment build only (tools/build_variant.sh ... -DBS_TRACE): clock64 stamps of CTA 0's warps over tiles [BS_TR_T0, BS_TR_T0 + 64)
#define BS_TR_T0 300u
#define BS_TR_SLOTS 64
__device__ unsigned long long *bs_trace_ptr;
#define BS_STAMP(tile, slot) do { if (blockIdx.x == 0 && lane == 0 && (tile) >= BS_TR_T0 && (tile) < BS_TR_T0 + 64u) bs_trace_ptr[((tile) - BS_TR_T0) * BS_TR_SLOTS + (slot)] = clock64(); } while (0)
#else
#define BS_STAMP(tile, slot) do { } while (0)
#endif

template <int METRIC>
__global__ void __launch_bounds__(BS_THREADS, 1)
rk_bscan_kernel(const __grid_constant__ CUtensorMap tmx, const __grid_constant__ CUtensorMap tmq16,
                const __grid_constant__ CUtensorMap tmq32, const __grid_constant__ CUtensorMap tmq64,
                const __grid_constant__ CUtensorMap tmq128, const RkBScanParams p)
{
    extern __shared__ unsigned char smem_raw[];
    const BsSmem L = bs_smem_layout(p.mask_words);
    // the swizzled stage ring must sit on a 1024-byte boundary
    unsigned char *smem = smem_raw + ((1024u - (rk_smem_u32(smem_raw) & 1023u)) & 1023u);
    uint64_t *full = reinterpret_cast<uint64_t *>(smem + L.off_bar);
    uint64_t *empty = full + BS_STAGES;
    uint64_t *tfull = empty + BS_STAGES;      // [2] accumulator ready
    uint64_t *tempty = tfull + 2;             // [2] accumulator drained
    uint32_t *s_tmem = reinterpret_cast<uint32_t *>(smem + L.off_tmem);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (threadIdx.x == 0) {
        for (int s = 0; s < BS_STAGES; s++) { rk_mbar_init(&full[s], 1); rk_mbar_init(&empty[s], 1); }
        for (int b = 0; b < 2; b++) { rk_mbar_init(&tfull[b], 1); rk_mbar_init(&tempty[b], BS_EW); }
        rk_fence_barrier_init();
    }
    if (warp == 1) bs_tmem_alloc(s_tmem, 2 * BS_N);
    bs_fence_before();
    __syncthreads();
    bs_fence_after();
    const uint32_t tmem_base = *s_tmem;

    if (warp == 0) {
        // ===================== TMA producer =====================
        int stage = 0;
        uint32_t phase = 0;
        uint32_t ptile = 0; (void) ptile;
        for (uint32_t ui = p.cta_off[blockIdx.x]; ui < p.cta_off[blockIdx.x + 1]; ui++) {
            const RkBUnit un = p.units[p.cta_units[ui]];
            // the query box is only as tall as the unit needs (16 / 32 / 64 / 128 rows)
            const uint32_t ncol = (un.nq + 15u) & ~15u;
            const CUtensorMap *tq = ncol <= 16 ? &tmq16 : ncol <= 32 ? &tmq32 : ncol <= 64 ? &tmq64 : &tmq128;
            const uint32_t qbox = ncol <= 16 ? 16u : ncol <= 32 ? 32u : ncol <= 64 ? 64u : 128u;
            const uint32_t tx = 16384u + qbox * 128u;
            const uint32_t ntiles = bs_unit_tiles(un);
            for (uint32_t i0 = 0; i0 < ntiles; i0 += 32) {
                // first rows of the next 32 tiles in one load (keeps the list lookup off the per-tile
                // critical path)
                const uint32_t tmine = un.tile_begin + (i0 + (uint32_t) lane) * un.tile_step;
                uint32_t rmine = 0;
                if (i0 + (uint32_t) lane < ntiles) rmine = p.tile_rows ? p.tile_rows[un.tl_off + tmine] : tmine * BS_M;
                const uint32_t nt = min(32u, ntiles - i0);
                for (uint32_t j = 0; j < nt; j++) {
                    const int row0 = (int) __shfl_sync(0xffffffffu, rmine, (int) j);
                    if (lane == 0) {
                        for (int kc = 0; kc < p.kchunks; kc++) {
                            rk_mbar_wait(&empty[stage], phase ^ 1u);
                            BS_STAMP(ptile, 0 + kc);                 // producer: stage free, about to issue k-chunk kc
                            unsigned char *st = smem + L.off_stage + (size_t) stage * BS_STAGE_BYTES;
                            rk_mbar_arrive_expect_tx(&full[stage], tx);
                            bs_tma_2d(st, &tmx, kc * BS_KC, row0, &full[stage]);
                            bs_tma_2d(st + 16384, tq, kc * BS_KC, (int) un.q0, &full[stage]);
                            if (++stage == BS_STAGES) { stage = 0; phase ^= 1u; }
                        }
                    }
                    __syncwarp();
                    ptile++;
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        int stage = 0;
        uint32_t phase = 0, tcount = 0;
        for (uint32_t ui = p.cta_off[blockIdx.x]; ui < p.cta_off[blockIdx.x + 1]; ui++) {
            const RkBUnit un = p.units[p.cta_units[ui]];
            uint32_t ncol = (un.nq + 15u) & ~15u;
            if (ncol < 16) ncol = 16;
            // c=F32, a=b=TF32, K-major both, N>>3, M>>4
            const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((ncol >> 3) << 17) | ((BS_M >> 4) << 24);
            const uint32_t ntiles = bs_unit_tiles(un);
            for (uint32_t ti = 0; ti < ntiles; ti++, tcount++) {
                const uint32_t buf = tcount & 1u, tph = (tcount >> 1) & 1u;
                BS_STAMP(tcount, 8);                                 // MMA warp: at the tile
                rk_mbar_wait(&tempty[buf], tph ^ 1u);
                BS_STAMP(tcount, 9);                                 // accumulator buffer free
                bs_fence_after();
                const uint32_t d_tmem = tmem_base + buf * BS_N;
                for (int kc = 0; kc < p.kchunks; kc++) {
                    rk_mbar_wait(&full[stage], phase);
                    BS_STAMP(tcount, 10 + kc);                       // k-chunk kc has landed
                    bs_fence_after();
                    if (lane == 0) {
                        const uint32_t a0 = rk_smem_u32(smem + L.off_stage + (size_t) stage * BS_STAGE_BYTES);
                        const uint64_t ad = bs_smem_desc(a0), bd = bs_smem_desc(a0 + 16384u);
#pragma unroll
                        for (int k = 0; k < BS_KC / 8; k++)      // 8 tf32 = 32 bytes per MMA: +2 in 16-byte units
                            bs_mma_tf32(d_tmem, ad + (uint64_t) (2 * k), bd + (uint64_t) (2 * k), idesc, (kc | k) != 0);
                        bs_commit(&empty[stage]);
                        if (kc == p.kchunks - 1) bs_commit(&tfull[buf]);
                    }
                    __syncwarp();
                    if (++stage == BS_STAGES) { stage = 0; phase ^= 1u; }
                }
            }
        }
    } else {
        // ===================== epilogue warps =====================
        const int ew = warp - 2;                       // 0..3 (compaction work split)
        const int quarter = warp & 3;                  // TMEM lanes this warp may read: [32 * (warp % 4), +32)
        const int erow = quarter * 32 + lane;          // row of the tile == TMEM lane
        const int chalf = ew >> 2;                     // which 16-column chunks of a tile this warp scores: chalf, chalf + 2, ...
        const int et = ew * 32 + lane;                 // epilogue thread index; threads below BS_N also own column et's state
        BsCtx c;
        c.tau = reinterpret_cast<uint32_t *>(smem + L.off_tau);
        float *tauf = reinterpret_cast<float *>(smem + L.off_tauf);
        c.cnt = reinterpret_cast<uint32_t *>(smem + L.off_cnt);
        c.hist = reinterpret_cast<uint32_t *>(smem + L.off_hist) + ew * 256;
        c.sel = reinterpret_cast<uint32_t *>(smem + L.off_sel) + ew * 4;
        c.scratch = p.scratch + (size_t) blockIdx.x * BS_N * BS_CAP;
        c.sortbuf = reinterpret_cast<uint64_t *>(smem + L.off_sort) + ew * BS_SORT_MAX;
        c.lane = lane;
        c.loose_extra = p.loose_extra;
        int32_t **s_trunc = reinterpret_cast<int32_t **>(smem + L.off_trunc);     // each query's trunc flag
        uint4 *s_qa = reinterpret_cast<uint4 *>(smem + L.off_qa);
        uint64_t *s_um = reinterpret_cast<uint64_t *>(smem + L.off_um);
        const int W = p.mask_words;
        auto esync = [] { rk_named_bar_sync(2, BS_EW * 32); };
        uint32_t tcount = 0;
        uint32_t qid = 0, qt_next = RK_KEY_SENTINEL;   // this thread's query (column et) and its shared threshold as last fetched
        for (uint32_t ui = p.cta_off[blockIdx.x]; ui < p.cta_off[blockIdx.x + 1]; ui++) {
            const RkBUnit un = p.units[p.cta_units[ui]];
            esync();                                   // previous unit fully emitted
            if (et < BS_N) {
                const bool live = (uint32_t) et < un.nq;
                // a padded column keeps tau = 0: no key is ever <= 0, so it never collects anything
                uint32_t tau0 = live ? RK_KEY_SENTINEL : 0u;
                // seeded threshold: the kp-th smallest ordering value over a SAMPLE of the rows this query
                // sees bounds the kp-th smallest over all of them from above
                if (live && p.tau_key && p.tau_count[un.q0 + et] >= p.kp) tau0 = rk_float_to_key(p.tau_key[(size_t) (un.q0 + et) * p.kp + (p.kp - 1)]);
                qid = live ? (p.slot_q ? p.slot_q[un.q0 + et] : un.q0 + (uint32_t) et) : 0u;
                if (live && p.qtau) tau0 = min(tau0, __ldcg(p.qtau + qid));       // what the query's other units have proved so far
                qt_next = tau0;
                c.tau[et] = tau0;
                tauf[et] = bs_tau_float(tau0);
                c.cnt[et] = 0;
                s_trunc[et] = p.trunc + qid;
                uint64_t w0 = ~0ull, w1 = 0ull;        // no permission filter: every row passes
                if (p.use_mask) {
                    w0 = live ? p.qmask[(size_t) (un.q0 + et) * W] : 0ull;
                    w1 = (live && W > 1) ? p.qmask[(size_t) (un.q0 + et) * W + 1] : 0ull;
                }
                s_qa[et] = make_uint4((uint32_t) w0, (uint32_t) (w0 >> 32), (uint32_t) w1, (uint32_t) (w1 >> 32));
                for (int w = 2; w < W; w++) s_um[et * W + w] = live ? p.qmask[(size_t) (un.q0 + et) * W + w] : 0ull;
            }
            esync();
            const uint32_t ncol = (un.nq + 15u) & ~15u;
            // per-row inputs are fetched one tile ahead (row start two tiles ahead) so that no global
            // load latency sits between two tiles
            const uint32_t ntiles = bs_unit_tiles(un);
            auto tile_row0 = [&](uint32_t i) -> uint32_t {       // first row of the unit's i-th tile
                const uint32_t t = un.tile_begin + i * un.tile_step;
                return i < ntiles ? (p.tile_rows ? p.tile_rows[un.tl_off + t] : t * BS_M) : 0xFFFFFFF0u;
            };
            uint32_t r0_cur = tile_row0(0), r0_next = tile_row0(1);
            float xn_n = 0.f;
            uint64_t ma_n = 0, mb_n = 0;
            bool own_n = true;
            // atoms plan: the unit nominates only rows of ITS document class (rows a tile drags in from
            // other classes are some other unit's to nominate — no row is nominated twice)
            const uint64_t *umask = (un.mask_id && p.unit_mask) ? p.unit_mask + (size_t) (un.mask_id - 1u) * W : nullptr;
            const uint64_t um0 = umask ? umask[0] : 0ull, um1 = (umask && W > 1) ? umask[1] : 0ull;
            auto fetch_row = [&](uint32_t r0) {
                const uint64_t rr = (uint64_t) r0 + (uint64_t) erow;
                xn_n = 0.f; ma_n = 0; mb_n = 0;
                if (rr < p.nrows) {
                    if (METRIC == 0) xn_n = p.xnorm[rr];
                    if (p.use_mask) {
                        const ulonglong2 m2 = reinterpret_cast<const ulonglong2 *>(p.rowmask2)[rr];
                        ma_n = m2.x;
                        mb_n = m2.y;
                        // (the comparison with the unit's class waits until the values are used, one tile later: doing it
                        // here would expose the load's latency on every tile)
                        own_n = true;
                        if (umask) for (int w = 2; w < W && own_n; w++) own_n = p.rowmask[rr * W + w] == umask[w];
                    } else {
                        ma_n = 1ull;                    // matches the all-ones query word
                    }
                }
            };
            fetch_row(r0_cur);
            uint64_t uhit0 = 0ull, uhit1 = 0ull;        // atoms plan: per chunk parity, which queries see the unit's class
            if (umask) {
                const uint32_t a0 = (uint32_t) um0, a1 = (uint32_t) (um0 >> 32), b0 = (uint32_t) um1, b1 = (uint32_t) (um1 >> 32);
#pragma unroll 1
                for (int i = 0; i < 8; i++) {
                    const uint32_t c0 = 16u * (uint32_t) i;
                    if (c0 >= ncol) break;
                    const uint4 qa = s_qa[c0 + (uint32_t) (lane & 15)];
                    const uint32_t hit = (a0 & qa.x) | (a1 & qa.y) | (b0 & qa.z) | (b1 & qa.w);
                    const uint64_t bal = (uint64_t) (__ballot_sync(0xffffffffu, hit != 0) & 0xFFFFu) << (16 * (i >> 1));
                    if (i & 1) uhit1 |= bal; else uhit0 |= bal;
                }
            }
            for (uint32_t ti = 0; ti < ntiles; ti++, tcount++) {
                const uint32_t buf = tcount & 1u, tph = (tcount >> 1) & 1u;
                BS_STAMP(tcount, 16 + ew * 6 + 0);                   // epilogue warp ew: loop top
                const uint64_t row = (uint64_t) r0_cur + (uint64_t) erow;
                // the two warps of a lane quarter swap chunk parity every tile (an odd number of chunks evens out over two tiles)
                const int cpar = (chalf + (int) ti) & 1;
                if (p.qtau && (ti & 3u) == 3u && (uint32_t) et < un.nq) {
                    // the shared threshold fetched four tiles ago (no load latency on this path), then the next fetch
                    if (qt_next < c.tau[et]) { c.tau[et] = qt_next; tauf[et] = bs_tau_float(qt_next); }
                    qt_next = __ldcg(p.qtau + qid);
                }
                if (umask && !(own_n && ma_n == um0 && mb_n == um1)) { own_n = false; ma_n = 0; mb_n = 0; }
                const bool valid = row < p.nrows && own_n;
                const float xn = xn_n;
                // which queries may see each row: the 32 rows of a warp belong to a few documents, so the mask test runs
                // once per (distinct row mask, query) — lanes test 16 queries at a time and vote — not per (row, query)
                uint64_t hit_lo = 0ull;                     // 16 bits per chunk of 16 queries (this warp's chunks)
                if (umask) {
                    // atoms plan: every row the unit may nominate carries the unit's own mask, so which queries see it was
                    // settled once per unit (uhit0 / uhit1) — nothing to group or vote on per tile
                    if ((ma_n | mb_n) != 0ull) hit_lo = cpar ? uhit1 : uhit0;
                } else {
                    const uint64_t ma = ma_n, mb = mb_n;
                    const uint32_t grp = __match_any_sync(0xffffffffu, ma) & __match_any_sync(0xffffffffu, mb);
                    uint32_t todo = 0xffffffffu;
                    while (todo) {
                        const int ld = __ffs((int) todo) - 1;
                        const uint32_t g = __shfl_sync(0xffffffffu, grp, ld);
                        const uint64_t a = __shfl_sync(0xffffffffu, ma, ld), b = __shfl_sync(0xffffffffu, mb, ld);
                        todo &= ~g;
                        if ((a | b) == 0ull) continue;
                        const bool mine = (g >> lane) & 1u;
                        const uint32_t a0 = (uint32_t) a, a1 = (uint32_t) (a >> 32), b0 = (uint32_t) b, b1 = (uint32_t) (b >> 32);
#pragma unroll 1
                        for (int i = 0; i < 4; i++) {
                            const uint32_t c0 = 16u * (uint32_t) (cpar + 2 * i);
                            if (c0 >= ncol) break;
                            const uint4 qa = s_qa[c0 + (uint32_t) (lane & 15)];
                            const uint32_t hit = (a0 & qa.x) | (a1 & qa.y) | (b0 & qa.z) | (b1 & qa.w);
                            const uint32_t bal = __ballot_sync(0xffffffffu, hit != 0) & 0xFFFFu;
                            if (mine) hit_lo |= (uint64_t) bal << (16 * i);
                        }
                    }
                }
                // issue the next tile's loads now; they are consumed on the next iteration
                r0_cur = r0_next;
                r0_next = tile_row0(ti + 2);
                fetch_row(r0_cur);
                BS_STAMP(tcount, 16 + ew * 6 + 1);                   // masks settled, next tile's loads issued
                rk_mbar_wait(&tfull[buf], tph);
                BS_STAMP(tcount, 16 + ew * 6 + 2);                   // accumulators ready
                bs_fence_after();
                const uint32_t taddr = tmem_base + buf * BS_N + ((uint32_t) (quarter * 32) << 16);
#pragma unroll 1
                for (int i = 0; i < 4; i++) {
                    const uint32_t c0 = 16u * (uint32_t) (cpar + 2 * i);
                    if (c0 >= ncol) break;
                    const uint32_t h16 = (uint32_t) (hit_lo >> (16 * i)) & 0xFFFFu;
                    if (W <= 2 && !__any_sync(0xffffffffu, h16 != 0)) continue;      // none of the 32 rows is visible to these 16 queries
                    uint32_t r[16];
                    bs_tmem_ld16(taddr + c0, r);
                    // straight-line: per column one FMA, a quarter of a 16-byte threshold load, one compare
                    uint32_t pass = 0;
#pragma unroll
                    for (int j4 = 0; j4 < 4; j4++) {
                        const float4 t4 = reinterpret_cast<const float4 *>(tauf + c0)[j4];
                        const float tj[4] = {t4.x, t4.y, t4.z, t4.w};
#pragma unroll
                        for (int jj = 0; jj < 4; jj++) {
                            const int j = j4 * 4 + jj;
                            const float v = __uint_as_float(r[j]);
                            const float dist = METRIC == 0 ? fmaf(-2.0f, v, xn) : -v;
                            r[j] = __float_as_uint(dist);
                            pass |= (uint32_t) (dist <= tj[jj]) << j;
                        }
                    }
                    pass &= h16;
                    if (W > 2 && p.use_mask && valid) {
                        // more than 128 roles: the remaining mask words (rare, slow path)
#pragma unroll
                        for (int j = 0; j < 16; j++) {
                            const uint32_t q = c0 + j;
                            uint64_t hit = 0;
                            for (int w = 2; w < W; w++) hit |= p.rowmask[row * W + w] & s_um[q * W + w];
                            if (hit != 0 && __uint_as_float(r[j]) <= tauf[q]) pass |= 1u << j;
                        }
                    }
                    if (__any_sync(0xffffffffu, pass != 0)) {
#pragma unroll
                        for (int j = 0; j < 16; j++) {
                            if (pass & (1u << j)) {
                                const uint32_t q = c0 + j;
                                const uint32_t slot = atomicAdd(&c.cnt[q], 1u);
                                __stcg(&c.scratch[(size_t) q * BS_CAP + slot], make_uint2(rk_float_to_key(__uint_as_float(r[j])), (uint32_t) row));
                            }
                        }
                    }
                }
                BS_STAMP(tcount, 16 + ew * 6 + 3);                   // chunks scored
                bs_fence_before();
                __syncwarp();
                if (lane == 0) rk_mbar_arrive(&tempty[buf]);
                // a tile adds at most 128 entries per query: compact what could overflow on the next one
                // (the reducing barrier also publishes this tile's appends)
                // (the count a thread reads here may still miss this tile's appends by the other
                // warps — at most 96 — hence the margin of two tiles: seen <= CAP - 256 means at most
                // CAP - 160 now and at most CAP - 32 after the next tile.  Checking every second tile with a trigger two
                // tiles lower was measured: the extra compactions cost more than the barrier saves, 2.31 vs 2.19 ms on C3.)
                const uint32_t trig = BS_CAP - 2 * BS_M;
                const bool need = rk_named_bar_red_or(2, BS_EW * 32, et < BS_N && c.cnt[et] > trig);
                BS_STAMP(tcount, 16 + ew * 6 + 4);                   // past the barrier
                if (need) {
                    for (uint32_t q = (uint32_t) ew; q < un.nq; q += BS_EW) {
                        const uint32_t n = c.cnt[q];
                        if (n > p.round_mark) {
                            uint32_t tau;
                            uint32_t m = bs_warp_compact(c, (int) q, n, (uint32_t) p.kp, &tau, false);
                            if (p.qtau && lane == 0) atomicMin(p.qtau + (uint32_t) (s_trunc[q] - p.trunc), tau);     // >= kp rows are <= tau
                            if (m > BS_CAP / 2 && tau > 0) {
                                // more than half the buffer ties at the kp-th key: drop that key altogether
                                // and record that the list is no longer complete (the query is re-run exactly)
                                tau -= 1u;
                                m = bs_warp_filter(c, (int) q, m, tau);
                                if (lane == 0) *s_trunc[q] = 1;
                            }
                            if (lane == 0) { c.cnt[q] = m; c.tau[q] = tau; tauf[q] = bs_tau_float(tau); }
                        }
                    }
                    esync();
                    BS_STAMP(tcount, 16 + ew * 6 + 5);               // compaction round done
                }
            }
            // unit end: every query's list, sorted by (approximate key, row), at most kp entries
            for (uint32_t q = (uint32_t) ew; q < un.nq; q += BS_EW) {
                uint32_t n = c.cnt[q];
                if (n > (uint32_t) p.kp) {
                    uint32_t tau;
                    n = bs_warp_compact(c, (int) q, n, (uint32_t) p.kp, &tau);
                    if (p.qtau && lane == 0) atomicMin(p.qtau + (uint32_t) (s_trunc[q] - p.trunc), tau);
                    if (n > BS_SORT_MAX && tau > 0) {
                        n = bs_warp_filter(c, (int) q, n, tau - 1u);
                        if (lane == 0) *s_trunc[q] = 1;
                    }
                    if (n > BS_SORT_MAX) n = BS_SORT_MAX;
                }
                const size_t list = (size_t) un.list_base + q;
                if (n <= (uint32_t) p.kp) bs_warp_emit_unsorted(c, (int) q, n, p.cand_key + list * p.kp, p.cand_row + list * p.kp, p.cand_count + list);
                else bs_warp_emit(c, (int) q, n, (uint32_t) p.kp, p.cand_key + list * p.kp, p.cand_row + list * p.kp, p.cand_count + list);
            }
        }
    }
    __syncthreads();
    if (warp == 1) bs_tmem_dealloc(tmem_base, 2 * BS_N);
}

// ---------------------------------------------------------------------------------------------
// exact re-scoring of the nominated rows + final top-k + sufficiency proof

#define RS_THREADS 256
#define RS_MAX 512

template <int METRIC>
__global__ void __launch_bounds__(RS_THREADS) rk_rescore_kernel(const RkRescoreParams p)
{
    extern __shared__ __align__(16) unsigned char rs_smem[];
    float4 *qs = reinterpret_cast<float4 *>(rs_smem);                                  // [dim4]
    uint64_t *s_tid = reinterpret_cast<uint64_t *>(rs_smem + (size_t) p.dim_pad * 4);   // [RS_MAX]
    uint32_t *s_key = reinterpret_cast<uint32_t *>(s_tid + RS_MAX);
    uint32_t *s_row = s_key + RS_MAX;
    const int q = blockIdx.x, t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const int dim4 = p.dim_pad >> 2;
    const float4 *q4 = reinterpret_cast<const float4 *>(p.q + (size_t) q * p.dim_pad);
    for (int i = t; i < dim4; i += RS_THREADS) qs[i] = q4[i];
    const int n = p.sel_count[q];
    int P = 32;
    while (P < n) P <<= 1;
    for (int i = n + t; i < P; i += RS_THREADS) { s_key[i] = RK_KEY_SENTINEL; s_tid[i] = ~0ull; s_row[i] = 0xFFFFFFFFu; }
    __syncthreads();
    // the scan kernel's summation order: 4 rows per warp step, 8 lanes per row — or, for wide rows, one row per warp step
    if (p.wide_rows) {
        for (int i = warp; i < n; i += RS_THREADS / 32) {
            const uint32_t row = p.sel_row[(size_t) q * p.kp + i];
            const float4 *xr = reinterpret_cast<const float4 *>(p.x + (size_t) row * p.dim_pad);
            float s = 0.f;
            if (!p.null_query) s = rk_row_sum32<METRIC>(xr, qs, dim4, lane);
            if (lane == 0) {
                s_key[i] = rk_float_to_key(METRIC == 0 ? s : -s);
                s_row[i] = row;
                s_tid[i] = p.tid[row];
            }
        }
    } else {
        const int g = lane >> 3, l8 = lane & 7;
        for (int i0 = warp * 4; i0 < n; i0 += (RS_THREADS / 32) * 4) {
            const int i = i0 + g;
            const bool v = i < n;
            const uint32_t row = p.sel_row[(size_t) q * p.kp + (v ? i : 0)];
            const float4 *xr = reinterpret_cast<const float4 *>(p.x + (size_t) row * p.dim_pad);
            float s = 0.f;
            if (!p.null_query) s = rk_row_sum8<METRIC>(xr, qs, dim4, l8);
            if (v && l8 == 0) {
                s_key[i] = rk_float_to_key(METRIC == 0 ? s : -s);
                s_row[i] = row;
                s_tid[i] = p.tid[row];
            }
        }
    }
    __syncthreads();
    // bitonic sort by (key, tid)
    for (int size = 2; size <= P; size <<= 1) {
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
            for (int i = t; i < (P >> 1); i += RS_THREADS) {
                const int lo = 2 * i - (i & (stride - 1)), hi = lo + stride;
                const bool up = (lo & size) == 0;
                const uint32_t ka = s_key[lo], kb = s_key[hi];
                const uint64_t ta = s_tid[lo], tb = s_tid[hi];
                const bool a_lt_b = ka < kb || (ka == kb && ta < tb), b_lt_a = kb < ka || (ka == kb && tb < ta);
                if (up ? b_lt_a : a_lt_b) {
                    s_key[lo] = kb; s_key[hi] = ka; s_tid[lo] = tb; s_tid[hi] = ta;
                    const uint32_t ra = s_row[lo], rb = s_row[hi];
                    s_row[lo] = rb; s_row[hi] = ra;
                }
            }
            __syncthreads();
        }
    }
    const int k = p.k;
    const int cnt = n < k ? n : k;
    const int oq = p.slot_q ? (int) p.slot_q[q] : q;      // the caller's query index
    const size_t ob = (size_t) oq * k;
    for (int i = t; i < k; i += RS_THREADS) {
        const bool ok = i < cnt;
        const float ord = ok ? rk_key_to_float(s_key[i]) : __int_as_float(0x7f800000);
        if (p.out_tid) p.out_tid[ob + i] = ok ? s_tid[i] : ~0ull;
        if (p.out_key) p.out_key[ob + i] = ord;
        if (p.out_doc) p.out_doc[ob + i] = ok ? p.doc_ids[p.rowdoc[s_row[i]]] : 0xFFFFFFFFu;
        if (p.out_dist) {
            double v;
            if (!ok) v = __longlong_as_double(0x7ff0000000000000ll);
            else if (p.null_query) v = 0.0;
            else if (METRIC == 0) v = sqrt((double) ord);      // vector.c:577
            else v = (double) ord;                             // vector.c:635
            p.out_dist[ob + i] = v;
        }
    }
    if (t == 0) {
        if (p.out_count) p.out_count[oq] = cnt;
        // Sufficiency: a row outside the candidate set has approximate value >= a_last (the largest
        // approximate value kept, when the list is saturated).  Its exact value is then >= a_last - eps.
        // The answer is proven iff that exceeds the exact k-th value (strictly, so that no tie on the
        // k-th value with a smaller tuple id can hide outside).
        int uncertain = p.trunc ? p.trunc[q] : 0;
        // fp32 overflow (vector_type.out:387-391, 413-417: distances come out as Infinity): the tensor-core value may have
        // been inf - inf = NaN, which no threshold test admits — such rows were never nominated.  Leave it to the scan.
        if (!(p.xnorm_max < 3.0e38f) || !(p.qnorm[q] < 3.0e38f)) uncertain = 1;
        if (n >= p.kp && cnt > 0) {
            // the nominated ordering value leaves the query's own norm out (L2): add it back
            const float a_last = p.sel_key[(size_t) q * p.kp + (n - 1)] + (METRIC == 0 ? p.qnorm[q] : 0.f);
            const float e_k = rk_key_to_float(s_key[cnt - 1]);
            float eps = 0.f;
            float eps_scale = p.eps_scale;
            if (p.probe) {
                // exact iff operands are small integers (exact in TF32, products exact in fp32) and every partial sum
                // stays below 2^24: |x|^2 + |q|^2 + 2 |x.q| <= (|x| + |q|)^2 by Cauchy-Schwarz
                const float nx = sqrtf(p.xnorm_max), nq = sqrtf(__uint_as_float((uint32_t) (p.probe[4] & 0xFFFFFFFFull)));
                const float qabs = __uint_as_float((uint32_t) (p.probe[1] & 0xFFFFFFFFull));       // largest |value| of the queries
                if (p.x_integer && p.probe[0] == 0ull && (nx + nq) * (nx + nq) < 16700000.0f && fmaxf(p.x_absmax, qabs) <= p.exact_absmax) eps_scale = 0.f;
            }
            if (eps_scale > 0.f) eps = eps_scale * sqrtf(p.xnorm_max * p.qnorm[q]) + 1e-6f * (p.xnorm_max + p.qnorm[q]);
            if (cnt < k || !(a_last - eps > e_k)) uncertain = 1;
        }
        p.uncertain[oq] = uncertain;
    }
}

// ---------------------------------------------------------------------------------------------
// preparation kernels: row norms, per-row masks, TF32-exactness probe

__global__ void rk_row_norms_kernel(const float *__restrict__ x, size_t n, int dim_pad, float *__restrict__ out)
{
    const int lane = threadIdx.x & 31;
    const size_t warp = (blockIdx.x * (size_t) blockDim.x + threadIdx.x) >> 5;
    const size_t nw = ((size_t) gridDim.x * blockDim.x) >> 5;
    const int dim4 = dim_pad >> 2;
    for (size_t r = warp; r < n; r += nw) {
        const float4 *p4 = reinterpret_cast<const float4 *>(x + r * (size_t) dim_pad);
        float s = 0.f;
        for (int c = lane; c < dim4; c += 32) { const float4 v = p4[c]; s = fmaf(v.x, v.x, s); s = fmaf(v.y, v.y, s); s = fmaf(v.z, v.z, s); s = fmaf(v.w, v.w, s); }
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        if (lane == 0) out[r] = s;
    }
}

__global__ void rk_rowmask_kernel(const uint32_t *__restrict__ rowdoc, const uint64_t *__restrict__ doc_mask, size_t n, int W,
                                  uint64_t *__restrict__ out)
{
    const size_t total = n * (size_t) W;
    for (size_t i = blockIdx.x * (size_t) blockDim.x + threadIdx.x; i < total; i += (size_t) gridDim.x * blockDim.x) {
        const size_t r = i / W;
        const int w = (int) (i % W);
        out[i] = doc_mask[(size_t) rowdoc[r] * W + w];
    }
}

// First two mask words of every row as one 16-byte record (what the tensor kernel stages per tile);
// rows [n, n_pad) get an empty mask so that a tile reaching past the store never passes the test.
__global__ void rk_rowmask2_kernel(const uint32_t *__restrict__ rowdoc, const uint64_t *__restrict__ doc_mask, size_t n, size_t n_pad,
                                   int W, ulonglong2 *__restrict__ out)
{
    for (size_t r = blockIdx.x * (size_t) blockDim.x + threadIdx.x; r < n_pad; r += (size_t) gridDim.x * blockDim.x) {
        ulonglong2 v = make_ulonglong2(0ull, 0ull);
        if (r < n) {
            const uint64_t *m = doc_mask + (size_t) rowdoc[r] * W;
            v.x = m[0];
            if (W > 1) v.y = m[1];
        }
        out[r] = v;
    }
}

// out[0] = number of elements that are not small integers (|v| <= 2047, v == rint(v)); out[1] = max |v| bits
__global__ void rk_tf32_probe_kernel(const float *__restrict__ x, size_t n, unsigned long long *out)
{
    unsigned bad = 0;
    unsigned mx = 0;
    for (size_t i = blockIdx.x * (size_t) blockDim.x + threadIdx.x; i < n; i += (size_t) gridDim.x * blockDim.x) {
        const float v = x[i];
        const float a = fabsf(v);
        bad += !(a <= 2047.0f && v == rintf(v));
        mx = max(mx, __float_as_uint(a));
    }
    bad = __reduce_add_sync(0xffffffffu, bad);
    mx = __reduce_max_sync(0xffffffffu, mx);
    if ((threadIdx.x & 31) == 0) {
        if (bad) atomicAdd(&out[0], (unsigned long long) bad);
        atomicMax(&out[1], (unsigned long long) mx);
    }
}

__global__ void rk_max_f32_kernel(const float *__restrict__ x, size_t n, unsigned *out)
{
    unsigned mx = 0;
    for (size_t i = blockIdx.x * (size_t) blockDim.x + threadIdx.x; i < n; i += (size_t) gridDim.x * blockDim.x)
        mx = max(mx, __float_as_uint(fabsf(x[i])));
    mx = __reduce_max_sync(0xffffffffu, mx);
    if ((threadIdx.x & 31) == 0) atomicMax(out, mx);
}

// fp32 rows [n][dim_pad] -> bf16 rows [n][dim_b] (round to nearest even; dim_b >= dim_pad, the tail is zero)
__global__ void rk_rows_to_bf16_kernel(const float *__restrict__ x, size_t n, int dim_pad, int dim_b, __nv_bfloat16 *__restrict__ out)
{
    const size_t total = n * (size_t) dim_b;
    for (size_t i = blockIdx.x * (size_t) blockDim.x + threadIdx.x; i < total; i += (size_t) gridDim.x * blockDim.x) {
        const size_t r = i / (size_t) dim_b;
        const int c = (int) (i - r * (size_t) dim_b);
        out[i] = __float2bfloat16_rn(c < dim_pad ? x[r * (size_t) dim_pad + c] : 0.f);
    }
}

cudaError_t rk_launch_rows_to_bf16(const float *x, size_t n, int dim_pad, int dim_b, void *out, cudaStream_t stream)
{
    if (n) rk_rows_to_bf16_kernel<<<148 * 16, 256, 0, stream>>>(x, n, dim_pad, dim_b, reinterpret_cast<__nv_bfloat16 *>(out));
    return cudaGetLastError();
}

cudaError_t rk_launch_row_norms(const float *x, size_t n, int dim_pad, float *out, cudaStream_t stream)
{
    if (n) rk_row_norms_kernel<<<148 * 8, 256, 0, stream>>>(x, n, dim_pad, out);
    return cudaGetLastError();
}
cudaError_t rk_launch_rowmask(const uint32_t *rowdoc, const uint64_t *doc_mask, size_t n, int W, uint64_t *out, cudaStream_t stream)
{
    if (n) rk_rowmask_kernel<<<148 * 8, 256, 0, stream>>>(rowdoc, doc_mask, n, W, out);
    return cudaGetLastError();
}
cudaError_t rk_launch_rowmask2(const uint32_t *rowdoc, const uint64_t *doc_mask, size_t n, size_t n_pad, int W, uint64_t *out,
                               cudaStream_t stream)
{
    if (n_pad) rk_rowmask2_kernel<<<148 * 8, 256, 0, stream>>>(rowdoc, doc_mask, n, n_pad, W, reinterpret_cast<ulonglong2 *>(out));
    return cudaGetLastError();
}
cudaError_t rk_launch_tf32_probe(const float *x, size_t n, unsigned long long *d_out2, cudaStream_t stream)
{
    if (n) rk_tf32_probe_kernel<<<148 * 8, 256, 0, stream>>>(x, n, d_out2);
    return cudaGetLastError();
}
cudaError_t rk_launch_max_f32(const float *x, size_t n, unsigned *d_out, cudaStream_t stream)
{
    if (n) rk_max_f32_kernel<<<148 * 4, 256, 0, stream>>>(x, n, d_out);
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// host launchers

size_t rk_bscan_smem_bytes(int mask_words) { return bs_smem_layout(mask_words).total; }
size_t rk_bscan_scratch_bytes(int grid) { return (size_t) grid * BS_N * BS_CAP * sizeof(uint2); }
int rk_bscan_lists_per_query() { return 1; }

typedef CUresult (*rk_encode_fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                 const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                 CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// 2-D tensor [rows][row_elems] of fp32 (box = 32 elements x box_rows) or bf16 (box = 64 elements x box_rows): 128 bytes per box
// row either way, 128-byte swizzle, zero fill out of bounds.
cudaError_t rk_make_tensor_map_2d_t(CUtensorMap *tm, const void *base, uint64_t rows, int row_elems, uint32_t box_rows, bool bf16);
cudaError_t rk_make_tensor_map_2d(CUtensorMap *tm, const float *base, uint64_t rows, int dim_pad, uint32_t box_rows)
{
    return rk_make_tensor_map_2d_t(tm, base, rows, dim_pad, box_rows, false);
}

cudaError_t rk_make_tensor_map_2d_t(CUtensorMap *tm, const void *base, uint64_t rows, int row_elems, uint32_t box_rows, bool bf16)
{
    static rk_encode_fn encode = nullptr;
    if (!encode) {
        void *fn = nullptr;
        cudaDriverEntryPointQueryResult qres;
        cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
        if (e != cudaSuccess) return e;
        if (!fn || qres != cudaDriverEntryPointSuccess) return cudaErrorNotSupported;
        encode = reinterpret_cast<rk_encode_fn>(fn);
    }
    cuuint64_t gdim[2] = {(cuuint64_t) row_elems, (cuuint64_t) (rows ? rows : 1)};
    cuuint64_t gstride[1] = {(cuuint64_t) row_elems * (bf16 ? 2u : 4u)};
    cuuint32_t box[2] = {bf16 ? 2u * BS_KC : BS_KC, box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = encode(tm, bf16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void *>(base), gdim, gstride, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS ? cudaSuccess : cudaErrorInvalidValue;
}

#ifdef BS_TRACE
#include <cstdio>
#include <cstdlib>
static unsigned long long *bs_trace_host = nullptr;
static void bs_trace_dump()
{
    const char *path = getenv("BS_TRACE_OUT");
    FILE *f = fopen(path ? path : "gpurun_out/bs_trace.txt", "w");
    if (!f || !bs_trace_host) return;
    for (int t = 0; t < 64; t++) {
        fprintf(f, "%d", t);
        for (int s = 0; s < BS_TR_SLOTS; s++) fprintf(f, " %llu", bs_trace_host[t * BS_TR_SLOTS + s]);
        fprintf(f, "\n");
    }
    fclose(f);
}
#endif

cudaError_t rk_launch_bscan(const RkBScanParams &p, const float *x, const float *q, uint64_t nq, int grid, cudaStream_t stream)
{
#ifdef BS_TRACE
    if (!bs_trace_host) {
        cudaHostAlloc(&bs_trace_host, 64 * BS_TR_SLOTS * 8, cudaHostAllocMapped);
        unsigned long long *dp = nullptr;
        cudaHostGetDevicePointer(&dp, bs_trace_host, 0);
        cudaMemcpyToSymbol(bs_trace_ptr, &dp, sizeof dp);
        atexit(bs_trace_dump);
    }
    memset(bs_trace_host, 0, 64 * BS_TR_SLOTS * 8);
#endif
    CUtensorMap tmx, tmq[4];
    cudaError_t e = rk_make_tensor_map_2d(&tmx, x, p.nrows, p.dim_pad, BS_M);
    if (e != cudaSuccess) return e;
    for (int i = 0; i < 4; i++) {
        e = rk_make_tensor_map_2d(&tmq[i], q, nq, p.dim_pad, 16u << i);
        if (e != cudaSuccess) return e;
    }
    const size_t smem = rk_bscan_smem_bytes(p.mask_words);
    auto kern = p.metric == 0 ? rk_bscan_kernel<0> : rk_bscan_kernel<1>;
    // (function attributes are per device: rk_multi runs one index per GPU in one process)
    static bool attr_set[64][2];
    int dev_ = 0;
    cudaGetDevice(&dev_);
    if (!attr_set[dev_ & 63][p.metric & 1]) {
        e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
        if (e != cudaSuccess) return e;
        attr_set[dev_ & 63][p.metric & 1] = true;
    }
    kern<<<grid, BS_THREADS, smem, stream>>>(tmx, tmq[0], tmq[1], tmq[2], tmq[3], p);
    return cudaGetLastError();
}

cudaError_t rk_launch_rescore(const RkRescoreParams &p, int nq, cudaStream_t stream)
{
    if (nq <= 0) return cudaSuccess;
    const size_t smem = (size_t) p.dim_pad * 4 + RS_MAX * 16;
    if (p.metric == 0) rk_rescore_kernel<0><<<nq, RS_THREADS, smem, stream>>>(p);
    else rk_rescore_kernel<1><<<nq, RS_THREADS, smem, stream>>>(p);
    return cudaGetLastError();
}
