// rk_bscan2.cu — batched scoring on CTA pairs (tcgen05 cta_group::2).
//
// Same contract as rk_bscan_kernel (rk_bscan.cu): D = X · Qᵀ on the tensor cores
//   L2:  ‖x − q‖² = ‖x‖² − 2·D + ‖q‖²        (pgvector/src/vector.c:549-563)
//   IP:  −x·q     = −D                         (pgvector/src/vector.c:596-606, :626-636)
// with the permission predicate (row_level_security.py:55-65) as a role-mask test in the epilogue, and
// only NOMINATES candidates for rk_rescore_kernel.  What changes is how the operands reach the tensor
// cores.  The one-CTA kernel stages 128 rows of X and 128 queries per 128x128 output tile: 32 flop per
// staged byte, and ncu shows it bound by the L2 -> shared-memory feed (8.8 TB/s of TMA traffic, tensor
// pipe 26 % busy).  Here the two CTAs of a cluster (one TPC) issue ONE
// `tcgen05.mma.cta_group::2.kind::tf32` of M = 256 (128 rows of X per CTA), N = up to 256 queries:
// each CTA stages its own 128 rows of X and only HALF of the query tile, the pair's tensor cores read
// both halves — 64 flop per staged byte, half the L2 -> shared traffic and half the shared-memory
// read bandwidth per flop.
//
// Shape: persistent, one cluster of two 320-thread CTAs per TPC, static schedule of unit PAIRS (units
// 2p and 2p+1 share their query tile; rank r of the cluster walks unit 2p + r, i.e. every other tile
// of the range).
//   warp 0     TMA producer (both CTAs): per k-chunk of 32 floats one 128x32 box of X and one
//              (N/2)x32 box of Q (cp.async.bulk.tensor.2d.cta_group::2, 128-byte swizzle), both
//              completing on the LEADER's full barrier.
//   warp 1     MMA issuer (leader CTA only): 4 x (M=256, N, K=8) per stage; tcgen05.commit multicast
//              releases the stage in both CTAs and publishes the accumulators (2 x 256 TMEM columns per
//              CTA, double buffered) to both epilogues.
//   warps 2-9  epilogue (each CTA on its own 128 rows): tcgen05.ld, ordering value, mask test,
//              threshold test, append to per-query candidate buffers, radix-select compaction.
//              Two warps per TMEM lane quarter split the columns.
#include "rk_bscan_dev.cuh"
#include "rk_kernels.h"

#define B2_M 128
#define B2_N 256
#define B2_KC 32
#define B2_THREADS 320
#define B2_EW 8                         // epilogue warps
#define B2_STAGE_BYTES 32768u           // X box 16 KB + half Q box <= 16 KB
#define B2_DUMMY_ROW 0x7FFFFF00         // a tile past the end of the store: TMA fills zeros, no row is valid

// ---------------------------------------------------------------------------------------------
// cluster / cta_group::2 PTX

__device__ __forceinline__ uint32_t b2_cluster_ctarank()
{
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}

__device__ __forceinline__ uint32_t b2_cluster_id()
{
    uint32_t r;
    asm volatile("mov.u32 %0, %%clusterid.x;" : "=r"(r));
    return r;
}

// shared::cluster address of the same variable in CTA `rank` of the cluster
__device__ __forceinline__ uint32_t b2_mapa(uint32_t smem_addr, uint32_t rank)
{
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(smem_addr), "r"(rank));
    return r;
}

__device__ __forceinline__ void b2_cluster_sync()
{
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}

__device__ __forceinline__ void b2_mbar_arrive_cluster(uint32_t cluster_addr)
{
    // (default semantics: release at CTA scope — the arrive only orders this warp's TMEM reads, already complete,
    // before the leader's next MMA; a cluster-scope release would also wait for the warp's global stores)
    asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}

// bounded wait: a protocol error traps instead of hanging the device
__device__ __forceinline__ void b2_mbar_wait(uint64_t *bar, uint32_t parity)
{
    const uint32_t a = rk_smem_u32(bar);
    uint32_t done = 0;
    for (uint32_t spin = 0; spin < (1u << 26); spin++) {
        asm volatile(
            "{\n\t"
            ".reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t"
            "}"
            : "=r"(done)
            : "r"(a), "r"(parity)
            : "memory");
        if (done) return;
    }
    __trap();
}

// 2-D tiled load into THIS CTA's shared memory, completing on a barrier of the pair's leader
__device__ __forceinline__ void b2_tma_2d(void *smem_dst, const CUtensorMap *tm, int c0, int c1, uint32_t leader_bar)
{
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
            rk_smem_u32(smem_dst)),
        "l"(tm), "r"(leader_bar), "r"(c0), "r"(c1)
        : "memory");
}

__device__ __forceinline__ void b2_tmem_alloc(uint32_t *smem_dst, uint32_t ncols)
{
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(rk_smem_u32(smem_dst)), "r"(ncols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}

__device__ __forceinline__ void b2_tmem_dealloc(uint32_t taddr, uint32_t ncols)
{
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}

// arrives on the barrier at this offset in both CTAs of the pair once all earlier MMAs have completed
__device__ __forceinline__ void b2_commit(uint64_t *bar)
{
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
                     rk_smem_u32(bar)),
                 "h"((uint16_t) 3)
                 : "memory");
}

// bf16 operands (kind::f16, a/b format bf16 in the instruction descriptor), K = 16 elements = 32 bytes per instruction
__device__ __forceinline__ void b2_mma_bf16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate)
{
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t"
        "}" ::"r"(tmem_d),
        "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}

__device__ __forceinline__ void b2_mma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate)
{
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::tf32 [%0], %1, %2, %3, p;\n\t"
        "}" ::"r"(tmem_d),
        "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}

// ---------------------------------------------------------------------------------------------

struct B2Smem {
    uint32_t off_stage, off_bar, off_tmem, off_tau, off_tauf, off_cnt, off_trunc, off_qa, off_um, off_hist, off_sel, off_sort, total;
    int stages;
};

__host__ __device__ inline B2Smem b2_smem_layout(int mask_words)
{
    B2Smem s;
    s.stages = mask_words <= 2 ? 5 : 4;      // the extra mask words of 256 queries take the fifth stage's room
    uint32_t o = 0;
    s.off_stage = o; o += (uint32_t) s.stages * B2_STAGE_BYTES;
    s.off_bar = o; o += 8u * (2 * 5 + 4);
    s.off_tmem = o; o += 16;
    s.off_tau = o; o += B2_N * 4;
    s.off_tauf = o; o += B2_N * 4;     // the same thresholds as floats (16-byte aligned: read four at a time)
    s.off_cnt = o; o += B2_N * 4;
    o = (o + 15) & ~15u;
    s.off_trunc = o; o += B2_N * 8;
    s.off_qa = o; o += B2_N * 16;
    s.off_um = o; o += mask_words > 2 ? B2_N * 8u * (uint32_t) mask_words : 0u;
    s.off_hist = o; o += B2_EW * 256 * 4;
    s.off_sel = o; o += B2_EW * 16;
    o = (o + 15) & ~15u;
    s.off_sort = o; o += B2_EW * BS_SORT_MAX * 8;
    s.total = o + 1024;
    return s;
}

__device__ __forceinline__ uint32_t b2_ncol(uint32_t nq)
{
    uint32_t n = 16;
    while (n < nq) n <<= 1;
    return n;
}

template <int METRIC>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(B2_THREADS, 1)
rk_bscan2_kernel(const __grid_constant__ CUtensorMap tmx, const __grid_constant__ CUtensorMap tmq8,
                 const __grid_constant__ CUtensorMap tmq16, const __grid_constant__ CUtensorMap tmq32,
                 const __grid_constant__ CUtensorMap tmq64, const __grid_constant__ CUtensorMap tmq128, const RkBScanParams p)
{
    extern __shared__ unsigned char smem_raw[];
    const B2Smem L = b2_smem_layout(p.mask_words);
    const int S = L.stages;
    unsigned char *smem = smem_raw + ((1024u - (rk_smem_u32(smem_raw) & 1023u)) & 1023u);
    uint64_t *full = reinterpret_cast<uint64_t *>(smem + L.off_bar);
    uint64_t *empty = full + 5;
    uint64_t *tfull = empty + 5;              // [2] accumulator ready (both CTAs, by multicast commit)
    uint64_t *tempty = tfull + 2;             // [2] accumulator drained (leader's: 16 epilogue warps of the pair arrive)
    uint32_t *s_tmem = reinterpret_cast<uint32_t *>(smem + L.off_tmem);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = b2_cluster_ctarank();
    const uint32_t cid = b2_cluster_id();
    const bool leader = rank == 0;

    if (threadIdx.x == 0) {
        for (int s = 0; s < S; s++) { rk_mbar_init(&full[s], 1); rk_mbar_init(&empty[s], 1); }
        for (int b = 0; b < 2; b++) { rk_mbar_init(&tfull[b], 1); rk_mbar_init(&tempty[b], 2 * B2_EW); }
        rk_fence_barrier_init();
    }
    if (warp == 1) b2_tmem_alloc(s_tmem, 2 * B2_N);
    bs_fence_before();
    __syncthreads();
    b2_cluster_sync();
    bs_fence_after();
    const uint32_t tmem_base = *s_tmem;

    const uint32_t u_begin = p.cta_off[cid], u_end = p.cta_off[cid + 1];

    if (warp == 0) {
        // ===================== TMA producer (both CTAs) =====================
        int stage = 0;
        uint32_t phase = 0;
        for (uint32_t ui = u_begin; ui < u_end; ui++) {
            const uint32_t pi = p.cta_units[ui];
            const RkBUnit un = p.units[2 * pi + rank];
            const uint32_t my_tiles = bs_unit_tiles(un);
            const uint32_t ntiles = max(my_tiles, bs_unit_tiles(p.units[2 * pi + (rank ^ 1u)]));
            const uint32_t ncol = b2_ncol(un.nq), half = ncol >> 1;
            const CUtensorMap *tq = half <= 8 ? &tmq8 : half <= 16 ? &tmq16 : half <= 32 ? &tmq32 : half <= 64 ? &tmq64 : &tmq128;
            const uint32_t tx = 2u * (16384u + half * 128u);      // both CTAs' boxes land on the leader's barrier
            const int qrow = (int) (un.q0 + rank * half);
            const int kc_elems = p.bf16 ? 2 * B2_KC : B2_KC;          // 128 bytes of a row per k-chunk either way
            for (uint32_t i0 = 0; i0 < ntiles; i0 += 32) {
                const uint32_t idx = i0 + (uint32_t) lane;
                uint32_t rmine = B2_DUMMY_ROW;
                if (idx < my_tiles) {
                    const uint32_t t = un.tile_begin + idx * un.tile_step;
                    rmine = p.tile_rows ? p.tile_rows[un.tl_off + t] : t * B2_M;
                }
                const uint32_t nt = min(32u, ntiles - i0);
                for (uint32_t j = 0; j < nt; j++) {
                    const int row0 = (int) __shfl_sync(0xffffffffu, rmine, (int) j);
                    if (lane == 0) {
                        for (int kc = 0; kc < p.kchunks; kc++) {
                            b2_mbar_wait(&empty[stage], phase ^ 1u);
                            unsigned char *st = smem + L.off_stage + (size_t) stage * B2_STAGE_BYTES;
                            if (leader) rk_mbar_arrive_expect_tx(&full[stage], tx);
                            const uint32_t lbar = b2_mapa(rk_smem_u32(&full[stage]), 0);
                            b2_tma_2d(st, &tmx, kc * kc_elems, row0, lbar);
                            b2_tma_2d(st + 16384, tq, kc * kc_elems, qrow, lbar);
                            if (++stage == S) { stage = 0; phase ^= 1u; }
                        }
                    }
                    __syncwarp();
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer (leader CTA) =====================
        if (leader) {
            int stage = 0;
            uint32_t phase = 0, tcount = 0;
            for (uint32_t ui = u_begin; ui < u_end; ui++) {
                const uint32_t pi = p.cta_units[ui];
                const RkBUnit ua = p.units[2 * pi], ub = p.units[2 * pi + 1];
                const uint32_t ncol = b2_ncol(ua.nq);
                // c=F32, a=b=TF32, K-major both, N>>3, M>>4 (M = 256 over the pair)
                // (bf16 nomination: a = b = BF16, kind::f16)
                const uint32_t fmt = p.bf16 ? 1u : 2u;
                const uint32_t idesc = (1u << 4) | (fmt << 7) | (fmt << 10) | ((ncol >> 3) << 17) | ((256u >> 4) << 24);
                const uint32_t ntiles = max(bs_unit_tiles(ua), bs_unit_tiles(ub));
                for (uint32_t ti = 0; ti < ntiles; ti++, tcount++) {
                    const uint32_t buf = tcount & 1u, tph = (tcount >> 1) & 1u;
                    b2_mbar_wait(&tempty[buf], tph ^ 1u);
                    bs_fence_after();
                    const uint32_t d_tmem = tmem_base + buf * B2_N;
                    for (int kc = 0; kc < p.kchunks; kc++) {
                        b2_mbar_wait(&full[stage], phase);
                        bs_fence_after();
                        if (lane == 0) {
                            const uint32_t a0 = rk_smem_u32(smem + L.off_stage + (size_t) stage * B2_STAGE_BYTES);
                            const uint64_t ad = bs_smem_desc(a0), bd = bs_smem_desc(a0 + 16384u);
                            // four MMAs per k-chunk, 32 bytes of K each: 8 TF32 or 16 bf16 elements (+2 in 16-byte units)
                            if (p.bf16) {
#pragma unroll
                                for (int k = 0; k < 4; k++)
                                    b2_mma_bf16(d_tmem, ad + (uint64_t) (2 * k), bd + (uint64_t) (2 * k), idesc, (kc | k) != 0);
                            } else {
#pragma unroll
                                for (int k = 0; k < 4; k++)
                                    b2_mma_tf32(d_tmem, ad + (uint64_t) (2 * k), bd + (uint64_t) (2 * k), idesc, (kc | k) != 0);
                            }
                            b2_commit(&empty[stage]);
                            if (kc == p.kchunks - 1) b2_commit(&tfull[buf]);
                        }
                        __syncwarp();
                        if (++stage == S) { stage = 0; phase ^= 1u; }
                    }
                }
            }
        }
    } else {
        // ===================== epilogue warps (each CTA: its own 128 rows) =====================
        const int ew = warp - 2;                       // 0..7
        const int quarter = warp & 3;                  // TMEM lanes this warp may read: [32 * (warp % 4), +32)
        const int chalf = ew >> 2;                     // which 16-column chunks of the tile: chalf, chalf + 2, ...
        const int erow = quarter * 32 + lane;
        const int et = ew * 32 + lane;                 // 0..255
        BsCtx c;
        c.tau = reinterpret_cast<uint32_t *>(smem + L.off_tau);
        c.cnt = reinterpret_cast<uint32_t *>(smem + L.off_cnt);
        c.hist = reinterpret_cast<uint32_t *>(smem + L.off_hist) + ew * 256;
        c.sel = reinterpret_cast<uint32_t *>(smem + L.off_sel) + ew * 4;
        c.scratch = p.scratch + (size_t) blockIdx.x * B2_N * BS_CAP;
        c.sortbuf = reinterpret_cast<uint64_t *>(smem + L.off_sort) + ew * BS_SORT_MAX;
        c.lane = lane;
        c.loose_extra = p.loose_extra;
        float *tauf = reinterpret_cast<float *>(smem + L.off_tauf);
        int32_t **s_trunc = reinterpret_cast<int32_t **>(smem + L.off_trunc);
        uint4 *s_qa = reinterpret_cast<uint4 *>(smem + L.off_qa);
        uint64_t *s_um = reinterpret_cast<uint64_t *>(smem + L.off_um);
        const int W = p.mask_words;
        auto esync = [] { rk_named_bar_sync(2, B2_EW * 32); };
        const uint32_t te0 = b2_mapa(rk_smem_u32(&tempty[0]), 0), te1 = b2_mapa(rk_smem_u32(&tempty[1]), 0);
        uint32_t tcount = 0;
        uint32_t qid = 0, qt_next = RK_KEY_SENTINEL;   // this thread's query (column et) and its shared threshold as last fetched
        for (uint32_t ui = u_begin; ui < u_end; ui++) {
            const uint32_t pi = p.cta_units[ui];
            const RkBUnit un = p.units[2 * pi + rank];
            const uint32_t my_tiles = bs_unit_tiles(un);
            const uint32_t ntiles = max(my_tiles, bs_unit_tiles(p.units[2 * pi + (rank ^ 1u)]));
            esync();                                   // previous unit fully emitted
            {
                const bool live = (uint32_t) et < un.nq;
                uint32_t tau0 = live ? RK_KEY_SENTINEL : 0u;      // a padded column never collects anything
                if (live && p.tau_key && p.tau_count[un.q0 + et] >= p.kp) tau0 = rk_float_to_key(p.tau_key[(size_t) (un.q0 + et) * p.kp + (p.kp - 1)]);
                qid = live ? (p.slot_q ? p.slot_q[un.q0 + et] : un.q0 + (uint32_t) et) : 0u;
                if (live && p.qtau) tau0 = min(tau0, __ldcg(p.qtau + qid));       // what the query's other units have proved so far (RkBScanParams::qtau)
                qt_next = tau0;
                c.tau[et] = tau0;
                tauf[et] = bs_tau_float(tau0);
                c.cnt[et] = 0;
                s_trunc[et] = p.trunc + qid;
                uint64_t w0 = ~0ull, w1 = 0ull;
                if (p.use_mask) {
                    w0 = live ? p.qmask[(size_t) (un.q0 + et) * W] : 0ull;
                    w1 = (live && W > 1) ? p.qmask[(size_t) (un.q0 + et) * W + 1] : 0ull;
                }
                s_qa[et] = make_uint4((uint32_t) w0, (uint32_t) (w0 >> 32), (uint32_t) w1, (uint32_t) (w1 >> 32));
                for (int w = 2; w < W; w++) s_um[et * W + w] = live ? p.qmask[(size_t) (un.q0 + et) * W + w] : 0ull;
            }
            esync();
            const uint32_t ncol = b2_ncol(un.nq);
            auto tile_row0 = [&](uint32_t i) -> uint32_t {
                const uint32_t t = un.tile_begin + i * un.tile_step;
                return i < my_tiles ? (p.tile_rows ? p.tile_rows[un.tl_off + t] : t * B2_M) : 0xFFFFFFF0u;
            };
            uint32_t r0_cur = tile_row0(0), r0_next = tile_row0(1);
            float xn_n = 0.f;
            uint64_t ma_n = 0, mb_n = 0;
            bool own_n = true;
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
                        ma_n = 1ull;
                    }
                }
            };
            fetch_row(r0_cur);
            for (uint32_t ti = 0; ti < ntiles; ti++, tcount++) {
                const uint32_t buf = tcount & 1u, tph = (tcount >> 1) & 1u;
                const uint64_t row = (uint64_t) r0_cur + (uint64_t) erow;
                if (p.qtau && (ti & 3u) == 3u && (uint32_t) et < un.nq) {
                    // the shared threshold fetched four tiles ago (no load latency on this path), then the next fetch
                    if (qt_next < c.tau[et]) { c.tau[et] = qt_next; tauf[et] = bs_tau_float(qt_next); }
                    qt_next = __ldcg(p.qtau + qid);
                }
                if (umask && !(own_n && ma_n == um0 && mb_n == um1)) { own_n = false; ma_n = 0; mb_n = 0; }
                const bool valid = row < p.nrows && own_n;
                const float xn = xn_n;
                // which of this warp's queries may see each row: the 32 rows of a warp belong to a few documents, so
                // the mask test runs once per (distinct row mask, query) — lanes test 16 queries at a time and vote —
                // instead of once per (row, query)
                uint64_t hit_lo = 0ull, hit_hi = 0ull;      // 16 bits per chunk of 16 queries
                {
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
                        for (int i = 0; i < 8; i++) {
                            const uint32_t c0 = 16u * (uint32_t) (chalf + 2 * i);
                            if (c0 >= ncol) break;
                            const uint4 qa = s_qa[c0 + (uint32_t) (lane & 15)];
                            const uint32_t hit = (a0 & qa.x) | (a1 & qa.y) | (b0 & qa.z) | (b1 & qa.w);
                            const uint32_t bal = __ballot_sync(0xffffffffu, hit != 0) & 0xFFFFu;
                            if (mine) { if (i < 4) hit_lo |= (uint64_t) bal << (16 * i); else hit_hi |= (uint64_t) bal << (16 * (i - 4)); }
                        }
                    }
                }
                r0_cur = r0_next;
                r0_next = tile_row0(ti + 2);
                fetch_row(r0_cur);
                b2_mbar_wait(&tfull[buf], tph);
                bs_fence_after();
                const uint32_t taddr = tmem_base + buf * B2_N + ((uint32_t) (quarter * 32) << 16);
#pragma unroll 1
                for (int i = 0; i < 8; i++) {
                    const uint32_t c0 = 16u * (uint32_t) (chalf + 2 * i);
                    if (c0 >= ncol) break;
                    const uint32_t h16 = (uint32_t) (i < 4 ? hit_lo >> (16 * i) : hit_hi >> (16 * (i - 4))) & 0xFFFFu;
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
                bs_fence_before();
                __syncwarp();
                if (lane == 0) b2_mbar_arrive_cluster(buf ? te1 : te0);
                // a tile adds at most 128 entries per query (only the four warps of one column half append
                // to it); the count read here may miss up to 96 of this tile's appends by the other three
                // — hence the margin of two tiles, as in rk_bscan_kernel
                const bool need = rk_named_bar_red_or(2, B2_EW * 32, c.cnt[et] > BS_CAP - 2 * B2_M);
                if (need) {
                    for (uint32_t q = (uint32_t) ew; q < un.nq; q += B2_EW) {
                        const uint32_t n = c.cnt[q];
                        if (n > p.round_mark) {
                            uint32_t tau;
                            uint32_t m = bs_warp_compact(c, (int) q, n, (uint32_t) p.kp, &tau, false);
                            if (p.qtau && lane == 0) atomicMin(p.qtau + (uint32_t) (s_trunc[q] - p.trunc), tau);     // >= kp rows are <= tau
                            if (m > BS_CAP / 2 && tau > 0) {
                                tau -= 1u;
                                m = bs_warp_filter(c, (int) q, m, tau);
                                if (lane == 0) *s_trunc[q] = 1;
                            }
                            if (lane == 0) { c.cnt[q] = m; c.tau[q] = tau; tauf[q] = bs_tau_float(tau); }
                        }
                    }
                    esync();
                }
            }
            for (uint32_t q = (uint32_t) ew; q < un.nq; q += B2_EW) {
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
    bs_fence_before();
    __syncthreads();
    b2_cluster_sync();          // no remote arrive / multicast commit may target a CTA that has left
    if (warp == 1) b2_tmem_dealloc(tmem_base, 2 * B2_N);
}

// ---------------------------------------------------------------------------------------------
// host launcher

size_t rk_bscan2_smem_bytes(int mask_words) { return b2_smem_layout(mask_words).total; }
size_t rk_bscan2_scratch_bytes(int grid) { return (size_t) grid * B2_N * BS_CAP * sizeof(uint2); }

cudaError_t rk_launch_bscan2(const RkBScanParams &p, const void *x, const void *q, uint64_t nq, int grid, cudaStream_t stream, int row_elems_b)
{
    CUtensorMap tmx, tmq[5];
    const bool bf = p.bf16 != 0;
    const int row_elems = bf ? row_elems_b : p.dim_pad;
    cudaError_t e = rk_make_tensor_map_2d_t(&tmx, x, p.nrows, row_elems, B2_M, bf);
    if (e != cudaSuccess) return e;
    for (int i = 0; i < 5; i++) {
        e = rk_make_tensor_map_2d_t(&tmq[i], q, nq, row_elems, 8u << i, bf);
        if (e != cudaSuccess) return e;
    }
    const size_t smem = rk_bscan2_smem_bytes(p.mask_words);
    auto kern = p.metric == 0 ? rk_bscan2_kernel<0> : rk_bscan2_kernel<1>;
    // (function attributes are per device: rk_multi runs one index per GPU in one process)
    static bool attr_set[64][2];
    int dev_ = 0;
    cudaGetDevice(&dev_);
    if (!attr_set[dev_ & 63][p.metric & 1]) {
        e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
        if (e != cudaSuccess) return e;
        attr_set[dev_ & 63][p.metric & 1] = true;
    }
    if (grid & 1) return cudaErrorInvalidConfiguration;
    kern<<<grid, B2_THREADS, smem, stream>>>(tmx, tmq[0], tmq[1], tmq[2], tmq[3], tmq[4], p);
    return cudaGetLastError();
}
