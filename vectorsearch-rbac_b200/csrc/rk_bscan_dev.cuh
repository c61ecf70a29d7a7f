// rk_bscan_dev.cuh — device helpers shared by the two tensor-core scoring kernels (rk_bscan.cu: one CTA
// per tile; rk_bscan2.cu: CTA pairs, tcgen05 cta_group::2): PTX wrappers, the per-query candidate
// buffers' radix-select compaction and the unit-end sort.
#pragma once
#include <cuda.h>

#include "rk_common.cuh"

#define BS_CAP 1024                    // entries of one query's candidate buffer
#define BS_SORT_MAX 512                // entries one warp sorts at unit end

// ---------------------------------------------------------------------------------------------
// PTX wrappers

__device__ __forceinline__ void bs_tma_2d(void *smem_dst, const CUtensorMap *tm, int c0, int c1, uint64_t *bar)
{
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
                     rk_smem_u32(smem_dst)),
                 "l"(tm), "r"(rk_smem_u32(bar)), "r"(c0), "r"(c1)
                 : "memory");
}

__device__ __forceinline__ void bs_tmem_alloc(uint32_t *smem_dst, uint32_t ncols)
{
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(rk_smem_u32(smem_dst)), "r"(ncols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}

__device__ __forceinline__ void bs_tmem_dealloc(uint32_t taddr, uint32_t ncols)
{
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}

__device__ __forceinline__ void bs_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void bs_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

__device__ __forceinline__ void bs_commit(uint64_t *bar)
{
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(rk_smem_u32(bar)) : "memory");
}

__device__ __forceinline__ void bs_mma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate)
{
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t"
        "}" ::"r"(tmem_d),
        "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}

__device__ __forceinline__ void bs_tmem_ld16(uint32_t taddr, uint32_t (&r)[16])
{
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
          "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr)
        : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// K-major, 128-byte swizzle, 8-row atoms 1024 bytes apart (what the 2-D TMA box above writes).
__device__ __forceinline__ uint64_t bs_smem_desc(uint32_t smem_addr)
{
    return (uint64_t) ((smem_addr >> 4) & 0x3FFFu) | (64ull << 32) | (1ull << 46) | (2ull << 61);
}


// A threshold key as the float the epilogue compares ordering values with: `v <= bs_tau_float(tau)` holds for every
// v with rk_float_to_key(v) <= tau (it may also admit +0.0 against a threshold of -0.0: an extra candidate, dropped by
// the next compaction).  The open threshold admits everything but NaN; key 0 (a padded column) admits nothing.
__device__ __forceinline__ float bs_tau_float(uint32_t tau)
{
    return tau == RK_KEY_SENTINEL ? __int_as_float(0x7f800000) : rk_key_to_float(tau);
}

struct BsCtx {
    uint32_t *tau, *cnt, *hist, *sel;
    uint2 *scratch;     // this CTA's [BS_N][BS_CAP]
    uint64_t *sortbuf;  // this warp's [BS_SORT_MAX]
    int lane;
    uint32_t loose_extra;
};

// kp-th smallest key of one query's buffer by a warp-level radix select (8 bits per pass), then in-place
// compaction key <= tau.  n <= BS_CAP.  Returns the new count; *tau_out receives the threshold.
// exact: all four passes, tau = the kp-th smallest key.  Otherwise the select stops after two or three passes as soon
// as the bucket holding the kp-th key is small (tau = the bucket's upper end: at most kp + 128 entries survive, and at
// most BS_CAP / 2) — a mid-unit compaction only has to make room and tighten the threshold, not hit kp exactly.
__device__ __forceinline__ uint32_t bs_warp_compact(const BsCtx &c, int q, uint32_t n, uint32_t kp, uint32_t *tau_out, bool exact = true)
{
    uint2 *buf = c.scratch + (size_t) q * BS_CAP;
    uint32_t key[BS_CAP / 32];
#pragma unroll
    for (int j = 0; j < BS_CAP / 32; j++) {
        const uint32_t i = (uint32_t) c.lane + 32u * j;
        key[j] = i < n ? __ldcg(&buf[i].x) : RK_KEY_SENTINEL;
    }
    uint32_t prefix = 0, mask = 0, kk = kp, below = 0, tau = 0;
    const uint32_t loose = min(kp + c.loose_extra, (uint32_t) BS_CAP / 2u);
#pragma unroll 1
    for (int shift = 24; shift >= 0; shift -= 8) {
#pragma unroll
        for (int b = 0; b < 8; b++) c.hist[c.lane * 8 + b] = 0;
        __syncwarp();
#pragma unroll
        for (int j = 0; j < BS_CAP / 32; j++) {
            const uint32_t i = (uint32_t) c.lane + 32u * j;
            if (i < n && (key[j] & mask) == prefix) atomicAdd(&c.hist[(key[j] >> shift) & 255u], 1u);
        }
        __syncwarp();
        uint32_t h[8], tot = 0;
#pragma unroll
        for (int b = 0; b < 8; b++) { h[b] = c.hist[c.lane * 8 + b]; tot += h[b]; }
        uint32_t inc = tot;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const uint32_t u = __shfl_up_sync(0xffffffffu, inc, o); if (c.lane >= o) inc += u; }
        uint32_t run = inc - tot;
#pragma unroll
        for (int b = 0; b < 8; b++) {
            if (run < kk && kk <= run + h[b]) { c.sel[0] = (uint32_t) (c.lane * 8 + b); c.sel[1] = kk - run; c.sel[2] = run; c.sel[3] = h[b]; }
            run += h[b];
        }
        __syncwarp();
        prefix |= c.sel[0] << shift;
        mask |= 0xFFu << shift;
        kk = c.sel[1];
        below += c.sel[2];                        // keys below the bucket that holds the kp-th
        const uint32_t in_bucket = c.sel[3];
        __syncwarp();
        tau = prefix | ((1u << shift) - 1u);      // the bucket's largest key
        if (!exact && shift <= 16 && below + in_bucket <= loose) break;
    }
    // compaction: rows of the survivors are read before anything is overwritten
    uint32_t row[BS_CAP / 32];
    uint32_t mine = 0;
#pragma unroll
    for (int j = 0; j < BS_CAP / 32; j++) {
        const uint32_t i = (uint32_t) c.lane + 32u * j;
        const bool keep = i < n && key[j] <= tau;
        row[j] = keep ? __ldcg(&buf[i].y) : 0u;
        mine += keep;
    }
    uint32_t inc = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const uint32_t u = __shfl_up_sync(0xffffffffu, inc, o); if (c.lane >= o) inc += u; }
    const uint32_t total = __shfl_sync(0xffffffffu, inc, 31);
    uint32_t pos = inc - mine;
    __syncwarp();
#pragma unroll
    for (int j = 0; j < BS_CAP / 32; j++) {
        const uint32_t i = (uint32_t) c.lane + 32u * j;
        if (i < n && key[j] <= tau) __stcg(&buf[pos++], make_uint2(key[j], row[j]));
    }
    __syncwarp();
    *tau_out = tau;
    return total;
}

// Keep the entries with key <= bound (in place).  Returns the new count.
__device__ __forceinline__ uint32_t bs_warp_filter(const BsCtx &c, int q, uint32_t n, uint32_t bound)
{
    uint2 *buf = c.scratch + (size_t) q * BS_CAP;
    uint2 e[BS_CAP / 32];
    uint32_t mine = 0;
#pragma unroll
    for (int j = 0; j < BS_CAP / 32; j++) {
        const uint32_t i = (uint32_t) c.lane + 32u * j;
        e[j] = i < n ? __ldcg(&buf[i]) : make_uint2(RK_KEY_SENTINEL, 0u);
        mine += i < n && e[j].x <= bound;
    }
    uint32_t inc = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const uint32_t u = __shfl_up_sync(0xffffffffu, inc, o); if (c.lane >= o) inc += u; }
    const uint32_t total = __shfl_sync(0xffffffffu, inc, 31);
    uint32_t pos = inc - mine;
    __syncwarp();
#pragma unroll
    for (int j = 0; j < BS_CAP / 32; j++) {
        const uint32_t i = (uint32_t) c.lane + 32u * j;
        if (i < n && e[j].x <= bound) __stcg(&buf[pos++], e[j]);
    }
    __syncwarp();
    return total;
}

// Sort the first n (<= BS_SORT_MAX) entries of a query's buffer by (key, row) in this warp's shared
// scratch and write the first min(n, kp) to the unit's candidate list.
__device__ __forceinline__ void bs_warp_emit(const BsCtx &c, int q, uint32_t n, uint32_t kp, uint32_t *out_key, uint32_t *out_row,
                                             int32_t *out_count)
{
    const uint2 *buf = c.scratch + (size_t) q * BS_CAP;
    uint32_t P = 32;
    while (P < n) P <<= 1;
    for (uint32_t i = c.lane; i < P; i += 32) {
        uint64_t v = ~0ull;
        if (i < n) { const uint2 e = __ldcg(&buf[i]); v = ((uint64_t) e.x << 32) | e.y; }
        c.sortbuf[i] = v;
    }
    __syncwarp();
    for (uint32_t size = 2; size <= P; size <<= 1) {
        for (uint32_t stride = size >> 1; stride > 0; stride >>= 1) {
            for (uint32_t i = c.lane; i < (P >> 1); i += 32) {
                const uint32_t lo = 2 * i - (i & (stride - 1)), hi = lo + stride;
                const bool up = (lo & size) == 0;
                const uint64_t a = c.sortbuf[lo], b = c.sortbuf[hi];
                if (up ? (b < a) : (a < b)) { c.sortbuf[lo] = b; c.sortbuf[hi] = a; }
            }
            __syncwarp();
        }
    }
    const uint32_t m = n < kp ? n : kp;
    for (uint32_t i = c.lane; i < m; i += 32) {
        const uint64_t v = c.sortbuf[i];
        out_key[i] = (uint32_t) (v >> 32);
        out_row[i] = (uint32_t) v;
    }
    if (c.lane == 0) *out_count = (int32_t) m;
    __syncwarp();
}

// The first n <= kp entries of a query's buffer as the unit's candidate list, UNSORTED except that the largest
// (key, row) goes last — all the per-query merge (rk_select_kernel, `unsorted`) needs: a full list's last key bounds
// the kp-th best from above.  Keeps the sort out of the scoring kernel's tail.
__device__ __forceinline__ void bs_warp_emit_unsorted(const BsCtx &c, int q, uint32_t n, uint32_t *out_key, uint32_t *out_row,
                                                      int32_t *out_count)
{
    const uint2 *buf = c.scratch + (size_t) q * BS_CAP;
    uint64_t best = 0;
    uint32_t besti = 0;
    // eight of a lane's entries are fetched before any is used (one memory round trip per 256 entries; n <= kp <= 512)
#pragma unroll 1
    for (uint32_t b0 = 0; b0 < n; b0 += 256) {
        uint2 e[8];
#pragma unroll
        for (int j = 0; j < 8; j++) {
            const uint32_t i = b0 + (uint32_t) c.lane + 32u * j;
            e[j] = i < n ? __ldcg(&buf[i]) : make_uint2(0u, 0u);
        }
#pragma unroll
        for (int j = 0; j < 8; j++) {
            const uint32_t i = b0 + (uint32_t) c.lane + 32u * j;
            if (i < n) {
                const uint64_t v = ((uint64_t) e[j].x << 32) | e[j].y;
                if (v >= best) { best = v; besti = i; }
                out_key[i] = e[j].x;
                out_row[i] = e[j].y;
            }
        }
    }
    uint64_t wbest = best;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { const uint64_t u = __shfl_xor_sync(0xffffffffu, wbest, o); wbest = u > wbest ? u : wbest; }
    __syncwarp();
    if (n > 1) {
        // (key, row) pairs of one list are distinct: exactly one lane owns the largest
        const bool owner = best == wbest && (uint32_t) c.lane < n;
        const uint32_t oi = __shfl_sync(0xffffffffu, besti, __ffs((int) __ballot_sync(0xffffffffu, owner)) - 1);
        if (c.lane == 0 && oi != n - 1) {
            const uint2 last = __ldcg(&buf[n - 1]);
            out_key[oi] = last.x; out_row[oi] = last.y;
            out_key[n - 1] = (uint32_t) (wbest >> 32); out_row[n - 1] = (uint32_t) wbest;
        }
    }
    if (c.lane == 0) *out_count = (int32_t) n;
    __syncwarp();
}

// Tiles a unit visits: tile_begin, tile_begin + step, ... below tile_end.
__device__ __forceinline__ uint32_t bs_unit_tiles(const RkBUnit &un)
{
    return un.tile_end > un.tile_begin ? (un.tile_end - un.tile_begin + un.tile_step - 1u) / un.tile_step : 0u;
}

// 2-D fp32 tensor map [rows][dim_pad], box = 32 floats x box_rows, 128-byte swizzle, zero fill out of bounds (rk_bscan.cu)
cudaError_t rk_make_tensor_map_2d(CUtensorMap *tm, const float *base, uint64_t rows, int dim_pad, uint32_t box_rows);
cudaError_t rk_make_tensor_map_2d_t(CUtensorMap *tm, const void *base, uint64_t rows, int row_elems, uint32_t box_rows, bool bf16);
