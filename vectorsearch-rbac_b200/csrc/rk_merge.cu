// rk_merge.cu — per-query k-way select over the candidate lists the scan kernel (or an all-gather
// across GPUs) produced.
//
// Stands in for the harness merge: "sort by distance, drop duplicates on (document_id, block_id),
// stop at k" (controller/baseline/prefilter/prefilter_role.py:174-189;
// controller/dynamic_partition/search.py:347-364) and for the final `[:topk]` of the ground truth
// (basic_benchmark/common_function.py:713-714, 1114-1124).  Also converts the fp32 sum to what the
// SQL operator returns: sqrt((double) sum) for `<->` (pgvector/src/vector.c:577), -(double) sum for
// `<#>` (vector.c:635).
#include "rk_common.cuh"
#include "rk_select.cuh"
#include "rk_kernels.h"

#define RK_SEL_THREADS 512
#define RK_SEL_CHUNK 1024
#define RK_SEL_LISTS 2048       // non-empty list ids staged per round

__global__ void __launch_bounds__(RK_SEL_THREADS) rk_select_kernel(const RkSelectParams p)
{
    __shared__ __align__(16) uint64_t s_tid[RK_CAND_CAP];
    __shared__ uint32_t s_key[RK_CAND_CAP];
    __shared__ uint32_t s_row[RK_CAND_CAP];
    __shared__ uint32_t s_count, s_tau, s_scan[RK_SEL_THREADS / 32];

    const int q = blockIdx.x;
    const int t = threadIdx.x;
    RkCandBuf cb{s_key, s_row, s_tid};
    auto sync = [] { __syncthreads(); };

    uint32_t l0, l1;
    const bool by_id = p.lid0 != nullptr || p.list_ids != nullptr;
    const uint32_t id0 = p.lid0 ? p.lid0[q] : 0u, idstep = p.lid0 ? p.lstep[q] : 0u;
    const uint32_t *ids = p.list_ids ? p.list_ids + p.qseg[q] : nullptr;       // explicit list ids of this query
    if (p.list_ids) { l0 = 0; l1 = p.qseg[q + 1] - p.qseg[q]; }
    else if (by_id) { l0 = 0; l1 = p.nlists[q]; }
    else if (p.lists_per_query > 0) { l0 = 0; l1 = (uint32_t) p.lists_per_query; }
    else { l0 = p.qseg[q]; l1 = p.qseg[q + 1]; }
    const uint32_t nl = l1 - l0;
    const int k = p.k;

    // In fixed-list mode list j of query q lives at shard j's block + q * k; blocks are dense
    // ([G][nq][k]) or shard_stride_bytes apart (packed all-gather buffer).  b4 / b8 index 4- and
    // 8-byte arrays respectively.
    const size_t dense = (size_t) gridDim.x * (size_t) p.list_stride;
    const size_t st4 = p.shard_stride_bytes ? p.shard_stride_bytes / 4 : dense;
    const size_t st8 = p.shard_stride_bytes ? p.shard_stride_bytes / 8 : dense;
    auto b4 = [&](uint32_t j) -> size_t {
        if (by_id) return (size_t) (ids ? ids[j] : id0 + j * idstep) * (size_t) p.list_stride;
        return p.lists_per_query > 0 ? (size_t) j * st4 + (size_t) q * p.list_stride : (size_t) (l0 + j) * (size_t) p.list_stride;
    };
    auto b8 = [&](uint32_t j) -> size_t {
        if (by_id) return (size_t) (ids ? ids[j] : id0 + j * idstep) * (size_t) p.list_stride;
        return p.lists_per_query > 0 ? (size_t) j * st8 + (size_t) q * p.list_stride : (size_t) (l0 + j) * (size_t) p.list_stride;
    };
    const uint32_t cstride = p.count_stride ? (uint32_t) p.count_stride : gridDim.x;
    auto list_count = [&](uint32_t j) -> int {
        if (by_id) return p.cand_count[ids ? ids[j] : id0 + j * idstep];
        return p.cand_count ? p.cand_count[p.lists_per_query > 0 ? j * cstride + q : l0 + j] : k;
    };
    // lists without tuple ids (tensor-engine nominations) order ties by row
    auto load_tid = [&](size_t a8, size_t a4) -> uint64_t { return p.cand_tid ? p.cand_tid[a8] : (uint64_t) p.cand_row[a4]; };

    auto load_key = [&](size_t a) -> uint32_t {
        const uint32_t raw = p.cand_key[a];
        return p.key_is_float ? rk_float_to_key(__uint_as_float(raw)) : raw;
    };
    if (t == 0) { s_count = 0; s_tau = RK_KEY_SENTINEL; }
    if (p.wait_n > 0) {
        // lists pushed by the peers: wait for every peer's arrival flag of this step
        if (t < p.wait_n && t != p.wait_self && !rk_wait_flag_sys(p.wait_flag + t, p.wait_seq)) *p.xerr = 1u;
    }
    __syncthreads();
    // upper bound of the k-th best: any list holding k entries proves tau <= its last key
    for (uint32_t j = t; j < nl; j += RK_SEL_THREADS) {
        if (list_count(j) >= k) atomicMin(&s_tau, load_key(b4(j) + k - 1));
    }
    __syncthreads();

    // Fast path: every list is sorted ascending, so the entries with key <= tau form a prefix of
    // each list.  Walk the prefixes and gather them; only if they do not fit the buffer (more than
    // RK_CAND_CAP entries tie below tau) fall back to the chunked scan with compactions.
    __shared__ uint32_t s_overflow, s_nlists;
    __shared__ uint32_t s_hist[256], s_sel[2];
    __shared__ uint32_t s_lists[RK_SEL_LISTS];
    if (t == 0) s_overflow = 0;
    __syncthreads();
    // One warp per list, 32 consecutive entries per step (one coalesced load instead of a chain of dependent ones).
    // Unsorted lists (tensor-engine nominations: only the last entry is known to be the largest) are read whole.
    {
        const uint32_t tau = s_tau;
        const int warp = t >> 5, lane = t & 31;
        // non-empty lists first (all counts fetched in one parallel round): a single query cut into a thousand
        // segments leaves most lists empty once CTAs carry their buffer across segments
        for (uint32_t lbase = 0; lbase < nl; lbase += RK_SEL_LISTS) {
        __syncthreads();
        if (t == 0) s_nlists = 0;
        __syncthreads();
        for (uint32_t j = lbase + (uint32_t) t; j < nl && j < lbase + RK_SEL_LISTS; j += RK_SEL_THREADS)
            if (list_count(j) > 0) s_lists[atomicAdd(&s_nlists, 1u)] = j;
        __syncthreads();
        const uint32_t nlive = s_nlists;
        for (uint32_t li = (uint32_t) warp; li < nlive; li += RK_SEL_THREADS / 32) {
            const uint32_t j = s_lists[li];
            const int cnt = list_count(j);
            const size_t a4 = b4(j), a8 = b8(j);
            // keys (and rows) of up to eight 32-entry steps are fetched together: one memory round trip per eight steps
            for (int b0 = 0; b0 < cnt; b0 += 256) {
                uint32_t kreg[8], rreg[8];
#pragma unroll
                for (int s8 = 0; s8 < 8; s8++) {
                    const int i = b0 + s8 * 32 + lane;
                    kreg[s8] = i < cnt ? load_key(a4 + i) : RK_KEY_SENTINEL;
                    rreg[s8] = i < cnt ? p.cand_row[a4 + i] : 0u;
                }
                bool stop = false;
#pragma unroll
                for (int s8 = 0; s8 < 8; s8++) {
                    const int i = b0 + s8 * 32 + lane;
                    if (!stop && b0 + s8 * 32 < cnt) {          // warp-uniform
                        bool take = false;
                        const uint32_t key = kreg[s8];
                        uint64_t tid = 0;
                        if (i < cnt && key <= tau && key != RK_KEY_SENTINEL) {
                            tid = p.cand_tid ? p.cand_tid[a8 + i] : (uint64_t) rreg[s8];
                            take = !(tid == 0xFFFFFFFFFFFFFFFFull && key == 0xFF800000u);   // exchange-form padding
                        }
                        const uint32_t bal = __ballot_sync(0xffffffffu, take);
                        if (bal) {
                            uint32_t base = 0;
                            if (lane == 0) base = atomicAdd(&s_count, (uint32_t) __popc(bal));
                            base = __shfl_sync(0xffffffffu, base, 0);
                            if (take) {
                                const uint32_t slot = base + (uint32_t) __popc(bal & ((1u << lane) - 1u));
                                if (slot >= RK_CAND_CAP) s_overflow = 1;
                                else { s_key[slot] = key; s_tid[slot] = tid; s_row[slot] = rreg[s8]; }
                            }
                        }
                        // sorted list: the first entry that fails ends the prefix
                        if (!p.unsorted && __any_sync(0xffffffffu, i < cnt && !take)) stop = true;
                    }
                }
                if (stop) break;
            }
        }
        }
    }
    __syncthreads();
    const bool overflow = s_overflow != 0;
    __syncthreads();
    if (overflow && t == 0) s_count = 0;
    __syncthreads();

    // keeps the entries with key <= (k-th smallest key) at the front of the buffer; returns how many
    auto sel_reduce = [&](int n) -> int {
        uint32_t prefix = 0, mask = 0, kk = (uint32_t) k;
#pragma unroll 1
        for (int shift = 24; shift >= 0; shift -= 8) {
            if (t < 256) s_hist[t] = 0;
            __syncthreads();
            for (int i = t; i < n; i += RK_SEL_THREADS) {
                const uint32_t kv = s_key[i];
                if ((kv & mask) == prefix) atomicAdd(&s_hist[(kv >> shift) & 255u], 1u);
            }
            __syncthreads();
            if (t < 32) {
                uint32_t h[8], tot = 0;
#pragma unroll
                for (int b = 0; b < 8; b++) { h[b] = s_hist[t * 8 + b]; tot += h[b]; }
                uint32_t inc = tot;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) { const uint32_t u = __shfl_up_sync(0xffffffffu, inc, o); if (t >= o) inc += u; }
                uint32_t run = inc - tot;
#pragma unroll
                for (int b = 0; b < 8; b++) {
                    if (run < kk && kk <= run + h[b]) { s_sel[0] = (uint32_t) (t * 8 + b); s_sel[1] = kk - run; }
                    run += h[b];
                }
            }
            __syncthreads();
            prefix |= s_sel[0] << shift;
            mask |= 0xFFu << shift;
            kk = s_sel[1];
        }
        const uint32_t kth = prefix;
        constexpr int PER = RK_CAND_CAP / RK_SEL_THREADS;      // 4
        uint32_t mk[PER], mr[PER];
        uint64_t mt[PER];
        int mc = 0;
#pragma unroll
        for (int i = 0; i < PER; i++) {
            const int idx = t * PER + i;
            if (idx < n && s_key[idx] <= kth) { mk[mc] = s_key[idx]; mr[mc] = s_row[idx]; mt[mc] = s_tid[idx]; mc++; }
        }
        int inc = mc;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int v = __shfl_up_sync(0xffffffffu, inc, o); if ((t & 31) >= o) inc += v; }
        if ((t & 31) == 31) s_scan[t >> 5] = (uint32_t) inc;
        __syncthreads();                      // also orders the reads above before the writes below
        int woff = 0, tot = 0;
        for (int w = 0; w < RK_SEL_THREADS / 32; w++) { if (w < (t >> 5)) woff += (int) s_scan[w]; tot += (int) s_scan[w]; }
        const int pos = woff + inc - mc;
#pragma unroll
        for (int i = 0; i < PER; i++)
            if (i < mc) { s_key[pos + i] = mk[i]; s_row[pos + i] = mr[i]; s_tid[pos + i] = mt[i]; }
        __syncthreads();
        return tot;
    };

    bool need_chunks = overflow;
    // With one staging round (nl <= RK_SEL_LISTS) s_lists still names the non-empty lists: a single query cut into many
    // segments leaves most of them empty (CTAs carry their buffer across segments), and only the live ones count here.
    const bool live_only = nl <= RK_SEL_LISTS;
    const uint32_t nsrc = live_only ? s_nlists : nl;
    if (overflow && !p.dedup && (size_t) nsrc * (size_t) k <= 32u * RK_SEL_THREADS) {
        // More than RK_CAND_CAP entries lie below the bound the full lists gave (many unsorted lists of similar
        // quality: the tensor engine's unit lists).  Instead of sorting them chunk by chunk, find the k-th smallest key
        // over ALL the lists exactly: every thread takes up to 32 entries into registers (all loads independent: a couple
        // of memory round trips), a 4-pass radix select runs over the registers, and only the entries up to the k-th key
        // are gathered and sorted.
        const uint32_t flat = nsrc * (uint32_t) k;
        uint32_t kr[32];
#pragma unroll
        for (int r = 0; r < 32; r++) {
            const uint32_t e = (uint32_t) t + (uint32_t) RK_SEL_THREADS * r;
            uint32_t kv = RK_KEY_SENTINEL;
            if (e < flat) {
                const uint32_t jj = e / (uint32_t) k, i = e - jj * (uint32_t) k;
                const uint32_t j = live_only ? s_lists[jj] : jj;
                if ((int) i < list_count(j)) {
                    kv = load_key(b4(j) + i);
                    if (kv == 0xFF800000u && load_tid(b8(j) + i, b4(j) + i) == 0xFFFFFFFFFFFFFFFFull) kv = RK_KEY_SENTINEL;   // padding
                }
            }
            kr[r] = kv;
        }
        uint32_t prefix = 0, mask = 0, kk = (uint32_t) k;
#pragma unroll 1
        for (int shift = 24; shift >= 0; shift -= 8) {
            if (t < 256) s_hist[t] = 0;
            __syncthreads();
#pragma unroll
            for (int r = 0; r < 32; r++)
                if (kr[r] != RK_KEY_SENTINEL && (kr[r] & mask) == prefix) atomicAdd(&s_hist[(kr[r] >> shift) & 255u], 1u);
            __syncthreads();
            if (t < 32) {
                uint32_t h[8], tot = 0;
#pragma unroll
                for (int b = 0; b < 8; b++) { h[b] = s_hist[t * 8 + b]; tot += h[b]; }
                uint32_t inc = tot;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) { const uint32_t u = __shfl_up_sync(0xffffffffu, inc, o); if (t >= o) inc += u; }
                uint32_t run = inc - tot;
#pragma unroll
                for (int b = 0; b < 8; b++) {
                    if (run < kk && kk <= run + h[b]) { s_sel[0] = (uint32_t) (t * 8 + b); s_sel[1] = kk - run; }
                    run += h[b];
                }
            }
            __syncthreads();
            prefix |= s_sel[0] << shift;
            mask |= 0xFFu << shift;
            kk = s_sel[1];
        }
        const uint32_t kth = prefix;
        if (t == 0) { s_overflow = 0; if (kth < s_tau) s_tau = kth; }
        __syncthreads();
#pragma unroll
        for (int r = 0; r < 32; r++) {
            if (kr[r] != RK_KEY_SENTINEL && kr[r] <= kth) {
                const uint32_t e = (uint32_t) t + (uint32_t) RK_SEL_THREADS * r;
                const uint32_t jj = e / (uint32_t) k, i = e - jj * (uint32_t) k;
                const uint32_t j = live_only ? s_lists[jj] : jj;
                const uint32_t slot = atomicAdd(&s_count, 1u);
                if (slot >= RK_CAND_CAP) s_overflow = 1;       // more than RK_CAND_CAP ties at the k-th key
                else { s_key[slot] = kr[r]; s_tid[slot] = load_tid(b8(j) + i, b4(j) + i); s_row[slot] = p.cand_row[b4(j) + i]; }
            }
        }
        __syncthreads();
        need_chunks = s_overflow != 0;
        __syncthreads();
        if (need_chunks && t == 0) s_count = 0;      // the chunked pass below redoes it with the tight bound
        __syncthreads();
    }

    const size_t total = need_chunks ? (size_t) nl * k : 0;
    for (size_t c0 = 0; c0 < total || c0 == 0; c0 += RK_SEL_CHUNK) {
        const uint32_t tau = s_tau;
        for (size_t idx = c0 + t; idx < total && idx < c0 + RK_SEL_CHUNK; idx += RK_SEL_THREADS) {
            const uint32_t j = (uint32_t) (idx / k);
            const int i = (int) (idx % k);
            if (i < list_count(j)) {
                const size_t a = b4(j) + i;
                const uint32_t key = load_key(a);
                const uint64_t tid = load_tid(b8(j) + i, a);
                // padding: sentinel key, or (+Inf, UINT64_MAX) in exchange form
                const bool pad = key == RK_KEY_SENTINEL || (tid == 0xFFFFFFFFFFFFFFFFull && key == 0xFF800000u);
                if (key <= tau && !pad) {
                    const uint32_t slot = atomicAdd(&s_count, 1u);
                    s_key[slot] = key; s_tid[slot] = tid; s_row[slot] = p.cand_row[a];
                }
            }
        }
        __syncthreads();
        int n = (int) s_count;
        const bool lastc = c0 + RK_SEL_CHUNK >= total;
        if (lastc || n > RK_CAND_CAP - RK_SEL_CHUNK) {
            // Without dedup only the k smallest keys can reach the output: cut the buffer down to them (exact 4-pass radix
            // select of the k-th smallest key, ties kept) before sorting, so the sort runs over ~k entries instead of up to
            // RK_CAND_CAP.  (With dedup a duplicate may sit among the k smallest, so everything gathered is sorted.)
            const bool setonly = p.out_maxlast && !p.dedup && lastc;
            if (!p.dedup && ((n > 2 * k && n > 64) || (setonly && n > k))) n = sel_reduce(n);
            if (setonly && n <= k) {
                // the k smallest are all there is: no order needed, only the largest (key, tid) moved to the end
                uint32_t bk = 0; uint64_t bt = 0; int bi = -1;
                for (int i = t; i < n; i += RK_SEL_THREADS)
                    if (bi < 0 || rk_cand_less(bk, bt, s_key[i], s_tid[i])) { bk = s_key[i]; bt = s_tid[i]; bi = i; }
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) {
                    const uint32_t ok = __shfl_xor_sync(0xffffffffu, bk, o);
                    const uint64_t ot = __shfl_xor_sync(0xffffffffu, bt, o);
                    const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
                    if (oi >= 0 && (bi < 0 || rk_cand_less(bk, bt, ok, ot))) { bk = ok; bt = ot; bi = oi; }
                }
                __syncthreads();                       // s_hist is free again
                if ((t & 31) == 0) {
                    s_hist[(t >> 5) * 4 + 0] = bk; s_hist[(t >> 5) * 4 + 1] = (uint32_t) bi;
                    s_hist[(t >> 5) * 4 + 2] = (uint32_t) bt; s_hist[(t >> 5) * 4 + 3] = (uint32_t) (bt >> 32);
                }
                __syncthreads();
                if (t == 0 && n > 1) {
                    for (int w = 0; w < RK_SEL_THREADS / 32; w++) {
                        const uint32_t ok = s_hist[w * 4]; const int oi = (int) s_hist[w * 4 + 1];
                        const uint64_t ot = (uint64_t) s_hist[w * 4 + 2] | ((uint64_t) s_hist[w * 4 + 3] << 32);
                        if (oi >= 0 && (bi < 0 || rk_cand_less(bk, bt, ok, ot))) { bk = ok; bt = ot; bi = oi; }
                    }
                    if (bi >= 0 && bi != n - 1) {
                        const uint32_t lk = s_key[n - 1], lr = s_row[n - 1]; const uint64_t lt = s_tid[n - 1];
                        s_key[n - 1] = s_key[bi]; s_row[n - 1] = s_row[bi]; s_tid[n - 1] = s_tid[bi];
                        s_key[bi] = lk; s_row[bi] = lr; s_tid[bi] = lt;
                    }
                }
                __syncthreads();
            } else {
                rk_bitonic_sort(cb, n, t, RK_SEL_THREADS, sync);
            }
            int m = n;
            if (p.dedup && n > 1) {
                // drop entries equal to their predecessor (same tuple reached through two partitions)
                // stable compaction with a block-wide exclusive scan over keep flags
                const int per = (n + RK_SEL_THREADS - 1) / RK_SEL_THREADS;
                const int b0 = t * per, b1 = min(n, b0 + per);
                int cnt = 0;
                for (int i = b0; i < b1; i++) cnt += !(i > 0 && s_tid[i] == s_tid[i - 1] && s_key[i] == s_key[i - 1]);
                int inc = cnt;
                for (int o = 1; o < 32; o <<= 1) { int v = __shfl_up_sync(0xffffffffu, inc, o); if ((t & 31) >= o) inc += v; }
                if ((t & 31) == 31) s_scan[t >> 5] = (uint32_t) inc;
                __syncthreads();
                int woff = 0, tot = 0;
                for (int w = 0; w < RK_SEL_THREADS / 32; w++) { if (w < (t >> 5)) woff += (int) s_scan[w]; tot += (int) s_scan[w]; }
                int pos = woff + inc - cnt;
                // read my block into registers is unbounded; instead write through a second pass:
                // entries only move left (pos <= i), and blocks are processed in index order per thread,
                // but different threads may race, so stage keep flags first, then move after a barrier.
                uint64_t mt[8]; uint32_t mk[8], mr[8]; int mp[8]; int mc = 0;   // per <= 4 since n <= 2048, 512 threads
                for (int i = b0; i < b1; i++) {
                    bool keepf = !(i > 0 && s_tid[i] == s_tid[i - 1] && s_key[i] == s_key[i - 1]);
                    if (keepf) { mt[mc] = s_tid[i]; mk[mc] = s_key[i]; mr[mc] = s_row[i]; mp[mc] = pos++; mc++; }
                }
                __syncthreads();
                for (int i = 0; i < mc; i++) { s_tid[mp[i]] = mt[i]; s_key[mp[i]] = mk[i]; s_row[mp[i]] = mr[i]; }
                __syncthreads();
                m = tot;
            }
            const int keep = m < k ? m : k;
            if (t == 0) {
                s_count = (uint32_t) keep;
                if (m >= k) { uint32_t kk = s_key[k - 1]; if (kk < s_tau) s_tau = kk; }
            }
            __syncthreads();
        }
    }

    const int cnt = (int) s_count;
    const size_t ob = (size_t) q * k;
    for (int i = t; i < k; i += RK_SEL_THREADS) {
        const bool ok = i < cnt;
        const uint32_t key = ok ? s_key[i] : RK_KEY_SENTINEL;
        const float ord = ok ? rk_key_to_float(key) : __int_as_float(0x7f800000);
        if (p.out_tid) p.out_tid[ob + i] = ok ? s_tid[i] : 0xFFFFFFFFFFFFFFFFull;
        if (p.out_row) p.out_row[ob + i] = ok ? s_row[i] : 0xFFFFFFFFu;
        if (p.out_key) p.out_key[ob + i] = ord;
        if (p.out_doc) {
            uint32_t d = 0xFFFFFFFFu;
            if (ok) d = p.rowdoc ? p.doc_ids[p.rowdoc[s_row[i]]] : s_row[i];
            p.out_doc[ob + i] = d;
        }
        if (p.out_dist) {
            double v;
            if (!ok) v = __longlong_as_double(0x7ff0000000000000ll);
            else if (p.null_query) v = 0.0;
            else if (p.metric == 0) v = sqrt((double) ord);     // vector.c:577
            else v = (double) ord;                              // ord already holds -sum (vector.c:635)
            p.out_dist[ob + i] = v;
        }
    }
    if (t == 0 && p.out_count) p.out_count[q] = cnt;
    if (p.ack_n > 0) {
        // every block has read its lists: tell the peers that this ring slot may be overwritten
        __syncthreads();
        if (t == 0) {
            __threadfence();
            if (atomicAdd(p.ack_done, 1u) == gridDim.x - 1) {
                *p.ack_done = 0;
                __threadfence_system();
                for (int g = 0; g < p.ack_n; g++) rk_st_release_sys(p.ack_flag[g], p.ack_seq);
            }
        }
    }
}

cudaError_t rk_launch_select(const RkSelectParams &p, int nq, cudaStream_t stream)
{
    if (nq <= 0) return cudaSuccess;
    rk_select_kernel<<<nq, RK_SEL_THREADS, 0, stream>>>(p);
    return cudaGetLastError();
}
