// rk_select.cuh — shared-memory candidate buffer + bitonic top-k used by the scan and select kernels.
//
// Stands in for PostgreSQL's bounded top-N heapsort under `ORDER BY distance LIMIT k`
// (driven by controller/baseline/prefilter/prefilter_role.py:162-172) and for pgvector's
// tuplesort of probed tuples (pgvector/src/ivfscan.c:155-171).  Order: (key, tid) ascending, where
// key is the order-preserving uint image of the fp32 ordering value.
#pragma once
#include "rk_common.cuh"

#define RK_CAND_CAP 2048

struct RkCandBuf {
    uint32_t *key;   // [RK_CAND_CAP]
    uint32_t *row;   // [RK_CAND_CAP]
    uint64_t *tid;   // [RK_CAND_CAP]
};

__device__ __forceinline__ bool rk_cand_less(uint32_t ka, uint64_t ta, uint32_t kb, uint64_t tb)
{
    return (ka < kb) || (ka == kb && ta < tb);
}

// Sort entries [0, n) ascending by (key, tid) with a block-cooperative bitonic network.
// `t` = thread index within the cooperating group of `nt` threads; Sync() is the group's barrier.
// n <= RK_CAND_CAP.  On return entries [0, n) are sorted; [n, P) hold sentinels.
template <typename Sync>
__device__ __forceinline__ void rk_bitonic_sort(RkCandBuf b, int n, int t, int nt, Sync sync)
{
    int P = 32;
    while (P < n) P <<= 1;
    for (int i = n + t; i < P; i += nt) {
        b.key[i] = RK_KEY_SENTINEL;
        b.tid[i] = 0xFFFFFFFFFFFFFFFFull;
        b.row[i] = 0xFFFFFFFFu;
    }
    sync();
    for (int size = 2; size <= P; size <<= 1) {
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
            for (int i = t; i < (P >> 1); i += nt) {
                int lo = 2 * i - (i & (stride - 1));
                int hi = lo + stride;
                bool up = (lo & size) == 0;
                uint32_t ka = b.key[lo], kb = b.key[hi];
                uint64_t ta = b.tid[lo], tb = b.tid[hi];
                bool swap = up ? rk_cand_less(kb, tb, ka, ta) : rk_cand_less(ka, ta, kb, tb);
                if (swap) {
                    b.key[lo] = kb; b.key[hi] = ka;
                    b.tid[lo] = tb; b.tid[hi] = ta;
                    uint32_t ra = b.row[lo], rb = b.row[hi];
                    b.row[lo] = rb; b.row[hi] = ra;
                }
            }
            sync();
        }
    }
}
