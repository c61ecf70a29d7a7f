// rk_ivf_kernels.cu — IVFFlat device code: nearest-centre assignment, k-means updates, list probe
// selection and the on-device work-list builder.
//
// Reference semantics (pgvector, paths relative to the reference checkout):
//   assignment ...... pgvector/src/ivfbuild.c:141-198  nearest centre by opclass FUNCTION 1, first minimum wins
//   k-means ......... pgvector/src/ivfkmeans.c:21-89, 258-494 (Lloyd iterations here; centres are not
//                     reproducible across implementations — pg_prng state — so parity is recall-level)
//   list choice ..... pgvector/src/ivfscan.c:36-107    distance to ALL centres, `probes` nearest, nearest first
//   list scan ....... pgvector/src/ivfscan.c:112-176   every tuple of the probed lists scored (rk_scan_kernel)
#include "rk_common.cuh"
#include "rk_kernels.h"

// ---------------------------------------------------------------------------------------------
// assignment: out_list[r] = argmin_c value(x_r, c), value = L2 squared sum or -inner product; ties
// go to the lowest centre index.  64 rows x 64 centres per block step, 4x4 micro-tile per thread.

#define AS_BM 64
#define AS_BN 64
#define AS_BK 16

template <int METRIC>
__global__ void __launch_bounds__(256) rk_assign_kernel(const float *__restrict__ X, size_t nrows, const float *__restrict__ Cn,
                                                        int lists, int dim_pad, uint32_t *__restrict__ out_list,
                                                        float *__restrict__ out_val)
{
    __shared__ float Xs[AS_BK][AS_BM + 4];
    __shared__ float Cs[AS_BK][AS_BN + 4];
    const int t = threadIdx.x;
    const int tx = t & 15, ty = t >> 4;
    const size_t row0 = (size_t) blockIdx.x * AS_BM;
    const int lr = t >> 2, lc = (t & 3) * 4;          // loader mapping: row lr, columns lc..lc+3 of the k-chunk

    float best[4];
    uint32_t besti[4];
#pragma unroll
    for (int i = 0; i < 4; i++) { best[i] = __int_as_float(0x7f800000); besti[i] = 0; }

    for (int n0 = 0; n0 < lists; n0 += AS_BN) {
        float acc[4][4];
#pragma unroll
        for (int i = 0; i < 4; i++)
#pragma unroll
            for (int j = 0; j < 4; j++) acc[i][j] = 0.f;
        for (int k0 = 0; k0 < dim_pad; k0 += AS_BK) {
            float4 xv = make_float4(0.f, 0.f, 0.f, 0.f), cv = xv;
            if (k0 + lc < dim_pad) {
                if (row0 + lr < nrows) xv = *reinterpret_cast<const float4 *>(X + (row0 + lr) * (size_t) dim_pad + k0 + lc);
                if (n0 + lr < lists) cv = *reinterpret_cast<const float4 *>(Cn + (size_t) (n0 + lr) * dim_pad + k0 + lc);
            }
            __syncthreads();
            Xs[lc + 0][lr] = xv.x; Xs[lc + 1][lr] = xv.y; Xs[lc + 2][lr] = xv.z; Xs[lc + 3][lr] = xv.w;
            Cs[lc + 0][lr] = cv.x; Cs[lc + 1][lr] = cv.y; Cs[lc + 2][lr] = cv.z; Cs[lc + 3][lr] = cv.w;
            __syncthreads();
#pragma unroll
            for (int k = 0; k < AS_BK; k++) {
                const float4 a = *reinterpret_cast<const float4 *>(&Xs[k][ty * 4]);
                const float4 b = *reinterpret_cast<const float4 *>(&Cs[k][tx * 4]);
                const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
                for (int i = 0; i < 4; i++)
#pragma unroll
                    for (int j = 0; j < 4; j++) {
                        if (METRIC == 0) { const float d = av[i] - bv[j]; acc[i][j] = fmaf(d, d, acc[i][j]); }
                        else acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
                    }
            }
        }
        // best of this centre tile per row: my 4 centres, then the 16 tx lanes of the row group
#pragma unroll
        for (int i = 0; i < 4; i++) {
            float bv = __int_as_float(0x7f800000);
            uint32_t bi = 0xFFFFFFFFu;
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const uint32_t c = (uint32_t) (n0 + tx * 4 + j);
                const float v = METRIC == 0 ? acc[i][j] : -acc[i][j];
                if (c < (uint32_t) lists && (v < bv || (v == bv && c < bi))) { bv = v; bi = c; }
            }
#pragma unroll
            for (int o = 8; o > 0; o >>= 1) {
                const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
                const uint32_t oi = __shfl_xor_sync(0xffffffffu, bi, o);
                if (ov < bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
            }
            // centre tiles are visited in ascending order: strict < keeps the first minimum
            if (bi != 0xFFFFFFFFu && (bv < best[i] || (bv == best[i] && bi < besti[i]))) { best[i] = bv; besti[i] = bi; }
        }
    }
    if (tx == 0) {
#pragma unroll
        for (int i = 0; i < 4; i++) {
            const size_t r = row0 + ty * 4 + i;
            if (r < nrows) {
                out_list[r] = besti[i];
                if (out_val) out_val[r] = best[i];
            }
        }
    }
}

cudaError_t rk_launch_assign(const float *X, size_t nrows, const float *C, int lists, int dim_pad, int metric,
                             uint32_t *out_list, float *out_val, cudaStream_t stream)
{
    if (nrows == 0) return cudaSuccess;
    const unsigned grid = (unsigned) ((nrows + AS_BM - 1) / AS_BM);
    if (metric == 0) rk_assign_kernel<0><<<grid, 256, 0, stream>>>(X, nrows, C, lists, dim_pad, out_list, out_val);
    else rk_assign_kernel<1><<<grid, 256, 0, stream>>>(X, nrows, C, lists, dim_pad, out_list, out_val);
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// k-means helpers (Lloyd): accumulate, finalize (+ spherical normalisation for inner product), changes

__global__ void rk_km_accumulate_kernel(const float *__restrict__ S, size_t ns, int dim_pad, const uint32_t *__restrict__ label,
                                        float *__restrict__ sums, uint32_t *__restrict__ counts)
{
    const int lane = threadIdx.x & 31;
    const size_t warp = (blockIdx.x * (size_t) blockDim.x + threadIdx.x) >> 5;
    const size_t nw = ((size_t) gridDim.x * blockDim.x) >> 5;
    for (size_t r = warp; r < ns; r += nw) {
        const uint32_t l = label[r];
        for (int c = lane; c < dim_pad; c += 32) atomicAdd(&sums[(size_t) l * dim_pad + c], S[r * (size_t) dim_pad + c]);
        if (lane == 0) atomicAdd(&counts[l], 1u);
    }
}

// one warp per centre: mean of its members; an empty cluster is re-seeded from a sample row
__global__ void rk_km_finalize_kernel(float *__restrict__ C, const float *__restrict__ sums, const uint32_t *__restrict__ counts,
                                      int lists, int dim_pad, const float *__restrict__ S, size_t ns, uint64_t seed, int normalize)
{
    const int lane = threadIdx.x & 31;
    const int c = (int) ((blockIdx.x * (size_t) blockDim.x + threadIdx.x) >> 5);
    if (c >= lists) return;
    const uint32_t n = counts[c];
    float nrm = 0.f;
    if (n) {
        const float inv = 1.0f / (float) n;
        for (int j = lane; j < dim_pad; j += 32) {
            const float v = sums[(size_t) c * dim_pad + j] * inv;
            C[(size_t) c * dim_pad + j] = v;
            nrm = fmaf(v, v, nrm);
        }
    } else {
        uint64_t h = (seed + (uint64_t) c) * 0x9E3779B97F4A7C15ull;
        h ^= h >> 29; h *= 0xBF58476D1CE4E5B9ull; h ^= h >> 32;
        const size_t r = (size_t) (h % (uint64_t) ns);
        for (int j = lane; j < dim_pad; j += 32) {
            const float v = S[r * (size_t) dim_pad + j];
            C[(size_t) c * dim_pad + j] = v;
            nrm = fmaf(v, v, nrm);
        }
    }
    if (normalize) {
        for (int o = 16; o > 0; o >>= 1) nrm += __shfl_xor_sync(0xffffffffu, nrm, o);
        if (nrm > 0.f) {
            const float s = rsqrtf(nrm);
            for (int j = lane; j < dim_pad; j += 32) C[(size_t) c * dim_pad + j] *= s;
        }
    }
}

__global__ void rk_km_normalize_rows_kernel(float *__restrict__ S, size_t ns, int dim_pad)
{
    const int lane = threadIdx.x & 31;
    const size_t warp = (blockIdx.x * (size_t) blockDim.x + threadIdx.x) >> 5;
    const size_t nw = ((size_t) gridDim.x * blockDim.x) >> 5;
    for (size_t r = warp; r < ns; r += nw) {
        float nrm = 0.f;
        for (int j = lane; j < dim_pad; j += 32) { const float v = S[r * (size_t) dim_pad + j]; nrm = fmaf(v, v, nrm); }
        for (int o = 16; o > 0; o >>= 1) nrm += __shfl_xor_sync(0xffffffffu, nrm, o);
        if (nrm > 0.f) {
            const float s = rsqrtf(nrm);
            for (int j = lane; j < dim_pad; j += 32) S[r * (size_t) dim_pad + j] *= s;
        }
    }
}

__global__ void rk_count_changed_kernel(const uint32_t *__restrict__ a, const uint32_t *__restrict__ b, size_t n, unsigned long long *out)
{
    unsigned c = 0;
    for (size_t i = blockIdx.x * (size_t) blockDim.x + threadIdx.x; i < n; i += (size_t) gridDim.x * blockDim.x) c += a[i] != b[i];
    c = __reduce_add_sync(0xffffffffu, c);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd(out, (unsigned long long) c);
}

cudaError_t rk_launch_km_accumulate(const float *S, size_t ns, int dim_pad, const uint32_t *label, float *sums, uint32_t *counts,
                                    cudaStream_t stream)
{
    rk_km_accumulate_kernel<<<148 * 4, 256, 0, stream>>>(S, ns, dim_pad, label, sums, counts);
    return cudaGetLastError();
}
cudaError_t rk_launch_km_finalize(float *C, const float *sums, const uint32_t *counts, int lists, int dim_pad, const float *S,
                                  size_t ns, uint64_t seed, int normalize, cudaStream_t stream)
{
    rk_km_finalize_kernel<<<(lists * 32 + 255) / 256, 256, 0, stream>>>(C, sums, counts, lists, dim_pad, S, ns, seed, normalize);
    return cudaGetLastError();
}
cudaError_t rk_launch_km_normalize_rows(float *S, size_t ns, int dim_pad, cudaStream_t stream)
{
    rk_km_normalize_rows_kernel<<<148 * 4, 256, 0, stream>>>(S, ns, dim_pad);
    return cudaGetLastError();
}
cudaError_t rk_launch_count_changed(const uint32_t *a, const uint32_t *b, size_t n, unsigned long long *d_out, cudaStream_t stream)
{
    rk_count_changed_kernel<<<148 * 2, 256, 0, stream>>>(a, b, n, d_out);
    return cudaGetLastError();
}

// dst[i] = src[perm[i]] for 4- and 8-byte elements
__global__ void rk_gather_u32_kernel(const uint32_t *__restrict__ src, uint32_t *__restrict__ dst, const uint32_t *__restrict__ perm, size_t n)
{
    for (size_t i = blockIdx.x * (size_t) blockDim.x + threadIdx.x; i < n; i += (size_t) gridDim.x * blockDim.x) dst[i] = src[perm[i]];
}
__global__ void rk_gather_u64_kernel(const uint64_t *__restrict__ src, uint64_t *__restrict__ dst, const uint32_t *__restrict__ perm, size_t n)
{
    for (size_t i = blockIdx.x * (size_t) blockDim.x + threadIdx.x; i < n; i += (size_t) gridDim.x * blockDim.x) dst[i] = src[perm[i]];
}
cudaError_t rk_launch_gather_u32(const uint32_t *src, uint32_t *dst, const uint32_t *perm, size_t n, cudaStream_t stream)
{
    if (n) rk_gather_u32_kernel<<<148 * 4, 256, 0, stream>>>(src, dst, perm, n);
    return cudaGetLastError();
}
cudaError_t rk_launch_gather_u64(const uint64_t *src, uint64_t *dst, const uint32_t *perm, size_t n, cudaStream_t stream)
{
    if (n) rk_gather_u64_kernel<<<148 * 4, 256, 0, stream>>>(src, dst, perm, n);
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// list probe: GetScanLists (ivfscan.c:36-107).  One block per query: value(q, centre) for every
// list, full sort of (key, list) in shared memory (ties -> lower list number), then the slice
// [probe_offset, probe_offset + probes) is written nearest first.

#define PR_THREADS 256

template <int METRIC>
__global__ void __launch_bounds__(PR_THREADS) rk_ivf_probe_kernel(const float *__restrict__ Q, const float *__restrict__ Cn, int lists,
                                                                  int dim_pad, int null_query, int probe_offset, int probes,
                                                                  uint32_t *__restrict__ out_list)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint64_t *ent = reinterpret_cast<uint64_t *>(smem_raw);                 // [P] (key << 32 | list)
    int P = 32;
    while (P < lists) P <<= 1;
    float4 *qs = reinterpret_cast<float4 *>(ent + P);                        // [dim_pad / 4]
    const int q = blockIdx.x;
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const int dim4 = dim_pad >> 2;
    const float4 *q4 = reinterpret_cast<const float4 *>(Q + (size_t) q * dim_pad);
    for (int i = t; i < dim4; i += PR_THREADS) qs[i] = q4[i];
    for (int i = lists + t; i < P; i += PR_THREADS) ent[i] = ~0ull;
    __syncthreads();
    // one warp per centre
    for (int c = warp; c < lists; c += PR_THREADS / 32) {
        const float4 *c4 = reinterpret_cast<const float4 *>(Cn + (size_t) c * dim_pad);
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        if (!null_query) {
            for (int j = lane; j < dim4; j += 32) {
                const float4 x = c4[j], y = qs[j];
                if (METRIC == 0) {
                    const float d0 = x.x - y.x, d1 = x.y - y.y, d2 = x.z - y.z, d3 = x.w - y.w;
                    acc.x = fmaf(d0, d0, acc.x); acc.y = fmaf(d1, d1, acc.y); acc.z = fmaf(d2, d2, acc.z); acc.w = fmaf(d3, d3, acc.w);
                } else {
                    acc.x = fmaf(x.x, y.x, acc.x); acc.y = fmaf(x.y, y.y, acc.y); acc.z = fmaf(x.z, y.z, acc.z); acc.w = fmaf(x.w, y.w, acc.w);
                }
            }
        }
        float s = (acc.x + acc.y) + (acc.z + acc.w);
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        if (lane == 0) ent[c] = ((uint64_t) rk_float_to_key(METRIC == 0 ? s : -s) << 32) | (uint32_t) c;
    }
    __syncthreads();
    for (int size = 2; size <= P; size <<= 1) {
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
            for (int i = t; i < (P >> 1); i += PR_THREADS) {
                const int lo = 2 * i - (i & (stride - 1)), hi = lo + stride;
                const bool up = (lo & size) == 0;
                const uint64_t a = ent[lo], b = ent[hi];
                if (up ? (b < a) : (a < b)) { ent[lo] = b; ent[hi] = a; }
            }
            __syncthreads();
        }
    }
    for (int j = t; j < probes; j += PR_THREADS) {
        const int src = probe_offset + j;
        out_list[(size_t) q * probes + j] = src < lists ? (uint32_t) (ent[src] & 0xFFFFFFFFu) : 0xFFFFFFFFu;
    }
}

size_t rk_ivf_probe_smem(int lists, int dim_pad)
{
    int P = 32;
    while (P < lists) P <<= 1;
    return (size_t) P * 8 + (size_t) dim_pad * 4;
}

cudaError_t rk_launch_ivf_probe(const float *Q, int nq, const float *C, int lists, int dim_pad, int metric, int null_query,
                                int probe_offset, int probes, uint32_t *out_list, cudaStream_t stream)
{
    const size_t smem = rk_ivf_probe_smem(lists, dim_pad);
    auto kern = metric == 0 ? rk_ivf_probe_kernel<0> : rk_ivf_probe_kernel<1>;
    static bool attr_set[2];
    if (smem > 48 * 1024 && !attr_set[metric & 1]) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
        if (e != cudaSuccess) return e;
        attr_set[metric & 1] = true;
    }
    kern<<<nq, PR_THREADS, smem, stream>>>(Q, C, lists, dim_pad, null_query, probe_offset, probes, out_list);
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// work list: one segment per (query, probe, chunk) over the list-sorted store.  The segment count
// is fixed by the host (nq * probes * chunks); chunks beyond a short list come out empty.

__global__ void rk_ivf_plan_kernel(const uint32_t *__restrict__ probe_list, const uint32_t *__restrict__ list_start, int nq, int probes,
                                   int chunks, uint32_t chunk_rows, uint32_t kind, RkSegment *__restrict__ segs)
{
    const size_t total = (size_t) nq * probes * chunks;
    for (size_t i = blockIdx.x * (size_t) blockDim.x + threadIdx.x; i < total; i += (size_t) gridDim.x * blockDim.x) {
        const int c = (int) (i % chunks);
        const size_t qj = i / chunks;
        const int q = (int) (qj / probes);
        const uint32_t l = probe_list[qj];
        RkSegment s;
        s.query = (uint32_t) q; s.kind = (uint16_t) kind; s.flags = 0; s.ext_lo = 0; s.ext_hi = 0;
        uint64_t b = 0, e = 0;
        if (l != 0xFFFFFFFFu) {
            const uint64_t l0 = list_start[l], l1 = list_start[l + 1];
            b = l0 + (uint64_t) c * chunk_rows;
            e = b + chunk_rows;
            if (e > l1) e = l1;
            if (b > l1) b = l1;
        }
        s.begin = b; s.end = e;
        segs[i] = s;
    }
}

cudaError_t rk_launch_ivf_plan(const uint32_t *probe_list, const uint32_t *list_start, int nq, int probes, int chunks,
                               uint32_t chunk_rows, uint32_t kind, RkSegment *segs, cudaStream_t stream)
{
    const size_t total = (size_t) nq * probes * chunks;
    if (total == 0) return cudaSuccess;
    const unsigned grid = (unsigned) std::min<size_t>((total + 255) / 256, 148 * 8);
    rk_ivf_plan_kernel<<<grid, 256, 0, stream>>>(probe_list, list_start, nq, probes, chunks, chunk_rows, kind, segs);
    return cudaGetLastError();
}
