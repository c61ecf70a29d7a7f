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
#include <algorithm>

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
                                                        float *__restrict__ out_val, const uint32_t *__restrict__ idx)
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
                // (idx: the rows to assign are X[idx[0 .. nrows)], results go to out[idx[.]] — the tensor-core path's refused rows)
                if (row0 + lr < nrows) xv = *reinterpret_cast<const float4 *>(X + (idx ? (size_t) idx[row0 + lr] : row0 + lr) * (size_t) dim_pad + k0 + lc);
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
                const size_t ro = idx ? (size_t) idx[r] : r;
                out_list[ro] = besti[i];
                if (out_val) out_val[ro] = best[i];
            }
        }
    }
}

cudaError_t rk_launch_assign(const float *X, size_t nrows, const float *C, int lists, int dim_pad, int metric,
                             uint32_t *out_list, float *out_val, cudaStream_t stream, const uint32_t *idx)
{
    if (nrows == 0) return cudaSuccess;
    const unsigned grid = (unsigned) ((nrows + AS_BM - 1) / AS_BM);
    if (metric == 0) rk_assign_kernel<0><<<grid, 256, 0, stream>>>(X, nrows, C, lists, dim_pad, out_list, out_val, idx);
    else rk_assign_kernel<1><<<grid, 256, 0, stream>>>(X, nrows, C, lists, dim_pad, out_list, out_val, idx);
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// k-means++ seeding (pgvector/src/ivfkmeans.c:21-89, InitCenters).  The reference's loop, per new centre i: every sample's
// distance to centre i (the operator class's k-means distance: l2_distance, or the spherical distance acos(x.c) / pi on
// normalised rows), weight[j] = min(weight[j], distance^2) (distance a double, weight a float), sum of the weights (double);
// then choice = sum * RandomDouble() and the next centre is the first sample whose running weight total reaches it.
// Here: two small kernels per centre, no host round trip (the chosen row number stays on the device).  Block b owns the
// sample rows [b R, (b + 1) R); it writes the double sum of its rows' weights, always added in the same order, so a build is
// reproducible for a seed.  The reference draws from PostgreSQL's PRNG; here u_i = splitmix64(seed, i) / 2^53.

#define KPP_THREADS 256

__device__ __forceinline__ double rk_kpp_uniform(uint64_t seed, uint32_t i)
{
    uint64_t h = (seed + 0x9E3779B97F4A7C15ull * (uint64_t) (i + 1));
    h ^= h >> 30; h *= 0xBF58476D1CE4E5B9ull; h ^= h >> 27; h *= 0x94D049BB133111EBull; h ^= h >> 31;
    return (double) (h >> 11) * (1.0 / 9007199254740992.0);
}

template <int SPHERICAL>
__global__ void __launch_bounds__(KPP_THREADS) rk_kpp_dist_kernel(const float *__restrict__ S, uint32_t ns, int dim_pad, uint32_t rows_per_block,
                                                                   const uint32_t *__restrict__ chosen, uint32_t i, float *__restrict__ weight,
                                                                   double *__restrict__ blocksum)
{
    extern __shared__ float kpp_c[];                 // the new centre
    __shared__ double wsum[KPP_THREADS / 32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const float4 *c4g = reinterpret_cast<const float4 *>(S + (size_t) chosen[i] * dim_pad);
    float4 *c4 = reinterpret_cast<float4 *>(kpp_c);
    const int dim4 = dim_pad >> 2;
    for (int k = threadIdx.x; k < dim4; k += KPP_THREADS) c4[k] = c4g[k];
    __syncthreads();
    const uint32_t r0 = blockIdx.x * rows_per_block, r1 = min(ns, r0 + rows_per_block);
    double acc = 0.0;
    for (uint32_t r = r0 + (uint32_t) warp; r < r1; r += KPP_THREADS / 32) {
        const float4 *x4 = reinterpret_cast<const float4 *>(S + (size_t) r * dim_pad);
        float s = 0.f;
        for (int k = lane; k < dim4; k += 32) {
            const float4 a = x4[k], b = c4[k];
            if (SPHERICAL) { s = fmaf(a.x, b.x, s); s = fmaf(a.y, b.y, s); s = fmaf(a.z, b.z, s); s = fmaf(a.w, b.w, s); }
            else {
                float d = a.x - b.x; s = fmaf(d, d, s);
                d = a.y - b.y; s = fmaf(d, d, s);
                d = a.z - b.z; s = fmaf(d, d, s);
                d = a.w - b.w; s = fmaf(d, d, s);
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        if (lane == 0) {
            double d;
            if (SPHERICAL) {                         // vector_spherical_distance (vector.c): acos of the clamped dot, over pi
                double dp = (double) s;
                dp = dp > 1.0 ? 1.0 : (dp < -1.0 ? -1.0 : dp);
                d = acos(dp) / 3.14159265358979323846;
            } else {
                d = sqrt((double) s);                // l2_distance (vector.c:568-578)
            }
            d *= d;
            float w = weight[r];
            if (d < (double) w) { w = (float) d; weight[r] = w; }
            acc += (double) w;
        }
    }
    if (lane == 0) wsum[warp] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int w = 0; w < KPP_THREADS / 32; w++) t += wsum[w];
        blocksum[blockIdx.x] = t;
    }
}

// One warp: choice = (sum of the weights) * u_i; chosen[i + 1] = the first sample whose running total reaches it.
__global__ void rk_kpp_pick_kernel(const float *__restrict__ weight, const double *__restrict__ blocksum, uint32_t nblocks, uint32_t ns,
                                   uint32_t rows_per_block, uint64_t seed, uint32_t i, uint32_t *__restrict__ chosen)
{
    const uint32_t lane = threadIdx.x;
    // level 1: which block
    const uint32_t per = (nblocks + 31) / 32;
    const uint32_t b0 = min(nblocks, lane * per), b1 = min(nblocks, b0 + per);
    double loc = 0.0;
    for (uint32_t b = b0; b < b1; b++) loc += blocksum[b];
    double inc = loc;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const double u = __shfl_up_sync(0xffffffffu, inc, o); if ((int) lane >= o) inc += u; }
    const double total = __shfl_sync(0xffffffffu, inc, 31);
    double choice = total * rk_kpp_uniform(seed, i);
    uint32_t hit = __ballot_sync(0xffffffffu, b1 > b0 && inc >= choice);
    uint32_t blk = nblocks - 1;
    double rem = 0.0;
    bool found = hit != 0;
    if (found) {
        const int L = __ffs((int) hit) - 1;
        double c = choice - (inc - loc);              // what is left when lane L's blocks begin
        uint32_t bb = b1 - 1;
        for (uint32_t b = b0; b < b1; b++) { const double v = blocksum[b]; if (c <= v) { bb = b; break; } c -= v; }
        blk = __shfl_sync(0xffffffffu, bb, L);
        rem = __shfl_sync(0xffffffffu, c, L);
    }
    uint32_t pick = ns - 1;                           // the reference's loop ends on the last sample when the total is never reached
    if (found) {
        // level 2: which row of the block
        const uint32_t r0 = blk * rows_per_block, r1 = min(ns, r0 + rows_per_block), n = r1 - r0;
        const uint32_t per2 = (n + 31) / 32;
        const uint32_t a0 = min(n, lane * per2), a1 = min(n, a0 + per2);
        double l2 = 0.0;
        for (uint32_t j = a0; j < a1; j++) l2 += (double) weight[r0 + j];
        double in2 = l2;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const double u = __shfl_up_sync(0xffffffffu, in2, o); if ((int) lane >= o) in2 += u; }
        const uint32_t hit2 = __ballot_sync(0xffffffffu, a1 > a0 && in2 >= rem);
        uint32_t jj = r1 - 1;
        if (hit2) {
            const int L2 = __ffs((int) hit2) - 1;
            double c = rem - (in2 - l2);
            uint32_t j2 = a1 - 1;
            for (uint32_t j = a0; j < a1; j++) { const double v = (double) weight[r0 + j]; if (c <= v) { j2 = j; break; } c -= v; }
            jj = r0 + __shfl_sync(0xffffffffu, j2, L2);
        }
        pick = jj;
    }
    if (lane == 0) chosen[i + 1] = pick;
}

// chosen[lists] (device) = the sample rows k-means++ seeds the centres with.  weight: [ns] floats, blocksum: [nblocks] doubles
// (nblocks = rk_kpp_blocks(ns)).  Asynchronous on `stream`.
uint32_t rk_kpp_blocks(uint32_t ns) { return std::min<uint32_t>(1184u, (ns + 63u) / 64u); }

__global__ void rk_kpp_init_kernel(float *weight, uint32_t ns, uint32_t *chosen, uint64_t seed)
{
    for (uint32_t j = blockIdx.x * blockDim.x + threadIdx.x; j < ns; j += gridDim.x * blockDim.x) weight[j] = 3.402823466e38f;   // FLT_MAX
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        // "Choose an initial center uniformly at random": RandomInt() % samples->length
        uint64_t h = seed * 0x9E3779B97F4A7C15ull + 0x632BE59BD9B4E019ull;
        h ^= h >> 30; h *= 0xBF58476D1CE4E5B9ull; h ^= h >> 27; h *= 0x94D049BB133111EBull; h ^= h >> 31;
        chosen[0] = (uint32_t) ((h >> 32) % (uint64_t) ns);
    }
}

cudaError_t rk_launch_kmeanspp(const float *S, uint32_t ns, int dim_pad, int lists, int spherical, uint64_t seed, float *weight,
                               double *blocksum, uint32_t *chosen, cudaStream_t stream)
{
    const uint32_t nb = rk_kpp_blocks(ns), rpb = (ns + nb - 1) / nb;
    rk_kpp_init_kernel<<<148, 256, 0, stream>>>(weight, ns, chosen, seed);
    const size_t smem = (size_t) dim_pad * 4;
    for (int i = 0; i + 1 < lists; i++) {
        if (spherical) rk_kpp_dist_kernel<1><<<nb, KPP_THREADS, smem, stream>>>(S, ns, dim_pad, rpb, chosen, (uint32_t) i, weight, blocksum);
        else rk_kpp_dist_kernel<0><<<nb, KPP_THREADS, smem, stream>>>(S, ns, dim_pad, rpb, chosen, (uint32_t) i, weight, blocksum);
        rk_kpp_pick_kernel<<<1, 32, 0, stream>>>(weight, blocksum, nb, ns, rpb, seed, (uint32_t) i, chosen);
    }
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
// (PR_QB queries per block; launched with 1: the general path — any probe count, single queries.  Batches take the
// two-kernel path below.)

template <int METRIC, int PR_QB>
__global__ void __launch_bounds__(PR_THREADS) rk_ivf_probe_kernel(const float *__restrict__ Q, int nq, const float *__restrict__ Cn, int lists,
                                                                  int dim_pad, int null_query, int probe_offset, int probes,
                                                                  uint32_t *__restrict__ out_list)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    int P = 32;
    while (P < lists) P <<= 1;
    uint64_t *ent = reinterpret_cast<uint64_t *>(smem_raw);                 // [PR_QB][P] (key << 32 | list)
    float4 *qs = reinterpret_cast<float4 *>(ent + (size_t) PR_QB * P);       // [PR_QB][dim_pad / 4]
    const int q0 = blockIdx.x * PR_QB;
    const int nqb = min(PR_QB, nq - q0);
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const int dim4 = dim_pad >> 2;
    for (int i = t; i < PR_QB * dim4; i += PR_THREADS) {
        const int b = i / dim4, j = i - b * dim4;
        qs[i] = b < nqb ? reinterpret_cast<const float4 *>(Q + (size_t) (q0 + b) * dim_pad)[j] : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    for (int b = 0; b < PR_QB; b++)
        for (int i = lists + t; i < P; i += PR_THREADS) ent[(size_t) b * P + i] = ~0ull;
    __syncthreads();
    // one warp per centre, PR_QB queries at a time; per (query, centre) the arithmetic is what a single-query pass does
    // (lane j takes float4 columns j, j + 32, ...; (x + y) + (z + w); xor-shuffles 16 .. 1)
    for (int c = warp; c < lists; c += PR_THREADS / 32) {
        const float4 *c4 = reinterpret_cast<const float4 *>(Cn + (size_t) c * dim_pad);
        float4 acc[PR_QB];
#pragma unroll
        for (int b = 0; b < PR_QB; b++) acc[b] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (!null_query) {
            for (int j = lane; j < dim4; j += 32) {
                const float4 x = c4[j];
#pragma unroll
                for (int b = 0; b < PR_QB; b++) {
                    const float4 y = qs[b * dim4 + j];
                    if (METRIC == 0) {
                        const float d0 = x.x - y.x, d1 = x.y - y.y, d2 = x.z - y.z, d3 = x.w - y.w;
                        acc[b].x = fmaf(d0, d0, acc[b].x); acc[b].y = fmaf(d1, d1, acc[b].y);
                        acc[b].z = fmaf(d2, d2, acc[b].z); acc[b].w = fmaf(d3, d3, acc[b].w);
                    } else {
                        acc[b].x = fmaf(x.x, y.x, acc[b].x); acc[b].y = fmaf(x.y, y.y, acc[b].y);
                        acc[b].z = fmaf(x.z, y.z, acc[b].z); acc[b].w = fmaf(x.w, y.w, acc[b].w);
                    }
                }
            }
        }
#pragma unroll
        for (int b = 0; b < PR_QB; b++) {
            float sv = (acc[b].x + acc[b].y) + (acc[b].z + acc[b].w);
            for (int o = 16; o > 0; o >>= 1) sv += __shfl_xor_sync(0xffffffffu, sv, o);
            if (lane == 0) ent[(size_t) b * P + c] = ((uint64_t) rk_float_to_key(METRIC == 0 ? sv : -sv) << 32) | (uint32_t) c;
        }
    }
    __syncthreads();
    // the PR_QB orderings: bitonic sorts side by side (P / 2 compare-exchanges per query and stage)
    for (int size = 2; size <= P; size <<= 1) {
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
            for (int i = t; i < nqb * (P >> 1); i += PR_THREADS) {
                const int b = i / (P >> 1), ii = i - b * (P >> 1);
                const int lo = 2 * ii - (ii & (stride - 1)), hi = lo + stride;
                const bool up = (lo & size) == 0;
                uint64_t *e = ent + (size_t) b * P;
                const uint64_t x = e[lo], y = e[hi];
                if (up ? (y < x) : (x < y)) { e[lo] = y; e[hi] = x; }
            }
            __syncthreads();
        }
    }
    for (int i = t; i < nqb * probes; i += PR_THREADS) {
        const int b = i / probes, j = i - b * probes;
        const int src = probe_offset + j;
        out_list[(size_t) (q0 + b) * probes + j] = src < lists ? (uint32_t) (ent[(size_t) b * P + src] & 0xFFFFFFFFu) : 0xFFFFFFFFu;
    }
}

// ---------------------------------------------------------------------------------------------
// list probe, batched (the path every multi-query call takes): two kernels.
//
//   rk_ivf_keys_kernel   value(q, centre) for a tile of queries x a range of centres -> keys[q][centre].  A warp keeps
//                        QW queries in registers (lane j holds float4 columns j, j + 32, ... of each) and walks the
//                        centres of a tile staged in shared memory by bulk copies (one per centre row, two buffers on mbarriers); a block of PK_WARPS warps
//                        shares every centre tile, so a centre is fetched from L2 once per PK_WARPS * QW queries instead
//                        of once per query.  Per (query, centre) the arithmetic is the one-block-per-query kernel's, in
//                        the same order (columns j, j + 32, ... into x / y / z / w sums, (x + y) + (z + w), xor-shuffles
//                        16 .. 1; out-of-range columns contribute fma(0, 0, acc) = acc), so the keys are bit-identical.
//   rk_ivf_order_kernel  one block per query: the `need` = probe_offset + probes smallest (key, list) pairs by radix
//                        select over the keys (bytes of key - min, so the bins spread), gathered in list order (ties at
//                        the threshold go to the lower list numbers, as the full sort would place them), sorted.

#define PK_WARPS 4
#define PK_THREADS (PK_WARPS * 32)
#define PO_THREADS 256
#define PO_MAX_NEED 1024

template <int METRIC, int NJ, int QW, int TC>
__global__ void __launch_bounds__(PK_THREADS) rk_ivf_keys_kernel(const float *__restrict__ Q, int nq, const float *__restrict__ Cn, int lists,
                                                                 int dim_pad, int null_query, int centres_per_block,
                                                                 uint32_t *__restrict__ keys)
{
    constexpr int RS = NJ * 32;                                             // float4 per staged centre (columns >= dim4 stay zero)
    constexpr int NV = QW * TC;                                             // (query, centre) values of a warp per tile: 32 or 8
    static_assert(NV == 32 || NV == 8, "the reduction below is written for 32 or 8 values per tile");
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float4 *tile = reinterpret_cast<float4 *>(smem_raw);                    // [2][TC][RS]
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const int dim4 = dim_pad >> 2;
    const int qbase = (blockIdx.x * PK_WARPS + warp) * QW;
    const int c_begin = blockIdx.y * centres_per_block;
    const int c_end = min(lists, c_begin + centres_per_block);
    const int ntiles = (c_end - c_begin + TC - 1) / TC;
    const float4 zero4 = make_float4(0.f, 0.f, 0.f, 0.f);

    // zero both buffers once: the column tail of every staged row stays zero (the copies never touch it)
    __shared__ __align__(8) uint64_t full[2];
    for (int i = t; i < 2 * TC * RS; i += PK_THREADS) tile[i] = zero4;
    if (t == 0) { rk_mbar_init(&full[0], 1); rk_mbar_init(&full[1], 1); rk_fence_barrier_init(); }
    __syncthreads();
    // one bulk copy per centre row, issued by the first lanes of warp 0, landing on the buffer's mbarrier
    auto load_tile = [&](int ti, int buf) {
        if (warp != 0) return;
        const int c0 = c_begin + ti * TC;
        const int nr = min(TC, c_end - c0);
        if (lane == 0) rk_mbar_arrive_expect_tx(&full[buf], (uint32_t) nr * (uint32_t) dim_pad * 4u);
        __syncwarp();
        if (lane < nr)
            rk_bulk_g2s(tile + ((size_t) buf * TC + lane) * RS, Cn + (size_t) (c0 + lane) * dim_pad, (uint32_t) dim_pad * 4u, &full[buf]);
    };
    if (ntiles > 0) load_tile(0, 0);

    float4 qr[QW][NJ];
#pragma unroll
    for (int b = 0; b < QW; b++)
#pragma unroll
        for (int jj = 0; jj < NJ; jj++) {
            const int j = lane + 32 * jj;
            qr[b][jj] = (qbase + b < nq && j < dim4) ? reinterpret_cast<const float4 *>(Q + (size_t) (qbase + b) * dim_pad)[j] : zero4;
        }

    for (int ti = 0; ti < ntiles; ti++) {
        if (ti + 1 < ntiles) load_tile(ti + 1, (ti + 1) & 1);
        rk_mbar_wait(&full[ti & 1], (uint32_t) (ti >> 1) & 1u);
        const float4 *tb = tile + (size_t) (ti & 1) * TC * RS + lane;
        const int c0 = c_begin + ti * TC;
        float v[NV];                                                        // v[b * TC + c]: this lane's part of value(query b, centre c0 + c)
#pragma unroll
        for (int c = 0; c < TC; c++) {
            float4 acc[QW];
#pragma unroll
            for (int b = 0; b < QW; b++) acc[b] = zero4;
            if (!null_query) {
#pragma unroll
                for (int jj = 0; jj < NJ; jj++) {
                    const float4 x = tb[c * RS + 32 * jj];
#pragma unroll
                    for (int b = 0; b < QW; b++) {
                        const float4 y = qr[b][jj];
                        if (METRIC == 0) {
                            const float d0 = x.x - y.x, d1 = x.y - y.y, d2 = x.z - y.z, d3 = x.w - y.w;
                            acc[b].x = fmaf(d0, d0, acc[b].x); acc[b].y = fmaf(d1, d1, acc[b].y);
                            acc[b].z = fmaf(d2, d2, acc[b].z); acc[b].w = fmaf(d3, d3, acc[b].w);
                        } else {
                            acc[b].x = fmaf(x.x, y.x, acc[b].x); acc[b].y = fmaf(x.y, y.y, acc[b].y);
                            acc[b].z = fmaf(x.z, y.z, acc[b].z); acc[b].w = fmaf(x.w, y.w, acc[b].w);
                        }
                    }
                }
            }
#pragma unroll
            for (int b = 0; b < QW; b++) v[b * TC + c] = (acc[b].x + acc[b].y) + (acc[b].z + acc[b].w);
        }
        // The xor-shuffle sums (16, 8, 4, 2, 1) of all NV values at once: at every step a lane keeps one half of its values
        // and hands the other half to its partner, which adds the same two numbers the full butterfly would add in both
        // lanes (a + b == b + a), so each value's sum is bit-identical to `for (o = 16; o; o >>= 1) v += shfl_xor(v, o)`.
        // Afterwards lane L holds value L * NV / 32.
#pragma unroll
        for (int st = 0; st < 5; st++) {
            const int o = 16 >> st;
            const int n_in = NV >> st;
            if (n_in > 1) {
                const bool up = (lane & o) != 0;
#pragma unroll
                for (int i = 0; i < n_in / 2; i++) {
                    const float keep = up ? v[i + n_in / 2] : v[i];
                    const float send = up ? v[i] : v[i + n_in / 2];
                    v[i] = keep + __shfl_xor_sync(0xffffffffu, send, o);
                }
            } else {
                v[0] += __shfl_xor_sync(0xffffffffu, v[0], o);
            }
        }
        {
            constexpr int LPV = 32 / NV;                                    // lanes holding the same value
            const int idx = lane / LPV;
            const int b = idx / TC, c = idx - b * TC;
            if ((lane % LPV) == 0 && c0 + c < c_end && qbase + b < nq)
                keys[(size_t) (qbase + b) * lists + c0 + c] = rk_float_to_key(METRIC == 0 ? v[0] : -v[0]);
        }
        __syncthreads();                                                    // the buffer is refilled by the next iteration's copy
                                                                            // (per-warp `empty` mbarriers instead: measured no faster)
    }
}

__global__ void __launch_bounds__(PO_THREADS) rk_ivf_order_kernel(const uint32_t *__restrict__ keys, int lists, int probe_offset, int probes,
                                                                  uint32_t *__restrict__ out_list)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint64_t *sel = reinterpret_cast<uint64_t *>(smem_raw);                 // [N2] chosen (key << 32 | list)
    const int need = probe_offset + probes;                                  // <= lists, <= PO_MAX_NEED
    int N2 = 32;
    while (N2 < need) N2 <<= 1;
    uint32_t *ks = reinterpret_cast<uint32_t *>(sel + N2);                  // [lists] key - min
    __shared__ uint32_t hist[256];
    __shared__ uint32_t red[2][PO_THREADS / 32];
    __shared__ uint32_t s_prefix, s_remaining, s_taken_lt, s_taken_eq;
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const uint32_t *kq = keys + (size_t) blockIdx.x * lists;

    uint32_t mn = 0xFFFFFFFFu, mx = 0;
    for (int i = t; i < lists; i += PO_THREADS) {
        const uint32_t v = kq[i];
        ks[i] = v;
        mn = min(mn, v); mx = max(mx, v);
    }
    mn = __reduce_min_sync(0xffffffffu, mn); mx = __reduce_max_sync(0xffffffffu, mx);
    if (lane == 0) { red[0][warp] = mn; red[1][warp] = mx; }
    __syncthreads();
    mn = red[0][0]; mx = red[1][0];
    for (int w = 1; w < PO_THREADS / 32; w++) { mn = min(mn, red[0][w]); mx = max(mx, red[1][w]); }
    const uint32_t range = mx - mn;
    const int bits = 32 - __clz(range | 1u);
    const int npass = (bits + 7) >> 3;
    if (t == 0) { s_prefix = 0; s_remaining = (uint32_t) need; s_taken_lt = 0; s_taken_eq = 0; }
    for (int i = t; i < lists; i += PO_THREADS) ks[i] -= mn;                // own entries: no barrier needed in between
    // radix select of the need-th smallest of ks: one byte per pass, high to low
    for (int pass = npass - 1; pass >= 0; pass--) {
        const int shift = pass * 8;
        hist[t] = 0;
        __syncthreads();
        const uint32_t prefix = s_prefix;
        const uint32_t himask = pass == npass - 1 ? 0u : (0xFFFFFFFFu << (shift + 8));
        for (int i = t; i < lists; i += PO_THREADS) {
            const uint32_t v = ks[i];
            if ((v & himask) == prefix) atomicAdd(&hist[(v >> shift) & 255u], 1u);
        }
        __syncthreads();
        if (warp == 0) {
            // bin holding the remaining-th entry: 8 bins per lane, warp scan of the lane sums
            uint32_t h[8], sum = 0;
#pragma unroll
            for (int i = 0; i < 8; i++) { h[i] = hist[lane * 8 + i]; sum += h[i]; }
            uint32_t inc = sum;
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t v = __shfl_up_sync(0xffffffffu, inc, o);
                if (lane >= o) inc += v;
            }
            const uint32_t rem = s_remaining;
            const uint32_t before = inc - sum;
            if (before < rem && rem <= inc) {
                uint32_t cum = before;
#pragma unroll
                for (int i = 0; i < 8; i++) {
                    if (cum < rem && rem <= cum + h[i]) {
                        s_prefix = prefix | ((uint32_t) (lane * 8 + i) << shift);
                        s_remaining = rem - cum;
                    }
                    cum += h[i];
                }
            }
        }
        __syncthreads();
    }
    // threshold = s_prefix; s_remaining of the entries equal to it are wanted (lowest list numbers first)
    const uint32_t thr = s_prefix;
    const uint32_t want_eq = s_remaining;
    const uint32_t want_lt = (uint32_t) need - want_eq;
    for (int i = need + t; i < N2; i += PO_THREADS) sel[i] = ~0ull;
    for (int base = 0; base < lists; base += PO_THREADS) {
        const int i = base + t;
        const uint32_t v = i < lists ? ks[i] : 0xFFFFFFFFu;
        const bool lt = i < lists && v < thr, eq = i < lists && v == thr;
        const uint32_t blt = __ballot_sync(0xffffffffu, lt), beq = __ballot_sync(0xffffffffu, eq);
        if (lane == 0) { red[0][warp] = __popc(blt); red[1][warp] = __popc(beq); }
        __syncthreads();
        uint32_t plt = s_taken_lt, peq = s_taken_eq, tlt = 0, teq = 0;
        for (int w = 0; w < PO_THREADS / 32; w++) {
            if (w < warp) { plt += red[0][w]; peq += red[1][w]; }
            tlt += red[0][w]; teq += red[1][w];
        }
        plt += __popc(blt & ((1u << lane) - 1u));
        peq += __popc(beq & ((1u << lane) - 1u));
        const uint64_t e = ((uint64_t) (v + mn) << 32) | (uint32_t) i;
        if (lt) sel[plt] = e;                                               // all of them: fewer than need
        if (eq && peq < want_eq) sel[want_lt + peq] = e;
        __syncthreads();
        if (t == 0) { s_taken_lt += tlt; s_taken_eq += teq; }
        // (the next round reads the counters after its own barrier)
        __syncthreads();
        if (s_taken_lt >= want_lt && s_taken_eq >= want_eq) break;
    }
    __syncthreads();
    for (int size = 2; size <= N2; size <<= 1) {
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
            for (int i = t; i < (N2 >> 1); i += PO_THREADS) {
                const int lo = 2 * i - (i & (stride - 1)), hi = lo + stride;
                const bool up = (lo & size) == 0;
                const uint64_t x = sel[lo], y = sel[hi];
                if (up ? (y < x) : (x < y)) { sel[lo] = y; sel[hi] = x; }
            }
            __syncthreads();
        }
    }
    for (int j = t; j < probes; j += PO_THREADS)
        out_list[(size_t) blockIdx.x * probes + j] = (uint32_t) (sel[probe_offset + j] & 0xFFFFFFFFu);
}

size_t rk_ivf_probe_smem(int lists, int dim_pad)
{
    int P = 32;
    while (P < lists) P <<= 1;
    return (size_t) P * 8 + (size_t) dim_pad * 4;
}

size_t rk_ivf_probe_keys_bytes(int nq, int lists) { return (size_t) nq * lists * 4; }

// cudaFuncSetAttribute is per device: remember per (kernel, device) that the shared-memory limit was raised
static cudaError_t pk_raise_smem(const void *kern, bool (&done)[64], int bytes)
{
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (done[dev & 63]) return cudaSuccess;
    e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (e == cudaSuccess) done[dev & 63] = true;
    return e;
}

template <int METRIC, int NJ, int QW, int TC>
static cudaError_t launch_keys(const float *Q, int nq, const float *C, int lists, int dim_pad, int null_query, uint32_t *keys, cudaStream_t stream)
{
    auto kern = rk_ivf_keys_kernel<METRIC, NJ, QW, TC>;
    const size_t smem = (size_t) 2 * TC * NJ * 32 * 16;
    static bool attr_set[64];
    if (smem + 1024 > 48 * 1024) {                                         // (+ the kernel's static shared memory)
        cudaError_t e = pk_raise_smem(reinterpret_cast<const void *>(kern), attr_set, (int) smem);
        if (e != cudaSuccess) return e;
    }
    const int qblocks = (nq + PK_WARPS * QW - 1) / (PK_WARPS * QW);
    // centre ranges: about three full waves of resident blocks (the last wave as full as the others), at least four tiles each
    static int resident = 0;                                               // (one GPU model per process)
    if (resident == 0) {
        int per_sm = 0, dev = 0, sms = 148;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, PK_THREADS, smem) != cudaSuccess || per_sm < 1) per_sm = 1;
        resident = per_sm * sms;
    }
    const int max_splits = std::max(1, (lists + 4 * TC - 1) / (4 * TC));
    int splits = std::max(1, std::min(max_splits, 3 * resident / qblocks));
    int cpb = (lists + splits - 1) / splits;
    cpb = (cpb + TC - 1) / TC * TC;
    splits = (lists + cpb - 1) / cpb;
    kern<<<dim3(qblocks, splits), PK_THREADS, smem, stream>>>(Q, nq, C, lists, dim_pad, null_query, cpb, keys);
    return cudaGetLastError();
}

template <int METRIC>
static cudaError_t launch_keys_m(const float *Q, int nq, const float *C, int lists, int dim_pad, int null_query, uint32_t *keys, cudaStream_t stream)
{
    const int nj = ((dim_pad >> 2) + 31) / 32;
    if (nj <= 1) return launch_keys<METRIC, 1, 4, 8>(Q, nq, C, lists, dim_pad, null_query, keys, stream);
    if (nj <= 2) return launch_keys<METRIC, 2, 4, 8>(Q, nq, C, lists, dim_pad, null_query, keys, stream);
    if (nj <= 4) return launch_keys<METRIC, 4, 4, 8>(Q, nq, C, lists, dim_pad, null_query, keys, stream);
    if (nj <= 6) return launch_keys<METRIC, 6, 4, 8>(Q, nq, C, lists, dim_pad, null_query, keys, stream);
    if (nj <= 8) return launch_keys<METRIC, 8, 4, 8>(Q, nq, C, lists, dim_pad, null_query, keys, stream);
    if (nj <= 12) return launch_keys<METRIC, 12, 2, 4>(Q, nq, C, lists, dim_pad, null_query, keys, stream);
    return launch_keys<METRIC, 16, 2, 4>(Q, nq, C, lists, dim_pad, null_query, keys, stream);   // dim <= 2000 (ivfflat.h:24)
}

// keys_ws: nq * lists uint32 of scratch (rk_ivf_probe_keys_bytes) for the batched path, or null.
cudaError_t rk_launch_ivf_probe(const float *Q, int nq, const float *C, int lists, int dim_pad, int metric, int null_query,
                                int probe_offset, int probes, uint32_t *out_list, uint32_t *keys_ws, cudaStream_t stream)
{
    if (nq <= 0 || probes <= 0) return cudaSuccess;
    const int need = probe_offset + probes;
    const bool batched = keys_ws != nullptr && nq >= 4 && need <= PO_MAX_NEED && need <= lists && (dim_pad >> 2) <= 512;
    if (batched) {
        cudaError_t e = metric == 0 ? launch_keys_m<0>(Q, nq, C, lists, dim_pad, null_query, keys_ws, stream)
                                    : launch_keys_m<1>(Q, nq, C, lists, dim_pad, null_query, keys_ws, stream);
        if (e != cudaSuccess) return e;
        int N2 = 32;
        while (N2 < need) N2 <<= 1;
        const size_t smem = (size_t) N2 * 8 + (size_t) lists * 4;
        static bool attr_set[64];
        if (smem + 2048 > 48 * 1024) {
            e = pk_raise_smem(reinterpret_cast<const void *>(rk_ivf_order_kernel), attr_set, 128 * 1024);
            if (e != cudaSuccess) return e;
        }
        rk_ivf_order_kernel<<<nq, PO_THREADS, smem, stream>>>(keys_ws, lists, probe_offset, probes, out_list);
        return cudaGetLastError();
    }
    const size_t smem = rk_ivf_probe_smem(lists, dim_pad);
    auto kern = metric == 0 ? rk_ivf_probe_kernel<0, 1> : rk_ivf_probe_kernel<1, 1>;
    static bool attr_set[2][64];
    if (smem > 48 * 1024) {
        cudaError_t e = pk_raise_smem(reinterpret_cast<const void *>(kern), attr_set[metric & 1], 227 * 1024);
        if (e != cudaSuccess) return e;
    }
    kern<<<nq, PR_THREADS, smem, stream>>>(Q, nq, C, lists, dim_pad, null_query, probe_offset, probes, out_list);
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
