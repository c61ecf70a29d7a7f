// rk_assign_tc.cu — nearest-centre assignment of the IVFFlat build on the tensor cores.
//
// pgvector assigns every sampled row (each Lloyd / Elkan iteration, pgvector/src/ivfkmeans.c:258-494) and finally every heap
// row (pgvector/src/ivfbuild.c:141-198) to its nearest centre: rows x lists x dim multiply-adds — 3.1e13 for C4 (10 M rows,
// 4096 lists, 768 dimensions), 2.6 s as fp32 FMA (rk_assign_kernel).  It is a dense contraction, so here:
//
//   1. rk_assign_tc_kernel   D = X · Cᵀ in TF32 on tcgen05 (fp32 rows straight from HBM, accumulators in TMEM), one 128-row
//      tile of X against all centres, 256 at a time.  One TMEM lane = one row, and a row's candidates never leave the
//      thread that owns its lane: each epilogue thread keeps the 4 best approximate scores (|c|² − 2 x·c, or −x·c) of its
//      half of the columns in registers — no shared-memory buffers, no atomics — and writes them out when its tile is done:
//      8 nominations per row.
//   2. rk_assign_rescore_kernel recomputes those 8 distances in exact fp32 with rk_assign_kernel's own arithmetic (one
//      sequential fmaf chain per (row, centre)), picks the smallest (ties: lower list number) and PROVES that no centre
//      outside the nominations can beat it: every such centre has an approximate score >= the smaller of the two 4th-best
//      nominations, and an approximate score is within eps = 2⁻⁹ |x| |c| (x 2 for L2) of the truth.  A row whose proof
//      fails is flagged ...
//   3. ... and re-assigned by rk_assign_kernel itself (through a row index list).
// So the result is, for every row, exactly what rk_assign_kernel returns — on integer-valued data the first minimum in list
// order, as pgvector's loop finds it — whatever the tensor cores rounded.
#include "rk_bscan_dev.cuh"
#include "rk_kernels.h"

#define AT_M 128
#define AT_N 256
#define AT_KC 32
#define AT_STAGES 4
#define AT_EW 8
#define AT_THREADS (64 + AT_EW * 32)
#define AT_STAGE_BYTES 49152u            // X box 128 x 128 B + centre box 256 x 128 B
#define AT_NOM 8                         // nominations per row (4 per column half)

__device__ __forceinline__ void at_insert(float v, uint32_t id, float (&bv)[4], uint32_t (&bi)[4])
{
    // keep the 4 smallest (score, id), ascending; called only when v beats the current 4th
    if (v < bv[2] || (v == bv[2] && id < bi[2])) {
        bv[3] = bv[2]; bi[3] = bi[2];
        if (v < bv[1] || (v == bv[1] && id < bi[1])) {
            bv[2] = bv[1]; bi[2] = bi[1];
            if (v < bv[0] || (v == bv[0] && id < bi[0])) { bv[1] = bv[0]; bi[1] = bi[0]; bv[0] = v; bi[0] = id; }
            else { bv[1] = v; bi[1] = id; }
        } else { bv[2] = v; bi[2] = id; }
    } else { bv[3] = v; bi[3] = id; }
}

struct RkAssignTcParams {
    uint64_t nrows;
    int lists, dim_pad, kchunks, ntiles, nct;      // nct = centre tiles of 256
    const float *cnorm;                            // [nct * 256] squared centre norms, +inf beyond `lists` (L2)
    float *nom_val;                                // [nrows][8]
    uint32_t *nom_id;                              // [nrows][8]
};

template <int METRIC>
__global__ void __launch_bounds__(AT_THREADS, 1)
rk_assign_tc_kernel(const __grid_constant__ CUtensorMap tmx, const __grid_constant__ CUtensorMap tmc, const RkAssignTcParams p)
{
    extern __shared__ unsigned char smem_raw[];
    unsigned char *smem = smem_raw + ((1024u - (rk_smem_u32(smem_raw) & 1023u)) & 1023u);
    uint64_t *full = reinterpret_cast<uint64_t *>(smem + AT_STAGES * AT_STAGE_BYTES);
    uint64_t *empty = full + AT_STAGES;
    uint64_t *tfull = empty + AT_STAGES;
    uint64_t *tempty = tfull + 2;
    uint32_t *s_tmem = reinterpret_cast<uint32_t *>(tempty + 2);
    float *s_cn = reinterpret_cast<float *>(s_tmem + 4);          // [nct * 256]
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (threadIdx.x == 0) {
        for (int s = 0; s < AT_STAGES; s++) { rk_mbar_init(&full[s], 1); rk_mbar_init(&empty[s], 1); }
        for (int b = 0; b < 2; b++) { rk_mbar_init(&tfull[b], 1); rk_mbar_init(&tempty[b], AT_EW); }
        rk_fence_barrier_init();
    }
    if (METRIC == 0)
        for (int i = threadIdx.x; i < p.nct * AT_N; i += AT_THREADS) s_cn[i] = p.cnorm[i];
    if (warp == 1) bs_tmem_alloc(s_tmem, 2 * AT_N);
    bs_fence_before();
    __syncthreads();
    bs_fence_after();
    const uint32_t tmem_base = *s_tmem;

    if (warp == 0) {
        // ===================== TMA producer =====================
        int stage = 0;
        uint32_t phase = 0;
        if (lane == 0) {
            for (int t = blockIdx.x; t < p.ntiles; t += gridDim.x)
                for (int ct = 0; ct < p.nct; ct++)
                    for (int kc = 0; kc < p.kchunks; kc++) {
                        rk_mbar_wait(&empty[stage], phase ^ 1u);
                        unsigned char *st = smem + (size_t) stage * AT_STAGE_BYTES;
                        rk_mbar_arrive_expect_tx(&full[stage], AT_STAGE_BYTES);
                        bs_tma_2d(st, &tmx, kc * AT_KC, t * AT_M, &full[stage]);
                        bs_tma_2d(st + 16384, &tmc, kc * AT_KC, ct * AT_N, &full[stage]);
                        if (++stage == AT_STAGES) { stage = 0; phase ^= 1u; }
                    }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        int stage = 0;
        uint32_t phase = 0, tcount = 0;
        // c = F32, a = b = TF32, K-major both, N >> 3, M >> 4
        const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((AT_N >> 3) << 17) | ((AT_M >> 4) << 24);
        for (int t = blockIdx.x; t < p.ntiles; t += gridDim.x)
            for (int ct = 0; ct < p.nct; ct++, tcount++) {
                const uint32_t buf = tcount & 1u, tph = (tcount >> 1) & 1u;
                rk_mbar_wait(&tempty[buf], tph ^ 1u);
                bs_fence_after();
                const uint32_t d_tmem = tmem_base + buf * AT_N;
                for (int kc = 0; kc < p.kchunks; kc++) {
                    rk_mbar_wait(&full[stage], phase);
                    bs_fence_after();
                    if (lane == 0) {
                        const uint32_t a0 = rk_smem_u32(smem + (size_t) stage * AT_STAGE_BYTES);
                        const uint64_t ad = bs_smem_desc(a0), bd = bs_smem_desc(a0 + 16384u);
#pragma unroll
                        for (int k = 0; k < AT_KC / 8; k++)
                            bs_mma_tf32(d_tmem, ad + (uint64_t) (2 * k), bd + (uint64_t) (2 * k), idesc, (kc | k) != 0);
                        bs_commit(&empty[stage]);
                        if (kc == p.kchunks - 1) bs_commit(&tfull[buf]);
                    }
                    __syncwarp();
                    if (++stage == AT_STAGES) { stage = 0; phase ^= 1u; }
                }
            }
    } else {
        // ===================== epilogue warps: per-lane top-4 in registers =====================
        const int ew = warp - 2, quarter = warp & 3, chalf = ew >> 2;
        uint32_t tcount = 0;
        const float INF = __int_as_float(0x7f800000);
        for (int t = blockIdx.x; t < p.ntiles; t += gridDim.x) {
            float bv[4] = {INF, INF, INF, INF};
            uint32_t bi[4] = {0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu};
            for (int ct = 0; ct < p.nct; ct++, tcount++) {
                const uint32_t buf = tcount & 1u, tph = (tcount >> 1) & 1u;
                rk_mbar_wait(&tfull[buf], tph);
                bs_fence_after();
                const uint32_t taddr = tmem_base + buf * AT_N + ((uint32_t) (quarter * 32) << 16);
#pragma unroll 1
                for (int i = 0; i < 8; i++) {
                    const uint32_t c0 = 16u * (uint32_t) (chalf + 2 * i);
                    const uint32_t col0 = (uint32_t) ct * AT_N + c0;
                    if (col0 >= (uint32_t) p.lists) break;
                    uint32_t r[16];
                    bs_tmem_ld16(taddr + c0, r);
#pragma unroll
                    for (int j = 0; j < 16; j++) {
                        const float d = __uint_as_float(r[j]);
                        float v = METRIC == 0 ? fmaf(-2.0f, d, s_cn[col0 + j]) : -d;
                        if (METRIC != 0 && col0 + j >= (uint32_t) p.lists) v = INF;        // (L2: the padded norms are +inf)
                        if (v < bv[3] || (v == bv[3] && col0 + j < bi[3])) at_insert(v, col0 + j, bv, bi);
                    }
                }
                bs_fence_before();
                __syncwarp();
                if (lane == 0) rk_mbar_arrive(&tempty[buf]);
            }
            const uint64_t row = (uint64_t) t * AT_M + (uint64_t) (quarter * 32 + lane);
            if (row < p.nrows) {
                float4 *ov = reinterpret_cast<float4 *>(p.nom_val + row * AT_NOM + chalf * 4);
                uint4 *oi = reinterpret_cast<uint4 *>(p.nom_id + row * AT_NOM + chalf * 4);
                *ov = make_float4(bv[0], bv[1], bv[2], bv[3]);
                *oi = make_uint4(bi[0], bi[1], bi[2], bi[3]);
            }
        }
    }
    __syncthreads();
    if (warp == 1) bs_tmem_dealloc(tmem_base, 2 * AT_N);
}

// ---------------------------------------------------------------------------------------------
// squared centre norms, padded with +inf to a multiple of 256 (approximate scores only: any summation order will do)

__global__ void rk_assign_cnorm_kernel(const float *__restrict__ C, int lists, int lists_pad, int dim_pad, float *__restrict__ out)
{
    const int lane = threadIdx.x & 31;
    const int c = (int) ((blockIdx.x * (size_t) blockDim.x + threadIdx.x) >> 5);
    if (c >= lists_pad) return;
    float s = 0.f;
    if (c < lists) {
        for (int j = lane; j < dim_pad; j += 32) { const float v = C[(size_t) c * dim_pad + j]; s = fmaf(v, v, s); }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    } else {
        s = __int_as_float(0x7f800000);
    }
    if (lane == 0) out[c] = s;
}

// ---------------------------------------------------------------------------------------------
// exact re-score of the 8 nominations of every row + proof; 8 lanes per row, one (row, centre) chain per lane

template <int METRIC>
__global__ void __launch_bounds__(256) rk_assign_rescore_kernel(const float *__restrict__ X, size_t nrows, const float *__restrict__ C,
                                                                 int dim_pad, const float *__restrict__ nom_val,
                                                                 const uint32_t *__restrict__ nom_id, const unsigned *__restrict__ cn_max_bits,
                                                                 uint32_t *__restrict__ out_list, float *__restrict__ out_val,
                                                                 uint32_t *__restrict__ unc_idx, uint32_t *__restrict__ unc_count)
{
    const size_t row = (blockIdx.x * (size_t) blockDim.x + threadIdx.x) >> 3;
    const int l8 = threadIdx.x & 7;
    const bool live = row < nrows;
    const size_t r = live ? row : 0;
    const uint32_t id = live ? nom_id[r * AT_NOM + l8] : 0xFFFFFFFFu;
    const float INF = __int_as_float(0x7f800000);
    float v = INF, xn = 0.f;
    if (id != 0xFFFFFFFFu) {
        // rk_assign_kernel's arithmetic: ONE accumulator, k ascending, fused multiply-add
        const float4 *x4 = reinterpret_cast<const float4 *>(X + r * (size_t) dim_pad);
        const float4 *c4 = reinterpret_cast<const float4 *>(C + (size_t) id * dim_pad);
        float acc = 0.f;
        for (int k = 0; k < dim_pad / 4; k++) {
            const float4 a = x4[k], b = __ldg(&c4[k]);
            if (METRIC == 0) {
                float d = a.x - b.x; acc = fmaf(d, d, acc);
                d = a.y - b.y; acc = fmaf(d, d, acc);
                d = a.z - b.z; acc = fmaf(d, d, acc);
                d = a.w - b.w; acc = fmaf(d, d, acc);
            } else {
                acc = fmaf(a.x, b.x, acc); acc = fmaf(a.y, b.y, acc); acc = fmaf(a.z, b.z, acc); acc = fmaf(a.w, b.w, acc);
            }
            xn = fmaf(a.x, a.x, xn); xn = fmaf(a.y, a.y, xn); xn = fmaf(a.z, a.z, xn); xn = fmaf(a.w, a.w, xn);
        }
        v = METRIC == 0 ? acc : -acc;
    }
    float bv = v;
    uint32_t bi = id;
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) {
        const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
        const uint32_t oi = __shfl_xor_sync(0xffffffffu, bi, o);
        if (ov < bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
        xn = fmaxf(xn, __shfl_xor_sync(0xffffffffu, xn, o));
    }
    if (live && l8 == 0) {
        // every centre outside the nominations has an approximate score >= the smaller 4th-best of the two column halves
        const float a_bound = fminf(nom_val[r * AT_NOM + 3], nom_val[r * AT_NOM + 7]);
        const float cn_max = __uint_as_float(*cn_max_bits);
        // TF32 truncates both operands (< 2^-10 each): |x.c| off by < 2^-9 |x||c|, + the fp32 accumulation of dim products
        // (5 % on top); the L2 score carries -2 x.c.  slack: rounding of the exact chains themselves (dim * 2^-24 relative)
        const float eps = (METRIC == 0 ? 2.f : 1.f) * (1.05f / 512.f) * sqrtf(xn * cn_max);
        const float slack = 1e-4f * (xn + cn_max);
        const float others = METRIC == 0 ? a_bound + xn : a_bound;       // lower bound of every other centre's true score, before eps
        const bool certain = a_bound == INF || bv < others - eps - slack;   // (a_bound == inf: fewer than 4 centres in a half, all nominated)
        out_list[r] = bi;
        if (out_val) out_val[r] = bv;
        if (!certain) unc_idx[atomicAdd(unc_count, 1u)] = (uint32_t) r;
    }
}

// ---------------------------------------------------------------------------------------------
// host

size_t rk_assign_tc_smem(int lists)
{
    const int nct = (lists + AT_N - 1) / AT_N;
    return (size_t) AT_STAGES * AT_STAGE_BYTES + 8 * (2 * AT_STAGES + 4) + 16 + (size_t) nct * AT_N * 4 + 1024;
}

bool rk_assign_tc_supported(size_t nrows, int lists, int dim_pad)
{
    return nrows >= 4096 && lists >= 256 && dim_pad % 4 == 0 && rk_assign_tc_smem(lists) <= 227 * 1024;
}

// scratch bytes for chunks of `chunk` rows
size_t rk_assign_tc_scratch_bytes(size_t chunk, int lists)
{
    const size_t nct = (size_t) (lists + AT_N - 1) / AT_N;
    return 256 + nct * AT_N * 4 + chunk * AT_NOM * 8 + chunk * 4 + 256;
}

// out_list[r] (and out_val[r]) for rows [0, nrows) — identical to rk_launch_assign's.  `scratch` holds
// rk_assign_tc_scratch_bytes(chunk, lists); *n_fallback receives the rows the proof sent to rk_assign_kernel.
// Synchronises `stream` once per chunk.
cudaError_t rk_assign_tc(const float *X, size_t nrows, const float *C, int lists, int dim_pad, int metric, uint32_t *out_list, float *out_val,
                         void *scratch, size_t chunk, int sm_count, cudaStream_t stream, size_t *n_fallback)
{
    cudaError_t e;
    const int nct = (lists + AT_N - 1) / AT_N;
    unsigned char *sb = reinterpret_cast<unsigned char *>(scratch);
    unsigned *d_ctl = reinterpret_cast<unsigned *>(sb);                   // [0] max centre norm bits, [1] uncertain count
    float *d_cn = reinterpret_cast<float *>(sb + 256);
    float *d_nv = reinterpret_cast<float *>(sb + 256 + (size_t) nct * AT_N * 4);
    uint32_t *d_ni = reinterpret_cast<uint32_t *>(d_nv + chunk * AT_NOM);
    uint32_t *d_unc = d_ni + chunk * AT_NOM;
    if ((e = cudaMemsetAsync(d_ctl, 0, 256, stream)) != cudaSuccess) return e;
    rk_assign_cnorm_kernel<<<(nct * AT_N * 32 + 255) / 256, 256, 0, stream>>>(C, lists, nct * AT_N, dim_pad, d_cn);
    if ((e = rk_launch_max_f32(d_cn, (size_t) lists, d_ctl, stream)) != cudaSuccess) return e;
    CUtensorMap tmc;
    if ((e = rk_make_tensor_map_2d(&tmc, C, (uint64_t) lists, dim_pad, AT_N)) != cudaSuccess) return e;
    const size_t smem = rk_assign_tc_smem(lists);
    auto kern = metric == 0 ? rk_assign_tc_kernel<0> : rk_assign_tc_kernel<1>;
    static bool attr_set[64][2];
    int dev_ = 0;
    cudaGetDevice(&dev_);
    if (!attr_set[dev_ & 63][metric & 1]) {
        if ((e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024)) != cudaSuccess) return e;
        attr_set[dev_ & 63][metric & 1] = true;
    }
    size_t fallback = 0;
    for (size_t r0 = 0; r0 < nrows; r0 += chunk) {
        const size_t n = std::min(chunk, nrows - r0);
        const float *Xc = X + r0 * (size_t) dim_pad;
        CUtensorMap tmx;
        if ((e = rk_make_tensor_map_2d(&tmx, Xc, (uint64_t) n, dim_pad, AT_M)) != cudaSuccess) return e;
        RkAssignTcParams p;
        p.nrows = n; p.lists = lists; p.dim_pad = dim_pad; p.kchunks = (dim_pad + AT_KC - 1) / AT_KC;
        p.ntiles = (int) ((n + AT_M - 1) / AT_M); p.nct = nct;
        p.cnorm = d_cn; p.nom_val = d_nv; p.nom_id = d_ni;
        kern<<<std::min(sm_count, p.ntiles), AT_THREADS, smem, stream>>>(tmx, tmc, p);
        if ((e = cudaGetLastError()) != cudaSuccess) return e;
        if ((e = cudaMemsetAsync(d_ctl + 1, 0, 4, stream)) != cudaSuccess) return e;
        const unsigned rgrid = (unsigned) ((n * 8 + 255) / 256);
        if (metric == 0) rk_assign_rescore_kernel<0><<<rgrid, 256, 0, stream>>>(Xc, n, C, dim_pad, d_nv, d_ni, d_ctl, out_list + r0, out_val ? out_val + r0 : nullptr, d_unc, d_ctl + 1);
        else rk_assign_rescore_kernel<1><<<rgrid, 256, 0, stream>>>(Xc, n, C, dim_pad, d_nv, d_ni, d_ctl, out_list + r0, out_val ? out_val + r0 : nullptr, d_unc, d_ctl + 1);
        if ((e = cudaGetLastError()) != cudaSuccess) return e;
        unsigned cnt = 0;
        if ((e = cudaMemcpyAsync(&cnt, d_ctl + 1, 4, cudaMemcpyDeviceToHost, stream)) != cudaSuccess) return e;
        if ((e = cudaStreamSynchronize(stream)) != cudaSuccess) return e;
        if (cnt) {
            // the rows the proof refused: rk_assign_kernel itself, through the index list
            if ((e = rk_launch_assign(Xc, cnt, C, lists, dim_pad, metric, out_list + r0, out_val ? out_val + r0 : nullptr, stream, d_unc)) != cudaSuccess) return e;
            fallback += cnt;
        }
    }
    if (n_fallback) *n_fallback = fallback;
    return cudaSuccess;
}
