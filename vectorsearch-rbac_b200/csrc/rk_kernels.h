// rk_kernels.h — internal launch API between the host side (rk_api.cu) and the kernels.
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include "rk_common.cuh"

int rk_device_sm_count(int device);
// Pipeline geometry of the scan kernel for a padded dimension.
void rk_scan_geometry(int dim_pad, int *tile_rows, int *nstages, size_t *smem);
size_t rk_scan_smem_bytes(int row_bytes, int tile_rows, int nstages);
// inline_q_host != NULL: fused single-query launch whose query (p.row_bytes <= RK_INLINE_Q * 4 bytes, host memory) rides in
// the kernel parameters
#define RK_INLINE_Q 1024
cudaError_t rk_launch_scan(const RkScanParams &p, int device, cudaStream_t stream, int grid_override,
                           const float *inline_q_host = nullptr);
cudaError_t rk_launch_select(const RkSelectParams &p, int nq, cudaStream_t stream);

// small utility kernels (rk_util.cu)
cudaError_t rk_launch_gather_rows(const float *src, float *dst, const uint32_t *perm, size_t n, int dim_pad,
                                  cudaStream_t stream);
cudaError_t rk_launch_scatter_chunks(const void *src, void *dst, const uint32_t *idx, size_t n, int words, cudaStream_t stream);
cudaError_t rk_launch_pad_rows(const float *src, float *dst, size_t n, int dim, int dim_pad, cudaStream_t stream);
// counts NaN / Inf elements of a device array (synchronous; result in *out)
cudaError_t rk_count_nonfinite(const float *d_x, size_t n, unsigned long long *out, cudaStream_t stream);

cudaError_t rk_launch_flag_nonfinite(const float *d_x, size_t n, uint32_t *d_flag, cudaStream_t stream);

// IVFFlat (rk_ivf_kernels.cu)
cudaError_t rk_launch_assign(const float *X, size_t nrows, const float *C, int lists, int dim_pad, int metric,
                             uint32_t *out_list, float *out_val, cudaStream_t stream, const uint32_t *idx = nullptr);
// the same assignment through the tensor cores (rk_assign_tc.cu): TF32 nomination, exact re-score + proof, refused rows through
// rk_launch_assign — identical results
bool rk_assign_tc_supported(size_t nrows, int lists, int dim_pad);
size_t rk_assign_tc_scratch_bytes(size_t chunk, int lists);
cudaError_t rk_assign_tc(const float *X, size_t nrows, const float *C, int lists, int dim_pad, int metric, uint32_t *out_list, float *out_val,
                         void *scratch, size_t chunk, int sm_count, cudaStream_t stream, size_t *n_fallback);
// k-means++ seeding (ivfkmeans.c:21-89): chosen[lists] = sample rows to start the centres from
uint32_t rk_kpp_blocks(uint32_t ns);
cudaError_t rk_launch_kmeanspp(const float *S, uint32_t ns, int dim_pad, int lists, int spherical, uint64_t seed, float *weight,
                               double *blocksum, uint32_t *chosen, cudaStream_t stream);
cudaError_t rk_launch_km_accumulate(const float *S, size_t ns, int dim_pad, const uint32_t *label, float *sums, uint32_t *counts,
                                    cudaStream_t stream);
cudaError_t rk_launch_km_finalize(float *C, const float *sums, const uint32_t *counts, int lists, int dim_pad, const float *S,
                                  size_t ns, uint64_t seed, int normalize, cudaStream_t stream);
cudaError_t rk_launch_km_normalize_rows(float *S, size_t ns, int dim_pad, cudaStream_t stream);
cudaError_t rk_launch_count_changed(const uint32_t *a, const uint32_t *b, size_t n, unsigned long long *d_out, cudaStream_t stream);
cudaError_t rk_launch_gather_u32(const uint32_t *src, uint32_t *dst, const uint32_t *perm, size_t n, cudaStream_t stream);
cudaError_t rk_launch_gather_u64(const uint64_t *src, uint64_t *dst, const uint32_t *perm, size_t n, cudaStream_t stream);
size_t rk_ivf_probe_smem(int lists, int dim_pad);
cudaError_t rk_launch_ivf_probe(const float *Q, int nq, const float *C, int lists, int dim_pad, int metric, int null_query,
                                int probe_offset, int probes, uint32_t *out_list, uint32_t *keys_ws, cudaStream_t stream);
size_t rk_ivf_probe_keys_bytes(int nq, int lists);
cudaError_t rk_launch_ivf_plan(const uint32_t *probe_list, const uint32_t *list_start, int nq, int probes, int chunks,
                               uint32_t chunk_rows, uint32_t kind, RkSegment *segs, cudaStream_t stream);

// batched tensor-core engine (rk_bscan.cu)
size_t rk_bscan_smem_bytes(int mask_words);
int rk_bscan_lists_per_query();   // candidate lists one unit writes per query (one per epilogue set)
size_t rk_bscan_scratch_bytes(int grid);
cudaError_t rk_launch_bscan(const RkBScanParams &p, const float *x, const float *q, uint64_t nq, int grid, cudaStream_t stream);
// CTA-pair variant (rk_bscan2.cu): units come in pairs (2p, 2p + 1) sharing a tile of up to 256 queries;
// cta_off / cta_units schedule PAIRS onto clusters; grid = 2 x clusters
size_t rk_bscan2_smem_bytes(int mask_words);
size_t rk_bscan2_scratch_bytes(int grid);
// x / q: fp32 [rows][dim_pad], or with p.bf16 their bf16 shadow copies [rows][row_elems_b]
cudaError_t rk_launch_bscan2(const RkBScanParams &p, const void *x, const void *q, uint64_t nq, int grid, cudaStream_t stream, int row_elems_b = 0);
cudaError_t rk_launch_rows_to_bf16(const float *x, size_t n, int dim_pad, int dim_b, void *out, cudaStream_t stream);
cudaError_t rk_launch_rescore(const RkRescoreParams &p, int nq, cudaStream_t stream);
cudaError_t rk_launch_row_norms(const float *x, size_t n, int dim_pad, float *out, cudaStream_t stream);
cudaError_t rk_launch_rowmask(const uint32_t *rowdoc, const uint64_t *doc_mask, size_t n, int W, uint64_t *out, cudaStream_t stream);
cudaError_t rk_launch_rowmask2(const uint32_t *rowdoc, const uint64_t *doc_mask, size_t n, size_t n_pad, int W, uint64_t *out,
                               cudaStream_t stream);
cudaError_t rk_launch_tf32_probe(const float *x, size_t n, unsigned long long *d_out2, cudaStream_t stream);
cudaError_t rk_launch_max_f32(const float *x, size_t n, unsigned *d_out, cudaStream_t stream);
