// rk_kernels.h — internal launch API between the host side (rk_api.cu) and the kernels.
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include "rk_common.cuh"

int rk_device_sm_count(int device);
// Pipeline geometry of the scan kernel for a padded dimension.
void rk_scan_geometry(int dim_pad, int *tile_rows, int *nstages, size_t *smem);
size_t rk_scan_smem_bytes(int row_bytes, int tile_rows, int nstages);
cudaError_t rk_launch_scan(const RkScanParams &p, int device, cudaStream_t stream, int grid_override);
cudaError_t rk_launch_select(const RkSelectParams &p, int nq, cudaStream_t stream);

// small utility kernels (rk_util.cu)
cudaError_t rk_launch_gather_rows(const float *src, float *dst, const uint32_t *perm, size_t n, int dim_pad,
                                  cudaStream_t stream);
cudaError_t rk_launch_pad_rows(const float *src, float *dst, size_t n, int dim, int dim_pad, cudaStream_t stream);
// counts NaN / Inf elements of a device array (synchronous; result in *out)
cudaError_t rk_count_nonfinite(const float *d_x, size_t n, unsigned long long *out, cudaStream_t stream);
