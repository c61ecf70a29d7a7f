// rk_util.cu — layout helpers run once at load / index-build time (not on the query path).
#include "rk_common.cuh"
#include <algorithm>

#include "rk_kernels.h"

// dst[i][:] = src[perm[i]][:]  (rows of dim_pad floats, float4 granularity)
__global__ void rk_gather_rows_kernel(const float4 *__restrict__ src, float4 *__restrict__ dst,
                                      const uint32_t *__restrict__ perm, size_t n, int dim4)
{
    const size_t total = n * (size_t) dim4;
    for (size_t i = blockIdx.x * (size_t) blockDim.x + threadIdx.x; i < total; i += (size_t) gridDim.x * blockDim.x) {
        const size_t r = i / dim4;
        const int c = (int) (i % dim4);
        dst[i] = src[(size_t) perm[r] * dim4 + c];
    }
}

// dst[n][dim_pad] = src[n][dim] zero-padded
__global__ void rk_pad_rows_kernel(const float *__restrict__ src, float *__restrict__ dst, size_t n, int dim, int dim_pad)
{
    const size_t total = n * (size_t) dim_pad;
    for (size_t i = blockIdx.x * (size_t) blockDim.x + threadIdx.x; i < total; i += (size_t) gridDim.x * blockDim.x) {
        const size_t r = i / dim_pad;
        const int c = (int) (i % dim_pad);
        dst[i] = c < dim ? src[r * dim + c] : 0.0f;
    }
}

cudaError_t rk_launch_gather_rows(const float *src, float *dst, const uint32_t *perm, size_t n, int dim_pad,
                                  cudaStream_t stream)
{
    if (n == 0) return cudaSuccess;
    rk_gather_rows_kernel<<<148 * 8, 256, 0, stream>>>(reinterpret_cast<const float4 *>(src),
                                                       reinterpret_cast<float4 *>(dst), perm, n, dim_pad / 4);
    return cudaGetLastError();
}

// dst[idx[i]][:] = src[i][:]  (chunks of `words` 32-bit words)
__global__ void rk_scatter_chunks_kernel(const uint32_t *__restrict__ src, uint32_t *__restrict__ dst, const uint32_t *__restrict__ idx,
                                         size_t n, int words)
{
    const size_t total = n * (size_t) words;
    for (size_t i = blockIdx.x * (size_t) blockDim.x + threadIdx.x; i < total; i += (size_t) gridDim.x * blockDim.x) {
        const size_t r = i / words;
        dst[(size_t) idx[r] * words + (i - r * words)] = src[i];
    }
}

cudaError_t rk_launch_scatter_chunks(const void *src, void *dst, const uint32_t *idx, size_t n, int words, cudaStream_t stream)
{
    if (n == 0 || words == 0) return cudaSuccess;
    rk_scatter_chunks_kernel<<<(unsigned) std::min<size_t>(148 * 4, (n * words + 255) / 256), 256, 0, stream>>>(
        reinterpret_cast<const uint32_t *>(src), reinterpret_cast<uint32_t *>(dst), idx, n, words);
    return cudaGetLastError();
}

cudaError_t rk_launch_pad_rows(const float *src, float *dst, size_t n, int dim, int dim_pad, cudaStream_t stream)
{
    if (n == 0) return cudaSuccess;
    rk_pad_rows_kernel<<<148 * 8, 256, 0, stream>>>(src, dst, n, dim, dim_pad);
    return cudaGetLastError();
}

__global__ void rk_count_nonfinite_kernel(const float *__restrict__ x, size_t n, unsigned long long *out)
{
    unsigned long long c = 0;
    for (size_t i = blockIdx.x * (size_t) blockDim.x + threadIdx.x; i < n; i += (size_t) gridDim.x * blockDim.x)
        c += !isfinite(x[i]);
    c = __reduce_add_sync(0xffffffffu, (unsigned) c);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd(out, c);
}

cudaError_t rk_count_nonfinite(const float *d_x, size_t n, unsigned long long *out, cudaStream_t stream)
{
    unsigned long long *d = nullptr;
    cudaError_t e = cudaMalloc(&d, sizeof *d);
    if (e != cudaSuccess) return e;
    cudaMemsetAsync(d, 0, sizeof *d, stream);
    rk_count_nonfinite_kernel<<<148 * 8, 256, 0, stream>>>(d_x, n, d);
    e = cudaMemcpyAsync(out, d, sizeof *d, cudaMemcpyDeviceToHost, stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(stream);
    cudaFree(d);
    return e;
}

// Asynchronous variant for the device-pointer query entry points: sets *flag when any element is NaN / Inf; the host
// reads the flag at the next rk_synchronize (vector_in's check, vector.c:102-113, without a sync on the search path).
__global__ void rk_flag_nonfinite_kernel(const float *__restrict__ x, size_t n, uint32_t *flag)
{
    unsigned c = 0;
    for (size_t i = blockIdx.x * (size_t) blockDim.x + threadIdx.x; i < n; i += (size_t) gridDim.x * blockDim.x)
        c |= (__float_as_uint(x[i]) & 0x7f800000u) == 0x7f800000u;
    if (__any_sync(0xffffffffu, c != 0) && (threadIdx.x & 31) == 0) atomicOr(flag, 1u);
}

cudaError_t rk_launch_flag_nonfinite(const float *d_x, size_t n, uint32_t *d_flag, cudaStream_t stream)
{
    if (n == 0) return cudaSuccess;
    const unsigned grid = (unsigned) std::min<size_t>((n + 2047) / 2048, 148 * 4);
    rk_flag_nonfinite_kernel<<<grid, 256, 0, stream>>>(d_x, n, d_flag);
    return cudaGetLastError();
}
