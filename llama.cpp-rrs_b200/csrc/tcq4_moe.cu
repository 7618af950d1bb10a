// MUL_MAT_ID (mixture-of-experts routing) around the TCQ4 MUL_MAT kernels: the row gather / scatter either side of the
// per-expert GEMMs.  Replaces the two get_rows_cuda calls of the reference's generic path (ggml-cuda.cu:2376-2380,
// 2431-2435); the routing itself (ids -> rows sorted by expert) is host code in tcq4_api.cu, as it is there.
//
// HBM-bound copies: one float4 per thread and step, rows addressed through an int32 index on either side.
#include "tcq4_kernels.h"

namespace tcq4 {

// dst[dst_idx ? dst_idx[r] : r][0..cols) = src[src_idx ? src_idx[r] : r][0..cols)      (cols % 4 == 0, 16-byte aligned rows)
__global__ void __launch_bounds__(256) tcq4_copy_rows_kernel(const float *__restrict__ src, const int32_t *__restrict__ src_idx,
                                                             float *__restrict__ dst, const int32_t *__restrict__ dst_idx,
                                                             long long rows, int cols4) {
    const long long total = rows * cols4;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const long long r = i / cols4;
        const int c = (int)(i - r * cols4);
        const long long rs = src_idx ? (long long)__ldg(src_idx + r) : r;
        const long long rd = dst_idx ? (long long)__ldg(dst_idx + r) : r;
        // plain cached load: with one activation row per token, the row is read once per expert the token uses
        const float4 v = __ldg(reinterpret_cast<const float4 *>(src) + rs * cols4 + c);
        reinterpret_cast<float4 *>(dst)[rd * cols4 + c] = v;
    }
}

cudaError_t launch_copy_rows(const float *src, const int32_t *src_idx, float *dst, const int32_t *dst_idx, int64_t rows,
                             int64_t cols, int num_sms, cudaStream_t stream) {
    if (rows <= 0) return cudaSuccess;
    if (cols <= 0 || (cols & 3)) return cudaErrorInvalidValue;
    const long long total = (long long)rows * (cols / 4);
    long long blocks = (total + 255) / 256;
    const long long cap = (long long)num_sms * 8;          // grid-stride above 8 resident CTAs per SM
    if (blocks > cap) blocks = cap;
    tcq4_copy_rows_kernel<<<(unsigned)blocks, 256, 0, stream>>>(src, src_idx, dst, dst_idx, (long long)rows, (int)(cols / 4));
    return cudaGetLastError();
}

}  // namespace tcq4
