// MUL_MAT_ID (mixture-of-experts routing) around the TCQ4 MUL_MAT kernels: the row gather / scatter either side of the
// per-expert GEMMs.  Replaces the two get_rows_cuda calls of the reference's generic path (ggml-cuda.cu:2376-2380,
// 2431-2435); the routing itself (ids -> rows sorted by expert) is host code in tcq4_api.cu, as it is there -- or, for the
// capturable path of round 2, the one-CTA kernel at the end of this file.
//
// HBM-bound copies: one float4 per thread and step, rows addressed through an int32 index on either side.
#include "tcq4_common.cuh"
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

// ---- device-side routing for MUL_MAT_ID with many tokens (round 2) ----------------------------------------------------
// The reference reads the ids on the host and sorts rows by expert there (ggml-cuda.cu:2349-2368: a stream synchronise, not
// capturable).  Here ONE CTA does it on the device: rows (token t, slot s) are binned by expert into CHUNKS of up to 64 rows
// -- the token width of the batched-decode GEMM (tcq4_gemm4.cu) -- laid out at 64-aligned VIRTUAL rows, so that a chunk is
// half of one 128-row activation unit of K1.  Outputs (workspace):
//   n_chunks            int
//   chunk_expert[c]     expert of chunk c           (c < n_chunks <= n_expert + R / 64)
//   src_map[v]          row of x for virtual row v  (t * x_slots + s % x_slots), -1: empty
//   dst_map[v]          row of y for virtual row v  (t * n_used + s),            -1: empty
// The position of a row inside its expert's chunks comes from an atomic counter: it varies from run to run, the RESULT of a
// row does not (every row is computed on its own; blocks are never cut in K in this mode).
constexpr int MOE_ROUTE_THREADS = 1024;
__global__ void __launch_bounds__(MOE_ROUTE_THREADS)
tcq4_moe_route_kernel(const int32_t *__restrict__ ids, long long ids_stride, int n_used, long long n_tokens, int n_expert,
                      int x_slots, int max_chunks, int *__restrict__ n_chunks, int *__restrict__ chunk_expert,
                      int32_t *__restrict__ src_map, int32_t *__restrict__ dst_map, int32_t *__restrict__ pos_tmp,
                      int *__restrict__ err_flag) {
    extern __shared__ int sm_i[];
    int *cnt = sm_i;                    // [n_expert] rows per expert
    int *chunk0 = sm_i + n_expert;      // [n_expert + 1] first chunk of each expert
    pdl_launch_dependents();
    pdl_wait();                         // the ids come from the router kernel in front of this one
    const long long R = n_tokens * n_used;
    for (int e = threadIdx.x; e < n_expert; e += blockDim.x) cnt[e] = 0;
    __syncthreads();
    for (long long r = threadIdx.x; r < R; r += blockDim.x) {
        const long long t = r / n_used;
        const int s = (int)(r - t * n_used);
        const int e = ids[t * ids_stride + s];
        int p = -1;
        if (e >= 0 && e < n_expert) p = atomicAdd(&cnt[e], 1);
        else if (err_flag) *err_flag = 1;          // as the decode path does: flag it, skip the row
        pos_tmp[r] = p;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        int c = 0;
        for (int e = 0; e < n_expert; e++) {
            chunk0[e] = c;
            c += (cnt[e] + 63) >> 6;
        }
        chunk0[n_expert] = c;
        *n_chunks = c < max_chunks ? c : max_chunks;      // (c <= n_expert + R / 64 = max_chunks by construction)
    }
    __syncthreads();
    const int nc = chunk0[n_expert];
    for (long long v = threadIdx.x; v < (long long)nc * 64; v += blockDim.x) {
        src_map[v] = -1;
        dst_map[v] = -1;
    }
    for (int e = threadIdx.x; e < n_expert; e += blockDim.x)
        for (int c = chunk0[e]; c < chunk0[e + 1]; c++) chunk_expert[c] = e;
    __syncthreads();
    for (long long r = threadIdx.x; r < R; r += blockDim.x) {
        const int p = pos_tmp[r];
        if (p < 0) continue;
        const long long t = r / n_used;
        const int s = (int)(r - t * n_used);
        const int e = ids[t * ids_stride + s];
        const long long v = ((long long)chunk0[e] + (p >> 6)) * 64 + (p & 63);
        src_map[v] = (int32_t)(t * x_slots + s % x_slots);
        dst_map[v] = (int32_t)r;
    }
}

cudaError_t launch_moe_route(const int32_t *ids, int64_t ids_stride, int64_t n_used, int64_t n_tokens, int64_t n_expert,
                             int64_t x_slots, int max_chunks, int *n_chunks, int *chunk_expert, int32_t *src_map,
                             int32_t *dst_map, int32_t *pos_tmp, int *err_flag, cudaStream_t stream) {
    const size_t smem = (size_t)(2 * n_expert + 1) * sizeof(int);
    if (smem > 48 * 1024) return cudaErrorInvalidValue;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(1);
    cfg.blockDim = dim3(MOE_ROUTE_THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    return cudaLaunchKernelEx(&cfg, tcq4_moe_route_kernel, ids, (long long)ids_stride, (int)n_used, (long long)n_tokens,
                              (int)n_expert, (int)x_slots, max_chunks, n_chunks, chunk_expert, src_map, dst_map, pos_tmp, err_flag);
}

}  // namespace tcq4
