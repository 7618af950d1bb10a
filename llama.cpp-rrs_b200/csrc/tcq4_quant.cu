// tcq4_quant.cu -- the two other producers / consumers of GGML_TYPE_TCQ4_K32 tensors:
//
//  K4  tcq4_quantize_tiles_kernel -- device weight quantiser, byte-identical to quantize_tcq4_tile
//      (/root/reference/ggml/src/ggml-quants.c:1528-1639) with apply_channel_perm_to_weights
//      (/root/reference/src/llama-quant.cpp:26-40) fused as a gather.  F32 [nrow][K] -> [ceil(nrow/8)][K/256] tiles.
//  K5  tcq4_get_rows_kernel -- GET_ROWS for a TCQ4 tensor: tile dequantisation of the selected rows
//      (k_get_rows_tcq4_tile, /root/reference/ggml/src/ggml-cuda/getrows.cu:96-171) fused with the inverse FWHT-256
//      the fork runs afterwards (getrows.cu:348-380; CPU: ggml/src/ggml-cpu/ops.cpp:4800-4824), one pass, no
//      intermediate row in global memory.
//
// Both are HBM-bound streaming kernels (K4: 4 B read + 0.578 B written per weight; K5: 0.578 B read + 4 B written per
// element) built on the same warp butterfly as the activation stage: one warp owns one 256-block, lane l holds elements
// l*8 .. l*8+7, stages h = 1, 2, 4 in registers and h = 8 .. 128 by shuffle, same (j, j+h) -> (x+y, x-y) pairs and stage
// order as ggml_fwht_impl (/root/reference/ggml/src/ggml-cpu/rrs.c:103-121), so every float is bit-identical to the
// scalar code.  No fast-math, no FMA contraction (-fmad=false).
#include "tcq4_common.cuh"
#include "tcq4_kernels.h"

namespace tcq4 {

// unnormalised FWHT-256 of one block held by a warp (layout: lane l, register r = element l*8 + r), then * 1/16.
// ggml_fwht_impl multiplies inside the last stage, (x +- y) * 0.0625f: the same single rounding-free scaling.
__device__ __forceinline__ void fwht256_warp(float (&v)[8], int lane) {
#pragma unroll
    for (int h = 1; h < 8; h <<= 1) {
#pragma unroll
        for (int r = 0; r < 8; r++) {
            if ((r & h) == 0) {
                const float a = v[r], b = v[r + h];
                v[r]     = a + b;
                v[r + h] = a - b;
            }
        }
    }
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const float sgn = (lane & d) ? -1.0f : 1.0f;       // upper lane: lower - upper = o + (-v)
#pragma unroll
        for (int r = 0; r < 8; r++) {
            const float o = __shfl_xor_sync(0xffffffffu, v[r], d);
            v[r] = __fmaf_rn(v[r], sgn, o);                // exact: v * (+-1) + o rounds once, like the add
        }
    }
#pragma unroll
    for (int r = 0; r < 8; r++) v[r] = v[r] * 0.0625f;
}

// roundf (half away from zero) from one FRND: rintf differs from roundf only on exact .5 ties (it goes to even), and
// x - rintf(x) is exact for the |x| < 2^22 that occur here, so the tie is detected exactly and repaired.
__device__ __forceinline__ float round_half_away(float x) {
    const float r = rintf(x);
    const float d = x - r;
    return (fabsf(d) == 0.5f) ? x + copysignf(0.5f, x) : r;      // tie: the neighbour away from zero
}

// ---------------------------------------------------------------------------------------------------------------
// K4: one CTA of 8 warps builds one tile (warp c = channel c) in shared memory and writes it out as 74 x 16 bytes.
// ---------------------------------------------------------------------------------------------------------------
constexpr int QZ_THREADS = 256;

__global__ void __launch_bounds__(QZ_THREADS)
tcq4_quantize_tiles_kernel(const float *__restrict__ w, const int32_t *__restrict__ perm, uint8_t *__restrict__ out,
                           long long nrow, int K, long long ntiles) {
    __shared__ __align__(16) Tile tile;
    const int lane = threadIdx.x & 31, c = threadIdx.x >> 5;
    const int g = lane >> 2, s = lane & 3;
    const int KT = K / TILE_K;

    // tile counter kept as (n-tile, k-tile) and advanced without a division per tile
    long long nt = blockIdx.x / KT;
    int kt = (int)(blockIdx.x - nt * KT);
    const long long nt_step = gridDim.x / KT;
    const int kt_step = (int)(gridDim.x - nt_step * KT);
    for (long long t = blockIdx.x; t < ntiles; t += gridDim.x) {
        const long long row = nt * TILE_CH + c;
        float v[8];
        if (row < nrow) {                                          // rows past the end are zero (ggml-quants.c:1565-1568)
            const float *src = w + (size_t)row * K;
            const int base = kt * TILE_K + lane * 8;
            if (perm == nullptr) {
                const float4 a = __ldcs(reinterpret_cast<const float4 *>(src + base));
                const float4 b = __ldcs(reinterpret_cast<const float4 *>(src + base) + 1);
                v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w;
                v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
            } else {                                               // W'[n][j] = W[n][perm[j]]  (llama-quant.cpp:36)
                const int4 p0 = __ldg(reinterpret_cast<const int4 *>(perm + base));
                const int4 p1 = __ldg(reinterpret_cast<const int4 *>(perm + base) + 1);
                v[0] = __ldg(src + p0.x); v[1] = __ldg(src + p0.y); v[2] = __ldg(src + p0.z); v[3] = __ldg(src + p0.w);
                v[4] = __ldg(src + p1.x); v[5] = __ldg(src + p1.y); v[6] = __ldg(src + p1.z); v[7] = __ldg(src + p1.w);
            }
        } else {
#pragma unroll
            for (int r = 0; r < 8; r++) v[r] = 0.0f;
        }
        fwht256_warp(v, lane);

        // per (channel, group) scale: absmax / 7, < 1e-10 -> 1  (ggml-quants.c:1576-1588); the lane's group is g
        float mx = 0.0f;
#pragma unroll
        for (int r = 0; r < 8; r++) {
            const float a = fabsf(v[r]);
            if (a > mx) mx = a;
        }
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
        float scale = __fdiv_rn(mx, 7.0f);
        if (scale < 1e-10f) scale = 1.0f;
        // S = max over the 8 groups (ggml-quants.c:1594-1600); scales are positive, so the "> 0 else 1" rule is moot
        float S = scale;
        S = fmaxf(S, __shfl_xor_sync(0xffffffffu, S, 4));
        S = fmaxf(S, __shfl_xor_sync(0xffffffffu, S, 8));
        S = fmaxf(S, __shfl_xor_sync(0xffffffffu, S, 16));
        const float sc_f = __fdiv_rn(scale, S) * 127.0f;
        const int sc = (int)fmaxf(-127.0f, fminf(127.0f, round_half_away(sc_f)));
        const float inv = __fdiv_rn(1.0f, scale);
        uint32_t reg = 0;
#pragma unroll
        for (int r = 0; r < 8; r++) {
            const int q = (int)fmaxf(-8.0f, fminf(7.0f, round_half_away(v[r] * inv)));
            reg |= ((uint32_t)q & 0xFu) << (4 * r);
        }
        // fragment order (ggml-quants.c:1381-1400): word c*4 + s of group g holds elements s*8 .. s*8+7 of channel c
        reinterpret_cast<uint32_t *>(tile.tiles[g])[c * 4 + s] = reg;
        if (s == 0) {
            tile.sc[c][g] = (int8_t)sc;
            tile.zc[c][g] = 0;
        }
        if (lane == 0) {
            tile.S[c] = __float2half_rn(S);
            tile.Z[c] = __float2half_rn(0.0f);
        }
        __syncthreads();
        if (threadIdx.x < TILE_BYTES / 16)
            __stcs(reinterpret_cast<uint4 *>(out + (size_t)t * TILE_BYTES) + threadIdx.x,
                   reinterpret_cast<const uint4 *>(&tile)[threadIdx.x]);
        __syncthreads();
        nt += nt_step;
        kt += kt_step;
        if (kt >= KT) { kt -= KT; nt++; }
    }
}

cudaError_t launch_quantize_tiles(const float *w, const int32_t *perm, void *tiles, int64_t nrow, int64_t K, int num_sms,
                                  cudaStream_t stream) {
    const int64_t ntiles = ((nrow + TILE_CH - 1) / TILE_CH) * (K / TILE_K);
    if (ntiles == 0) return cudaSuccess;
    const int64_t cap = (int64_t)num_sms * 8;                     // 8 CTAs x 8 warps per SM, then grid-stride
    const int grid = (int)(ntiles < cap ? ntiles : cap);
    tcq4_quantize_tiles_kernel<<<grid, QZ_THREADS, 0, stream>>>(w, perm, reinterpret_cast<uint8_t *>(tiles),
                                                                (long long)nrow, (int)K, (long long)ntiles);
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------------------------
// K5: one warp per (selected row, 256-block): the lane's 8 elements are ONE 32-bit word of the tile
// (channel c, group l/4, slice l%4), dequantised as scale * q + zero with scale = S*sc/127, zero = Z*zc/127
// (ggml-quants.c:1425-1463, getrows.cu:137-163), then the FWHT-256 and a coalesced 32-byte store per lane.
// A row index outside [0, nrows_src) yields a zero row and raises the error flag (the reference reads out of bounds).
// ---------------------------------------------------------------------------------------------------------------
constexpr int GR_WARPS = 8;

__global__ void __launch_bounds__(GR_WARPS * 32)
tcq4_get_rows_kernel(const uint8_t *__restrict__ tiles, const int32_t *__restrict__ rows, float *__restrict__ dst,
                     long long n_idx, long long nrows_src, int K, long long dst_stride, int *__restrict__ err_flag) {
    const int lane = threadIdx.x & 31;
    const int KT = K / TILE_K;
    const long long nblocks = n_idx * KT;
    const long long warp0 = (long long)blockIdx.x * GR_WARPS + (threadIdx.x >> 5);
    const long long nwarps = (long long)gridDim.x * GR_WARPS;
    const int g = lane >> 2, s = lane & 3;

    for (long long b = warp0; b < nblocks; b += nwarps) {
        const long long i = b / KT;
        const int kt = (int)(b - i * KT);
        const long long r = rows[i];
        float v[8];
        if (r < 0 || r >= nrows_src) {
            if (lane == 0 && err_flag) atomicExch(err_flag, 1);
#pragma unroll
            for (int j = 0; j < 8; j++) v[j] = 0.0f;
        } else {
            const int c = (int)(r & 7);
            const uint8_t *tp = tiles + ((size_t)(r >> 3) * KT + kt) * TILE_BYTES;
            const uint32_t word = __ldg(reinterpret_cast<const uint32_t *>(tp) + g * 32 + c * 4 + s);
            const float S = __half2float(__ldg(reinterpret_cast<const __half *>(tp + TILE_OFF_S) + c));
            const float Z = __half2float(__ldg(reinterpret_cast<const __half *>(tp + TILE_OFF_Z) + c));
            const int sc = (int)(int8_t)__ldg(tp + TILE_OFF_SC + c * 8 + g);
            const int zc = (int)(int8_t)__ldg(tp + TILE_OFF_ZC + c * 8 + g);
            const float scale = __fdiv_rn(S * (float)sc, 127.0f);
            const float zero = __fdiv_rn(Z * (float)zc, 127.0f);
#pragma unroll
            for (int j = 0; j < 8; j++) {
                const int q = ((int)(word << (28 - 4 * j))) >> 28;           // sign-extended nibble j
                v[j] = scale * (float)q + zero;
            }
            fwht256_warp(v, lane);
        }
        float4 *o = reinterpret_cast<float4 *>(dst + (size_t)i * dst_stride + (size_t)kt * TILE_K + lane * 8);
        __stcs(o, make_float4(v[0], v[1], v[2], v[3]));
        __stcs(o + 1, make_float4(v[4], v[5], v[6], v[7]));
    }
}

cudaError_t launch_get_rows(const void *tiles, const int32_t *rows, float *dst, int64_t n_idx, int64_t nrows_src,
                            int64_t K, int64_t dst_stride, int *err_flag, int num_sms, cudaStream_t stream) {
    const int64_t nblocks = n_idx * (K / TILE_K);
    if (nblocks == 0) return cudaSuccess;
    const int64_t ctas = (nblocks + GR_WARPS - 1) / GR_WARPS;
    const int64_t cap = (int64_t)num_sms * 8;
    tcq4_get_rows_kernel<<<(int)(ctas < cap ? ctas : cap), GR_WARPS * 32, 0, stream>>>(
        reinterpret_cast<const uint8_t *>(tiles), rows, dst, (long long)n_idx, (long long)nrows_src, (int)K,
        (long long)dst_stride, err_flag);
    return cudaGetLastError();
}

}  // namespace tcq4
