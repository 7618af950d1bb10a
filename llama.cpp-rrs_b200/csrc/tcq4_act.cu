// tcq4_act.cu -- K1: Rotated-Runtime-Smooth activation stage (perm gather -> FWHT-256 -> 1/16 ->
// per-block max-scale -> int4 codes + group sums) as a warp-shuffle butterfly over coalesced
// 128-bit loads.  One warp owns one 256-block: 8 floats per lane (two LDG.128), FWHT stages
// h=1,2,4 in registers and h=8..128 by __shfl_xor_sync; no shared memory, no block barrier.
//
// Replaces tcq4_rrs_fused_activation_tc_kernel (/root/reference/ggml/src/ggml-cuda/tcq4_k32.cu:181-277,
// one 256-thread CTA + 1 KB smem + ~14 __syncthreads per block).
//
// Two output formats:
//   MODE_REF   : block_rrs_int4_tc [M][KT] (160 B), byte-compatible with the reference
//   MODE_STAGE : s8 codes in the UMMA K-major layout the tcgen05 GEMM bulk-copies (tcq4_common.cuh)
#include <cstdlib>

#include "tcq4_common.cuh"
#include "tcq4_kernels.h"

namespace tcq4 {

constexpr int ACT_WARPS_PER_CTA = 8;

enum { MODE_REF = 0, MODE_STAGE = 1, MODE_NIB = 2 };

// one element of the fused input transform (see ActInput): x is what K1 would have read from the op's output
template <int IN>
__device__ __forceinline__ float act_input_value(const ActInput &in, float xv, const float *__restrict__ x2_row, float rs, int idx) {
    if (IN == ACT_IN_NORM) return (rs * xv) * __ldg(in.col_weight + idx);          // scale * x * mul, left to right (norm.cu:143-144)
    if (IN == ACT_IN_SWIGLU) return (xv / (1.0f + expf(-xv))) * __ldcg(x2_row + idx);      // op_silu(x) * g (unary.cu:266)
    return xv;
}

template <int MODE, int IN = ACT_IN_NONE>
__global__ void __launch_bounds__(ACT_WARPS_PER_CTA * 32)
tcq4_act_quant_kernel(const float *__restrict__ x, const int32_t *__restrict__ perm, int M, int K,
                      ActBlockTC *__restrict__ out_ref, ActStage st, const __grid_constant__ ZeroList zl,
                      const int32_t *__restrict__ row_map, const int *__restrict__ n_chunks, const ActInput in) {
    const int lane = threadIdx.x & 31;
    const int KT = K / TILE_K;
    const long long nblocks = (long long)M * KT;
    const long long warp0 = (long long)blockIdx.x * ACT_WARPS_PER_CTA + (threadIdx.x >> 5);
    const long long nwarps = (long long)gridDim.x * ACT_WARPS_PER_CTA;

    __shared__ __align__(16) float scratch_all[ACT_WARPS_PER_CTA * 256];
    float *scratch = scratch_all + (threadIdx.x >> 5) * 256;

    pdl_launch_dependents();     // the GEMM behind this kernel may set itself up (it waits before reading the codes)
    pdl_wait();                  // x and the staging arena belong to the previous kernel until here

    if (MODE == MODE_NIB && zl.n > 0) {      // the GEMM behind this launch accumulates into its outputs (K split) / counts arrivals
        const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int r = 0; r < zl.n; r++) {
            const long long c4 = zl.cols[r] >> 2, n4 = zl.rows[r] * c4;
            for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
                const long long row = i / c4;
                reinterpret_cast<float4 *>(zl.p[r] + row * zl.ld[r])[i - row * c4] = z;
            }
        }
    }

    // MUL_MAT_ID (device-routed, tcq4_moe.cu): m is a VIRTUAL row -- 64-row chunks of one expert each -- whose source row of x is
    // row_map[m] (-1: empty slot of a chunk); only the first *n_chunks chunks exist
    long long nblocks_rt = nblocks;
    if (MODE == MODE_NIB && row_map != nullptr) {
        const long long lim = (long long)__ldcg(n_chunks) * 64 * KT;
        if (lim < nblocks_rt) nblocks_rt = lim;
    }
    for (long long b = warp0; b < nblocks_rt; b += nwarps) {
        const int m = (int)(b / KT);
        const int kt = (int)(b - (long long)m * KT);
        long long m_src = m;
        if (MODE == MODE_NIB && row_map != nullptr) {
            m_src = __ldcg(row_map + m);
            if (m_src < 0) continue;
        }
        float v[8];
        // ld.global.cg: every element is read once, and under programmatic stream serialization a stale L1 line of x
        // could survive from an overlapping predecessor
        if (IN == ACT_IN_NONE) {
            rrs_load_block_cg(x + (size_t)m_src * K, perm, kt, lane, v);
        } else {      // the producing op fused into the load: same values the op's own kernel would have written
            const float *x_row = x + (size_t)m_src * K;
            const float *x2_row = IN == ACT_IN_SWIGLU ? in.x2 + (size_t)m_src * K : nullptr;
            const float rs = IN == ACT_IN_NORM ? __ldcg(in.row_scale + m_src) : 0.0f;
            const int base = kt * TILE_K + lane * 8;
            int idx[8];
            if (perm == nullptr) {
#pragma unroll
                for (int r = 0; r < 8; r++) idx[r] = base + r;
            } else {
                const int4 p0 = __ldg(reinterpret_cast<const int4 *>(perm + base));
                const int4 p1 = __ldg(reinterpret_cast<const int4 *>(perm + base) + 1);
                idx[0] = p0.x; idx[1] = p0.y; idx[2] = p0.z; idx[3] = p0.w;
                idx[4] = p1.x; idx[5] = p1.y; idx[6] = p1.z; idx[7] = p1.w;
            }
#pragma unroll
            for (int r = 0; r < 8; r++) v[r] = act_input_value<IN>(in, __ldcg(x_row + idx[r]), x2_row, rs, idx[r]);
        }
        const ActQuantT a = rrs_quantize_block_t<true>(v, lane, scratch);
        const uint2 e = rrs_codes_layout_a(a, lane, scratch);        // byte r = code of element lane*8 + r
        int gs = 0;                                                  // lane g < 8 keeps the sum of group g
#pragma unroll
        for (int g = 0; g < 8; g++) gs = (lane == g) ? a.gsum[g] : gs;

        if (MODE == MODE_REF) {
            // qs[t] = q[2t] | q[2t+1] << 4  (tcq4_k32.cu:266-270): lane l owns bytes l*4 .. l*4+3
            ActBlockTC *blk = out_ref + b;
            const uint32_t lo = e.x & 0x0F0F0F0Fu, hi = e.y & 0x0F0F0F0Fu;       // nibbles of elements 0-3 / 4-7
            const uint32_t p0 = lo | (lo >> 4);                                  // bytes 0,2: e0|e1<<4, e2|e3<<4
            const uint32_t p1 = hi | (hi >> 4);
            const uint32_t w = (p0 & 0xFFu) | ((p0 >> 8) & 0xFF00u) | ((p1 & 0xFFu) << 16) | ((p1 << 8) & 0xFF000000u);
            reinterpret_cast<uint32_t *>(blk->qs)[lane] = w;
            if (lane == 0) blk->smooth_scale = a.max_abs;
            if (lane < 8) blk->sum_q[lane] = (int16_t)gs;
            if (lane < 3) reinterpret_cast<uint32_t *>(blk->pad)[lane] = 0u;
        } else if (MODE == MODE_NIB) {
            // u = q + 8 as nibbles, element i of the lane's 8-slice at bits [4i, 4i+4): word (lane & 3) of group lane >> 2
            const uint32_t t0 = (e.x ^ 0x08080808u) & 0x0F0F0F0Fu, t1 = (e.y ^ 0x08080808u) & 0x0F0F0F0Fu;
            const uint32_t p0 = t0 | (t0 >> 4), p1 = t1 | (t1 >> 4);        // bytes 0 and 2 hold two nibbles each
            const uint32_t w = __byte_perm(p0, p1, 0x6420);
            const int mb = m >> 7, r = m & 127;
            int8_t *unit = st.codes + ((size_t)mb * KT + kt) * ACT_NIB_UNIT_BYTES;
            *reinterpret_cast<uint32_t *>(unit + (lane >> 2) * 2048 + r * 16 + (lane & 3) * 4) = w;
            if (lane == 0) {
                const float a7 = __fdiv_rn(a.max_abs, 7.0f);
                st.a7[b] = a7;
                st.a7t[((size_t)mb * KT + kt) * 128 + r] = a7;
            }
            if (lane < 8) st.sumq[b * 8 + lane] = (int16_t)gs;
        } else {
            // s8 codes, sigma order {0,2,4,6,1,3,5,7}; lane pair (2j, 2j+1) forms 16-B K-chunk j
            const uint32_t w0 = __byte_perm(e.x, e.y, 0x6420);       // e0 e2 e4 e6
            const uint32_t w1 = __byte_perm(e.x, e.y, 0x7531);       // e1 e3 e5 e7
            const int mb = m >> 7, r = m & 127;
            int8_t *unit = st.codes + ((size_t)mb * KT + kt) * ACT_UNIT_BYTES;
            uint2 *dst = reinterpret_cast<uint2 *>(unit + (lane >> 1) * ACT_LBO + r * 16 + (lane & 1) * 8);
            *dst = make_uint2(w0, w1);
            if (lane == 0) st.a7[b] = __fdiv_rn(a.max_abs, 7.0f);
            if (lane < 8) st.sumq[b * 8 + lane] = (int16_t)gs;
        }
    }
}

static thread_local const ActInput *tl_pending_in = nullptr;
void set_pending_act_input(const ActInput *in) { tl_pending_in = in; }

static int act_grid(int64_t nblocks, int num_sms) {
    const int64_t ctas = (nblocks + ACT_WARPS_PER_CTA - 1) / ACT_WARPS_PER_CTA;
    const int64_t cap = (int64_t)num_sms * 8;     // 8 CTAs x 8 warps = full occupancy, then grid-stride
    return (int)(ctas < cap ? ctas : cap);
}

template <typename Kern>
static cudaError_t launch_act(Kern kern, int grid, cudaStream_t stream, const float *x, const int32_t *perm, int M, int K,
                              ActBlockTC *out_ref, ActStage st, int which = 0, const ZeroList *zlp = nullptr,
                              const int32_t *row_map = nullptr, const int *n_chunks = nullptr) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3(ACT_WARPS_PER_CTA * 32);
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    // Dependent launch is OFF by default (measured round 1: no gain for prefill, -8 % for the decode chain once the
    // kernels' own prologues had been shortened); TCQ4_PDL=1 turns it on for experiments -- then x must be read
    // with ld.global.cg (rrs_load_block_cg).
    // The nibble-staging launch (which == 2) feeds tcq4_gemm3.cu, whose prologue and first weight copies run under this
    // kernel's tail: dependent launch is on for that pair unless TCQ4_NO_PDL is set.
    static const bool pdl = getenv("TCQ4_PDL") != nullptr, no_pdl = getenv("TCQ4_NO_PDL") != nullptr;
    cfg.numAttrs = (pdl || (which == 2 && !no_pdl)) ? 1 : 0;
    // same L1/shared split as the GEMM that follows (no SM reconfiguration between the two launches)
    static bool carve_ref = false, carve_stage = false, carve_nib = false;
    bool &done = (out_ref != nullptr) ? carve_ref : (which == 2 ? carve_nib : carve_stage);
    if (!done) {
        cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, (int)cudaSharedmemCarveoutMaxShared);
        done = true;
    }
    ZeroList zl{};
    if (zlp) zl = *zlp;
    ActInput in{};
    return cudaLaunchKernelEx(&cfg, kern, x, perm, M, K, out_ref, st, zl, row_map, n_chunks, in);
}

// the same launch with the pending input transform (stage / nibble outputs only)
template <int MODE>
static cudaError_t launch_act_in(int grid, cudaStream_t stream, const float *x, const int32_t *perm, int M, int K, ActStage st,
                                 int which, const ZeroList *zlp) {
    const ActInput in = *tl_pending_in;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3(ACT_WARPS_PER_CTA * 32);
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    static const bool pdl = getenv("TCQ4_PDL") != nullptr, no_pdl = getenv("TCQ4_NO_PDL") != nullptr;
    cfg.numAttrs = (pdl || (which == 2 && !no_pdl)) ? 1 : 0;
    ZeroList zl{};
    if (zlp) zl = *zlp;
    if (in.mode == ACT_IN_NORM)
        return cudaLaunchKernelEx(&cfg, tcq4_act_quant_kernel<MODE, ACT_IN_NORM>, x, perm, M, K, (ActBlockTC *)nullptr, st, zl,
                                  (const int32_t *)nullptr, (const int *)nullptr, in);
    return cudaLaunchKernelEx(&cfg, tcq4_act_quant_kernel<MODE, ACT_IN_SWIGLU>, x, perm, M, K, (ActBlockTC *)nullptr, st, zl,
                              (const int32_t *)nullptr, (const int *)nullptr, in);
}

cudaError_t launch_act_quant_ref(const float *x, const int32_t *perm, int64_t M, int64_t K, void *out_blocks,
                                 int num_sms, cudaStream_t stream) {
    const int64_t nblocks = M * (K / TILE_K);
    if (nblocks == 0) return cudaSuccess;
    ActStage none{nullptr, nullptr, nullptr, nullptr};
    return launch_act(tcq4_act_quant_kernel<MODE_REF>, act_grid(nblocks, num_sms), stream, x, perm, (int)M, (int)K,
                      reinterpret_cast<ActBlockTC *>(out_blocks), none);
}

cudaError_t launch_act_quant_stage(const float *x, const int32_t *perm, int64_t M, int64_t K, ActStage st,
                                   int num_sms, cudaStream_t stream) {
    const int64_t nblocks = M * (K / TILE_K);
    if (nblocks == 0) return cudaSuccess;
    if (tl_pending_in && tl_pending_in->mode != ACT_IN_NONE)
        return launch_act_in<MODE_STAGE>(act_grid(nblocks, num_sms), stream, x, perm, (int)M, (int)K, st, 0, nullptr);
    return launch_act(tcq4_act_quant_kernel<MODE_STAGE>, act_grid(nblocks, num_sms), stream, x, perm, (int)M, (int)K,
                      (ActBlockTC *)nullptr, st);
}

cudaError_t launch_act_quant_nib_zl(const float *x, const int32_t *perm, int64_t M, int64_t K, ActStage st, int num_sms,
                                    cudaStream_t stream, const ZeroList *zl) {
    const int64_t nblocks = M * (K / TILE_K);
    if (nblocks == 0) return cudaSuccess;
    if (tl_pending_in && tl_pending_in->mode != ACT_IN_NONE)
        return launch_act_in<MODE_NIB>(act_grid(nblocks, num_sms), stream, x, perm, (int)M, (int)K, st, 2, zl);
    return launch_act(tcq4_act_quant_kernel<MODE_NIB>, act_grid(nblocks, num_sms), stream, x, perm, (int)M, (int)K,
                      (ActBlockTC *)nullptr, st, 2, zl);
}
// MUL_MAT_ID: max_rows virtual rows (64 per chunk), of which the first *n_chunks * 64 exist; row r reads x[row_map[r]]
cudaError_t launch_act_quant_nib_rows(const float *x, const int32_t *perm, int64_t max_rows, int64_t K, ActStage st, int num_sms,
                                      cudaStream_t stream, const int32_t *row_map, const int *n_chunks) {
    const int64_t nblocks = max_rows * (K / TILE_K);
    if (nblocks == 0) return cudaSuccess;
    return launch_act(tcq4_act_quant_kernel<MODE_NIB>, act_grid(nblocks, num_sms), stream, x, perm, (int)max_rows, (int)K,
                      (ActBlockTC *)nullptr, st, 2, nullptr, row_map, n_chunks);
}
cudaError_t launch_act_quant_nib(const float *x, const int32_t *perm, int64_t M, int64_t K, ActStage st, int num_sms,
                                 cudaStream_t stream, float *zero_y, int64_t zero_floats) {
    ZeroList zl{};
    if (zero_y != nullptr && zero_floats >= 4) {
        zl.n = 1;
        zl.p[0] = zero_y;
        zl.rows[0] = 1;
        zl.cols[0] = zero_floats & ~(int64_t)3;
        zl.ld[0] = zl.cols[0];
    }
    return launch_act_quant_nib_zl(x, perm, M, K, st, num_sms, stream, &zl);
}

// ---- the input transforms on their own ------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) tcq4_apply_input_kernel(const ActInput in, const float *__restrict__ x, float *__restrict__ out,
                                                               long long M, int K) {
    const long long total = M * K;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const long long m = i / K;
        const int k = (int)(i - m * K);
        const float xv = x[i];
        if (in.mode == ACT_IN_NORM) out[i] = act_input_value<ACT_IN_NORM>(in, xv, nullptr, in.row_scale[m], k);
        else if (in.mode == ACT_IN_SWIGLU) out[i] = act_input_value<ACT_IN_SWIGLU>(in, xv, in.x2 + m * K, 0.0f, k);
        else out[i] = xv;
    }
}
cudaError_t launch_apply_input(const ActInput &in, const float *x, float *out, int64_t M, int64_t K, int num_sms, cudaStream_t stream) {
    const long long total = (long long)M * K;
    if (total == 0) return cudaSuccess;
    long long grid = (total + 255) / 256;
    if (grid > (long long)num_sms * 8) grid = (long long)num_sms * 8;
    tcq4_apply_input_kernel<<<(unsigned)grid, 256, 0, stream>>>(in, x, out, (long long)M, (int)K);
    return cudaGetLastError();
}

// rsqrt(mean(x^2) + eps) of one row per CTA, in the order of rms_norm_f32 (norm.cu:125-137): per-thread partial sums over
// col = tid, tid + BS, ..., a shuffle-xor tree (16 .. 1) inside each warp, the warps' sums through shared memory and the same
// tree again
template <int BS>
__global__ void __launch_bounds__(BS) tcq4_rms_row_scale_kernel(const float *__restrict__ x, float *__restrict__ row_scale, int K, float eps) {
    __shared__ float s_sum[32];
    const float *xr = x + (size_t)blockIdx.x * K;
    float tmp = 0.0f;
    for (int col = threadIdx.x; col < K; col += BS) {
        const float xi = xr[col];
        tmp += xi * xi;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) tmp += __shfl_xor_sync(0xffffffffu, tmp, o);
    if (BS > 32) {
        if ((threadIdx.x & 31) == 0) s_sum[threadIdx.x >> 5] = tmp;
        __syncthreads();
        tmp = (threadIdx.x & 31) < BS / 32 ? s_sum[threadIdx.x & 31] : 0.0f;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) tmp += __shfl_xor_sync(0xffffffffu, tmp, o);
    }
    if (threadIdx.x == 0) row_scale[blockIdx.x] = rsqrtf(tmp / K + eps);
}
cudaError_t launch_rms_row_scale(const float *x, float *row_scale, int64_t M, int64_t K, float eps, cudaStream_t stream) {
    if (M == 0) return cudaSuccess;
    if (K < 1024) tcq4_rms_row_scale_kernel<256><<<(unsigned)M, 256, 0, stream>>>(x, row_scale, (int)K, eps);
    else tcq4_rms_row_scale_kernel<1024><<<(unsigned)M, 1024, 0, stream>>>(x, row_scale, (int)K, eps);
    return cudaGetLastError();
}

__global__ void tcq4_zero_kernel(float4 *__restrict__ y, long long n4) {
    pdl_launch_dependents();
    pdl_wait();
    const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) y[i] = z;
}
__global__ void tcq4_zero_list_kernel(const __grid_constant__ ZeroList zl) {
    pdl_launch_dependents();
    pdl_wait();
    const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int r = 0; r < zl.n; r++) {
        const long long c4 = zl.cols[r] >> 2, n4 = zl.rows[r] * c4;
        for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
            const long long row = i / c4;
            reinterpret_cast<float4 *>(zl.p[r] + row * zl.ld[r])[i - row * c4] = z;
        }
    }
}
cudaError_t launch_zero_list(const ZeroList &zl, int num_sms, cudaStream_t stream) {
    if (zl.n == 0) return cudaSuccess;
    tcq4_zero_list_kernel<<<(unsigned)num_sms, 256, 0, stream>>>(zl);
    return cudaGetLastError();
}
cudaError_t launch_zero_floats(float *y, int64_t n_floats, int num_sms, cudaStream_t stream) {
    const long long n4 = n_floats / 4;
    if (n4 == 0) return cudaSuccess;
    long long grid = (n4 + 255) / 256;
    if (grid > (long long)num_sms * 8) grid = (long long)num_sms * 8;
    tcq4_zero_kernel<<<(unsigned)grid, 256, 0, stream>>>(reinterpret_cast<float4 *>(y), n4);
    return cudaGetLastError();
}

}  // namespace tcq4
