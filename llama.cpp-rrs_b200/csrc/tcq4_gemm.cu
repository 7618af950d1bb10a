// tcq4_gemm.cu -- K2: W4A4 prefill GEMM on the 5th-gen tensor cores (tcgen05.mma kind::i8,
// int32 accumulators in TMEM), fed by the activation kernel's s8 staging (tcq4_act.cu).
//
// Replaces tcq4_rrs_gemm_imma_kernel (/root/reference/ggml/src/ggml-cuda/tcq4_k32.cu:1217-1412,
// mma.sync.m16n8k32.s4 -- emulated as unpack + IMMA.16816.S8 on sm_100a) and, for M > 4, the live
// scalar kernel tcq4_rrs_fused_gemm_smallM (tcq4_k32.cu:733-889).
//
// Exact integer formulation.  kind::i8 has K = 32 per instruction = one scale group, and
// the weight scale S*sc[n][g]/127 changes every group.  Instead of draining TMEM every 32-deep step,
// the int8 scale code is folded into the weight: p = sc*q in [-1016, 1016] is split exactly as
// p = 16*hi + lo, hi in [-64, 63], lo in [0, 15] (both s8).  The B operand of ONE MMA per group is
// 256 rows: 128 channels' hi bytes followed by the same channels' lo bytes (a tcgen05.mma takes
// ~128 cycles whatever N <= 256 is -- measured, tests/gpu_umma_peak.py -- so N = 256 is the only
// full-rate shape for cta_group::1).  TMEM columns [0,128) accumulate D_hi = A*hi and [128,256)
// D_lo = A*lo over the whole 256-block, so  16*D_hi + D_lo = sum_g sc_g * dot_g  exactly in int32,
// and the fp32 epilogue runs once per 256-block with  a7[m][kt] * S[n][kt] / 127.
//
// Weight unpack (CUDA cores): unpack_chunk_split in tcq4_common.cuh -- the group scale is split as sc = 16*sh + sl
// and folded into the nibbles four bytes at a time ("hi" rows = sh*q, "lo" rows = sl*q, both s8).  The resulting
// K byte order inside an 8-slice is {0,2,4,6,1,3,5,7}; the activation kernel writes its s8 codes in the same order.
//
// CTA = 128 tokens x 128 channels, 18 warps:
//   warp 0       bulk-copy producer: per k-tile one 32 KB activation unit + 16 raw 1184-B weight tiles
//   warp 1       TMEM allocator + single-thread MMA issuer (8 MMAs 128x256x32 per k-tile)
//   warps 2-9    weight unpack (raw tiles -> K-major no-swizzle UMMA operand, hi rows then lo rows)
//   warps 10-17  epilogue: tcgen05.ld, combine hi/lo, scale, fp32 accumulate in registers, store
// TMEM: 2 accumulator stages x (128 hi + 128 lo) columns, so MMA(kt+1) overlaps epilogue(kt).
#include "tcq4_common.cuh"
#include "tcq4_kernels.h"

namespace tcq4 {

constexpr int GM_BM = 128;                  // tokens per CTA
constexpr int GM_BN = 128;                  // channels per CTA
constexpr int GM_NTL = GM_BN / TILE_CH;     // 16 n-tiles per CTA
constexpr int GM_UNPACK_WARP0 = 2, GM_UNPACK_WARPS = 8;
constexpr int GM_EPI_WARP0 = GM_UNPACK_WARP0 + GM_UNPACK_WARPS, GM_EPI_WARPS = 8;
constexpr int GM_THREADS = (GM_EPI_WARP0 + GM_EPI_WARPS) * 32;
constexpr int GM_SN_WARPS = 4;              // unpack warps 0-3 also publish S[] for the epilogue

constexpr int GM_A_BYTES = ACT_UNIT_BYTES;                 // 32768 per stage
constexpr int GM_RAW_BYTES = GM_NTL * TILE_BYTES;          // 18944 per stage
constexpr int GM_B_LBO = 2 * GM_NTL * 128;                 // 4096: K-chunk stride (32 row-groups x 128 B)
constexpr int GM_BGROUP_BYTES = 2 * GM_B_LBO;              // 8192: one 32-group = 2 K-chunks
constexpr int GM_BHALF_BYTES = 4 * GM_BGROUP_BYTES;        // 32768: 4 groups
constexpr int GM_SN_RING = 4;

struct GemmSmem {
    static constexpr int off_a = 0;
    static constexpr int off_raw = off_a + 2 * GM_A_BYTES;
    static constexpr int off_b = off_raw + 2 * GM_RAW_BYTES;           // 103424, 128-aligned
    static constexpr int off_sn = off_b + 2 * GM_BHALF_BYTES;
    static constexpr int off_zflag = off_sn + GM_SN_RING * GM_BN * 4;
    static constexpr int off_bar = off_zflag + GM_SN_RING * GM_SN_WARPS * 4;
    static constexpr int n_bar = 12 + GM_SN_RING;
    static constexpr int off_tmem = off_bar + n_bar * 8;
    static constexpr int total = off_tmem + 16;
};
static_assert(GemmSmem::off_b % 128 == 0, "B operand must be 128-byte aligned");
static_assert(GemmSmem::off_bar % 8 == 0, "barriers must be 8-byte aligned");

// Zero-point term of one output element for one k-tile (tcq4_k32.cu:850,872-873):
//   a7 * sum_g sum_q[m][kt][g] * Z[n]*zc[n][g]/127.
// The reference quantiser always writes Z = 0 (ggml-quants.c:1603), so this is a cold path.
__device__ __noinline__ float zero_point_term(const uint8_t *__restrict__ w, const int16_t *__restrict__ sumq, int m,
                                              int n, int kt, int KT, float a7) {
    const uint8_t *tp = w + ((size_t)(n >> 3) * KT + kt) * TILE_BYTES;
    const int c = n & 7;
    const float zv = __half2float(*reinterpret_cast<const __half *>(tp + TILE_OFF_Z + 2 * c));
    float zs = 0.0f;
    for (int g = 0; g < 8; g++) {
        const float bz = __fdiv_rn(zv * (float)*reinterpret_cast<const int8_t *>(tp + TILE_OFF_ZC + c * 8 + g), 127.0f);
        zs += ((float)sumq[((size_t)m * KT + kt) * 8 + g] * a7) * bz;
    }
    return zs;
}

template <bool DEBUG>
__global__ void __launch_bounds__(GM_THREADS, 1)
tcq4_gemm_kernel(const uint8_t *__restrict__ w, ActStage st, float *__restrict__ y, int M, int N, int K,
                 int32_t *__restrict__ dbg_isum) {
    extern __shared__ __align__(128) uint8_t smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int KT = K / TILE_K, NT = N / TILE_CH;
    const int mb = blockIdx.x;                      // m-block fastest: CTAs sharing weights run together
    const int nt0 = blockIdx.y * GM_NTL;            // first n-tile of this CTA
    const int ntl = min(GM_NTL, NT - nt0);          // n-tiles actually present

    uint8_t *sA = smem + GemmSmem::off_a;
    uint8_t *sRaw = smem + GemmSmem::off_raw;
    uint8_t *sB = smem + GemmSmem::off_b;
    float *sSn = reinterpret_cast<float *>(smem + GemmSmem::off_sn);
    int *sZflag = reinterpret_cast<int *>(smem + GemmSmem::off_zflag);
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem + GemmSmem::off_bar);
    uint64_t *full = bars + 0, *empty = bars + 2, *bfull = bars + 4, *bfree = bars + 6, *accfull = bars + 8,
             *accempty = bars + 10, *snfull = bars + 12;
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(smem + GemmSmem::off_tmem);

    if (threadIdx.x == 0) {
        for (int i = 0; i < 2; i++) {
            mbar_init(&full[i], 1);
            mbar_init(&empty[i], 1 + GM_UNPACK_WARPS);
            mbar_init(&bfull[i], GM_UNPACK_WARPS);
            mbar_init(&bfree[i], 1);
            mbar_init(&accfull[i], 1);
            mbar_init(&accempty[i], GM_EPI_WARPS);
        }
        for (int i = 0; i < GM_SN_RING; i++) mbar_init(&snfull[i], GM_SN_WARPS);
        fence_mbar_init();
    }
    if (warp == 1) {
        tmem_alloc(tmem_slot, 512);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;

    if (warp == 0) {
        // ===================== producer =====================
        if (lane == 0) {
            const uint64_t pol_w = l2_policy_evict_first();
            for (int kt = 0; kt < KT; kt++) {
                const int s = kt & 1, u = kt >> 1;
                if (u > 0) mbar_wait(&empty[s], (uint32_t)((u - 1) & 1));
                mbar_arrive_expect_tx(&full[s], (uint32_t)(GM_A_BYTES + ntl * TILE_BYTES));
                bulk_g2s(sA + s * GM_A_BYTES, st.codes + ((size_t)mb * KT + kt) * ACT_UNIT_BYTES, GM_A_BYTES, &full[s]);
                for (int j = 0; j < ntl; j++)
                    bulk_g2s_stream(sRaw + s * GM_RAW_BYTES + j * TILE_BYTES,
                                    w + ((size_t)(nt0 + j) * KT + kt) * TILE_BYTES, TILE_BYTES, &full[s], pol_w);
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        if (lane == 0) {
            constexpr uint32_t idesc = umma_idesc_i8(GM_BM, 2 * GM_BN, 1, 1);     // 128 x 256 x 32, s8 x s8
            const uint32_t a_base = smem_u32(sA), b_base = smem_u32(sB);
            for (int kt = 0; kt < KT; kt++) {
                const int s = kt & 1, u = kt >> 1, as = kt & 1;
                mbar_wait(&full[s], (uint32_t)(u & 1));
                if (u > 0) mbar_wait(&accempty[as], (uint32_t)((u - 1) & 1));
                const uint32_t d = tmem + as * 256;
                for (int h = 0; h < 2; h++) {
                    mbar_wait(&bfull[h], (uint32_t)(kt & 1));
                    tc_fence_after();
#pragma unroll
                    for (int gl = 0; gl < 4; gl++) {
                        const int g = h * 4 + gl;
                        const uint64_t a_desc =
                            umma_desc_kmajor_noswz(a_base + s * GM_A_BYTES + (2 * g) * ACT_LBO, ACT_LBO, ACT_SBO);
                        const uint64_t b_desc =
                            umma_desc_kmajor_noswz(b_base + h * GM_BHALF_BYTES + gl * GM_BGROUP_BYTES, GM_B_LBO, 128);
                        umma_i8(d, a_desc, b_desc, idesc, g > 0 ? 1u : 0u);
                    }
                    umma_commit(&bfree[h]);
                }
                umma_commit(&empty[s]);
                umma_commit(&accfull[as]);
            }
        }
    } else if (warp < GM_EPI_WARP0) {
        // ===================== weight unpack =====================
        const int uw = warp - GM_UNPACK_WARP0;
        const int ut = threadIdx.x - GM_UNPACK_WARP0 * 32;          // 0..255
        for (int kt = 0; kt < KT; kt++) {
            const int s = kt & 1, u = kt >> 1;
            mbar_wait(&full[s], (uint32_t)(u & 1));
            const uint8_t *raw = sRaw + s * GM_RAW_BYTES;
            if (uw < GM_SN_WARPS) {     // per-channel super-scale for the epilogue + zero-point presence flag
                const int j = ut >> 3, c = ut & 7;
                float sv = 0.0f;
                int zf = 0;
                if (j < ntl) {
                    const uint8_t *tp = raw + j * TILE_BYTES;
                    sv = __half2float(*reinterpret_cast<const __half *>(tp + TILE_OFF_S + 2 * c));
                    const float zv = __half2float(*reinterpret_cast<const __half *>(tp + TILE_OFF_Z + 2 * c));
                    const uint2 zc = *reinterpret_cast<const uint2 *>(tp + TILE_OFF_ZC + c * 8);
                    zf = (zv != 0.0f) && ((zc.x | zc.y) != 0u);
                }
                const int ring = kt & (GM_SN_RING - 1);
                sSn[ring * GM_BN + ut] = sv;
                const int anyz = __any_sync(0xffffffffu, zf);
                __syncwarp();
                if (lane == 0) {
                    sZflag[ring * GM_SN_WARPS + uw] = anyz;
                    mbar_arrive(&snfull[ring]);      // release: epilogue(kt) acquires before reading
                }
            }
            for (int h = 0; h < 2; h++) {
                if (kt > 0) mbar_wait(&bfree[h], (uint32_t)((kt - 1) & 1));
                uint8_t *bh = sB + h * GM_BHALF_BYTES;
#pragma unroll
                for (int it = 0; it < 2; it++) {
                    const int qid = it * 256 + ut;                  // (gl, j, c)
                    const int c = qid & 7, j = (qid >> 3) & 15, gl = qid >> 7;
                    if (j < ntl) {
                        const uint8_t *tp = raw + j * TILE_BYTES;
                        const int g = h * 4 + gl;
                        const uint4 wq = *reinterpret_cast<const uint4 *>(tp + g * 128 + c * 16);
                        const int sc = *reinterpret_cast<const int8_t *>(tp + TILE_OFF_SC + c * 8 + g);
                        uint8_t *row = bh + gl * GM_BGROUP_BYTES + j * 128 + c * 16;
                        unpack_chunk_split(wq, sc, row, row + GM_NTL * 128, GM_B_LBO);      // lo rows follow the 128 hi rows
                    }
                }
                fence_proxy_async_smem();
                __syncwarp();
                if (lane == 0) mbar_arrive(&bfull[h]);
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty[s]);
        }
    } else {
        // ===================== epilogue =====================
        const int ew = warp - GM_EPI_WARP0;                 // 0..7
        const int quarter = warp & 3;                       // TMEM lane quarter this warp may access
        const int colhalf = ew >> 2;                        // 0: columns [0,64), 1: [64,128)
        const int row = quarter * 32 + lane;
        const int m = mb * GM_BM + row;
        const bool mvalid = m < M;
        const int n_base = nt0 * TILE_CH + colhalf * 64;
        float acc[64];
#pragma unroll
        for (int i = 0; i < 64; i++) acc[i] = 0.0f;

        for (int kt = 0; kt < KT; kt++) {
            const int as = kt & 1, u = kt >> 1;
            const float a7 = mvalid ? st.a7[(size_t)m * KT + kt] : 0.0f;
            const float rm = a7 * (1.0f / 127.0f);          // row factor a7/127; column factor S[n]
            const int ring = kt & (GM_SN_RING - 1);
            mbar_wait(&snfull[ring], (uint32_t)((kt / GM_SN_RING) & 1));
            mbar_wait(&accfull[as], (uint32_t)(u & 1));
            tc_fence_after();
            const float *sn = sSn + ring * GM_BN + colhalf * 64;
            const int *zfl = sZflag + ring * GM_SN_WARPS;
            const int anyz = zfl[0] | zfl[1] | zfl[2] | zfl[3];
            const uint32_t t_hi = tmem + ((uint32_t)(quarter * 32) << 16) + as * 256 + colhalf * 64;
            const uint32_t t_lo = t_hi + GM_BN;
#pragma unroll
            for (int ch = 0; ch < 8; ch++) {
                uint32_t rh[8], rl[8];
                tmem_ld_32x8(t_hi + ch * 8, rh);
                tmem_ld_32x8(t_lo + ch * 8, rl);
                // S[] of this k-tile is read BEFORE the stage is handed back: accempty(kt) is what lets
                // MMA(kt+2) -> empty -> TMA(kt+4) -> the unpack warps overwrite ring slot (kt+4)&3 == kt&3.
                const float4 s0 = *reinterpret_cast<const float4 *>(sn + ch * 8);
                const float4 s1 = *reinterpret_cast<const float4 *>(sn + ch * 8 + 4);
                tmem_ld_wait();
                if (ch == 7) {               // all TMEM and ring reads of this stage are done: hand it back
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&accempty[as]);
                }
                const float sv[8] = {s0.x, s0.y, s0.z, s0.w, s1.x, s1.y, s1.z, s1.w};
#pragma unroll
                for (int i = 0; i < 8; i++) {
                    const int isum = ((int)rh[i] << 4) + (int)rl[i];         // = sum_g sc*dot, exact
                    const float f = (float)isum * rm;
                    acc[ch * 8 + i] = __fmaf_rn(f, sv[i], acc[ch * 8 + i]);
                    if (DEBUG) {
                        const int n = n_base + ch * 8 + i;
                        if (dbg_isum && mvalid && n < N) dbg_isum[((size_t)m * N + n) * KT + kt] = isum;
                    }
                }
            }
            if (anyz && mvalid) {
#pragma unroll
                for (int i = 0; i < 64; i++) {
                    const int n = n_base + i;
                    if (n < N) acc[i] += zero_point_term(w, st.sumq, m, n, kt, KT, a7);
                }
            }
        }
        if (mvalid) {
            float *yrow = y + (size_t)m * N + n_base;
#pragma unroll
            for (int i = 0; i < 64; i += 4) {
                if (n_base + i + 3 < N) {
                    *reinterpret_cast<float4 *>(yrow + i) = make_float4(acc[i], acc[i + 1], acc[i + 2], acc[i + 3]);
                }
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc(tmem, 512);
}

bool gemm_plan(int64_t M, int64_t N, int64_t K, int num_sms, size_t smem_optin, GemmPlan *plan) {
    (void)num_sms;
    if (M < 1 || (size_t)GemmSmem::total > smem_optin) return false;
    const int64_t KT = K / TILE_K, NT = N / TILE_CH;
    const int64_t MB = (M + GM_BM - 1) / GM_BM;
    plan->grid = dim3((unsigned)MB, (unsigned)((NT + GM_NTL - 1) / GM_NTL), 1);
    plan->smem_bytes = GemmSmem::total;
    plan->stage_codes_bytes = (size_t)round_up(MB * KT * (int64_t)ACT_UNIT_BYTES, 256);
    plan->stage_a7_bytes = (size_t)round_up(M * KT * (int64_t)sizeof(float), 256);
    plan->stage_sumq_bytes = (size_t)round_up(M * KT * 8 * (int64_t)sizeof(int16_t), 256);
    return true;
}

cudaError_t launch_gemm(const void *w_tiles, ActStage st, float *y, int64_t M, int64_t N, int64_t K,
                        const GemmPlan &plan, int32_t *dbg_isum, cudaStream_t stream) {
    if (dbg_isum) {
        tcq4_gemm_kernel<true><<<plan.grid, GM_THREADS, plan.smem_bytes, stream>>>((const uint8_t *)w_tiles, st, y, (int)M,
                                                                                   (int)N, (int)K, dbg_isum);
    } else {
        tcq4_gemm_kernel<false><<<plan.grid, GM_THREADS, plan.smem_bytes, stream>>>((const uint8_t *)w_tiles, st, y, (int)M,
                                                                                    (int)N, (int)K, nullptr);
    }
    return cudaGetLastError();
}

cudaError_t gemm_init_attributes(size_t smem_optin) {
    (void)smem_optin;
    cudaError_t e = cudaFuncSetAttribute(tcq4_gemm_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         GemmSmem::total);
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(tcq4_gemm_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, GemmSmem::total);
}

}  // namespace tcq4
