// tcq4_gemm2.cu -- K2 for M > 128: W4A4 prefill GEMM on CTA PAIRS (tcgen05.mma.cta_group::2 kind::i8).
//
// Same exact integer formulation as tcq4_gemm.cu (p = sc*q = 16*hi + lo folded into the weight operand, one
// 256-wide accumulation per 256-block, fp32 epilogue once per 256-block), re-tiled so that the CUDA-core work
// per tensor-core cycle halves:
//   * a cluster of two CTAs (one TPC) owns 256 tokens x (16*ntl) channels; each CTA stages its own 128 token
//     rows of s8 activation codes and unpacks only ITS half of the channels (ntl n-tiles = 8*ntl channels ->
//     16*ntl operand rows: hi rows then lo rows).  One M = 256, N = 32*ntl, K = 32 MMA per 32-group, issued by
//     the leader CTA, reads A and B from both CTAs' shared memory and writes 128 lanes x N columns of TMEM in
//     each.  Weight unpack per SM is half of the single-CTA kernel's for the same MMA time.
//   * ntl (1..8) is chosen by the planner so that the number of pair-tiles is a multiple of the 74 SM pairs
//     as nearly as possible (M = 512, N = 9728: ntl = 6 -> 204 tiles = 2.76 waves instead of 152 = 2.05).
//   * epilogue: hi/lo combine is one IMAD, the fp32 steps are packed f32x2 (FMUL2 / FFMA2).
// Cross-CTA signalling: the unpack warps of both CTAs arrive (release.cluster) on the LEADER's bfull barriers,
// the epilogue warps of both CTAs on the leader's accempty barriers; the leader's tcgen05.commit multicasts
// to both CTAs' bfree / empty / accfull barriers; per-channel super-scales S travel through DSMEM.
//
// Replaces tcq4_rrs_gemm_imma_kernel (/root/reference/ggml/src/ggml-cuda/tcq4_k32.cu:1217-1412) and, for these M,
// the live scalar kernel tcq4_rrs_fused_gemm_smallM (tcq4_k32.cu:733-889).
#include "tcq4_common.cuh"
#include "tcq4_kernels.h"

namespace tcq4 {

constexpr int G2_BM = 128;                  // tokens per CTA (256 per pair)
constexpr int G2_MAXNTL = 8;                // n-tiles per CTA (8 channels each)
constexpr int G2_UNPACK_WARP0 = 2, G2_UNPACK_WARPS = 8;
constexpr int G2_EPI_WARP0 = G2_UNPACK_WARP0 + G2_UNPACK_WARPS, G2_EPI_WARPS = 8;
constexpr int G2_THREADS = (G2_EPI_WARP0 + G2_EPI_WARPS) * 32;

constexpr int G2_A_BYTES = ACT_UNIT_BYTES;                   // 32768 per stage
constexpr int G2_RAW_BYTES = G2_MAXNTL * TILE_BYTES;         // 9472 per stage
constexpr int G2_BH_BYTES = 16384;                           // one half k-tile (4 groups) of B at ntl = 8
constexpr int G2_BH_STAGES = 4;
constexpr int G2_SN_RING = 4;

struct Gemm2Smem {
    static constexpr int off_a = 0;
    static constexpr int off_raw = off_a + 2 * G2_A_BYTES;                 // 65536
    static constexpr int off_b = off_raw + 2 * G2_RAW_BYTES;               // 84480
    static constexpr int off_sn = off_b + G2_BH_STAGES * G2_BH_BYTES;      // 150016
    static constexpr int off_zflag = off_sn + G2_SN_RING * 128 * 4;
    static constexpr int off_bar = off_zflag + G2_SN_RING * 4 * 4;
    static constexpr int n_bar = 2 + 2 + G2_BH_STAGES + G2_BH_STAGES + 2 + 2 + G2_SN_RING;
    static constexpr int off_tmem = off_bar + n_bar * 8;
    static constexpr int total = off_tmem + 16;
};
static_assert(Gemm2Smem::off_b % 128 == 0, "B operand must be 128-byte aligned");
static_assert(Gemm2Smem::off_bar % 8 == 0, "barriers must be 8-byte aligned");

// ---- cluster-scope helpers --------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t g2_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ uint32_t g2_mapa(uint32_t local_addr, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local_addr), "r"(rank));
    return r;
}
// arrive on a barrier anywhere in the cluster (address from g2_mapa), releasing this thread's prior writes
__device__ __forceinline__ void g2_arrive_cluster(uint32_t cluster_addr) {
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
// wait on a LOCAL barrier whose arrivals may come from the peer CTA
__device__ __forceinline__ void g2_wait_cluster(uint64_t *bar, uint32_t parity) {
    uint32_t ok;
    do {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(ok)
            : "r"(smem_u32(bar)), "r"(parity)
            : "memory");
    } while (!ok);
}
__device__ __forceinline__ void g2_cluster_sync() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void g2_st_remote_f32(uint32_t cluster_addr, float v) {
    asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(cluster_addr), "f"(v) : "memory");
}
__device__ __forceinline__ void g2_st_remote_b32(uint32_t cluster_addr, int v) {
    asm volatile("st.shared::cluster.b32 [%0], %1;" ::"r"(cluster_addr), "r"(v) : "memory");
}
// D[tmem, both CTAs] (+)= A[smem, both CTAs] * B[smem, both CTAs]; issued by one thread of the leader CTA
__device__ __forceinline__ void umma_i8_2cta(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                             uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::i8 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
        "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// arrive on the barrier at this shared-memory offset in BOTH CTAs when all prior MMAs of this thread completed
__device__ __forceinline__ void umma_commit_2cta(uint64_t *bar) {
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::
                     "r"(smem_u32(bar)), "h"((uint16_t)3)
                 : "memory");
}
__device__ __forceinline__ uint64_t g2_pack(float lo, float hi) {
    uint64_t r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void g2_unpack(uint64_t v, float &lo, float &hi) {
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ uint64_t g2_mul2(uint64_t a, uint64_t b) {
    uint64_t r;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ uint64_t g2_fma2(uint64_t a, uint64_t b, uint64_t c) {
    uint64_t r;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
    return r;
}

// Zero-point term of one output element for one k-tile (tcq4_k32.cu:850,872-873); cold (Z = 0 from the quantiser)
__device__ __noinline__ float g2_zero_point_term(const uint8_t *__restrict__ w, const int16_t *__restrict__ sumq, int m,
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

// Unpack one 16-byte chunk (channel c of group g: 32 nibbles) into 32 hi bytes + 32 lo bytes (see tcq4_gemm.cu).
//   out_hi / out_lo point at this channel's 16-byte row in K-chunk 0; K-chunk 1 is `lbo` bytes further.
__device__ __forceinline__ void g2_unpack_chunk(const uint4 wq, int sc, uint8_t *out_hi, uint8_t *out_lo, int lbo) {
    const uint32_t mul = (uint32_t)(16 * sc);
    const uint32_t bias = (uint32_t)(32768 - 128 * sc) * 0x00010001u;
    const uint32_t wv[4] = {wq.x, wq.y, wq.z, wq.w};
    uint32_t hi[8], lo[8];
#pragma unroll
    for (int s = 0; s < 4; s++) {
        uint32_t P[4];
#pragma unroll
        for (int i = 0; i < 4; i++) {
            const uint32_t up = ((wv[s] >> (4 * i)) & 0x000F000Fu) ^ 0x00080008u;    // u = q + 8 in two fields
            P[i] = up * mul + bias;                                                  // 16*sc*q + 32768
        }
        hi[s * 2 + 0] = __byte_perm(P[0], P[1], 0x7351) ^ 0x80808080u;               // elements 0,1,4,5
        hi[s * 2 + 1] = __byte_perm(P[2], P[3], 0x7351) ^ 0x80808080u;               // elements 2,3,6,7
        lo[s * 2 + 0] = __byte_perm(P[0], P[1], 0x6240) >> 4;
        lo[s * 2 + 1] = __byte_perm(P[2], P[3], 0x6240) >> 4;
    }
    *reinterpret_cast<uint4 *>(out_hi) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
    *reinterpret_cast<uint4 *>(out_hi + lbo) = make_uint4(hi[4], hi[5], hi[6], hi[7]);
    *reinterpret_cast<uint4 *>(out_lo) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
    *reinterpret_cast<uint4 *>(out_lo + lbo) = make_uint4(lo[4], lo[5], lo[6], lo[7]);
}

template <bool DEBUG>
__global__ void __launch_bounds__(G2_THREADS, 1)
tcq4_gemm2_kernel(const uint8_t *__restrict__ w, ActStage st, float *__restrict__ y, int M, int N, int K, int ntl,
                  int32_t *__restrict__ dbg_isum) {
    extern __shared__ __align__(128) uint8_t smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int KT = K / TILE_K, NT = N / TILE_CH;
    const uint32_t rank = g2_ctarank();
    const int mb = blockIdx.x;                                   // 128-row block of this CTA (pair = blocks 2p, 2p+1)
    const int nt_pair = blockIdx.y * (2 * ntl);                  // first n-tile of the pair
    const int nt0 = nt_pair + (int)rank * ntl;                   // first n-tile this CTA unpacks
    const int ntl_r = max(0, min(ntl, NT - nt0));                // n-tiles actually present
    const int lbo = 256 * ntl;                                   // K-chunk stride of B: (16*ntl rows) * 16 B
    const int bgroup = 2 * lbo;                                  // one 32-group = 2 K-chunks

    uint8_t *sA = smem + Gemm2Smem::off_a;
    uint8_t *sRaw = smem + Gemm2Smem::off_raw;
    uint8_t *sB = smem + Gemm2Smem::off_b;
    float *sSn = reinterpret_cast<float *>(smem + Gemm2Smem::off_sn);          // [ring][2 halves][64]
    int *sZflag = reinterpret_cast<int *>(smem + Gemm2Smem::off_zflag);        // [ring][2 halves][2 warps]
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem + Gemm2Smem::off_bar);
    uint64_t *full = bars + 0, *empty = bars + 2, *bfull = bars + 4, *bfree = bars + 4 + G2_BH_STAGES,
             *accfull = bars + 4 + 2 * G2_BH_STAGES, *accempty = accfull + 2, *snfull = accempty + 2;
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(smem + Gemm2Smem::off_tmem);

    if (threadIdx.x == 0) {
        for (int i = 0; i < 2; i++) {
            mbar_init(&full[i], 1);
            mbar_init(&empty[i], 1 + G2_UNPACK_WARPS);           // leader's commit (multicast) + local unpack warps
            mbar_init(&accfull[i], 1);                           // leader's commit (multicast)
            mbar_init(&accempty[i], 2 * G2_EPI_WARPS);           // epilogue warps of both CTAs (used in the leader)
        }
        for (int i = 0; i < G2_BH_STAGES; i++) {
            mbar_init(&bfull[i], 2 * G2_UNPACK_WARPS);           // unpack warps of both CTAs (used in the leader)
            mbar_init(&bfree[i], 1);                             // leader's commit (multicast)
        }
        for (int i = 0; i < G2_SN_RING; i++) mbar_init(&snfull[i], 4);     // 2 publishing warps per CTA
        fence_mbar_init();
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(512)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    g2_cluster_sync();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;

    if (warp == 0) {
        // ===================== producer (each CTA loads its own activation rows and its own weight tiles) =====
        if (lane == 0) {
            const uint64_t pol_w = l2_policy_evict_first();
            for (int kt = 0; kt < KT; kt++) {
                const int s = kt & 1, u = kt >> 1;
                if (u > 0) mbar_wait(&empty[s], (uint32_t)((u - 1) & 1));
                mbar_arrive_expect_tx(&full[s], (uint32_t)(G2_A_BYTES + ntl_r * TILE_BYTES));
                bulk_g2s(sA + s * G2_A_BYTES, st.codes + ((size_t)mb * KT + kt) * ACT_UNIT_BYTES, G2_A_BYTES, &full[s]);
                for (int j = 0; j < ntl_r; j++)
                    bulk_g2s_stream(sRaw + s * G2_RAW_BYTES + j * TILE_BYTES,
                                    w + ((size_t)(nt0 + j) * KT + kt) * TILE_BYTES, TILE_BYTES, &full[s], pol_w);
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer (leader CTA only) =====================
        if (rank == 0 && lane == 0) {
            const uint32_t idesc = umma_idesc_i8(2 * G2_BM, 32 * ntl, 1, 1);      // 256 x (32*ntl) x 32, s8 x s8
            const uint32_t a_base = smem_u32(sA), b_base = smem_u32(sB);
            for (int kt = 0; kt < KT; kt++) {
                const int s = kt & 1, u = kt >> 1, as = kt & 1;
                if (u > 0) g2_wait_cluster(&accempty[as], (uint32_t)((u - 1) & 1));
                tc_fence_after();
                const uint32_t d = tmem + as * 256;
                for (int h = 0; h < 2; h++) {
                    const int hq = 2 * kt + h, hs = hq & (G2_BH_STAGES - 1), v = hq / G2_BH_STAGES;
                    g2_wait_cluster(&bfull[hs], (uint32_t)(v & 1));
                    tc_fence_after();
#pragma unroll
                    for (int gl = 0; gl < 4; gl++) {
                        const int g = h * 4 + gl;
                        const uint64_t a_desc =
                            umma_desc_kmajor_noswz(a_base + s * G2_A_BYTES + (2 * g) * ACT_LBO, ACT_LBO, ACT_SBO);
                        const uint64_t b_desc =
                            umma_desc_kmajor_noswz(b_base + hs * G2_BH_BYTES + gl * bgroup, (uint32_t)lbo, 128);
                        umma_i8_2cta(d, a_desc, b_desc, idesc, g > 0 ? 1u : 0u);
                    }
                    umma_commit_2cta(&bfree[hs]);
                }
                umma_commit_2cta(&empty[s]);
                umma_commit_2cta(&accfull[as]);
            }
        }
    } else if (warp < G2_EPI_WARP0) {
        // ===================== weight unpack (this CTA's ntl n-tiles) =====================
        const int uw = warp - G2_UNPACK_WARP0;
        const int ut = threadIdx.x - G2_UNPACK_WARP0 * 32;          // 0..255
        // chunk of this thread inside a half k-tile: (gl, j, c) with 8*ntl chunks per group
        const int per_group = 8 * ntl;
        const int gl_t = ut / per_group, rem_t = ut - gl_t * per_group;
        const int j_t = rem_t >> 3, c_t = rem_t & 7;
        const bool has_chunk = (gl_t < 4) && (j_t < ntl_r);
        const uint32_t bfull_leader = g2_mapa(smem_u32(bfull), 0);
        const uint32_t peer = rank ^ 1u;
        const uint32_t sn_peer = g2_mapa(smem_u32(sSn), peer), zf_peer = g2_mapa(smem_u32(sZflag), peer),
                       snfull_peer = g2_mapa(smem_u32(snfull), peer);
        for (int kt = 0; kt < KT; kt++) {
            const int s = kt & 1, u = kt >> 1;
            mbar_wait(&full[s], (uint32_t)(u & 1));
            const uint8_t *raw = sRaw + s * G2_RAW_BYTES;
            if (uw < 2) {     // per-channel super-scale for both CTAs' epilogues + zero-point presence flag
                const int j = ut >> 3, c = ut & 7;
                float sv = 0.0f;
                int zf = 0;
                if (j < ntl_r) {
                    const uint8_t *tp = raw + j * TILE_BYTES;
                    sv = __half2float(*reinterpret_cast<const __half *>(tp + TILE_OFF_S + 2 * c));
                    const float zv = __half2float(*reinterpret_cast<const __half *>(tp + TILE_OFF_Z + 2 * c));
                    const uint2 zc = *reinterpret_cast<const uint2 *>(tp + TILE_OFF_ZC + c * 8);
                    zf = (zv != 0.0f) && ((zc.x | zc.y) != 0u);
                }
                const int ring = kt & (G2_SN_RING - 1);
                const int idx = (ring * 2 + (int)rank) * 64 + ut;
                sSn[idx] = sv;
                g2_st_remote_f32(sn_peer + (uint32_t)idx * 4, sv);
                const int anyz = __any_sync(0xffffffffu, zf);
                __syncwarp();
                if (lane == 0) {
                    const int zi = (ring * 2 + (int)rank) * 2 + uw;
                    sZflag[zi] = anyz;
                    g2_st_remote_b32(zf_peer + (uint32_t)zi * 4, anyz);
                    g2_arrive_cluster(g2_mapa(smem_u32(&snfull[ring]), rank));     // local, cluster-scope release
                    g2_arrive_cluster(snfull_peer + (uint32_t)ring * 8);
                }
            }
            for (int h = 0; h < 2; h++) {
                const int hq = 2 * kt + h, hs = hq & (G2_BH_STAGES - 1), v = hq / G2_BH_STAGES;
                if (v > 0) mbar_wait(&bfree[hs], (uint32_t)((v - 1) & 1));
                if (has_chunk) {
                    const uint8_t *tp = raw + j_t * TILE_BYTES;
                    const int g = h * 4 + gl_t;
                    const uint4 wq = *reinterpret_cast<const uint4 *>(tp + g * 128 + c_t * 16);
                    const int sc = *reinterpret_cast<const int8_t *>(tp + TILE_OFF_SC + c_t * 8 + g);
                    uint8_t *row = sB + hs * G2_BH_BYTES + gl_t * bgroup + j_t * 128 + c_t * 16;
                    g2_unpack_chunk(wq, sc, row, row + ntl * 128, lbo);      // lo rows follow the 8*ntl hi rows
                }
                fence_proxy_async_smem();
                __syncwarp();
                if (lane == 0) g2_arrive_cluster(bfull_leader + (uint32_t)hs * 8);
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty[s]);
        }
    } else {
        // ===================== epilogue =====================
        const int ew = warp - G2_EPI_WARP0;                 // 0..7
        const int quarter = warp & 3;                       // TMEM lane quarter this warp may access
        const int colhalf = ew >> 2;                        // 0: channels unpacked by CTA 0, 1: by CTA 1
        const int row = quarter * 32 + lane;
        const int m = mb * G2_BM + row;
        const bool mvalid = m < M;
        const int n_base = (nt_pair + colhalf * ntl) * TILE_CH;
        const uint32_t accempty_leader = g2_mapa(smem_u32(accempty), 0);
        uint64_t acc2[32];                                  // 64 fp32 accumulators as f32x2 pairs
#pragma unroll
        for (int i = 0; i < 32; i++) acc2[i] = 0ull;

        for (int kt = 0; kt < KT; kt++) {
            const int as = kt & 1, u = kt >> 1;
            const float a7 = mvalid ? st.a7[(size_t)m * KT + kt] : 0.0f;
            const float rm = a7 * (1.0f / 127.0f);          // row factor a7/127; column factor S[n]
            const uint64_t rm2 = g2_pack(rm, rm);
            const int ring = kt & (G2_SN_RING - 1);
            g2_wait_cluster(&snfull[ring], (uint32_t)((kt / G2_SN_RING) & 1));
            mbar_wait(&accfull[as], (uint32_t)(u & 1));
            tc_fence_after();
            const float *sn = sSn + (ring * 2 + colhalf) * 64;
            const int *zfl = sZflag + (ring * 2 + colhalf) * 2;
            const int anyz = zfl[0] | zfl[1];
            const uint32_t t_hi = tmem + ((uint32_t)(quarter * 32) << 16) + as * 256 + colhalf * (16 * ntl);
            const uint32_t t_lo = t_hi + 8 * ntl;
#pragma unroll
            for (int ch = 0; ch < G2_MAXNTL; ch++) {
                if (ch < ntl) {
                    uint32_t rh[8], rl[8];
                    tmem_ld_32x8(t_hi + ch * 8, rh);
                    tmem_ld_32x8(t_lo + ch * 8, rl);
                    // S[] of this k-tile is read BEFORE the stage is handed back (ring slot reuse, see tcq4_gemm.cu)
                    const ulonglong2 s01 = *reinterpret_cast<const ulonglong2 *>(sn + ch * 8);
                    const ulonglong2 s23 = *reinterpret_cast<const ulonglong2 *>(sn + ch * 8 + 4);
                    tmem_ld_wait();
                    if (ch == ntl - 1) {         // all TMEM and ring reads of this stage are done: hand it back
                        tc_fence_before();
                        __syncwarp();
                        if (lane == 0) g2_arrive_cluster(accempty_leader + (uint32_t)as * 8);
                    }
                    const uint64_t sv2[4] = {s01.x, s01.y, s23.x, s23.y};
#pragma unroll
                    for (int i = 0; i < 4; i++) {
                        const int is0 = (int)rh[2 * i] * 16 + (int)rl[2 * i];             // = sum_g sc*dot, exact
                        const int is1 = (int)rh[2 * i + 1] * 16 + (int)rl[2 * i + 1];
                        const uint64_t f2 = g2_mul2(g2_pack((float)is0, (float)is1), rm2);
                        acc2[ch * 4 + i] = g2_fma2(f2, sv2[i], acc2[ch * 4 + i]);
                        if (DEBUG) {
                            const int n = n_base + ch * 8 + 2 * i;
                            if (dbg_isum && mvalid && n < N) {
                                dbg_isum[((size_t)m * N + n) * KT + kt] = is0;
                                dbg_isum[((size_t)m * N + n + 1) * KT + kt] = is1;
                            }
                        }
                    }
                }
            }
            if (anyz && mvalid) {
#pragma unroll
                for (int i = 0; i < 32; i++) {
                    const int n = n_base + 2 * i;
                    if (2 * i < 8 * ntl && n < N) {
                        float a, b;
                        g2_unpack(acc2[i], a, b);
                        a += g2_zero_point_term(w, st.sumq, m, n, kt, KT, a7);
                        b += g2_zero_point_term(w, st.sumq, m, n + 1, kt, KT, a7);
                        acc2[i] = g2_pack(a, b);
                    }
                }
            }
        }
        if (mvalid) {
            float *yrow = y + (size_t)m * N + n_base;
#pragma unroll
            for (int i = 0; i < 32; i += 2) {
                if (2 * i < 8 * ntl && n_base + 2 * i + 3 < N) {
                    float a, b, c, d;
                    g2_unpack(acc2[i], a, b);
                    g2_unpack(acc2[i + 1], c, d);
                    *reinterpret_cast<float4 *>(yrow + 2 * i) = make_float4(a, b, c, d);
                }
            }
        }
    }

    tc_fence_before();
    g2_cluster_sync();
    if (warp == 1)
        asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512) : "memory");
}

bool gemm2_plan(int64_t M, int64_t N, int64_t K, int num_sms, size_t smem_optin, Gemm2Plan *plan) {
    if (M < 1 || (size_t)Gemm2Smem::total > smem_optin || num_sms < 2) return false;
    const int64_t NT = N / TILE_CH;
    const int64_t MB2 = (M + 2 * G2_BM - 1) / (2 * G2_BM);
    const int64_t pairs = num_sms / 2;
    // n-tiles per CTA: minimise (waves of pair-tiles) x (work per tile + a fixed fill/drain share)
    int best = G2_MAXNTL;
    int64_t best_cost = -1;
    for (int ntl = G2_MAXNTL; ntl >= 2; ntl--) {
        const int64_t tiles = MB2 * ((NT + 2 * ntl - 1) / (2 * ntl));
        const int64_t waves = (tiles + pairs - 1) / pairs;
        const int64_t cost = waves * (ntl + 1);
        if (best_cost < 0 || cost < best_cost) {
            best_cost = cost;
            best = ntl;
        }
    }
    plan->ntl = best;
    plan->grid = dim3((unsigned)(2 * MB2), (unsigned)((NT + 2 * best - 1) / (2 * best)), 1);
    plan->smem_bytes = Gemm2Smem::total;
    return true;
}

cudaError_t launch_gemm2(const void *w_tiles, ActStage st, float *y, int64_t M, int64_t N, int64_t K,
                         const Gemm2Plan &plan, int32_t *dbg_isum, cudaStream_t stream) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = plan.grid;
    cfg.blockDim = dim3(G2_THREADS);
    cfg.dynamicSmemBytes = plan.smem_bytes;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    if (dbg_isum)
        return cudaLaunchKernelEx(&cfg, tcq4_gemm2_kernel<true>, (const uint8_t *)w_tiles, st, y, (int)M, (int)N, (int)K,
                                  plan.ntl, dbg_isum);
    return cudaLaunchKernelEx(&cfg, tcq4_gemm2_kernel<false>, (const uint8_t *)w_tiles, st, y, (int)M, (int)N, (int)K,
                              plan.ntl, (int32_t *)nullptr);
}

cudaError_t gemm2_init_attributes(size_t smem_optin) {
    (void)smem_optin;
    cudaError_t e = cudaFuncSetAttribute(tcq4_gemm2_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         Gemm2Smem::total);
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(tcq4_gemm2_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, Gemm2Smem::total);
}

}  // namespace tcq4
