// tcq4_gemm4.cu -- K2c: batched-decode GEMM for 9 <= M <= 64 tokens (BASELINE configs[2], "decode batch 1-64"), fan-out groups of
// up to four matrices in one launch, and the grouped launch of MUL_MAT_ID with many tokens.
//
// At these batch sizes the MUL_MAT is bound by the weight stream (0.578 B per weight from HBM) and by the CUDA-core work of
// widening every nibble once, not by tensor throughput -- but the 256-token tiles of the prefill GEMMs make the tensor pipe
// (and their epilogues) pay for 256 rows of which at most 64 exist: 32-47 us per Llama-3-8B matrix against a 5 us memory
// floor.  This kernel turns the operands around:
//   * WEIGHTS are the A operand: 128 channels per CTA = the 128 TMEM lanes.  A thread owns one channel and half of the
//     k-tile: four 16-byte loads of nibbles, the magic-number widening of tcq4_gemm3.cu (LOP3 + HFMA2: p = sc*q exactly in
//     fp16), four tcgen05.st -- the widened weights never touch shared memory.
//   * TOKENS are the N dimension of tcgen05.mma.cta_group::1.kind::f16 with A in TMEM (ideal cost 128*N/256 cycles per
//     instruction; measured ~40-60 at N <= 64: the 128 rows of A set a floor).  The B operand (Mt x 256 fp16, Mt = 32 or 64) is
//     widened from K1's nibble units by the same warps into shared memory (8 KB of nibbles per k-tile at Mt = 64).
//   * D: 128 lanes (channels) x Mt columns (tokens) of fp32 holding the exact integer sums; the epilogue thread owns a channel:
//     acc[t] += float(isum) * (a7[t]/127) * S, S a per-thread scalar, a7/127 broadcast from shared memory -- the reference's
//     operation order, bit-identical to the other kernels' fp32 results.  Stores are coalesced (a warp = 32 channels of one token).
//   * Work = one line of (128-channel block, k-tile) units over all matrices of the group, cut into one contiguous range per CTA:
//       - more blocks than SMs: stream-K.  A range is at least one block long, so a block is cut by at most ONE boundary; its
//         two parts meet in y through red.global.add.f32 on an output K1 has zeroed (two addends commute: the result does not
//         depend on arrival order) and every SM works to the end (224 blocks on 148 SMs: 48.4 k-tiles each instead of 64);
//       - at most half as many blocks as SMs: the blocks' K halves (two addends again), or P >= 3 K parts per block that leave
//         their sums in workspace slots and count themselves in -- the LAST part to arrive adds the slots in part order and
//         stores (deterministic, no waiting, no zeroed output; the planner weighs a part's k-tiles against that reduction);
//       - otherwise whole blocks.
//     (An ordered fix-up where a block's owner WAITS for the other parts was built first and measured 5-20 us per launch:
//     every hand-over is two dependent L2 round trips under load, and a waiting CTA delays the SM's next launch.)
//   * MUL_MAT_ID: the line is (chunk, block) pairs, a chunk = up to 64 rows of one expert at 64-aligned virtual rows of K1's
//     units (tcq4_moe.cu); the expert's tiles come through a 3-D tensor map, the epilogue scatters rows through a map.
// Warps: 0 weight producer (TMA, HBM), 1 MMA issuer, 2 token producer (bulk copies, L2), 3 publisher (per-channel scales and
// row factors for the epilogue), 4-19 widening -- two teams of 8 on alternating k-tiles (lane quarter = warp % 4, K half =
// (warp / 4) % 2) so that the hand-over latencies of one team hide under the other's work -- 20-23 epilogue.  Rings: 6 weight
// stages (114 KB in flight: a weight stream is bound by latency x bandwidth per SM), 4 token stages, 2 operand stages (A in
// TMEM, B in shared memory), 2 accumulator stages.
//
// Replaces, for these M, the live kernel tcq4_rrs_fused_gemm_smallM (/root/reference/ggml/src/ggml-cuda/tcq4_k32.cu:733-889),
// which serves every M with one scalar loop (dispatch rrs.cu:317-323), and the per-expert loop of ggml_cuda_mul_mat_id
// (ggml-cuda.cu:2287-2436) for a TCQ4 src0.
#include <cuda.h>
#include <cuda_fp16.h>

#include <algorithm>
#include <cstdlib>
#include <cstring>

#include "tcq4_common.cuh"
#include "tcq4_kernels.h"

namespace tcq4 {

constexpr int G4_CH = 128;                                // channels per CTA (TMEM lanes)
constexpr int G4_MAXMT = 64;                              // tokens (MMA N), 32 or 64
constexpr int G4_CW_WARP0 = 4, G4_CW_WARPS = 16;          // widening warps
constexpr int G4_EPI_WARP0 = G4_CW_WARP0 + G4_CW_WARPS, G4_EPI_WARPS = 4;
constexpr int G4_THREADS = (G4_EPI_WARP0 + G4_EPI_WARPS) * 32;           // 768
// Two raw rings.  Weights come from HBM: what bounds a weight stream is the bytes in flight per SM (latency x bandwidth:
// ~60 KB per SM for 7 TB/s), so their ring is deep; the tokens' nibbles come from L2 and get a shallow one.
constexpr int G4_RAW_W = 16 * TILE_BYTES;                 // 18944: the 16 tiles of the block, whole (nibbles + metadata)
constexpr int G4_W_BYTES = ((G4_RAW_W + 127) / 128) * 128;                // 18944
constexpr int G4_W_STAGES = 6;
constexpr int G4_RAW_A = 8 * G4_MAXMT * 16;               // 8192: [8 groups][Mt rows][16 B] nibbles of the tokens
constexpr int G4_RAW_A7 = G4_MAXMT * 4;                   // 256
constexpr int G4_A_BYTES = G4_RAW_A + G4_RAW_A7;          // 8448
constexpr int G4_A_STAGES = 4;
constexpr int G4_B_BYTES = 32 * G4_MAXMT * 16;            // 32768: [32 K-chunks][Mt][16 B] fp16
constexpr int G4_RING = 8;
// TMEM: accumulators 2 x 64 columns, then two k-tile buffers of fp16 weights (256 K = 128 columns of packed pairs)
constexpr int G4_TM_A0 = 2 * G4_MAXMT;                    // 128
constexpr int G4_TM_A_COLS = 128;
static_assert(G4_TM_A0 + 2 * G4_TM_A_COLS <= 512, "TMEM budget");

struct Gemm4Smem {
    static constexpr int off_b = 0;                                            // 2 x 32 KB
    static constexpr int off_raw = off_b + 2 * G4_B_BYTES;                     // 65536: weight ring
    static constexpr int off_act = off_raw + G4_W_STAGES * G4_W_BYTES;         // token ring
    static constexpr int off_s = off_act + G4_A_STAGES * G4_A_BYTES;           // [ring][128] S
    static constexpr int off_rm = off_s + G4_RING * G4_CH * 4;                 // [ring][64] a7/127
    static constexpr int off_zf = off_rm + G4_RING * G4_MAXMT * 4;             // [ring][4]
    static constexpr int off_bar = off_zf + G4_RING * 4 * 4;
    static constexpr int n_bar = 2 * G4_W_STAGES + 2 * G4_A_STAGES + 2 + 2 + 2 + 2 + G4_RING;
    static constexpr int off_tmem = off_bar + n_bar * 8;
    static constexpr int total = off_tmem + 16;
};
static_assert(Gemm4Smem::off_raw % 128 == 0 && G4_W_BYTES % 128 == 0 && Gemm4Smem::off_act % 128 == 0 && G4_A_BYTES % 128 == 0 &&
                  Gemm4Smem::off_bar % 8 == 0, "alignment");
static_assert(Gemm4Smem::total <= 227 * 1024, "shared memory budget");

__device__ __forceinline__ void g4_wait(uint64_t *bar, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "G4_WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1, %2;\n\t"
        "@!p bra G4_WAIT_%=;\n\t}" ::"r"(smem_u32(bar)),
        "r"(parity), "r"(20000u)
        : "memory");
}
__device__ __forceinline__ void g4_tma_2d(void *dst_smem, const CUtensorMap *map, int c0, int c1, uint64_t *bar, uint64_t policy) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1, {%3, %4}], "
        "[%2], %5;" ::"r"(smem_u32(dst_smem)),
        "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "l"(policy)
        : "memory");
}
__device__ __forceinline__ void g4_tma_3d(void *dst_smem, const CUtensorMap *map, int c0, int c1, int c2, uint64_t *bar, uint64_t policy) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1, {%3, %4, %5}], "
        "[%2], %6;" ::"r"(smem_u32(dst_smem)),
        "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "l"(policy)
        : "memory");
}
// D[tmem] (+)= A[tmem] * B[smem], fp16 x fp16 -> fp32, one CTA
__device__ __forceinline__ void umma_f16_ts(uint32_t d_tmem, uint32_t a_tmem, uint32_t b_lo, uint32_t b_hi, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t.reg .b64 bd;\n\t"
        "setp.ne.b32 p, %5, 0;\n\t"
        "mov.b64 bd, {%2, %3};\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], bd, %4, p;\n\t}" ::"r"(d_tmem),
        "r"(a_tmem), "r"(b_lo), "r"(b_hi), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void g4_tmem_st_32x16(uint32_t taddr, const uint32_t (&r)[16]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};" ::
            "r"(taddr),
        "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]), "r"(r[10]),
        "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
        : "memory");
}
__device__ __forceinline__ uint32_t g4_hfma2(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t r;
    asm("fma.rn.f16x2 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
    return r;
}
constexpr uint32_t G4_MAGIC = 0x54005400u;      // (64.0, 64.0): see tcq4_gemm3.cu for the widening identities
__device__ __forceinline__ uint32_t g4_lop(uint32_t a) {
    uint32_t r;
    asm("lop3.b32 %0, %1, %2, %3, 0x6A;" : "=r"(r) : "r"(a), "r"(0x00F000F0u), "r"(G4_MAGIC | 0x00800080u));
    return r;
}
__device__ __forceinline__ uint4 g4_cvt_w(uint32_t w, uint32_t sc2, uint32_t c72) {      // two's-complement q, p = sc * q
    return make_uint4(g4_hfma2(g4_lop(w << 4), sc2, c72), g4_hfma2(g4_lop(w), sc2, c72), g4_hfma2(g4_lop(w >> 4), sc2, c72),
                      g4_hfma2(g4_lop(w >> 8), sc2, c72));
}
__device__ __forceinline__ uint4 g4_cvt_act(uint32_t w) {      // u = q + 8 stored by K1
    constexpr uint32_t K16 = 0x4C004C00u, KM1032 = 0xE408E408u, K1 = 0x3C003C00u, KM72 = 0xD480D480u;
    const uint32_t hi = w >> 8;
    return make_uint4(g4_hfma2((w & 0x000F000Fu) | G4_MAGIC, K16, KM1032), g4_hfma2((w & 0x00F000F0u) | G4_MAGIC, K1, KM72),
                      g4_hfma2((hi & 0x000F000Fu) | G4_MAGIC, K16, KM1032), g4_hfma2((hi & 0x00F000F0u) | G4_MAGIC, K1, KM72));
}
__device__ __forceinline__ uint64_t g4_pack(float lo, float hi) {
    uint64_t r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void g4_unpack(uint64_t v, float &lo, float &hi) {
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ uint64_t g4_mul2(uint64_t a, uint64_t b) {
    uint64_t r;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ uint64_t g4_fma2(uint64_t a, uint64_t b, uint64_t c) {
    uint64_t r;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
    return r;
}
__device__ __noinline__ float g4_zero_point_term(const uint8_t *__restrict__ w, const int16_t *__restrict__ sumq, int m, int n,
                                                 int kt, int KT, float a7) {
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

// Timeline stamps and what-if switches (diagnostics build only, TCQ4_G4_PROF / TCQ4_G4_EXP): SM clocks of the middle CTA,
// [role 4: MMA issuer, widening warp 0, widening warp 15, epilogue warp 0][k-tile 64][event 8]
#ifndef G4_NOZP
#define G4_NOZP 0
#endif
constexpr int G4_PROF_KT = 64;
enum { G4_WHOLE = 0, G4_TWO = 1, G4_PARTS = 2 };      // how the ranges of the blocks * KT line meet in y (Gemm4Plan::mode)
#ifdef TCQ4_DIAG
__device__ long long g_g4_prof[4 * G4_PROF_KT * 8];
#define G4_STAMP(role, kt, ev)                                                                                            \
    do {                                                                                                                  \
        if (prof_on && (kt) < G4_PROF_KT) g_g4_prof[((role) * G4_PROF_KT + (kt)) * 8 + (ev)] = clock64();                 \
    } while (0)
#define G4_EXP(bit) (exp & (bit))
#else
#define G4_STAMP(role, kt, ev) do { } while (0)
#define G4_EXP(bit) false
#endif

// the same summed over the k-tiles [kt0, kt1) of a segment (tiles without zero points contribute exact zeros)
__device__ __noinline__ float g4_zero_point_range(const uint8_t *__restrict__ w, const int16_t *__restrict__ sumq,
                                                  const float *__restrict__ a7, int m, int n, int kt0, int kt1, int KT) {
    float z = 0.0f;
    for (int kt = kt0; kt < kt1; kt++) z += g4_zero_point_term(w, sumq, m, n, kt, KT, a7[(size_t)m * KT + kt]);
    return z;
}

// up to four matrices over the same tokens (a fan-out group: Q/K/V, gate/up): their 128-channel blocks form one line
struct G4Params {
    CUtensorMap tm[GEMM4_MAX_SEG];
    const uint8_t *w[GEMM4_MAX_SEG];
    float *y[GEMM4_MAX_SEG];
    long long ldc[GEMM4_MAX_SEG];
    int N[GEMM4_MAX_SEG];
    int blk0[GEMM4_MAX_SEG + 1];      // first block of each segment (unused: INT_MAX)
    int nseg;
    // bias / residual fused into the epilogue (tcq4_b200_mul_mat_add; one segment only): the part of a block that holds k-tile 0
    // starts its accumulators from bias[n] + res[m][n] (same order as the pair GEMMs); res may alias y
    const float *bias;                // [N] or null
    const float *res;                 // [M][ldres] or null
    long long ldres;
    // MUL_MAT_ID, device-routed (tcq4_moe.cu): the line is (chunk, block) pairs -- a chunk = up to 64 rows of one expert at the
    // 64-aligned virtual rows of K1's units; tm[0] is then a 3-D map [expert][tile row][words], w[0] / N[0] / y[0] describe one
    // expert's matrix and the [row][N] output
    const int *id_nchunks;            // null: not this mode
    const int *id_chunk_expert;
    const int32_t *id_dst_map;        // virtual row -> row of y (-1: empty)
    long long id_stride;              // bytes between experts
    int id_blocks;                    // 128-channel blocks of one expert
};

// NCH = Mt / 32 (1 or 2).  ID: the MUL_MAT_ID walk over (chunk, block) pairs -- an instantiation of its own, because its
// scatter code inside the epilogue cost the plain kernel 7 % when it was a run-time branch (register allocation of the hot loop)
// EXTRA: the K-parts reduction and the bias / residual start values are compiled in (same reason: code the plain launch never runs)
template <int NCH, bool DEBUG, bool ID = false, bool EXTRA = false>
__global__ void __launch_bounds__(G4_THREADS, 1)
tcq4_gemm4_kernel(const __grid_constant__ G4Params P, ActStage st, int M, int K, int num_blocks, int mode, int parts, float *__restrict__ partials,
                  int *__restrict__ counters, int32_t *__restrict__ dbg_isum, int exp) {
    extern __shared__ __align__(128) uint8_t smem[];
#ifdef TCQ4_DIAG
    const bool prof_on = (exp & 0x100) && blockIdx.x == gridDim.x / 2 && (threadIdx.x & 31) == 0;
#endif
    constexpr int MT = 32 * NCH;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int KT = K / TILE_K;
    uint8_t *sB = smem + Gemm4Smem::off_b;
    uint8_t *sRaw = smem + Gemm4Smem::off_raw;
    uint8_t *sAct = smem + Gemm4Smem::off_act;
    float *sS = reinterpret_cast<float *>(smem + Gemm4Smem::off_s);
    float *sRm = reinterpret_cast<float *>(smem + Gemm4Smem::off_rm);
    int *sZf = reinterpret_cast<int *>(smem + Gemm4Smem::off_zf);
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem + Gemm4Smem::off_bar);
    uint64_t *wfull = bars, *wempty = wfull + G4_W_STAGES, *afull = wempty + G4_W_STAGES, *aempty = afull + G4_A_STAGES,
             *opfull = aempty + G4_A_STAGES, *opfree = opfull + 2, *accfull = opfree + 2, *accempty = accfull + 2, *snfull = accempty + 2;
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(smem + Gemm4Smem::off_tmem);

    if (threadIdx.x == 0) {
        for (int i = 0; i < G4_W_STAGES; i++) {
            mbar_init(&wfull[i], 1);
            mbar_init(&wempty[i], G4_CW_WARPS / 2 + 1);      // one team + the publisher
        }
        for (int i = 0; i < G4_A_STAGES; i++) {
            mbar_init(&afull[i], 1);
            mbar_init(&aempty[i], G4_CW_WARPS / 2 + 1);
        }
        for (int i = 0; i < 2; i++) {
            mbar_init(&opfull[i], G4_CW_WARPS / 2);
            mbar_init(&opfree[i], 1);
            mbar_init(&accfull[i], 1);
            mbar_init(&accempty[i], G4_EPI_WARPS);
        }
        for (int i = 0; i < G4_RING; i++) mbar_init(&snfull[i], 1);      // the publisher warp
        fence_mbar_init();
    }
    if (warp == 0 && lane < P.nseg) asm volatile("prefetch.tensormap [%0];" ::"l"(&P.tm[lane]) : "memory");
    if (warp == 1) {
        tmem_alloc(tmem_slot, 512);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;
    pdl_launch_dependents();
    // this CTA's range of the blocks * KT line, walked segment by segment (block nb, k-tiles [kt0, kt1)) by every role
    // (mode G4_WHOLE: ranges of whole blocks only -- strided outputs cannot be zeroed by K1)
    const bool whole_blocks = (mode == G4_WHOLE);
    constexpr bool id_mode = ID;
    if (id_mode) {      // the chunk table comes from the routing kernel two launches ahead: nothing may start before it is there
        pdl_wait();
        num_blocks = __ldcg(P.id_nchunks) * P.id_blocks;
    }
    const long long total_units = (long long)num_blocks * KT;
    const long long lin0 = whole_blocks ? ((long long)num_blocks * blockIdx.x / gridDim.x) * KT : total_units * blockIdx.x / gridDim.x;
    const long long lin1 = whole_blocks ? ((long long)num_blocks * (blockIdx.x + 1) / gridDim.x) * KT : total_units * (blockIdx.x + 1) / gridDim.x;
#define G4_FOR_SEGMENTS                                                                                              \
    for (long long lin = lin0, seg_len = 0; lin < lin1; lin += seg_len)                                             \
        if (const int nb = (int)(lin / KT); true)                                                                    \
            if (const int kt0 = (int)(lin - (long long)nb * KT); true)                                              \
                if (const int kt1 = (int)(((long long)KT - kt0 < lin1 - lin) ? KT : kt0 + (lin1 - lin)); (seg_len = kt1 - kt0), true)

    if (warp == 0) {
        asm volatile("setmaxnreg.dec.sync.aligned.u32 64;" ::: "memory");
        // ===================== weight producer (the weights do not depend on the kernel before this one: no griddepcontrol.wait) ==========
        if (lane == 0) {
            const uint64_t pol_w = l2_policy_evict_first();
            int s = 0, u = 0;
            G4_FOR_SEGMENTS {
                const int sg = (nb >= P.blk0[1]) + (nb >= P.blk0[2]) + (nb >= P.blk0[3]);
                int row16 = (nb - P.blk0[sg]) * 16, expert = 0;
                if (id_mode) {
                    const int c = nb / P.id_blocks;
                    row16 = (nb - c * P.id_blocks) * 16;
                    expert = __ldcg(P.id_chunk_expert + c);
                }
#pragma unroll 1
                for (int kt = kt0; kt < kt1; kt++) {
                    if (u > 0) g4_wait(&wempty[s], (uint32_t)((u - 1) & 1));
                    mbar_arrive_expect_tx(&wfull[s], (uint32_t)G4_RAW_W);
                    if (G4_EXP(16))      // what-if: the same bytes as one contiguous unit (is the tile layout what bounds the stream?)
                        bulk_g2s(sRaw + s * G4_W_BYTES, P.w[0] + ((size_t)nb * KT + kt) * G4_RAW_W, G4_RAW_W, &wfull[s]);
                    else if (id_mode)
                        g4_tma_3d(sRaw + s * G4_W_BYTES, &P.tm[0], kt * (TILE_BYTES / 8), row16, expert, &wfull[s], pol_w);
                    else
                    g4_tma_2d(sRaw + s * G4_W_BYTES, &P.tm[sg], kt * (TILE_BYTES / 8), row16, &wfull[s], pol_w);
                    if (++s == G4_W_STAGES) { s = 0; u++; }
                }
            }
        }
    } else if (warp == 2) {
        asm volatile("setmaxnreg.dec.sync.aligned.u32 64;" ::: "memory");
        // ===================== token producer: K1's nibbles and row factors of the k-tile (L2) =====================
        if (lane == 0) {
            pdl_wait();
            constexpr uint32_t total_tx = (uint32_t)(8 * MT * 16 + MT * 4);
            int s = 0, u = 0;
            G4_FOR_SEGMENTS {
#pragma unroll 1
                for (int kt = kt0; kt < kt1; kt++) {
                    if (u > 0) g4_wait(&aempty[s], (uint32_t)((u - 1) & 1));
                    uint8_t *stg = sAct + s * G4_A_BYTES;
                    mbar_arrive_expect_tx(&afull[s], total_tx);
                    // (MUL_MAT_ID: chunk c of the block is half `c & 1` of K1's 128-row unit c / 2)
                    const int chunk = id_mode ? nb / P.id_blocks : 0;
                    const size_t unit = (size_t)(chunk >> 1) * KT + kt;
                    const int row0 = (chunk & 1) * 64;
                    const uint8_t *a_src = reinterpret_cast<const uint8_t *>(st.codes) + unit * ACT_NIB_UNIT_BYTES + row0 * 16;
#pragma unroll
                    for (int g = 0; g < 8; g++) bulk_g2s(stg + g * (MT * 16), a_src + g * 2048, MT * 16, &afull[s]);
                    bulk_g2s(stg + 8 * MT * 16, st.a7t + unit * 128 + row0, MT * 4, &afull[s]);
                    if (++s == G4_A_STAGES) { s = 0; u++; }
                }
            }
        }
    } else if (warp == 1) {
        asm volatile("setmaxnreg.dec.sync.aligned.u32 64;" ::: "memory");
        // ===================== MMA issuer =====================
        constexpr uint32_t idesc = (1u << 4) | ((uint32_t)(MT >> 3) << 17) | ((uint32_t)(G4_CH >> 4) << 24);      // F32 acc, f16 x f16, M = 128, N = MT
        int gk = 0;
        G4_FOR_SEGMENTS {
            for (int kt = kt0; kt < kt1; kt++, gk++) {
                const int b = gk & 1, u2 = gk >> 1;
                G4_STAMP(0, gk, 0);
                if (u2 > 0) g4_wait(&accempty[b], (uint32_t)((u2 - 1) & 1));
                G4_STAMP(0, gk, 1);
                g4_wait(&opfull[b], (uint32_t)(u2 & 1));
                G4_STAMP(0, gk, 2);
                tc_fence_after();
                if (elect_one()) {
                    const uint32_t d = tmem + b * G4_MAXMT;
                    const uint32_t a_t = tmem + G4_TM_A0 + b * G4_TM_A_COLS;
                    // the descriptor's start-address field sits in the low word (address >> 4): a K step is a constant added to it
                    const uint64_t bd0 = umma_desc_kmajor_noswz(smem_u32(sB) + b * G4_B_BYTES, MT * 16, 128);
                    const uint32_t b_lo = (uint32_t)bd0, b_hi = (uint32_t)(bd0 >> 32);
                    if (!G4_EXP(8))
#pragma unroll
                    for (int ks = 0; ks < 16; ks++)      // K = 16: 8 TMEM columns of A, two 8-element K-chunks of B
                        umma_f16_ts(d, a_t + ks * 8, b_lo + (uint32_t)ks * ((2 * MT * 16) >> 4), b_hi, idesc, ks ? 1u : 0u);
                    umma_commit(&opfree[b]);
                    umma_commit(&accfull[b]);
                }
                __syncwarp();
                G4_STAMP(0, gk, 3);
            }
        }
    } else if (warp >= G4_CW_WARP0 && warp < G4_EPI_WARP0) {
        asm volatile("setmaxnreg.dec.sync.aligned.u32 64;" ::: "memory");
        // ===================== widening warps: two teams of 8, team t owns the k-tiles with (index & 1) == t =====================
        // Every step of a k-tile's hand-over (barrier waits, shared-memory loads, tcgen05.st completion, fences, arrives) is
        // latency, ~1 100 cycles of it per k-tile when all 16 warps walk the same k-tile in lockstep (measured).  Two
        // teams on alternating k-tiles overlap each other's latencies and amortise them over twice the work per warp.
        const int cw = warp - G4_CW_WARP0;                  // 0..15
        const int team = cw >> 3;                           // = operand stage (TMEM A buffer, B slot) this team fills
        const int hf = (cw >> 2) & 1;                       // K half: groups 4hf .. 4hf+3 of this thread's channel
        const int ch = (warp & 3) * 32 + lane;              // channel of the block = TMEM lane
        const int tt = (cw & 7) * 32 + lane;                // 0..255 within the team: token units (group, row)
        constexpr int A_UNITS = (8 * MT) / 256;             // token units per thread and k-tile: 1 (Mt = 32) or 2
        const uint32_t ta = tmem + ((uint32_t)((warp & 3) * 32) << 16) + G4_TM_A0 + team * G4_TM_A_COLS + hf * 64;
        const int w_off = (ch >> 3) * TILE_BYTES + (ch & 7) * 16 + hf * 512;
        const int sc_off = (ch >> 3) * TILE_BYTES + TILE_OFF_SC + (ch & 7) * 8 + hf * 4;
        uint8_t *sBt = sB + team * G4_B_BYTES;
        int gk = 0;
        G4_FOR_SEGMENTS {
            for (int kt = kt0; kt < kt1; kt++, gk++) {
                if ((gk & 1) != team) continue;
                const int s = gk % G4_W_STAGES, as = gk % G4_A_STAGES;
                const int prole = cw == 0 ? 1 : 2;
#ifdef TCQ4_DIAG
                const bool prof_on_cw = prof_on && (cw == 0 || cw == 15);
#define G4_CSTAMP(ev) do { if (prof_on_cw && gk < G4_PROF_KT) g_g4_prof[(prole * G4_PROF_KT + gk) * 8 + (ev)] = clock64(); } while (0)
#else
#define G4_CSTAMP(ev) do { } while (0)
#endif
                (void)prole;
                G4_CSTAMP(0);
                g4_wait(&wfull[s], (uint32_t)((gk / G4_W_STAGES) & 1));
                G4_CSTAMP(1);
                g4_wait(&afull[as], (uint32_t)((gk / G4_A_STAGES) & 1));
                G4_CSTAMP(2);
                const uint8_t *stg = sRaw + s * G4_W_BYTES;
                const uint8_t *sta = sAct + as * G4_A_BYTES;
                uint4 wv[4];
#pragma unroll
                for (int g = 0; g < 4; g++) wv[g] = *reinterpret_cast<const uint4 *>(stg + w_off + g * 128);
                const uint32_t sc4 = *reinterpret_cast<const uint32_t *>(stg + sc_off);      // the 4 int8 group scales
                uint4 wa[A_UNITS];
#pragma unroll
                for (int j = 0; j < A_UNITS; j++) {
                    const int un = tt + 256 * j, a_g = un / MT, a_row = un - a_g * MT;
                    wa[j] = *reinterpret_cast<const uint4 *>(sta + a_g * (MT * 16) + a_row * 16);
                }
                G4_CSTAMP(7);
                if (gk > 1) {
                    g4_wait(&opfree[team], (uint32_t)(((gk >> 1) - 1) & 1));
                    tc_fence_after();
                }
                G4_CSTAMP(3);
                // ---- weights: 4 groups = 16 words -> 16 operand cells -> 64 TMEM columns of this lane
                if (!G4_EXP(1)) {
#pragma unroll
                    for (int g = 0; g < 4; g++) {
                        const __half sh = __int2half_rn((int)(int8_t)(sc4 >> (8 * g)));
                        const __half2 s2 = __halves2half2(sh, sh);
                        const __half2 c2 = __hmul2(s2, __float2half2_rn(-72.0f));      // exact: 72*sc = 8 * (9*sc), 9*sc < 2048
                        const uint32_t su = *reinterpret_cast<const uint32_t *>(&s2), cu = *reinterpret_cast<const uint32_t *>(&c2);
                        const uint4 c0 = g4_cvt_w(wv[g].x, su, cu), c1 = g4_cvt_w(wv[g].y, su, cu), c2v = g4_cvt_w(wv[g].z, su, cu),
                                    c3 = g4_cvt_w(wv[g].w, su, cu);
                        const uint32_t r[16] = {c0.x, c0.y, c0.z, c0.w, c1.x, c1.y, c1.z, c1.w, c2v.x, c2v.y, c2v.z, c2v.w, c3.x, c3.y, c3.z, c3.w};
                        g4_tmem_st_32x16(ta + g * 16, r);
                    }
                }
                G4_CSTAMP(4);
                // ---- tokens: (group, row) units of 4 words -> 4 operand cells of the B operand each
                if (!G4_EXP(2)) {
#pragma unroll
                    for (int j = 0; j < A_UNITS; j++) {
                        const int un = tt + 256 * j, a_g = un / MT, a_row = un - a_g * MT;
                        uint8_t *dst = sBt + (a_g * 4) * (MT * 16) + a_row * 16;
                        *reinterpret_cast<uint4 *>(dst) = g4_cvt_act(wa[j].x);
                        *reinterpret_cast<uint4 *>(dst + 1 * (MT * 16)) = g4_cvt_act(wa[j].y);
                        *reinterpret_cast<uint4 *>(dst + 2 * (MT * 16)) = g4_cvt_act(wa[j].z);
                        *reinterpret_cast<uint4 *>(dst + 3 * (MT * 16)) = g4_cvt_act(wa[j].w);
                    }
                }
                G4_CSTAMP(5);
                fence_proxy_async_smem();
                asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
                tc_fence_before();
                __syncwarp();
                G4_CSTAMP(6);
                if (lane == 0) {
                    mbar_arrive(&opfull[team]);
                    mbar_arrive(&wempty[s]);
                    mbar_arrive(&aempty[as]);
                }
            }
        }
    } else if (warp < G4_CW_WARP0) {
        asm volatile("setmaxnreg.dec.sync.aligned.u32 64;" ::: "memory");
        // ===================== publisher (warp 3): what the epilogue needs of a k-tile, out of the rings =====================
        // per-channel super-scales S, the zero-point flag, the tokens' row factors a7 / 127 -- 8 rings slots ahead of the epilogue,
        // so that the raw stages can be recycled as soon as the widening team is done with them
        int gk = 0;
        G4_FOR_SEGMENTS {
            (void)nb;
            for (int kt = kt0; kt < kt1; kt++, gk++) {
                const int s = gk % G4_W_STAGES, as = gk % G4_A_STAGES, ring = gk & (G4_RING - 1);
                g4_wait(&wfull[s], (uint32_t)((gk / G4_W_STAGES) & 1));
                g4_wait(&afull[as], (uint32_t)((gk / G4_A_STAGES) & 1));
                const uint8_t *stg = sRaw + s * G4_W_BYTES;
                const uint8_t *sta = sAct + as * G4_A_BYTES;
                int zf = 0;
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    const int ch = j * 32 + lane;
                    const uint8_t *md = stg + (ch >> 3) * TILE_BYTES;
                    sS[ring * G4_CH + ch] = __half2float(*reinterpret_cast<const __half *>(md + TILE_OFF_S + 2 * (ch & 7)));
                    const float zv = __half2float(*reinterpret_cast<const __half *>(md + TILE_OFF_Z + 2 * (ch & 7)));
                    const uint2 zc = *reinterpret_cast<const uint2 *>(md + TILE_OFF_ZC + (ch & 7) * 8);
                    zf |= (zv != 0.0f) && ((zc.x | zc.y) != 0u);
                }
#pragma unroll
                for (int j = 0; j < NCH; j++) {
                    const int t = j * 32 + lane;
                    sRm[ring * G4_MAXMT + t] = *reinterpret_cast<const float *>(sta + 8 * MT * 16 + t * 4) * (1.0f / 127.0f);
                }
                const int anyz = __any_sync(0xffffffffu, zf);
                // (the ring slot is free: the epilogue finished k-tile gk - 8 before the MMA of gk - 6 could start, and that
                //  MMA's commit is what let the team widen k-tile gk - 4 ... the raw stage of gk itself was filled after that)
                __syncwarp();
                if (lane == 0) {
                    sZf[ring * 4] = anyz;
                    sZf[ring * 4 + 1] = 0;
                    sZf[ring * 4 + 2] = 0;
                    sZf[ring * 4 + 3] = 0;
                    mbar_arrive(&snfull[ring]);
                    mbar_arrive(&wempty[s]);
                    mbar_arrive(&aempty[as]);
                }
            }
        }
    } else {
        asm volatile("setmaxnreg.inc.sync.aligned.u32 152;" ::: "memory");
        // ===================== epilogue: thread = channel =====================
        pdl_wait();
        const int q4 = warp & 3;
        const int chl = q4 * 32 + lane;
        const uint32_t t_base = tmem + ((uint32_t)(q4 * 32) << 16);
        uint64_t acc2[MT / 2];
        int gk = 0;
        G4_FOR_SEGMENTS {
            const int sg = (nb >= P.blk0[1]) + (nb >= P.blk0[2]) + (nb >= P.blk0[3]);
            int n = (nb - P.blk0[sg]) * G4_CH + chl;      // channel within its matrix
            const int N = P.N[sg];
            float *__restrict__ y = P.y[sg];
            const long long ldc = P.ldc[sg];
            const uint8_t *__restrict__ w = P.w[sg];
            int vrow0 = 0;                                  // first (virtual) token row of this segment
            if (id_mode) {
                const int c = nb / P.id_blocks;
                n = (nb - c * P.id_blocks) * G4_CH + chl;
                w += (size_t)__ldcg(P.id_chunk_expert + c) * P.id_stride;
                vrow0 = c * 64;
            }
#pragma unroll
            for (int i = 0; i < MT / 2; i++) acc2[i] = 0ull;
            int zseen = 0;
            if (EXTRA && (P.bias != nullptr || P.res != nullptr) && kt0 == 0 && n < N) {
                const float bv = P.bias ? __ldg(P.bias + n) : 0.0f;
#pragma unroll
                for (int i = 0; i < MT / 2; i++) {
                    float a = bv, bb = bv;
                    if (P.res != nullptr) {      // ld.global.cg: the residual may be the output of the kernel before this one
                        if (2 * i < M) a += __ldcg(P.res + (size_t)(2 * i) * P.ldres + n);
                        if (2 * i + 1 < M) bb += __ldcg(P.res + (size_t)(2 * i + 1) * P.ldres + n);
                    }
                    acc2[i] = g4_pack(a, bb);
                }
            }
            for (int kt = kt0; kt < kt1; kt++, gk++) {
                const int b = gk & 1, u2 = gk >> 1, ring = gk & (G4_RING - 1);
#ifdef TCQ4_DIAG
                const bool prof_on_e = prof_on && warp == G4_EPI_WARP0;
#define G4_ESTAMP(ev) do { if (prof_on_e && gk < G4_PROF_KT) g_g4_prof[(3 * G4_PROF_KT + gk) * 8 + (ev)] = clock64(); } while (0)
#else
#define G4_ESTAMP(ev) do { } while (0)
#endif
                G4_ESTAMP(0);
                g4_wait(&snfull[ring], (uint32_t)((gk / G4_RING) & 1));
                G4_ESTAMP(1);
                g4_wait(&accfull[b], (uint32_t)(u2 & 1));
                G4_ESTAMP(2);
                tc_fence_after();
                const float S = sS[ring * G4_CH + chl];
                const uint64_t S2 = g4_pack(S, S);
                const float *rmp = sRm + ring * G4_MAXMT;
                const int anyz = sZf[ring * 4 + 0] | sZf[ring * 4 + 1] | sZf[ring * 4 + 2] | sZf[ring * 4 + 3];
                uint32_t r[NCH][32];
#pragma unroll
                for (int c = 0; c < NCH; c++) tmem_ld_32x32(t_base + b * G4_MAXMT + c * 32, r[c]);      // one wait for all of them
                tmem_ld_wait();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&accempty[b]);
                G4_ESTAMP(3);
                if (!G4_EXP(4))
#pragma unroll
                for (int c = 0; c < NCH; c++) {
#pragma unroll
                    for (int i = 0; i < 16; i++) {
                        const float f0 = __uint_as_float(r[c][2 * i]), f1 = __uint_as_float(r[c][2 * i + 1]);      // exact integer sums
                        const uint64_t rm2 = *reinterpret_cast<const uint64_t *>(rmp + c * 32 + 2 * i);
                        acc2[c * 16 + i] = g4_fma2(g4_mul2(g4_pack(f0, f1), rm2), S2, acc2[c * 16 + i]);
                        if (DEBUG) {
                            const int t = c * 32 + 2 * i;
                            if (dbg_isum && n < N) {
                                if (t < M) dbg_isum[((size_t)t * N + n) * KT + kt] = (int)f0;
                                if (t + 1 < M) dbg_isum[((size_t)(t + 1) * N + n) * KT + kt] = (int)f1;
                            }
                        }
                    }
                }
                G4_ESTAMP(4);
                zseen |= anyz;
            }
            // Zero points (cold: the quantiser writes Z = 0): their term of every k-tile of the segment is added here, after the
            // loop -- as 64 call sites inside the loop the same code cost the plain path 2-4 %.  The sum is the same set of terms in
            // another order (last-bit differences against the per-k-tile form of the other kernels; the oracle bar holds).
            if (!G4_NOZP && zseen && n < N) {
#pragma unroll
                for (int i = 0; i < MT / 2; i++) {
                    float a, bb;
                    g4_unpack(acc2[i], a, bb);
                    if (2 * i < M) a += g4_zero_point_range(w, st.sumq, st.a7, vrow0 + 2 * i, n, kt0, kt1, KT);
                    if (2 * i + 1 < M) bb += g4_zero_point_range(w, st.sumq, st.a7, vrow0 + 2 * i + 1, n, kt0, kt1, KT);
                    acc2[i] = g4_pack(a, bb);
                }
            }
            // ---- end of the segment.  A warp stores 32 consecutive channels of one token: coalesced.
            if (EXTRA && mode == G4_PARTS) {
                // The block is cut into `parts` K ranges, one CTA each (narrow, deep matrices: every SM gets a share of K).  Each
                // part leaves its sums in a workspace slot and counts itself in; whoever arrives LAST adds the slots in part order
                // (its own too, re-read: one code path) and stores -- deterministic, no waiting, no zeroed output.  Counters
                // are per (block, lane quarter): a warp needs only the same warp's columns of the other parts.
                float *slot = partials + (size_t)blockIdx.x * (MT * G4_CH) + chl;
#pragma unroll
                for (int i = 0; i < MT / 2; i++) {
                    float a, bb;
                    g4_unpack(acc2[i], a, bb);
                    if (2 * i < M) __stcg(slot + (2 * i) * G4_CH, a);
                    if (2 * i + 1 < M) __stcg(slot + (2 * i + 1) * G4_CH, bb);
                }
                __threadfence();
                __syncwarp();
                int old = 0;
                if (lane == 0) old = atomicAdd(counters + nb * 4 + q4, 1);
                old = __shfl_sync(0xffffffffu, old, 0);
                if (old != parts - 1) continue;
                __threadfence();
                if (lane == 0) counters[nb * 4 + q4] = 0;      // ready for the next launch
                const float *base = partials + (size_t)nb * parts * (MT * G4_CH) + chl;
                for (int t0 = 0; t0 < M; t0 += 8) {            // 8 tokens x parts loads in flight per round trip
                    float sum[8];
#pragma unroll
                    for (int i = 0; i < 8; i++) sum[i] = 0.0f;
#pragma unroll 1
                    for (int p0 = 0; p0 < parts; p0 += 4) {
                        float pv[4][8];
#pragma unroll
                        for (int pp = 0; pp < 4; pp++)
#pragma unroll
                            for (int i = 0; i < 8; i++)
                                pv[pp][i] = (p0 + pp < parts && t0 + i < M) ? __ldcg(base + (size_t)(p0 + pp) * (MT * G4_CH) + (t0 + i) * G4_CH) : 0.0f;
#pragma unroll
                        for (int pp = 0; pp < 4; pp++)
#pragma unroll
                            for (int i = 0; i < 8; i++)
                                if (p0 + pp < parts) sum[i] += pv[pp][i];
                    }
                    if (n < N) {
#pragma unroll
                        for (int i = 0; i < 8; i++)
                            if (t0 + i < M) y[(size_t)(t0 + i) * ldc + n] = sum[i];
                    }
                }
                continue;
            }
            if (id_mode) {      // MUL_MAT_ID: token t of the chunk goes to row dst_map[vrow0 + t] of y (-1: an empty slot)
                if (n < N) {
#pragma unroll
                    for (int h = 0; h < MT / 16; h++) {
                        int drow[16];
#pragma unroll
                        for (int i = 0; i < 16; i++) drow[i] = __ldcg(P.id_dst_map + vrow0 + h * 16 + i);
#pragma unroll
                        for (int i = 0; i < 8; i++) {
                            float a, bb;
                            g4_unpack(acc2[h * 8 + i], a, bb);
                            if (drow[2 * i] >= 0) y[(size_t)drow[2 * i] * ldc + n] = a;
                            if (drow[2 * i + 1] >= 0) y[(size_t)drow[2 * i + 1] * ldc + n] = bb;
                        }
                    }
                }
                continue;
            }
            // A part of a block (stream-K, K halves) is added to the zeroed output: a block has at most two parts, so the sum does
            // not depend on their order.
            if (n < N) {
                const bool part = (kt0 > 0) || (kt1 < KT);
#pragma unroll
                for (int i = 0; i < MT / 2; i++) {
                    float a, bb;
                    g4_unpack(acc2[i], a, bb);
                    if (part) {
                        if (2 * i < M) atomicAdd(y + (size_t)(2 * i) * ldc + n, a);
                        if (2 * i + 1 < M) atomicAdd(y + (size_t)(2 * i + 1) * ldc + n, bb);
                    } else {
                        if (2 * i < M) y[(size_t)(2 * i) * ldc + n] = a;
                        if (2 * i + 1 < M) y[(size_t)(2 * i + 1) * ldc + n] = bb;
                    }
                }
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc(tmem, 512);
}

// ---- host side ---------------------------------------------------------------------------------------------------
typedef CUresult (*G4EncodeTiled3)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                   const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                   CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
bool gemm4_plan(int64_t M, const int64_t *Ns, int nseg, int64_t K, int num_sms, size_t smem_optin, Gemm4Plan *plan, bool no_atomic) {
    if (M < 1 || M > G4_MAXMT || (size_t)Gemm4Smem::total > smem_optin || num_sms < 1 || nseg < 1 || nseg > GEMM4_MAX_SEG) return false;
    int64_t blocks = 0;
    for (int i = 0; i < nseg; i++) blocks += (Ns[i] + G4_CH - 1) / G4_CH;
    const int64_t KT = K / TILE_K;
    static const bool no_split = getenv("TCQ4_G4_NOSPLIT") != nullptr;
    static const bool no_parts = getenv("TCQ4_G4_NOPARTS") != nullptr;
    int64_t grid = blocks < num_sms ? blocks : num_sms;
    int mode = G4_WHOLE, parts = 1;
    if (!no_split) {
        if (blocks > num_sms && blocks % num_sms != 0) {
            if (!no_atomic) mode = G4_TWO;      // stream-K: ranges >= one block, a block is cut at most once (strided outputs are cleared row by row)
        } else if (2 * blocks <= num_sms && KT >= 8) {
            // K parts.  Cost model in k-tile times (~1 us): a part's k-tiles + the last arriver's reduction (one atomic round
            // trip, then ceil(M/8) * ceil(P/4) load round trips); two parts meet through red.add instead and cost nothing extra
            const int64_t pmax = std::min<int64_t>({(int64_t)num_sms / blocks, (int64_t)GEMM4_MAX_PARTS, KT / 4});
            double best = (double)((KT + 1) / 2);
            int bp = 2;
            for (int64_t p = 3; p <= pmax && !no_parts; p++) {
                const double fix = 2.0 + 0.6 * (double)((M + 7) / 8) * (double)((p + 3) / 4);
                const double c = (double)((KT + p - 1) / p) + fix;
                if (c < best - 0.5) {      // (calibrated on forced part counts, TCQ4_G4_PARTS: N = 4096 / 2560 / 1024, K = 4096 .. 14336, M = 12 .. 64)
                    best = c;
                    bp = (int)p;
                }
            }
            static const int force_p = getenv("TCQ4_G4_PARTS") ? atoi(getenv("TCQ4_G4_PARTS")) : 0;      // experiments
            if (force_p >= 1 && force_p <= pmax) bp = force_p;
            parts = bp;
            grid = blocks * parts;
            // (no_atomic -- a fused residual may sit in y, which can then neither be cleared nor take two adds in either order:
            //  two parts go through the slots as well)
            mode = parts == 1 ? G4_WHOLE : (parts == 2 && !no_atomic) ? G4_TWO : G4_PARTS;
        }
    }
    plan->grid = (int)grid;
    plan->num_blocks = (int)blocks;
    plan->mode = mode;
    plan->parts = parts;
    plan->mt = M <= 32 ? 32 : 64;
    plan->smem_bytes = Gemm4Smem::total;
    plan->partial_bytes = mode == G4_PARTS ? (size_t)grid * plan->mt * G4_CH * 4 : 0;
    plan->counter_bytes = mode == G4_PARTS ? (size_t)blocks * 16 : 0;
    return true;
}

cudaError_t launch_gemm4(const Gemm4Seg *segs, int nseg, ActStage st, int64_t M, int64_t K, const Gemm4Plan &plan,
                         float *partials, int *counters, int32_t *dbg_isum, cudaStream_t stream, const float *res, const float *bias) {
    if (nseg < 1 || nseg > GEMM4_MAX_SEG) return cudaErrorInvalidValue;
    G4Params P;
    memset(&P, 0, sizeof(P));
    const int64_t KT = K / TILE_K;
    int blk = 0;
    for (int i = 0; i <= GEMM4_MAX_SEG; i++) P.blk0[i] = 0x7fffffff;
    for (int i = 0; i < nseg; i++) {
        if (!weight_tensor_map(&P.tm[i], segs[i].w, segs[i].N / TILE_CH, KT, TILE_BYTES / 8, 16)) return cudaErrorInvalidValue;
        P.w[i] = reinterpret_cast<const uint8_t *>(segs[i].w);
        P.y[i] = segs[i].y;
        P.ldc[i] = (long long)segs[i].ldc;
        P.N[i] = (int)segs[i].N;
        P.blk0[i] = blk;
        blk += (int)((segs[i].N + G4_CH - 1) / G4_CH);
    }
    P.nseg = nseg;
    if (blk != plan.num_blocks) return cudaErrorInvalidValue;
    if ((res || bias) && (nseg != 1 || plan.mode == G4_TWO)) return cudaErrorInvalidValue;
    P.bias = bias;
    P.res = res;
    P.ldres = (long long)segs[0].N;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)plan.grid);
    cfg.blockDim = dim3(G4_THREADS);
    cfg.dynamicSmemBytes = plan.smem_bytes;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    static const bool no_pdl = getenv("TCQ4_NO_PDL") != nullptr;
    cfg.numAttrs = no_pdl ? 0 : 1;
    int exp = 0;
#ifdef TCQ4_DIAG
    exp = (getenv("TCQ4_G4_EXP") ? atoi(getenv("TCQ4_G4_EXP")) : 0) | (getenv("TCQ4_G4_PROF") ? 0x100 : 0);
#endif
    if (dbg_isum && nseg != 1) return cudaErrorInvalidValue;
    const bool extra = plan.mode == G4_PARTS || res != nullptr || bias != nullptr;
#define G4_LAUNCH(NCH_, DBG_, EX_)                                                                                                  \
    cudaLaunchKernelEx(&cfg, tcq4_gemm4_kernel<NCH_, DBG_, false, EX_>, P, st, (int)M, (int)K, plan.num_blocks, plan.mode, plan.parts, \
                       partials, counters, dbg_isum, exp)
    if (plan.mt == 32) {
        if (dbg_isum) return extra ? G4_LAUNCH(1, true, true) : G4_LAUNCH(1, true, false);
        return extra ? G4_LAUNCH(1, false, true) : G4_LAUNCH(1, false, false);
    }
    if (dbg_isum) return extra ? G4_LAUNCH(2, true, true) : G4_LAUNCH(2, true, false);
    return extra ? G4_LAUNCH(2, false, true) : G4_LAUNCH(2, false, false);
#undef G4_LAUNCH
}

// MUL_MAT_ID with device-side routing: one launch over (chunk, block) pairs; the chunk count lives on the device, so the grid is
// sized for the bound and CTAs whose range is empty fall through
cudaError_t launch_gemm4_id(const void *w_experts, int64_t n_expert, size_t expert_stride_bytes, ActStage st, float *y, int64_t N,
                            int64_t K, const int *n_chunks, const int *chunk_expert, const int32_t *dst_map, int max_chunks,
                            int num_sms, cudaStream_t stream) {
    static G4EncodeTiled3 enc = [] {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult q = cudaDriverEntryPointSymbolNotFound;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess)
            p = nullptr;
        return (G4EncodeTiled3)p;
    }();
    if (!enc) return cudaErrorInvalidValue;
    G4Params P;
    memset(&P, 0, sizeof(P));
    const int64_t KT = K / TILE_K, NT = N / TILE_CH;
    const cuuint64_t gdim[3] = {(cuuint64_t)KT * (TILE_BYTES / 8), (cuuint64_t)NT, (cuuint64_t)n_expert};
    const cuuint64_t gstride[2] = {(cuuint64_t)KT * TILE_BYTES, (cuuint64_t)expert_stride_bytes};
    const cuuint32_t box[3] = {TILE_BYTES / 8, 16, 1};
    const cuuint32_t estr[3] = {1, 1, 1};
    if (enc(&P.tm[0], CU_TENSOR_MAP_DATA_TYPE_UINT64, 3, const_cast<void *>(w_experts), gdim, gstride, box, estr,
            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
        return cudaErrorInvalidValue;
    for (int i = 0; i <= GEMM4_MAX_SEG; i++) P.blk0[i] = 0x7fffffff;
    P.blk0[0] = 0;
    P.w[0] = reinterpret_cast<const uint8_t *>(w_experts);
    P.y[0] = y;
    P.ldc[0] = (long long)N;
    P.N[0] = (int)N;
    P.nseg = 1;
    P.id_nchunks = n_chunks;
    P.id_chunk_expert = chunk_expert;
    P.id_dst_map = dst_map;
    P.id_stride = (long long)expert_stride_bytes;
    P.id_blocks = (int)((N + G4_CH - 1) / G4_CH);
    const long long items = (long long)max_chunks * P.id_blocks;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(items < num_sms ? items : num_sms));
    cfg.blockDim = dim3(G4_THREADS);
    cfg.dynamicSmemBytes = Gemm4Smem::total;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    static const bool no_pdl = getenv("TCQ4_NO_PDL") != nullptr;
    cfg.numAttrs = no_pdl ? 0 : 1;
    return cudaLaunchKernelEx(&cfg, tcq4_gemm4_kernel<2, false, true>, P, st, 64, (int)K, 0, (int)G4_WHOLE, 1, (float *)nullptr, (int *)nullptr,
                              (int32_t *)nullptr, 0);
}

cudaError_t gemm4_read_prof(long long *host_out, size_t bytes) {
#ifdef TCQ4_DIAG
    return cudaMemcpyFromSymbol(host_out, g_g4_prof, bytes < sizeof(g_g4_prof) ? bytes : sizeof(g_g4_prof));
#else
    (void)host_out; (void)bytes;
    return cudaErrorNotSupported;
#endif
}

cudaError_t gemm4_init_attributes(size_t smem_optin) {
    if ((size_t)Gemm4Smem::total > smem_optin) return cudaErrorInvalidValue;
    cudaError_t e;
#define G4_ATTR(...)                                                                                                                \
    if ((e = cudaFuncSetAttribute(tcq4_gemm4_kernel<__VA_ARGS__>, cudaFuncAttributeMaxDynamicSharedMemorySize, Gemm4Smem::total)) != \
        cudaSuccess)                                                                                                                \
        return e;
    G4_ATTR(1, true, false, false) G4_ATTR(1, true, false, true) G4_ATTR(1, false, false, false) G4_ATTR(1, false, false, true)
    G4_ATTR(2, true, false, false) G4_ATTR(2, true, false, true) G4_ATTR(2, false, false, false) G4_ATTR(2, false, false, true)
    G4_ATTR(2, false, true, false)
#undef G4_ATTR
    return cudaSuccess;
}

}  // namespace tcq4
