// tcq4_gemm2.cu -- K2 for M > 128: W4A4 prefill GEMM on CTA PAIRS (tcgen05.mma.cta_group::2 kind::i8).
//
// Same exact integer formulation as tcq4_gemm.cu (p = sc*q = 16*(sh*q) + sl*q folded into the weight operand, one
// 256-wide accumulation per 256-block, fp32 epilogue once per 256-block), re-tiled so that the CUDA-core work
// per tensor-core cycle halves:
//   * a cluster of two CTAs (one TPC) owns 256 tokens x (16*NTL) channels; each CTA stages its own 128 token
//     rows of s8 activation codes and unpacks only ITS half of the channels (NTL n-tiles = 8*NTL channels ->
//     16*NTL operand rows: hi rows then lo rows).  One M = 256, N = 32*NTL, K = 32 MMA per 32-group, issued by
//     the leader CTA, reads A and B from both CTAs' shared memory and writes 128 lanes x N columns of TMEM in
//     each.  Weight unpack per SM is half of the single-CTA kernel's for the same MMA time.
//   * NTL (2..8) is a template parameter chosen by the planner (waves of pair-tiles x work per tile); the
//     kernel is bound by issue slots of the unpack + epilogue warps (ncu: profiles/), so every loop over
//     n-tiles is unrolled at compile time and nothing in the inner loops depends on a kernel argument.
//   * epilogue per output and 256-block: one IMAD (16*hi + lo), one I2F, and the two fp32 steps packed as
//     f32x2 (FMUL2 / FFMA2), in the reference's operation order.
// Pipeline: 3 stages of {activation unit 32 KB + NTL raw tiles + the peer tiles' 160 metadata bytes} filled by
// one bulk-copy thread, 4 half-k-tile stages of unpacked B, 2 TMEM accumulator stages.
// Cross-CTA signalling: the unpack warps of both CTAs arrive on the LEADER's bfull barriers (which also tells
// the leader that the peer's activation unit has landed, because those warps waited for the peer's stage),
// the epilogue warps of both CTAs on the leader's accempty barriers; the leader's tcgen05.commit multicasts
// to both CTAs' bfree / empty / accfull barriers.  No data crosses CTAs through generic loads/stores: each CTA
// bulk-copies the 160 metadata bytes of the peer's tiles for its own epilogue's per-channel super-scales.
//
// Replaces tcq4_rrs_gemm_imma_kernel (/root/reference/ggml/src/ggml-cuda/tcq4_k32.cu:1217-1412) and, for these M,
// the live scalar kernel tcq4_rrs_fused_gemm_smallM (tcq4_k32.cu:733-889).
#include <cuda.h>      // CUtensorMap (types only; the encoder is fetched through cudaGetDriverEntryPoint)

#include <cstdlib>
#include <type_traits>

#include "tcq4_common.cuh"
#include "tcq4_kernels.h"

namespace tcq4 {

constexpr int G2_BM = 128;                  // tokens per CTA (256 per pair)
constexpr int G2_MAXNTL = 8;                // n-tiles per CTA (8 channels each)
constexpr int G2_MINNTL = 2;
constexpr int G2_UNPACK_WARP0 = 4, G2_UNPACK_WARPS = 8;     // warps 0..3: producer, MMA issuer, two idle (warpgroup-aligned roles)
constexpr int G2_EPI_WARP0 = G2_UNPACK_WARP0 + G2_UNPACK_WARPS, G2_EPI_WARPS = 8;
constexpr int G2_THREADS = (G2_EPI_WARP0 + G2_EPI_WARPS) * 32;             // 20 warps = 5 warpgroups

constexpr int G2_A_BYTES = ACT_UNIT_BYTES;                                  // 32768
constexpr int G2_RAW_BYTES = G2_MAXNTL * TILE_BYTES;                        // 9472: own tiles
constexpr int G2_META_BYTES = 2 * G2_MAXNTL * 160;                          // 2560: S, Z, sc, zc of the pair's 2*NTL tiles
constexpr int G2_STAGE_BYTES = G2_A_BYTES + G2_RAW_BYTES + G2_META_BYTES;   // 44800
#ifndef TCQ4_G2_STAGES
#define TCQ4_G2_STAGES 3
#endif
#ifndef TCQ4_G2_OUT_BOXES
#define TCQ4_G2_OUT_BOXES 2      /* 2 load stages + 6 boxes was measured: 10 % slower (the load ring is the wrong place to save) */
#endif
constexpr int G2_STAGES = TCQ4_G2_STAGES;
constexpr int G2_OUT_BOXES = TCQ4_G2_OUT_BOXES;                             // 32 x 32 B store boxes per epilogue warp
constexpr int G2_BH_BYTES = 16384;                                          // one half k-tile (4 groups) of B at NTL = 8
constexpr int G2_BH_STAGES = 4;
constexpr int G2_SN_RING = 8;

struct Gemm2Smem {
    static constexpr int off_stage = 0;                                     // [stage]{A | raw | meta}
    static constexpr int off_b = off_stage + G2_STAGES * G2_STAGE_BYTES;    // 134400
    static constexpr int off_sn = off_b + G2_BH_STAGES * G2_BH_BYTES;       // 199936
    static constexpr int off_out = off_sn + G2_SN_RING * 128 * 4;           // [8 epilogue warps][boxes][32 rows x 32 B]
    static constexpr int off_zflag = off_out + G2_EPI_WARPS * G2_OUT_BOXES * 1024;
    static constexpr int off_bar = off_zflag + G2_SN_RING * 4 * 4;
    static constexpr int n_bar = 2 * G2_STAGES + 2 * G2_BH_STAGES + 2 + 2 + G2_SN_RING;
    static constexpr int off_tmem = off_bar + n_bar * 8;
    static constexpr int total = off_tmem + 16;
};
static_assert(G2_STAGE_BYTES % 128 == 0 && Gemm2Smem::off_b % 128 == 0 && Gemm2Smem::off_out % 128 == 0,
              "operands and store boxes must be 128-byte aligned");
static_assert(Gemm2Smem::off_bar % 8 == 0, "barriers must be 8-byte aligned");
static_assert(Gemm2Smem::total <= 227 * 1024, "shared memory budget");

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
// Arrive on a barrier anywhere in the cluster (address from g2_mapa).  Default semantics (release at CTA scope),
// exactly what CUTLASS's ClusterBarrier::arrive(cta_id) emits for its 2-SM pipelines: the data this signal covers
// is shared memory of the ARRIVING CTA read by the tensor core's async proxy (ordered by fence.proxy.async before
// the arrive) or TMEM (ordered by tcgen05.fence), never another CTA's generic loads -- a cluster-scope release
// would cost MEMBAR.ALL.GPU per arrive and CCTL.IVALL per wait (seen in SASS, and measured: -30 %).
__device__ __forceinline__ void g2_arrive_cluster(uint32_t cluster_addr) {
    asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
// Wait with a suspend-time hint: the thread sleeps in hardware until the phase completes or the hint (ns) expires,
// instead of re-issuing try_wait every few dozen cycles (polling was ~20 % of all issued instructions).
__device__ __forceinline__ void g2_wait(uint64_t *bar, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "G2_WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1, %2;\n\t"
        "@!p bra G2_WAIT_%=;\n\t}" ::"r"(smem_u32(bar)),
        "r"(parity), "r"(20000u)
        : "memory");
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
// same, the two 64-bit shared-memory descriptors given as {low, high} 32-bit halves
__device__ __forceinline__ void umma_i8_2cta_parts(uint32_t d_tmem, uint32_t a_lo, uint32_t a_hi, uint32_t b_lo,
                                                   uint32_t b_hi, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\t"
        "mov.b64 da, {%1, %2};\n\t"
        "mov.b64 db, {%3, %4};\n\t"
        "setp.ne.b32 p, %6, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::i8 [%0], da, db, %5, p;\n\t}" ::"r"(d_tmem),
        "r"(a_lo), "r"(a_hi), "r"(b_lo), "r"(b_hi), "r"(idesc), "r"(accumulate)
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

__device__ __forceinline__ int g2_mad16(uint32_t hi, uint32_t lo) {      // 16*hi + lo as ONE integer multiply-add
    int r;
    asm("mad.lo.s32 %0, %1, 16, %2;" : "=r"(r) : "r"(hi), "r"(lo));
    return r;
}

// Timeline stamps (PROF instantiation only; diagnostics): [cta 2][role 6][k-tile 16][event 8] SM clocks of one cluster
constexpr int G2_PROF_KT = 32;
__device__ long long g_g2_prof[2 * 6 * G2_PROF_KT * 8];      // [cta 2][role 6][global k-tile 32][event 8]
constexpr int G2_PROF_CTAS = 4096;
__device__ long long g_g2_cta[G2_PROF_CTAS * 6];      // per CTA: smid, entry, set up, main loop done, stored, exit (ns)
__device__ __forceinline__ long long g2_globaltimer() {
    long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
#define G2_CTA_STAMP(ev)                                                                           \
    do {                                                                                           \
        if (PROF && cta_id < G2_PROF_CTAS) g_g2_cta[cta_id * 6 + (ev)] = g2_globaltimer();         \
    } while (0)
#define G2_STAMP(role, kt, ev)                                                                     \
    do {                                                                                           \
        if (PROF && prof_on && (kt) < G2_PROF_KT)                                                  \
            g_g2_prof[(((int)rank * 6 + (role)) * G2_PROF_KT + (kt)) * 8 + (ev)] = clock64();      \
    } while (0)

// 2-D tiled bulk tensor copy global -> shared, completing on an mbarrier (coordinates: {inner, outer} in elements)
__device__ __forceinline__ void g2_tma_2d(void *dst_smem, const CUtensorMap *map, int c0, int c1, uint64_t *bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
            smem_u32(dst_smem)),
        "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
        : "memory");
}
__device__ __forceinline__ void g2_tma_2d_hint(void *dst_smem, const CUtensorMap *map, int c0, int c1, uint64_t *bar,
                                               uint64_t policy) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1, {%3, %4}], "
        "[%2], %5;" ::"r"(smem_u32(dst_smem)),
        "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "l"(policy)
        : "memory");
}

__device__ __forceinline__ void g2_tma_store_2d(const CUtensorMap *map, const void *src_smem, int c0, int c1) {
    asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];" ::"l"(map),
                 "r"(smem_u32(src_smem)), "r"(c0), "r"(c1)
                 : "memory");
}
__device__ __forceinline__ void g2_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N_PENDING>
__device__ __forceinline__ void g2_store_wait_read() {
    asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N_PENDING) : "memory");
}

// ---- activation stage inside the GEMM (FUSED mode, opt-in: TCQ4_G2_FUSE=1; measured slower, see tcq4_api.cu) --------
// The stand-alone activation kernel costs a launch, a one-wave latency chain and a drain in front of EVERY MUL_MAT
// (~4 of the 27 us of an average Qwen3-4B prefill matmul).  In fused mode the GEMM's own unpack and epilogue warps --
// idle until the first weight tiles have landed and the first accumulator is complete -- quantise the M * K/256 blocks
// (spread over all CTAs of the grid) into the same staging layout, then a grid-wide counter releases the producer's
// activation copies and the epilogue's scale reads.  All CTAs of the grid are co-resident (persistent, one per SM), so
// the barrier cannot deadlock; the last CTA to leave resets the counters (cnt[0] = warps done, cnt[1] = CTAs gone).
struct G2Fuse {
    const float   *x;        // null: the staging area was filled by tcq4_act_quant_kernel before this launch
    const int32_t *perm;
    int           *cnt;      // two zeroed ints in the workspace
    // epilogue fusion (independent of the above): accumulators start at bias[n] + res[m][n] instead of zero
    const float   *res;      // [M][N] contiguous or null; may alias y
    const float   *bias;     // [N] or null
};
__device__ __forceinline__ int g2_ld_acquire(const int *p) {
    int v;
    asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void g2_fused_act(const G2Fuse &fz, const ActStage &st, int M, int K, int wi, int nw, int lane,
                                             int *sm_cnt) {
    const int KT = K / TILE_K;
    const int nblocks = M * KT;
    float vn[8];
    if (wi < nblocks) rrs_load_block_cg(fz.x + (size_t)(wi / KT) * K, fz.perm, wi % KT, lane, vn);
    for (int b = wi; b < nblocks; b += nw) {
        const int m = b / KT, kt = b - m * KT;
        float v[8];
#pragma unroll
        for (int i = 0; i < 8; i++) v[i] = vn[i];
        if (b + nw < nblocks) rrs_load_block_cg(fz.x + (size_t)((b + nw) / KT) * K, fz.perm, (b + nw) % KT, lane, vn);
        const ActQuant a = rrs_quantize_block(v, lane);               // lane holds elements lane*8 .. lane*8+7
        // s8 codes, sigma order {0,2,4,6,1,3,5,7}; lane pair (2j, 2j+1) forms 16-B K-chunk j (as tcq4_act.cu MODE_STAGE)
        const uint32_t w0 = (uint32_t)(a.q[0] & 0xFF) | ((uint32_t)(a.q[2] & 0xFF) << 8) |
                            ((uint32_t)(a.q[4] & 0xFF) << 16) | ((uint32_t)(a.q[6] & 0xFF) << 24);
        const uint32_t w1 = (uint32_t)(a.q[1] & 0xFF) | ((uint32_t)(a.q[3] & 0xFF) << 8) |
                            ((uint32_t)(a.q[5] & 0xFF) << 16) | ((uint32_t)(a.q[7] & 0xFF) << 24);
        int8_t *unit = st.codes + ((size_t)(m >> 7) * KT + kt) * ACT_UNIT_BYTES;
        *reinterpret_cast<uint2 *>(unit + (lane >> 1) * ACT_LBO + (m & 127) * 16 + (lane & 1) * 8) = make_uint2(w0, w1);
        if (lane == 0) st.a7[b] = __fdiv_rn(a.max_abs, 7.0f);
        if ((lane & 3) == 0) st.sumq[b * 8 + (lane >> 2)] = (int16_t)a.gsum;
    }
    // publish: every lane's stores are fenced (generic and async proxy: the codes are read by bulk copies of OTHER SMs)
    __threadfence();
    asm volatile("fence.proxy.async;" ::: "memory");
    __syncwarp();
    // one global atomic per CTA (2 368 warps on one address cost ~10 us): the CTA's 16th warp reports for all of them
    if (lane == 0 && atomicAdd_block(sm_cnt, 1) == G2_UNPACK_WARPS + G2_EPI_WARPS - 1) {
        __threadfence();
        atomicAdd(fz.cnt, 1);
    }
}
// producer lane: wait until every CTA has reported, then open the CTA-local gate for the epilogue warps
__device__ __forceinline__ void g2_fused_wait(const G2Fuse &fz, int target, volatile int *sm_go) {
    while (g2_ld_acquire(fz.cnt) < target) __nanosleep(32);
    asm volatile("fence.proxy.async;" ::: "memory");
    *sm_go = 1;
}

// PERSISTENT: one cluster per SM pair walks the (row-pair, column-block) tiles t = cluster, cluster + C, ...
// (row-pairs fastest, so that clusters running at the same time share weight tiles in L2).  Every ring and both
// TMEM accumulator stages keep running across tile boundaries (all roles count k-tiles globally), so the
// producer / unpack / MMA warps are already inside the next tile while the epilogue warps finish and store the
// previous one; there is no per-tile launch, TMEM allocation, cluster barrier or drain (those cost ~40 % of the
// non-persistent version: profiles/).
struct G2YMaps {       // one store map per destination matrix (this GPU's C, or every GPU's C when N-sharded)
    CUtensorMap m[MAX_PEERS];
    int n;
    float *mc;             // non-null: multicast address of C (already offset to this call's first column)
    long long mc_ldc;
};

// OUTM: 0 = one destination (ggml's dst; the code path of the single-GPU product), 1 = several peer-mapped
// destinations (one tiled store per GPU), 2 = one NVSwitch multicast destination (multimem.st).  A template parameter:
// carrying the choice at run time cost the plain path 8-10 % (measured).
// FUSE: the opt-in fused activation stage; a template parameter because its run-time checks alone cost the default
// path 1.7 % (measured: 7.04 vs 6.92 ms per Qwen3-4B prefill step).
template <int NTL, bool DEBUG, bool PROF = false, int OUTM = 0, bool FUSE = false, bool ADD = false>
__global__ void __launch_bounds__(G2_THREADS, 1)
tcq4_gemm2_kernel(const __grid_constant__ CUtensorMap tm_tiles, const __grid_constant__ CUtensorMap tm_meta,
                  const __grid_constant__ G2YMaps tm_y, const uint8_t *__restrict__ w, ActStage st, const G2Fuse fz, int M, int N,
                  int K, int MB2, int num_tiles, int32_t *__restrict__ dbg_isum) {
    extern __shared__ __align__(128) uint8_t smem[];
    constexpr int LBO = 256 * NTL;                               // K-chunk stride of B: (16*NTL rows) * 16 B
    constexpr int BGROUP = 2 * LBO;                              // one 32-group = 2 K-chunks
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int KT = K / TILE_K, NT = N / TILE_CH;
    const uint32_t rank = g2_ctarank();
    const int cluster_id = blockIdx.x >> 1, num_clusters = gridDim.x >> 1;
    const bool prof_on = PROF && cluster_id == num_clusters / 2;
    const int cta_id = (int)blockIdx.x;
    if (PROF && threadIdx.x == G2_EPI_WARP0 * 32) {
        uint32_t smid;
        asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
        if (cta_id < G2_PROF_CTAS) g_g2_cta[cta_id * 6] = smid;
        G2_CTA_STAMP(1);
    }

    uint8_t *sStage = smem + Gemm2Smem::off_stage;
    uint8_t *sB = smem + Gemm2Smem::off_b;
    float *sSn = reinterpret_cast<float *>(smem + Gemm2Smem::off_sn);          // [ring][2 halves][64]
    int *sZflag = reinterpret_cast<int *>(smem + Gemm2Smem::off_zflag);        // [ring][2 halves][2 warps]
    uint8_t *sOut = smem + Gemm2Smem::off_out;                                 // [epilogue warp][2][32 rows x 32 B]
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem + Gemm2Smem::off_bar);
    uint64_t *full = bars, *empty = full + G2_STAGES, *bfull = empty + G2_STAGES, *bfree = bfull + G2_BH_STAGES,
             *accfull = bfree + G2_BH_STAGES, *accempty = accfull + 2, *snfull = accempty + 2;
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(smem + Gemm2Smem::off_tmem);

    if (threadIdx.x == 0) {
        for (int i = 0; i < G2_STAGES; i++) {
            mbar_init(&full[i], 1);
            mbar_init(&empty[i], 1 + G2_UNPACK_WARPS);           // leader's commit (multicast) + local unpack warps
        }
        for (int i = 0; i < G2_BH_STAGES; i++) {
            mbar_init(&bfull[i], 2 * G2_UNPACK_WARPS);           // unpack warps of both CTAs (used in the leader)
            mbar_init(&bfree[i], 1);                             // leader's commit (multicast)
        }
        for (int i = 0; i < 2; i++) {
            mbar_init(&accfull[i], 1);                           // leader's commit (multicast)
            mbar_init(&accempty[i], 2 * G2_EPI_WARPS);           // epilogue warps of both CTAs (used in the leader)
        }
        for (int i = 0; i < G2_SN_RING; i++) mbar_init(&snfull[i], 4);     // 4 local publishing warps
        tmem_slot[1] = 0;        // fused activation stage: warps of this CTA that have reported / gate for the epilogue
        tmem_slot[2] = 0;
        fence_mbar_init();
    }
    if (warp == 0 && lane == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(&tm_tiles) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&tm_meta) : "memory");
        if (OUTM != 2)
            for (int pi = 0; pi < (OUTM == 1 ? tm_y.n : 1); pi++)
                asm volatile("prefetch.tensormap [%0];" ::"l"(&tm_y.m[pi]) : "memory");
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
    // Register re-balance between the warpgroups (the kernel launches with 96 per thread): the epilogue keeps 64
    // fp32 accumulators and wants two TMEM chunks in flight, everybody else needs few.
    // (setmaxnreg is the first instruction of each warpgroup's role branch below)
    if (threadIdx.x == G2_EPI_WARP0 * 32) G2_CTA_STAMP(2);
    pdl_launch_dependents();     // the next kernel of the stream may set itself up behind this one
    pdl_wait();                  // codes / scales (activation kernel) and y belong to earlier kernels until here
    const int fused_nw = (G2_UNPACK_WARPS + G2_EPI_WARPS) * (int)gridDim.x;     // warps that share the activation stage
    const int fused_target = (int)gridDim.x;                                    // one report per CTA
    int *sm_cnt = reinterpret_cast<int *>(tmem_slot) + 1;                       // zeroed with the barriers above
    volatile int *sm_go = reinterpret_cast<volatile int *>(tmem_slot) + 2;

    if (warp < G2_UNPACK_WARP0) {
      if constexpr (NTL >= 7) asm volatile("setmaxnreg.dec.sync.aligned.u32 48;" ::: "memory");
      else asm volatile("setmaxnreg.dec.sync.aligned.u32 56;" ::: "memory");
      if (warp == 0) {
        // ===================== producer: three copies per k-tile =====================
        //   the CTA's activation unit (32 KB, contiguous), its NTL raw tiles and the 160 metadata bytes of all 2*NTL
        //   tiles of the pair.  The tensor is viewed as a 2-D array of 8-byte words [n-tile][KT*148]; one tiled
        //   copy with a box of {148 or 20 words, NTL or 2*NTL rows} replaces one bulk copy per tile (17 copies per
        //   k-tile issued by one thread cost ~2200 cycles of uniform-datapath work and were THE bottleneck of the
        //   first version of this kernel).  Rows beyond the tensor are zero-filled and counted in the byte total.
        if (lane == 0) {
            constexpr uint32_t total_tx = (uint32_t)(G2_A_BYTES + NTL * TILE_BYTES + 2 * NTL * 160);
            const uint64_t pol_w = l2_policy_evict_first();
            int s = 0, u = 0, gk = 0;
            for (int t = cluster_id; t < num_tiles; t += num_clusters) {
                const int mb = 2 * (t % MB2) + (int)rank, nt_pair = (t / MB2) * (2 * NTL), nt0 = nt_pair + (int)rank * NTL;
                const uint8_t *a_src = reinterpret_cast<const uint8_t *>(st.codes) + (size_t)mb * KT * ACT_UNIT_BYTES;
                for (int kt = 0; kt < KT; kt++, gk++) {
                    if (u > 0) g2_wait(&empty[s], (uint32_t)((u - 1) & 1));
                    G2_STAMP(0, gk, 0);
                    uint8_t *stg = sStage + s * G2_STAGE_BYTES;
                    mbar_arrive_expect_tx(&full[s], total_tx);
                    g2_tma_2d_hint(stg + G2_A_BYTES, &tm_tiles, kt * (TILE_BYTES / 8), nt0, &full[s], pol_w);
                    g2_tma_2d(stg + G2_A_BYTES + G2_RAW_BYTES, &tm_meta, kt * (TILE_BYTES / 8) + TILE_NIB / 8, nt_pair,
                              &full[s]);
                    // fused mode: the weights are already on their way; the codes exist once every warp has reported
                    if (FUSE && gk == 0) g2_fused_wait(fz, fused_target, sm_go);
                    bulk_g2s(stg, a_src, G2_A_BYTES, &full[s]);
                    a_src += ACT_UNIT_BYTES;
                    G2_STAMP(0, gk, 1);
                    if (++s == G2_STAGES) { s = 0; u++; }
                }
            }
        }
      } else if (warp == 1) {
        // ===================== MMA issuer (leader CTA only) =====================
        // The whole warp walks the loop converged and ONE elected lane issues (elect.sync, which lets ptxas keep the
        // operands in uniform registers without a per-instruction election loop); the descriptors are a constant high
        // word plus a low word that advances by compile-time constants.  This loop IS the kernel's critical path: the
        // first version (one lane under `lane == 0`, descriptors rebuilt from addresses, ~200 instructions per k-tile
        // at ~10 cycles each in a loaded SM) issued a k-tile's eight MMAs slower than the tensor pipe executes them.
        if (rank == 0) {
            constexpr uint32_t idesc = umma_idesc_i8(2 * G2_BM, 32 * NTL, 1, 1);      // 256 x (32*NTL) x 32, s8 x s8
            constexpr uint32_t a_hi = (uint32_t)(ACT_SBO >> 4) | (1u << 14);          // bits 32..: SBO, version bit 46
            constexpr uint32_t b_hi = (uint32_t)(128 >> 4) | (1u << 14);
            const uint32_t a_lo0 = (smem_u32(sStage) >> 4) | ((uint32_t)(ACT_LBO >> 4) << 16);
            const uint32_t b_lo0 = (smem_u32(sB) >> 4) | ((uint32_t)(LBO >> 4) << 16);
            int my_tiles = 0;
            for (int t = cluster_id; t < num_tiles; t += num_clusters) my_tiles++;
            const int total_k = my_tiles * KT;
            int s = 0;
            for (int gk = 0; gk < total_k; gk++) {
                const int as = gk & 1, u2 = gk >> 1;
                if (lane == 0) G2_STAMP(1, gk, 0);
                if (u2 > 0) g2_wait(&accempty[as], (uint32_t)((u2 - 1) & 1));
                if (lane == 0) G2_STAMP(1, gk, 1);
                const uint32_t d = tmem + as * 256;
                const uint32_t a_lo = a_lo0 + (uint32_t)s * (G2_STAGE_BYTES >> 4);
#pragma unroll
                for (int h = 0; h < 2; h++) {
                    const int hs = 2 * (gk & 1) + h, v = gk >> 1;
                    g2_wait(&bfull[hs], (uint32_t)(v & 1));      // also: both CTAs' stage s has landed
                    if (lane == 0) G2_STAMP(1, gk, 2 + h);
                    tc_fence_after();
                    if (elect_one()) {
                        const uint32_t b_lo = b_lo0 + (uint32_t)hs * (G2_BH_BYTES >> 4);
#pragma unroll
                        for (int gl = 0; gl < 4; gl++) {
                            const int g = h * 4 + gl;
                            umma_i8_2cta_parts(d, a_lo + (uint32_t)g * ((2 * ACT_LBO) >> 4), a_hi,
                                               b_lo + (uint32_t)gl * (BGROUP >> 4), b_hi, idesc, g > 0 ? 1u : 0u);
                        }
                        umma_commit_2cta(&bfree[hs]);
                        if (h == 1) {
                            umma_commit_2cta(&empty[s]);
                            umma_commit_2cta(&accfull[as]);
                        }
                    }
                    __syncwarp();
                }
                if (lane == 0) G2_STAMP(1, gk, 4);
                if (++s == G2_STAGES) s = 0;
            }
        }
      }
    } else if (warp < G2_EPI_WARP0) {
        if constexpr (NTL >= 7) asm volatile("setmaxnreg.dec.sync.aligned.u32 48;" ::: "memory");
        else asm volatile("setmaxnreg.dec.sync.aligned.u32 56;" ::: "memory");
        // ===================== weight unpack (this CTA's NTL n-tiles) =====================
        const int uw = warp - G2_UNPACK_WARP0;
        if (FUSE)
            g2_fused_act(fz, st, M, K, (int)blockIdx.x * (G2_UNPACK_WARPS + G2_EPI_WARPS) + uw, fused_nw, lane, sm_cnt);
        const int ut = threadIdx.x - G2_UNPACK_WARP0 * 32;          // 0..255
        // chunk of this thread inside a half k-tile: (gl, j, c) with 8*NTL chunks per group
        constexpr int PER_GROUP = 8 * NTL;
        const int gl_t = ut / PER_GROUP, rem_t = ut - gl_t * PER_GROUP;
        const int j_t = rem_t >> 3, c_t = rem_t & 7;
        const bool warp_has_chunk = (uw * 32) < 4 * PER_GROUP;      // warp-uniform: any lane of this warp may have one
        const uint32_t bfull_leader = g2_mapa(smem_u32(bfull), 0);
        const int tile_off = G2_A_BYTES + j_t * TILE_BYTES;
        const int b_off = gl_t * BGROUP + j_t * 128 + c_t * 16;
        // publishing role of warps 0..3: super-scale of channel (half, j, c) of the pair's 16*NTL channels
        const int p_half = ut >> 6, p_j = (ut & 63) >> 3, p_c = ut & 7;
        const bool p_valid = p_j < NTL;                             // absent tiles were zero-filled by the copy
        const int p_off = G2_A_BYTES + G2_RAW_BYTES + (p_half * NTL + p_j) * 160;
        int s = 0, u = 0, gk = 0;
        for (int t = cluster_id; t < num_tiles; t += num_clusters) {
            const int nt0 = (t / MB2) * (2 * NTL) + (int)rank * NTL;
            const bool has_chunk = (gl_t < 4) && (j_t < NT - nt0);  // n-tiles beyond the tensor: nothing to unpack
            for (int kt = 0; kt < KT; kt++, gk++) {
                const int prole = (lane == 0 && (uw == 0 || uw == 7)) ? (uw == 0 ? 2 : 3) : -1;
                if (prole >= 0) G2_STAMP(prole, gk, 0);
                g2_wait(&full[s], (uint32_t)(u & 1));
                if (prole >= 0) G2_STAMP(prole, gk, 1);
                const uint8_t *stg = sStage + s * G2_STAGE_BYTES;
                if (uw < 4) {     // per-channel super-scale of all 16*NTL channels + zero-point presence flag
                    float sv = 0.0f;
                    int zf = 0;
                    if (p_valid) {
                        const uint8_t *md = stg + p_off;
                        sv = __half2float(*reinterpret_cast<const __half *>(md + 2 * p_c));
                        const float zv = __half2float(*reinterpret_cast<const __half *>(md + 16 + 2 * p_c));
                        const uint2 zc = *reinterpret_cast<const uint2 *>(md + 96 + p_c * 8);
                        zf = (zv != 0.0f) && ((zc.x | zc.y) != 0u);
                    }
                    const int ring = gk & (G2_SN_RING - 1);
                    sSn[ring * 128 + ut] = sv;
                    const int anyz = __any_sync(0xffffffffu, zf);
                    __syncwarp();
                    if (lane == 0) {
                        sZflag[ring * 4 + uw] = anyz;
                        mbar_arrive(&snfull[ring]);      // release: epilogue(gk) acquires before reading
                    }
                }
                // both halves' raw words are fetched up front, and each half's arithmetic runs BEFORE the wait for its
                // destination stage: what remains behind the wait is four stores, the proxy fence and the arrive
                uint4 wq2[2] = {make_uint4(0, 0, 0, 0), make_uint4(0, 0, 0, 0)};
                int sc2[2] = {0, 0};
                if (warp_has_chunk && has_chunk) {
                    const uint8_t *tp = stg + tile_off;
#pragma unroll
                    for (int h = 0; h < 2; h++) {
                        const int g = h * 4 + gl_t;
                        wq2[h] = *reinterpret_cast<const uint4 *>(tp + g * 128 + c_t * 16);
                        sc2[h] = *reinterpret_cast<const int8_t *>(tp + TILE_OFF_SC + c_t * 8 + g);
                    }
                }
#pragma unroll
                for (int h = 0; h < 2; h++) {
                    const int hs = 2 * (gk & 1) + h, v = gk >> 1;
                    uint32_t uhi[8], ulo[8];
                    if (warp_has_chunk && has_chunk) unpack_chunk_compute(wq2[h], sc2[h], uhi, ulo);
                    if (v > 0) g2_wait(&bfree[hs], (uint32_t)((v - 1) & 1));
                    if (prole >= 0) G2_STAMP(prole, gk, 2 + 2 * h);
                    if (warp_has_chunk) {
                        if (has_chunk) {
                            uint8_t *row = sB + hs * G2_BH_BYTES + b_off;
                            unpack_chunk_store(uhi, ulo, row, row + NTL * 128, LBO);    // lo rows follow the 8*NTL hi rows
                        }
                    }
                    if (prole >= 0) G2_STAMP(prole, gk, 3 + 2 * h);
                }
                // ONE generic->async proxy fence and one hand-over per k-tile, not one per half: the fence costs a few hundred
                // cycles on the unpack warps' critical path (measured: Qwen3-4B prefill step 6.93 -> 6.72 ms, the
                // M=4096 N=9728 K=2560 launch 190 -> 184 us).  bfree stays per half so the next k-tile's first half can
                // be overwritten while the tensor core still reads the second.
                if (warp_has_chunk) fence_proxy_async_smem();
                __syncwarp();
                if (lane == 0) {
                    g2_arrive_cluster(bfull_leader + (uint32_t)(2 * (gk & 1)) * 8);
                    g2_arrive_cluster(bfull_leader + (uint32_t)(2 * (gk & 1) + 1) * 8);
                }
                if (lane == 0) mbar_arrive(&empty[s]);      // after the __syncwarp above: every lane has read its raw bytes
                if (++s == G2_STAGES) { s = 0; u++; }
            }
        }
    } else {
        // ===================== epilogue =====================
        // 8*NTL accumulators + two 32-register TMEM steps in flight
        if constexpr (NTL >= 7) asm volatile("setmaxnreg.inc.sync.aligned.u32 168;" ::: "memory");
        else asm volatile("setmaxnreg.inc.sync.aligned.u32 152;" ::: "memory");
        const int ew = warp - G2_EPI_WARP0;                 // 0..7
        const int quarter = warp & 3;                       // TMEM lane quarter this warp may access
        const int colhalf = ew >> 2;                        // 0: channels unpacked by CTA 0, 1: by CTA 1
        const int row = quarter * 32 + lane;
        const uint32_t accempty_leader = g2_mapa(smem_u32(accempty), 0);
        const uint32_t t_base = tmem + ((uint32_t)(quarter * 32) << 16) + colhalf * (16 * NTL);
        uint8_t *obuf = sOut + ew * (G2_OUT_BOXES * 1024);  // this warp's 32 x 32 B store boxes
        uint64_t acc2[4 * NTL];                             // 8*NTL fp32 accumulators as f32x2 pairs
        int gk = 0;
        int mb = 2 * (cluster_id % MB2) + (int)rank;
        float a7 = 0.0f;
        if (FUSE) {
            g2_fused_act(fz, st, M, K, (int)blockIdx.x * (G2_UNPACK_WARPS + G2_EPI_WARPS) + G2_UNPACK_WARPS + ew, fused_nw, lane,
                         sm_cnt);
            if (lane == 0) {                                    // the scales below come from other CTAs' warps
                while (*sm_go == 0) __nanosleep(32);
                (void)g2_ld_acquire(fz.cnt);
            }
            __syncwarp();
        }
        if (cluster_id < num_tiles && mb * G2_BM + row < M) a7 = __ldcg(st.a7 + (size_t)(mb * G2_BM + row) * KT);
        for (int t = cluster_id; t < num_tiles; t += num_clusters) {
            const int m = mb * G2_BM + row;
            const bool mvalid = m < M;
            const int n_base = ((t / MB2) * (2 * NTL) + colhalf * NTL) * TILE_CH;
            const float *a7p = st.a7 + (size_t)(mvalid ? m : 0) * KT;
            // first scale of the NEXT tile of this cluster (fetched during the last k-tile of this one)
            const int t_next = t + num_clusters;
            const int mb_next = 2 * (t_next % MB2) + (int)rank;
            const int m_next = mb_next * G2_BM + row;
            const float *a7_next = (t_next < num_tiles && m_next < M) ? st.a7 + (size_t)m_next * KT : nullptr;
#pragma unroll
            for (int i = 0; i < 4 * NTL; i++) acc2[i] = 0ull;
            if (ADD && (fz.res != nullptr || fz.bias != nullptr)) {      // a template parameter: its mere presence cost 0.3 %
                // fused bias / residual: the tile's accumulators start from them (their latency hides under the first k-tile)
#pragma unroll
                for (int ch = 0; ch < NTL; ch++) {
                    const int n = n_base + ch * 8;
                    if (n < N) {
                        float4 v0 = make_float4(0.f, 0.f, 0.f, 0.f), v1 = v0;
                        if (fz.bias != nullptr) {
                            v0 = __ldg(reinterpret_cast<const float4 *>(fz.bias + n));
                            v1 = __ldg(reinterpret_cast<const float4 *>(fz.bias + n + 4));
                        }
                        if (fz.res != nullptr && mvalid) {
                            const float4 r0 = *reinterpret_cast<const float4 *>(fz.res + (size_t)m * N + n);
                            const float4 r1 = *reinterpret_cast<const float4 *>(fz.res + (size_t)m * N + n + 4);
                            v0.x += r0.x; v0.y += r0.y; v0.z += r0.z; v0.w += r0.w;
                            v1.x += r1.x; v1.y += r1.y; v1.z += r1.z; v1.w += r1.w;
                        }
                        acc2[ch * 4 + 0] = g2_pack(v0.x, v0.y);
                        acc2[ch * 4 + 1] = g2_pack(v0.z, v0.w);
                        acc2[ch * 4 + 2] = g2_pack(v1.x, v1.y);
                        acc2[ch * 4 + 3] = g2_pack(v1.z, v1.w);
                    }
                }
            }
            for (int kt = 0; kt < KT; kt++, gk++) {
                const int as = gk & 1, u2 = gk >> 1;
                const float rm = a7 * (1.0f / 127.0f);          // row factor a7/127; column factor S[n]
                const float a7_cur = a7;
                if (kt + 1 < KT) {
                    if (mvalid) a7 = a7p[kt + 1];               // next block's scale: its latency hides under this block
                } else {
                    a7 = a7_next ? a7_next[0] : 0.0f;
                }
                const uint64_t rm2 = g2_pack(rm, rm);
                const int ring = gk & (G2_SN_RING - 1);
                const int erole = (lane == 0 && (ew == 0 || ew == 7)) ? (ew == 0 ? 4 : 5) : -1;
                if (erole >= 0) G2_STAMP(erole, gk, 0);
                g2_wait(&snfull[ring], (uint32_t)((gk / G2_SN_RING) & 1));
                g2_wait(&accfull[as], (uint32_t)(u2 & 1));
                if (erole >= 0) G2_STAMP(erole, gk, 1);
                tc_fence_after();
                const float *sn = sSn + ring * 128 + colhalf * 64;
                const int anyz = sZflag[ring * 4 + colhalf * 2] | sZflag[ring * 4 + colhalf * 2 + 1];
                const uint32_t t_hi = t_base + as * 256;
                // TMEM loads: 16 columns of hi and 16 of lo per step, the next step requested before the math of this one
                // (tcgen05.wait::ld waits for ALL loads of the thread, so a step costs max(load latency, math): with 8-column
                // steps the read phase of a k-tile was NTL x ~250 cycles of load latency -- longer than the k-tile's MMA time,
                // and the two accumulator stages stalled the tensor pipe).  An odd NTL ends with one 8-column step.
                constexpr int NP = NTL / 2;                     // 16-column steps
                uint32_t rh[2][16], rl[2][16];
                if (NP > 0) {
                    tmem_ld_32x16(t_hi, rh[0]);
                    tmem_ld_32x16(t_hi + 8 * NTL, rl[0]);
                } else {
                    tmem_ld_32x8(t_hi, *reinterpret_cast<uint32_t(*)[8]>(&rh[0][0]));
                    tmem_ld_32x8(t_hi + 8 * NTL, *reinterpret_cast<uint32_t(*)[8]>(&rl[0][0]));
                }
#pragma unroll
                for (int cp = 0; cp < NP + (NTL & 1); cp++) {
                    const bool tail = cp == NP;                  // the 8-column step of an odd NTL
                    const int nch = tail ? 1 : 2;                // n-tiles in this step
                    // S[] of this k-tile is read BEFORE the stage is handed back (ring slot reuse, see tcq4_gemm.cu)
                    ulonglong2 s01[2], s23[2];
#pragma unroll
                    for (int j = 0; j < 2; j++)
                        if (j < nch) {
                            s01[j] = *reinterpret_cast<const ulonglong2 *>(sn + (2 * cp + j) * 8);
                            s23[j] = *reinterpret_cast<const ulonglong2 *>(sn + (2 * cp + j) * 8 + 4);
                        }
                    tmem_ld_wait();                  // step cp has landed
                    if (cp + 1 < NP) {
                        tmem_ld_32x16(t_hi + (cp + 1) * 16, rh[(cp + 1) & 1]);
                        tmem_ld_32x16(t_hi + 8 * NTL + (cp + 1) * 16, rl[(cp + 1) & 1]);
                    } else if (cp + 1 == NP && (NTL & 1)) {
                        tmem_ld_32x8(t_hi + (cp + 1) * 16, *reinterpret_cast<uint32_t(*)[8]>(&rh[(cp + 1) & 1][0]));
                        tmem_ld_32x8(t_hi + 8 * NTL + (cp + 1) * 16, *reinterpret_cast<uint32_t(*)[8]>(&rl[(cp + 1) & 1][0]));
                    } else {                         // all TMEM reads of this stage are done: hand it back
                        tc_fence_before();
                        __syncwarp();
                        if (lane == 0) g2_arrive_cluster(accempty_leader + (uint32_t)as * 8);
                        if (erole >= 0) G2_STAMP(erole, gk, 2);
                    }
#pragma unroll
                    for (int j = 0; j < 2; j++) {
                        if (j >= nch) continue;
                        const int ch = 2 * cp + j;
                        const uint64_t sv2[4] = {s01[j].x, s01[j].y, s23[j].x, s23[j].y};
#pragma unroll
                        for (int i = 0; i < 4; i++) {
                            const int is0 = g2_mad16(rh[cp & 1][j * 8 + 2 * i], rl[cp & 1][j * 8 + 2 * i]);             // = sum_g sc*dot, exact
                            const int is1 = g2_mad16(rh[cp & 1][j * 8 + 2 * i + 1], rl[cp & 1][j * 8 + 2 * i + 1]);
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
                if (erole >= 0) G2_STAMP(erole, gk, 3);
                if (anyz && mvalid) {
#pragma unroll
                    for (int i = 0; i < 4 * NTL; i++) {
                        const int n = n_base + 2 * i;
                        if (n < N) {
                            float a, b;
                            g2_unpack(acc2[i], a, b);
                            a += g2_zero_point_term(w, st.sumq, m, n, kt, KT, a7_cur);
                            b += g2_zero_point_term(w, st.sumq, m, n + 1, kt, KT, a7_cur);
                            acc2[i] = g2_pack(a, b);
                        }
                    }
                }
            }
            // ---- the tile's 32 rows x 8*NTL channels of this warp: one 32 x 32 B box per 8 channels through shared
            //      memory and a tiled bulk store (rows >= M and channels >= N are clipped by the copy engine)
            if (threadIdx.x == G2_EPI_WARP0 * 32) G2_CTA_STAMP(3);
            if (lane == 0) g2_store_wait_read<0>();         // the boxes are free again (stores of the previous tile)
            __syncwarp();
            const int row0 = mb * G2_BM + quarter * 32;
#pragma unroll
            for (int ch = 0; ch < NTL; ch++) {
                if (OUTM == 2) {
                    // NVLS form: chunks go through the two boxes in PAIRS and are read back 8 rows x 64 B per instruction,
                    // then stored ONCE to the multicast address; the switch replicates every store into all GPUs' C
                    // (this one included).  Wider row segments matter: 32-byte multicast writes reach ~140 GB/s per GPU.
                    uint8_t *bx = obuf + (ch & 1) * 1024;
                    if ((ch & 1) == 0) __syncwarp();             // the previous pair has been read back
                    *reinterpret_cast<ulonglong2 *>(bx + lane * 32) = make_ulonglong2(acc2[ch * 4 + 0], acc2[ch * 4 + 1]);
                    *reinterpret_cast<ulonglong2 *>(bx + lane * 32 + 16) = make_ulonglong2(acc2[ch * 4 + 2], acc2[ch * 4 + 3]);
                    if ((ch & 1) == 1) {                         // both chunks of the pair are in the boxes
                        __syncwarp();
                        const int cpair = ch - 1;
                        if (n_base + cpair * 8 < N) {            // N % 8 == 0 but not % 16: the second chunk may be outside
                            const int q4 = lane & 3, rr = lane >> 2;
                            const uint8_t *src = obuf + (q4 >> 1) * 1024 + (q4 & 1) * 16;
                            const bool in2 = (q4 < 2) || (n_base + ch * 8 < N);
#pragma unroll
                            for (int k = 0; k < 4; k++) {
                                const int r = rr + 8 * k;
                                const float4 v = *reinterpret_cast<const float4 *>(src + r * 32);
                                if (in2 && row0 + r < M)
                                    multimem_st_v4(tm_y.mc + (size_t)(row0 + r) * tm_y.mc_ldc + n_base + cpair * 8 + q4 * 4, v);
                            }
                        }
                    } else if (ch == NTL - 1) {                  // odd NTL: the last chunk alone, 16 rows x 32 B per instruction
                        __syncwarp();
                        const float4 v0 = *reinterpret_cast<const float4 *>(bx + lane * 16);            // rows 0..15
                        const float4 v1 = *reinterpret_cast<const float4 *>(bx + 512 + lane * 16);      // rows 16..31
                        if (n_base + ch * 8 < N) {
                            float *dst = tm_y.mc + (size_t)(row0 + (lane >> 1)) * tm_y.mc_ldc + n_base + ch * 8 + (lane & 1) * 4;
                            if (row0 + (lane >> 1) < M) multimem_st_v4(dst, v0);
                            if (row0 + 16 + (lane >> 1) < M) multimem_st_v4(dst + (size_t)16 * tm_y.mc_ldc, v1);
                        }
                    }
                    continue;
                }
                uint8_t *box = obuf + (ch % G2_OUT_BOXES) * 1024;
                if (ch >= G2_OUT_BOXES) {
                    if (lane == 0) g2_store_wait_read<G2_OUT_BOXES - 1>();     // the store that used this box has read it
                    __syncwarp();
                }
                *reinterpret_cast<ulonglong2 *>(box + lane * 32) = make_ulonglong2(acc2[ch * 4 + 0], acc2[ch * 4 + 1]);
                *reinterpret_cast<ulonglong2 *>(box + lane * 32 + 16) = make_ulonglong2(acc2[ch * 4 + 2], acc2[ch * 4 + 3]);
                fence_proxy_async_smem();
                __syncwarp();
                if (lane == 0) {
                    if (n_base + ch * 8 < N && row0 < M) {
                        if (OUTM == 1) {
                            for (int pi = 0; pi < tm_y.n; pi++) g2_tma_store_2d(&tm_y.m[pi], box, n_base + ch * 8, row0);
                        } else {
                            g2_tma_store_2d(&tm_y.m[0], box, n_base + ch * 8, row0);
                        }
                    }
                    g2_store_commit();
                }
            }
            if (threadIdx.x == G2_EPI_WARP0 * 32) G2_CTA_STAMP(4);
            mb = mb_next;
        }
        if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");     // all stores complete before exit
    }

    tc_fence_before();
    g2_cluster_sync();
    if (FUSE && threadIdx.x == 0) {                          // last CTA out re-arms the counters for the next launch
        if (atomicAdd(fz.cnt + 1, 1) == (int)gridDim.x - 1) {
            fz.cnt[0] = 0;
            fz.cnt[1] = 0;
            __threadfence();
        }
    }
    if (threadIdx.x == G2_EPI_WARP0 * 32) G2_CTA_STAMP(5);
    if (warp == 1)
        asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512) : "memory");
}

static int64_t g2_tiles(int64_t M, int64_t N, int ntl) {
    const int64_t NT = N / TILE_CH, MB2 = (M + 2 * G2_BM - 1) / (2 * G2_BM);
    return MB2 * ((NT + 2 * ntl - 1) / (2 * ntl));
}

bool gemm2_plan(int64_t M, int64_t N, int64_t K, int num_sms, size_t smem_optin, Gemm2Plan *plan) {
    (void)K;
    if (M < 1 || (size_t)Gemm2Smem::total > smem_optin || num_sms < 2) return false;
    const int64_t pairs = num_sms / 2;
    // n-tiles per CTA: the persistent clusters each walk ceil(tiles / pairs) tiles of ntl n-tiles per CTA; a small
    // per-tile term stands for the output store and for the narrower MMA's lower rate
    int best = G2_MAXNTL;
    int64_t best_cost = -1;
    for (int ntl = G2_MAXNTL; ntl >= G2_MINNTL; ntl--) {
        const int64_t tiles = g2_tiles(M, N, ntl);
        const int64_t rounds = (tiles + pairs - 1) / pairs;
        const int64_t cost = rounds * (4 * ntl + 1);
        if (best_cost < 0 || cost < best_cost) {
            best_cost = cost;
            best = ntl;
        }
    }
    plan->ntl = best;
    const int64_t tiles = g2_tiles(M, N, best);
    plan->grid = dim3((unsigned)(2 * (tiles < pairs ? tiles : pairs)), 1, 1);
    plan->smem_bytes = Gemm2Smem::total;
    return true;
}

// ---- tensor maps: the TCQ4 tensor as [NT rows][KT*148 eight-byte words], boxes of whole tiles / of their metadata
typedef CUresult (*G2EncodeTiled)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static G2EncodeTiled g2_encoder() {
    static G2EncodeTiled fn = [] {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult q = cudaDriverEntryPointSymbolNotFound;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
            q != cudaDriverEntryPointSuccess)
            p = nullptr;
        return (G2EncodeTiled)p;
    }();
    return fn;
}
static bool g2_make_map(CUtensorMap *map, const void *w, int64_t NT, int64_t KT, uint32_t box_words, uint32_t box_rows) {
    G2EncodeTiled enc = g2_encoder();
    if (!enc) return false;
    const cuuint64_t gdim[2] = {(cuuint64_t)KT * (TILE_BYTES / 8), (cuuint64_t)NT};
    const cuuint64_t gstride[1] = {(cuuint64_t)KT * TILE_BYTES};
    const cuuint32_t box[2] = {box_words, box_rows};
    const cuuint32_t estr[2] = {1, 1};
    return enc(map, CU_TENSOR_MAP_DATA_TYPE_UINT64, 2, const_cast<void *>(w), gdim, gstride, box, estr,
               CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE,
               CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// the fp32 output [M][N] with boxes of 32 rows x 8 channels
static bool g2_make_y_map(CUtensorMap *map, float *y, int64_t M, int64_t N, int64_t ldc) {
    G2EncodeTiled enc = g2_encoder();
    if (!enc) return false;
    const cuuint64_t gdim[2] = {(cuuint64_t)N, (cuuint64_t)M};
    const cuuint64_t gstride[1] = {(cuuint64_t)ldc * sizeof(float)};
    const cuuint32_t box[2] = {8, 32};
    const cuuint32_t estr[2] = {1, 1};
    return enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, y, gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
               CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

template <int NTL>
static cudaError_t launch_gemm2_t(cudaLaunchConfig_t *cfg, const void *w_tiles, ActStage st, G2Fuse fz, YOut y, int64_t M,
                                  int64_t N, int64_t K, int32_t *dbg_isum) {
    CUtensorMap tm_tiles, tm_meta;
    G2YMaps tm_y;
    if (!g2_make_map(&tm_tiles, w_tiles, N / TILE_CH, K / TILE_K, TILE_BYTES / 8, NTL) ||
        !g2_make_map(&tm_meta, w_tiles, N / TILE_CH, K / TILE_K, 160 / 8, 2 * NTL))
        return cudaErrorInvalidValue;      // unaligned tensor base or no driver entry point
    if (y.n < 1 || y.n > MAX_PEERS) return cudaErrorInvalidValue;
    tm_y.n = y.n;
    tm_y.mc = y.mc;
    tm_y.mc_ldc = y.ldc;
    for (int pi = 0; pi < MAX_PEERS; pi++)
        if (!g2_make_y_map(&tm_y.m[pi], y.p[pi < y.n ? pi : 0], M, N, y.ldc)) return cudaErrorInvalidValue;
    const int MB2 = (int)((M + 2 * G2_BM - 1) / (2 * G2_BM));
    const int num_tiles = (int)g2_tiles(M, N, NTL);
    if (y.mc)
        return cudaLaunchKernelEx(cfg, tcq4_gemm2_kernel<NTL, false, false, 2>, tm_tiles, tm_meta, tm_y, (const uint8_t *)w_tiles, st, fz, (int)M, (int)N, (int)K, MB2, num_tiles, (int32_t *)nullptr);
    if (y.n > 1)
        return cudaLaunchKernelEx(cfg, tcq4_gemm2_kernel<NTL, false, false, 1>, tm_tiles, tm_meta, tm_y, (const uint8_t *)w_tiles, st, fz, (int)M, (int)N, (int)K, MB2, num_tiles, (int32_t *)nullptr);
#ifdef TCQ4_DIAG
    static const bool prof = getenv("TCQ4_G2_PROF") != nullptr;     // diagnostics build: timeline stamps of one cluster
    if (prof && !dbg_isum) {
        cudaFuncSetAttribute(tcq4_gemm2_kernel<NTL, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             Gemm2Smem::total);
        return cudaLaunchKernelEx(cfg, tcq4_gemm2_kernel<NTL, false, true>, tm_tiles, tm_meta, tm_y, (const uint8_t *)w_tiles, st, fz, (int)M, (int)N, (int)K, MB2, num_tiles, (int32_t *)nullptr);
    }
#endif
    if (fz.res != nullptr || fz.bias != nullptr) {      // tcq4_b200_mul_mat_add (single destination)
        cudaFuncSetAttribute(tcq4_gemm2_kernel<NTL, false, false, 0, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             Gemm2Smem::total);
        return cudaLaunchKernelEx(cfg, tcq4_gemm2_kernel<NTL, false, false, 0, false, true>, tm_tiles, tm_meta, tm_y,
                                  (const uint8_t *)w_tiles, st, fz, (int)M, (int)N, (int)K, MB2, num_tiles, (int32_t *)nullptr);
    }
    if (fz.x != nullptr) {      // TCQ4_G2_FUSE=1 (single destination only; checked by the caller)
        if (dbg_isum) {
            cudaFuncSetAttribute(tcq4_gemm2_kernel<NTL, true, false, 0, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 Gemm2Smem::total);
            return cudaLaunchKernelEx(cfg, tcq4_gemm2_kernel<NTL, true, false, 0, true>, tm_tiles, tm_meta, tm_y,
                                      (const uint8_t *)w_tiles, st, fz, (int)M, (int)N, (int)K, MB2, num_tiles, dbg_isum);
        }
        cudaFuncSetAttribute(tcq4_gemm2_kernel<NTL, false, false, 0, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             Gemm2Smem::total);
        return cudaLaunchKernelEx(cfg, tcq4_gemm2_kernel<NTL, false, false, 0, true>, tm_tiles, tm_meta, tm_y,
                                  (const uint8_t *)w_tiles, st, fz, (int)M, (int)N, (int)K, MB2, num_tiles, (int32_t *)nullptr);
    }
    if (dbg_isum)
        return cudaLaunchKernelEx(cfg, tcq4_gemm2_kernel<NTL, true>, tm_tiles, tm_meta, tm_y, (const uint8_t *)w_tiles, st, fz,
                                  (int)M, (int)N, (int)K, MB2, num_tiles, dbg_isum);
    return cudaLaunchKernelEx(cfg, tcq4_gemm2_kernel<NTL, false>, tm_tiles, tm_meta, tm_y, (const uint8_t *)w_tiles, st, fz,
                              (int)M, (int)N, (int)K, MB2, num_tiles, (int32_t *)nullptr);
}

cudaError_t launch_gemm2(const void *w_tiles, ActStage st, YOut y, int64_t M, int64_t N, int64_t K,
                         const Gemm2Plan &plan, int32_t *dbg_isum, cudaStream_t stream, const float *fuse_x,
                         const int32_t *fuse_perm, int *fuse_cnt, const float *res, const float *bias) {
    const G2Fuse fz{fuse_x, fuse_perm, fuse_cnt, res, bias};
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = plan.grid;
    cfg.blockDim = dim3(G2_THREADS);
    cfg.dynamicSmemBytes = plan.smem_bytes;
    cfg.stream = stream;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    static const bool pdl = getenv("TCQ4_PDL") != nullptr;      // opt-in, see tcq4_gemv.cu (no measurable gain here)
    cfg.numAttrs = pdl ? 2 : 1;
    static const int force_ntl = getenv("TCQ4_G2_NTL") ? atoi(getenv("TCQ4_G2_NTL")) : 0;     // experiments only
    int ntl = plan.ntl;
    if (force_ntl >= G2_MINNTL && force_ntl <= G2_MAXNTL) {
        ntl = force_ntl;
        int dev = 0, sms = 2;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        const int64_t tiles = g2_tiles(M, N, ntl), cap = sms / 2;
        cfg.gridDim.x = (unsigned)(2 * (tiles < cap ? tiles : cap));
    }
    switch (ntl) {
        case 2: return launch_gemm2_t<2>(&cfg, w_tiles, st, fz, y, M, N, K, dbg_isum);
        case 3: return launch_gemm2_t<3>(&cfg, w_tiles, st, fz, y, M, N, K, dbg_isum);
        case 4: return launch_gemm2_t<4>(&cfg, w_tiles, st, fz, y, M, N, K, dbg_isum);
        case 5: return launch_gemm2_t<5>(&cfg, w_tiles, st, fz, y, M, N, K, dbg_isum);
        case 6: return launch_gemm2_t<6>(&cfg, w_tiles, st, fz, y, M, N, K, dbg_isum);
        case 7: return launch_gemm2_t<7>(&cfg, w_tiles, st, fz, y, M, N, K, dbg_isum);
        case 8: return launch_gemm2_t<8>(&cfg, w_tiles, st, fz, y, M, N, K, dbg_isum);
        default: return cudaErrorInvalidValue;
    }
}

cudaError_t gemm2_read_prof(long long *host_out, size_t bytes) {
    // first the one-cluster timeline, then (if the caller's buffer is larger) the per-CTA stamps
    const size_t n1 = bytes < sizeof(g_g2_prof) ? bytes : sizeof(g_g2_prof);
    cudaError_t e = cudaMemcpyFromSymbol(host_out, g_g2_prof, n1);
    if (e != cudaSuccess || bytes <= sizeof(g_g2_prof)) return e;
    size_t n2 = bytes - sizeof(g_g2_prof);
    if (n2 > sizeof(g_g2_cta)) n2 = sizeof(g_g2_cta);
    return cudaMemcpyFromSymbol(reinterpret_cast<char *>(host_out) + sizeof(g_g2_prof), g_g2_cta, n2);
}

template <int NTL>
static cudaError_t gemm2_set_smem() {
    cudaError_t e = cudaFuncSetAttribute(tcq4_gemm2_kernel<NTL, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         Gemm2Smem::total);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(tcq4_gemm2_kernel<NTL, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, Gemm2Smem::total);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(tcq4_gemm2_kernel<NTL, false, false, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             Gemm2Smem::total);
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(tcq4_gemm2_kernel<NTL, false, false, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                Gemm2Smem::total);
}

cudaError_t gemm2_init_attributes(size_t smem_optin) {
    (void)smem_optin;
    cudaError_t e;
    if ((e = gemm2_set_smem<2>()) != cudaSuccess) return e;
    if ((e = gemm2_set_smem<3>()) != cudaSuccess) return e;
    if ((e = gemm2_set_smem<4>()) != cudaSuccess) return e;
    if ((e = gemm2_set_smem<5>()) != cudaSuccess) return e;
    if ((e = gemm2_set_smem<6>()) != cudaSuccess) return e;
    if ((e = gemm2_set_smem<7>()) != cudaSuccess) return e;
    return gemm2_set_smem<8>();
}

}  // namespace tcq4
