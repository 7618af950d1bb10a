// tcq4_gemm3.cu -- K2 (round 2): W4A4 prefill GEMM on CTA pairs with tcgen05.mma.cta_group::2 kind::f16.
//
// Why fp16 operands for an integer GEMM.  The per-group weight scale makes the B operand p = sc*q an 11-bit integer
// (|p| <= 1016).  The round-1 kernel (tcq4_gemm2.cu) split it into two s8 halves and ran kind::i8 on twice the columns;
// its epilogue then had to read two TMEM values per output, recombine them (IMAD) and convert (I2F) once per
// 256-block -- the kernel was bound by the issue slots of that epilogue and of the unpack, not by the tensor pipe
// (profiles/prof_gemm2_*_r01d).  fp16 holds every |p| <= 2048 and every activation code exactly, every product
// a*p is an integer below 2^13 and every partial sum of a 256-deep dot product an integer below 2^21, so the fp32
// accumulator of kind::f16 carries the SAME integer the int32 accumulator would (measured on the hardware:
// profiles/f16_exact_r02.txt, 0 mismatches incl. the all-maximal pattern).  kind::f16 runs at half the MAC rate of
// kind::i8 but needs half the columns: the same tensor-pipe time per output, with ONE TMEM value per output that is
// already a float.  Epilogue: 2 packed-fp32 instructions per 2 outputs instead of 7; unpack: one LOP3 + one HFMA2 per
// two weights (the magic-number trick below) instead of the byte-field multiplies.
//
// Data movement.  Operands travel narrow and are widened inside the SM.  fp16 activations from L2 would need 64 KB per
// CTA and k-tile = 50 B/clk/SM, above the measured L2->SM fabric peak (profiles/fabric_peak_r02.txt).  Instead:
//   * activations arrive as ONE byte per code, the high byte of fp16(q) (every integer in [-7, 7] has a zero low byte
//     in fp16), written by K1; a byte permute against zero widens two codes per instruction -- no arithmetic;
//   * weights arrive as the GGUF tiles themselves; a nibble masked into bits [4, 8) of a half and OR-ed with 0x5400 is
//     the fp16 number 64 + u (u = q + 8; the XOR that makes u from the two's-complement nibble rides in the same LOP3):
//         (64 + u) * sc - 72 * sc = (u - 8) * sc = q * sc   exactly
//     (one HFMA2 per two weights; 72*sc = 8*(9*sc) is exact in fp16).
// Both are done by the same 8 unpack warps, half a k-tile per hand-over (one generic->async proxy fence per half).
//
// Tiling: a cluster of two CTAs owns 256 tokens x (16*NTL) channels (NTL = 12: 192 channels, where the M = 256
// kind::f16 instruction reaches its full rate: <= 192 columns cost the same 98 cycles, profiles/umma_peak_r01.txt).
// Per 256-block 16 MMAs of 256 x 192 x 16 accumulate in 192 TMEM columns (two accumulator stages), the fp32 epilogue
// runs once per 256-block: acc += float(isum) * (a7[m]/127) * S[n], in the reference's operation order.
// Several weight matrices that share their activations (Q/K/V, gate/up) are ONE launch: the column blocks of all
// segments form one tile list (SEG).  Narrow matrices split K in two halves whose partial sums meet in y through a
// TMA reduce-add (y zeroed by the activation kernel; two addends commute, so the result is deterministic).
//
// Warp roles (20 warps): 0 producer (bulk / tiled TMA copies), 1 MMA issuer (leader CTA), 2-3 idle, 4-11 unpack
// (nibbles -> fp16 operands), 12-19 epilogue.  Rings: 4 raw half-k-tile stages, 4 fp16 quarter-k-tile stages,
// 2 TMEM accumulator stages; all run across tile boundaries (persistent, one cluster per SM pair).
//
// Replaces tcq4_rrs_gemm_imma_kernel (/root/reference/ggml/src/ggml-cuda/tcq4_k32.cu:1217-1412) and, for these M,
// the live scalar kernel tcq4_rrs_fused_gemm_smallM (tcq4_k32.cu:733-889).
#include <cuda.h>      // CUtensorMap (types only; the encoder is fetched through cudaGetDriverEntryPoint)
#include <cuda_fp16.h>

#include <cstdlib>
#include <cstring>
#include <mutex>
#include <unordered_map>

#include "tcq4_common.cuh"
#include "tcq4_kernels.h"

namespace tcq4 {

constexpr int G3_BM = 128;                               // tokens per CTA (256 per pair)
constexpr int G3_MAXNTL = 12;                            // n-tiles (8 channels) unpacked per CTA
constexpr int G3_AW_WARP0 = 4, G3_AW_WARPS = 8;          // activation warps: two per TMEM lane quarter (thread = token row)
constexpr int G3_UNPACK_WARP0 = G3_AW_WARP0 + G3_AW_WARPS, G3_UNPACK_WARPS = 12;     // weight warps: two (group, channel) units per thread and k-tile
constexpr int G3_EPI_WARP0 = G3_UNPACK_WARP0 + G3_UNPACK_WARPS, G3_EPI_WARPS = 8;
constexpr int G3_THREADS = (G3_EPI_WARP0 + G3_EPI_WARPS) * 32;          // 1024

// one raw stage = one k-tile of this CTA's inputs
constexpr int G3_RAW_A = ACT_NIB_UNIT_BYTES;             // 16384: activation nibbles [8 groups][128 rows][16 B]
constexpr int G3_RAW_W = G3_MAXNTL * 1024;               // 12288: the nibbles of this CTA's NTL tiles
constexpr int G3_RAW_META = 2 * G3_MAXNTL * 160;         // 3840: S, Z, sc, zc of the pair's 2*NTL tiles
constexpr int G3_RAW_A7 = G3_BM * 4;                     // 512: a7 of this CTA's 128 rows for the k-tile
constexpr int G3_RAW_BYTES = G3_RAW_A + G3_RAW_W + G3_RAW_META + G3_RAW_A7;      // 33024
// bytes in flight = bandwidth x latency: ~1.3 us of L2 latency under load at 33 KB per k-tile (0.8 us) needs two stages in
// flight besides the one being read (two stages of a fatter unit WAS the kernel's period: what-if runs, tests/gpu_g3_exp.py)
constexpr int G3_RAW_STAGES = 3;
constexpr int G3_F_Q = 8 * 8 * G3_MAXNTL * 16;           // 12288: quarter k-tile (64 K) of this CTA's 8*NTL channels in fp16
constexpr int G3_F_BYTES = 4 * G3_F_Q;                   // 49152: one k-tile of the fp16 B operand
constexpr int G3_F_STAGES = 2;                           // the weight warps run up to two k-tiles ahead of the tensor pipe
constexpr int G3_SN_RING = 8;
constexpr int G3_SN_STRIDE = 16 * G3_MAXNTL;             // floats per ring slot
constexpr int G3_OUT_BOXES = 2;
// TMEM columns: two fp32 accumulator stages of 16*NTL columns, then four quarter-k-tile buffers of fp16 activations
// (64 K = 32 columns of packed pairs each)
constexpr int G3_TM_D_STRIDE = 16 * G3_MAXNTL;           // 192
constexpr int G3_TM_A0 = 2 * G3_TM_D_STRIDE;             // 384
constexpr int G3_TM_A_COLS = 32;
static_assert(G3_TM_A0 + 4 * G3_TM_A_COLS <= 512, "TMEM budget");

struct Gemm3Smem {
    static constexpr int off_f = 0;
    static constexpr int off_raw = off_f + G3_F_STAGES * G3_F_BYTES;
    static constexpr int off_sn = off_raw + G3_RAW_STAGES * G3_RAW_BYTES;
    static constexpr int off_rm = off_sn + G3_SN_RING * G3_SN_STRIDE * 4;              // [ring][128 rows] a7 / 127
    static constexpr int off_out = off_rm + G3_SN_RING * G3_BM * 4;
    static constexpr int off_zflag = off_out + G3_EPI_WARPS * G3_OUT_BOXES * 1024;
    static constexpr int off_bar = off_zflag + G3_SN_RING * 8 * 4;
    static constexpr int n_bar = 2 * G3_RAW_STAGES + 2 * G3_F_STAGES + 2 + 2 + 8 + 8 + G3_SN_RING;
    static constexpr int off_tmem = off_bar + n_bar * 8;
    static constexpr int total = off_tmem + 16;
};
static_assert(G3_RAW_BYTES % 128 == 0 && G3_F_BYTES % 128 == 0 && G3_F_Q % 128 == 0 && Gemm3Smem::off_raw % 128 == 0 && Gemm3Smem::off_sn % 16 == 0 &&
                  Gemm3Smem::off_out % 128 == 0 && (G3_RAW_A + G3_RAW_W) % 128 == 0,
              "operands, copy destinations and store boxes must be 128-byte aligned");
static_assert(Gemm3Smem::off_bar % 8 == 0, "barriers must be 8-byte aligned");
static_assert(Gemm3Smem::total <= 227 * 1024, "shared memory budget");

// ---- cluster-scope helpers (same protocol as tcq4_gemm2.cu) ----------------------------------------------------
__device__ __forceinline__ uint32_t g3_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ uint32_t g3_mapa(uint32_t local_addr, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local_addr), "r"(rank));
    return r;
}
// default (CTA-scope release) semantics on purpose: what the signal covers is read by the tensor core's async proxy
// (ordered by fence.proxy.async before the arrive) or is TMEM (tcgen05.fence); see tcq4_gemm2.cu
__device__ __forceinline__ void g3_arrive_cluster(uint32_t cluster_addr) {
    asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
__device__ __forceinline__ void g3_wait(uint64_t *bar, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "G3_WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1, %2;\n\t"
        "@!p bra G3_WAIT_%=;\n\t}" ::"r"(smem_u32(bar)),
        "r"(parity), "r"(20000u)
        : "memory");
}
// Busy poll (no suspend): for the few threads whose wake-up latency is on the critical path of the tensor pipe
__device__ __forceinline__ void g3_spin(uint64_t *bar, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "G3_SPIN_%=:\n\t"
        "mbarrier.test_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@!p bra G3_SPIN_%=;\n\t}" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
__device__ __forceinline__ void g3_cluster_sync() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// D[tmem, both CTAs] (+)= A[smem, both CTAs] * B[smem, both CTAs], fp16 x fp16 -> fp32; one thread of the leader CTA
__device__ __forceinline__ void umma_f16_2cta_parts(uint32_t d_tmem, uint32_t a_lo, uint32_t a_hi, uint32_t b_lo,
                                                    uint32_t b_hi, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\t"
        "mov.b64 da, {%1, %2};\n\t"
        "mov.b64 db, {%3, %4};\n\t"
        "setp.ne.b32 p, %6, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], da, db, %5, p;\n\t}" ::"r"(d_tmem),
        "r"(a_lo), "r"(a_hi), "r"(b_lo), "r"(b_hi), "r"(idesc), "r"(accumulate)
        : "memory");
}
// same with the A operand in TENSOR MEMORY (written there by tcgen05.st): 128 lanes = this CTA's token rows, 8 columns of
// packed fp16 pairs per K = 16
__device__ __forceinline__ void umma_f16_2cta_ts(uint32_t d_tmem, uint32_t a_tmem, uint32_t b_lo, uint32_t b_hi,
                                                 uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t.reg .b64 db;\n\t"
        "mov.b64 db, {%2, %3};\n\t"
        "setp.ne.b32 p, %5, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], [%1], db, %4, p;\n\t}" ::"r"(d_tmem),
        "r"(a_tmem), "r"(b_lo), "r"(b_hi), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void g3_tmem_st_32x16(uint32_t taddr, const uint32_t (&r)[16]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};" ::
            "r"(taddr),
        "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]), "r"(r[10]),
        "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
        : "memory");
}
__device__ __forceinline__ void g3_tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void g3_commit_2cta(uint64_t *bar) {
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::
                     "r"(smem_u32(bar)), "h"((uint16_t)3)
                 : "memory");
}
__device__ __forceinline__ uint64_t g3_pack(float lo, float hi) {
    uint64_t r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void g3_unpack(uint64_t v, float &lo, float &hi) {
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ uint64_t g3_mul2(uint64_t a, uint64_t b) {
    uint64_t r;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ uint64_t g3_fma2(uint64_t a, uint64_t b, uint64_t c) {
    uint64_t r;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
    return r;
}
__device__ __forceinline__ uint32_t g3_hfma2(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t r;
    asm("fma.rn.f16x2 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
    return r;
}

// ---- narrow operands -> fp16 (exact) ------------------------------------------------------------------------------
// Weights: a 32-bit word holds 8 nibbles, element i at bits [4i, 4i+4).  Masking two of them into bits [4, 8) of the two
// halves of a register and OR-ing 0x5400 gives the fp16 pair (64 + u_a, 64 + u_b) (the mantissa's unit at exponent 6 is
// 1/16, so mantissa bit 4 is worth 1).  One HFMA2 removes the offset and applies the scale, rounding once on an exactly
// representable result.  Order of the 8 values inside the 16-byte operand cell: elements (0,4,1,5,2,6,3,7); the
// activation kernel writes its bytes in the same order, which is all the MMA needs.
constexpr uint32_t G3_MAGIC = 0x54005400u;      // (64.0, 64.0)
// activation nibbles (u = q + 8 stored by K1, element i at bits [4i, 4i+4)) -> 8 fp16 values q in the cell order
__device__ __forceinline__ uint4 g3_cvt_act(uint32_t w) {
    constexpr uint32_t K16 = 0x4C004C00u, KM1032 = 0xE408E408u, K1 = 0x3C003C00u, KM72 = 0xD480D480u;
    const uint32_t hi = w >> 8;
    const uint32_t t0 = (w & 0x000F000Fu) | G3_MAGIC;       // 64 + u/16 : elements 0, 4
    const uint32_t t1 = (w & 0x00F000F0u) | G3_MAGIC;       // 64 + u    : elements 1, 5
    const uint32_t t2 = (hi & 0x000F000Fu) | G3_MAGIC;      //             elements 2, 6
    const uint32_t t3 = (hi & 0x00F000F0u) | G3_MAGIC;      //             elements 3, 7
    return make_uint4(g3_hfma2(t0, K16, KM1032), g3_hfma2(t1, K1, KM72), g3_hfma2(t2, K16, KM1032), g3_hfma2(t3, K1, KM72));
}
// 16 activation bytes (two operand cells) -> 16 fp16: byte b becomes the half (b << 8)
__device__ __forceinline__ void g3_cvt_act16(const uint4 w, uint4 &c0, uint4 &c1) {
    c0 = make_uint4(__byte_perm(w.x, 0u, 0x1404), __byte_perm(w.x, 0u, 0x3424), __byte_perm(w.y, 0u, 0x1404), __byte_perm(w.y, 0u, 0x3424));
    c1 = make_uint4(__byte_perm(w.z, 0u, 0x1404), __byte_perm(w.z, 0u, 0x3424), __byte_perm(w.w, 0u, 0x1404), __byte_perm(w.w, 0u, 0x3424));
}
// ((a ^ c) & m) | (c & ~m) in one LOP3: mask two nibbles into bits [4, 8) of the halves, flip their top bit (two's
// complement -> biased) and OR the magic exponent
__device__ __forceinline__ uint32_t g3_lop(uint32_t a) {
    uint32_t r;
    asm("lop3.b32 %0, %1, %2, %3, 0x6A;" : "=r"(r) : "r"(a), "r"(0x00F000F0u), "r"(G3_MAGIC | 0x00800080u));
    return r;
}
__device__ __forceinline__ uint4 g3_cvt_w(uint32_t w, uint32_t sc2, uint32_t c72) {      // two's-complement q, p = sc * q
    const uint32_t t0 = g3_lop(w << 4);           // elements 0, 4
    const uint32_t t1 = g3_lop(w);                // elements 1, 5
    const uint32_t t2 = g3_lop(w >> 4);           // elements 2, 6
    const uint32_t t3 = g3_lop(w >> 8);           // elements 3, 7
    return make_uint4(g3_hfma2(t0, sc2, c72), g3_hfma2(t1, sc2, c72), g3_hfma2(t2, sc2, c72), g3_hfma2(t3, sc2, c72));
}

// Zero-point term of one output element for one k-tile (tcq4_k32.cu:850,872-873); cold (Z = 0 from the quantiser)
__device__ __noinline__ float g3_zero_point_term(const uint8_t *__restrict__ w, const int16_t *__restrict__ sumq, int m,
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

__device__ __forceinline__ void g3_tma_2d(void *dst_smem, const CUtensorMap *map, int c0, int c1, uint64_t *bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
            smem_u32(dst_smem)),
        "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
        : "memory");
}
__device__ __forceinline__ void g3_tma_2d_hint(void *dst_smem, const CUtensorMap *map, int c0, int c1, uint64_t *bar,
                                               uint64_t policy) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1, {%3, %4}], "
        "[%2], %5;" ::"r"(smem_u32(dst_smem)),
        "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "l"(policy)
        : "memory");
}
__device__ __forceinline__ void g3_tma_store_2d(const CUtensorMap *map, const void *src_smem, int c0, int c1) {
    asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];" ::"l"(map),
                 "r"(smem_u32(src_smem)), "r"(c0), "r"(c1)
                 : "memory");
}
// y += box (fp32), performed by the copy engine at L2: the K-split halves of a tile meet here
__device__ __forceinline__ void g3_tma_reduce_add_2d(const CUtensorMap *map, const void *src_smem, int c0, int c1) {
    asm volatile("cp.reduce.async.bulk.tensor.2d.global.shared::cta.add.tile.bulk_group [%0, {%2, %3}], [%1];" ::"l"(map),
                 "r"(smem_u32(src_smem)), "r"(c0), "r"(c1)
                 : "memory");
}
__device__ __forceinline__ void g3_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N_PENDING>
__device__ __forceinline__ void g3_store_wait_read() {
    asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N_PENDING) : "memory");
}

// ---- kernel parameters -------------------------------------------------------------------------------------------
constexpr int G3_MAX_SEG = 4;
struct G3Seg {
    CUtensorMap tm_tiles;      // [NT rows][KT*148 8-byte words], box {128 words = the 1024 bytes of nibbles, NTL rows}
    CUtensorMap tm_meta;       // same tensor, box {20 words = S, Z, sc, zc, 2*NTL rows}
    CUtensorMap tm_y;          // fp32 [M][ldc], box {8 channels, 32 rows}
};
struct G3Params {
    G3Seg seg[G3_MAX_SEG];
    const uint8_t *w[G3_MAX_SEG];
    int N[G3_MAX_SEG];
    int cb0[G3_MAX_SEG + 1];   // first column block of each segment in the launch's tile list
    int nseg;
    int ksplit;                // 1 or 2
    const float *res;          // ADD: [M][N] contiguous or null (single segment only); may alias y
    const float *bias;         // ADD: [N] or null
    int res_in_y;              // ADD with ksplit > 1 and res == y: y already holds the residual, nothing to pre-add
    int exp;                   // diagnostics (PROF instantiation only): bit 0 skip epilogue math, 1 skip activation conversion, 2 skip weight conversion
};

struct G3Tile {
    int sg, mb, cbl, kt0, kt1, ks;
};
__device__ __forceinline__ G3Tile g3_decode(const G3Params &P, int t, int MB2, int KT, uint32_t rank) {
    G3Tile r;
    r.ks = t % P.ksplit;
    const int t2 = t / P.ksplit;
    r.mb = 2 * (t2 % MB2) + (int)rank;
    const int cb = t2 / MB2;
    int sg = 0;
#pragma unroll
    for (int i = 1; i < G3_MAX_SEG; i++)
        if (i < P.nseg && cb >= P.cb0[i]) sg = i;
    r.sg = sg;
    r.cbl = cb - P.cb0[sg];
    r.kt0 = (int)(((long long)r.ks * KT) / P.ksplit);
    r.kt1 = (int)(((long long)(r.ks + 1) * KT) / P.ksplit);
    return r;
}

// Timeline stamps (diagnostics build, TCQ4_G3_PROF=1): SM clocks of one cluster's leader CTA, [role 4][k-tile 64][event 8]
constexpr int G3_PROF_KT = 64;
#ifdef TCQ4_DIAG
__device__ long long g_g3_prof[4 * G3_PROF_KT * 8];
#define G3_STAMP(role, kt, ev)                                                                          \
    do {                                                                                                \
        if (PROF && prof_on && (kt) < G3_PROF_KT) g_g3_prof[((role) * G3_PROF_KT + (kt)) * 8 + (ev)] = clock64(); \
    } while (0)
#else
#define G3_STAMP(role, kt, ev) do { } while (0)
#endif

template <int NTL, bool DEBUG, bool ADD, bool PROF = false>
__global__ void __launch_bounds__(G3_THREADS, 1)
tcq4_gemm3_kernel(const __grid_constant__ G3Params P, ActStage st, int M, int K, int MB2, int num_tiles,
                  int32_t *__restrict__ dbg_isum) {
    static_assert(NTL % 4 == 0 && NTL >= 4 && NTL <= G3_MAXNTL, "NTL: a multiple of 4, at most 12");
    extern __shared__ __align__(128) uint8_t smem[];
    constexpr int NCOL = 16 * NTL;                     // channels per pair-tile = MMA N = TMEM columns per stage
    constexpr int LBO_B = 8 * NTL * 16;                // K-chunk stride of the B operand
    constexpr int SN_WARPS = (NCOL + 31) / 32;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int KT = K / TILE_K;
    const uint32_t rank = g3_ctarank();
    const int cluster_id = blockIdx.x >> 1, num_clusters = gridDim.x >> 1;
    const bool prof_on = PROF && cluster_id == num_clusters / 2 && rank == 0 && lane == 0;
    (void)prof_on;

    uint8_t *sF = smem + Gemm3Smem::off_f;
    uint8_t *sRaw = smem + Gemm3Smem::off_raw;
    float *sSn = reinterpret_cast<float *>(smem + Gemm3Smem::off_sn);
    float *sRm = reinterpret_cast<float *>(smem + Gemm3Smem::off_rm);
    int *sZflag = reinterpret_cast<int *>(smem + Gemm3Smem::off_zflag);
    uint8_t *sOut = smem + Gemm3Smem::off_out;
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem + Gemm3Smem::off_bar);
    uint64_t *rawfull = bars, *rawempty = rawfull + G3_RAW_STAGES, *bfull = rawempty + G3_RAW_STAGES,
             *bfree = bfull + G3_F_STAGES, *accfull = bfree + G3_F_STAGES, *accempty = accfull + 2, *afull = accempty + 2,
             *afree = afull + 8, *snfull = afree + 8;      // afull / afree: [issuing warp = k-tile & 1][quarter]
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(smem + Gemm3Smem::off_tmem);

    if (threadIdx.x == 0) {
        for (int i = 0; i < G3_RAW_STAGES; i++) {
            mbar_init(&rawfull[i], 1);
            mbar_init(&rawempty[i], G3_UNPACK_WARPS + G3_AW_WARPS);      // every unpack warp releases the k-tile's raw stage once
        }
        for (int i = 0; i < G3_F_STAGES; i++) {
            mbar_init(&bfull[i], 2 * G3_UNPACK_WARPS);           // weight warps of both CTAs (used in the leader)
            mbar_init(&bfree[i], 1);                             // leader's commit (multicast)
        }
        for (int i = 0; i < 2; i++) {
            mbar_init(&accfull[i], 1);                           // leader's commit (multicast)
            mbar_init(&accempty[i], 2 * G3_EPI_WARPS);           // epilogue warps of both CTAs (used in the leader)
        }
        for (int i = 0; i < 8; i++) {
            mbar_init(&afull[i], G3_AW_WARPS);                   // the four activation warps that own this quarter, both CTAs (leader)
            mbar_init(&afree[i], 1);                             // leader's commit (multicast)
        }
        for (int i = 0; i < G3_SN_RING; i++) mbar_init(&snfull[i], SN_WARPS + 4);      // S by weight warps, a7/127 by four activation warps
        fence_mbar_init();
    }
    if (warp == 0 && lane == 0) {
        for (int i = 0; i < P.nseg; i++) {
            asm volatile("prefetch.tensormap [%0];" ::"l"(&P.seg[i].tm_tiles) : "memory");
            asm volatile("prefetch.tensormap [%0];" ::"l"(&P.seg[i].tm_meta) : "memory");
            asm volatile("prefetch.tensormap [%0];" ::"l"(&P.seg[i].tm_y) : "memory");
        }
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(512)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    g3_cluster_sync();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;
    pdl_launch_dependents();     // the next kernel of the stream may set itself up behind this one
    // griddepcontrol.wait is executed per role below, by the threads that touch what earlier kernels produced: the
    // producer before its first activation copy, the epilogue warps before they read a7 / res and before they store.
    // Weight tiles are constants of the model: their copies, the barrier set-up and the TMEM allocation above run
    // while the activation kernel in front of this launch is still finishing.

    if (warp < G3_AW_WARP0) {
      asm volatile("setmaxnreg.dec.sync.aligned.u32 48;" ::: "memory");
      if (warp == 0) {
        // ===================== producer: four copies per k-tile =====================
        if (lane == 0) {
            constexpr uint32_t total_tx = (uint32_t)(G3_RAW_A + NTL * 1024 + 2 * NTL * 160 + G3_RAW_A7);
            const uint64_t pol_w = l2_policy_evict_first();
            // pass 0: the weight copies of the first G3_RAW_STAGES k-tiles (nothing to wait for); then griddepcontrol.wait;
            // pass 1: everything else in ring order
#pragma unroll 1
            for (int pass = 0; pass < 2; pass++) {
                int s = 0, u = 0, i = 0;
                bool done = false;
                for (int t = cluster_id; t < num_tiles && !done; t += num_clusters) {
                    const G3Tile tl = g3_decode(P, t, MB2, KT, rank);
                    const int nt_pair = tl.cbl * (2 * NTL), nt0 = nt_pair + (int)rank * NTL;
                    const uint8_t *a_src = reinterpret_cast<const uint8_t *>(st.codes) + (size_t)tl.mb * KT * ACT_NIB_UNIT_BYTES;
                    const float *a7_src = st.a7t + (size_t)tl.mb * KT * G3_BM;
                    const CUtensorMap *mt = &P.seg[tl.sg].tm_tiles, *mm = &P.seg[tl.sg].tm_meta;
#pragma unroll 1
                    for (int kt = tl.kt0; kt < tl.kt1; kt++, i++) {
                        if (pass == 0 && i >= G3_RAW_STAGES) {
                            done = true;
                            break;
                        }
                        uint8_t *stg = sRaw + s * G3_RAW_BYTES;
                        if (pass == 0 || i >= G3_RAW_STAGES) {
                            if (u > 0) g3_wait(&rawempty[s], (uint32_t)((u - 1) & 1));
                            mbar_arrive_expect_tx(&rawfull[s], total_tx);
                            g3_tma_2d_hint(stg + G3_RAW_A, mt, kt * (TILE_BYTES / 8), nt0, &rawfull[s], pol_w);
                            g3_tma_2d(stg + G3_RAW_A + G3_RAW_W, mm, kt * (TILE_BYTES / 8) + TILE_NIB / 8, nt_pair, &rawfull[s]);
                        }
                        if (pass == 1) {
                            bulk_g2s(stg, a_src + (size_t)kt * ACT_NIB_UNIT_BYTES, G3_RAW_A, &rawfull[s]);
                            bulk_g2s(stg + G3_RAW_A + G3_RAW_W + G3_RAW_META, a7_src + (size_t)kt * G3_BM, G3_RAW_A7, &rawfull[s]);
                        }
                        if (++s == G3_RAW_STAGES) { s = 0; u++; }
                    }
                }
                if (pass == 0) pdl_wait();       // the activation codes belong to the kernel in front until here
            }
        }
      } else if (warp <= 2) {
        // ===================== MMA issuers (leader CTA only): warp 1 takes the even k-tiles, warp 2 the odd ones =====================
        // One issuing thread spends ~600 cycles per k-tile outside the MMA instructions themselves (three commits, loop,
        // the accumulator / weight / first-quarter waits: ~50 dependent instructions at ~10 cycles each in a loaded SM) --
        // longer than the 391 cycles of tensor-pipe work it has queued, so the pipe drained at every k-tile boundary
        // (timeline of the what-if run with all CUDA-core work off: tests/gpu_g3_exp.py).  Two issuers alternate: each owns
        // one accumulator stage, one slot of the weight ring and its own set of activation barriers, and has two k-tiles
        // of pipe time per k-tile of its own.  MMAs of different k-tiles write different accumulators, so their
        // interleaving in the pipe's queue is free; tcgen05.commit tracks the issuing thread's own MMAs.
        if (rank == 0) {
            constexpr uint32_t idesc = (1u << 4) | ((uint32_t)(NCOL >> 3) << 17) | ((uint32_t)((2 * G3_BM) >> 4) << 24);
            constexpr uint32_t b_hi = (uint32_t)(128 >> 4) | (1u << 14);              // SBO = 128 B, descriptor version 1
            const int as = warp - 1;                                                  // this issuer's accumulator stage / ring slot
            const uint32_t b_lo0 = ((smem_u32(sF) + as * G3_F_BYTES) >> 4) | ((uint32_t)(LBO_B >> 4) << 16);
            const uint32_t d = tmem + as * G3_TM_D_STRIDE;
            uint64_t *my_afull = afull + as * 4, *my_afree = afree + as * 4;
            int total_k = 0;
            for (int t = cluster_id; t < num_tiles; t += num_clusters) {
                const G3Tile tl = g3_decode(P, t, MB2, KT, rank);
                total_k += tl.kt1 - tl.kt0;
            }
            for (int gk = as, u2 = 0; gk < total_k; gk += 2, u2++) {
                if (as == 0) G3_STAMP(0, gk, 0);
                if (u2 > 0) g3_wait(&accempty[as], (uint32_t)((u2 - 1) & 1));
                if (as == 0) G3_STAMP(0, gk, 1);
                g3_wait(&bfull[as], (uint32_t)(u2 & 1));       // the whole k-tile of fp16 weights (ring of two k-tiles)
                if (as == 0) G3_STAMP(0, gk, 2);
#pragma unroll 1      // rolled on purpose: the unrolled form kept four sets of descriptors alive and spilled them into the issue path
                for (int q = 0; q < 4; q++) {
                    g3_wait(&my_afull[q], (uint32_t)(u2 & 1));      // both CTAs' activations of this quarter are in TMEM
                    if (as == 0) G3_STAMP(0, gk, 3 + q);
                    tc_fence_after();
                    if (elect_one()) {
                        const uint32_t b_lo = b_lo0 + (uint32_t)q * (G3_F_Q >> 4);
                        const uint32_t a_t = tmem + G3_TM_A0 + q * G3_TM_A_COLS;
#pragma unroll
                        for (int ks = 0; ks < 4; ks++)      // K = 16 per instruction: 8 TMEM columns of A, two 8-element K-chunks of B
                            umma_f16_2cta_ts(d, a_t + ks * 8, b_lo + (uint32_t)ks * ((2 * LBO_B) >> 4), b_hi, idesc, (q | ks) ? 1u : 0u);
                        g3_commit_2cta(&my_afree[q]);
                        if (q == 3) {
                            g3_commit_2cta(&bfree[as]);
                            g3_commit_2cta(&accfull[as]);
                        }
                    }
                    __syncwarp();
                }
            }
        }
      }
    } else if (warp < G3_UNPACK_WARP0) {
        asm volatile("setmaxnreg.dec.sync.aligned.u32 32;" ::: "memory");
        // ===================== activation warps: nibbles -> fp16 -> TMEM (the MMA's A operand) =====================
        // thread = token row = TMEM lane.  Going through registers and tcgen05.st instead of shared memory keeps 128 KB per
        // k-tile (the fp16 stores and the tensor core's operand reads) off the shared-memory pipe.  Two warps per lane
        // quarter: warp j of the pair owns quarters j and j + 2 of every k-tile (32 bytes of nibbles of the row -> 32 columns each);
        // the MMA frees a quarter buffer as soon as it has consumed it, so a warp has 3/4 of a k-tile to refill it.
        const int aw = warp - G3_AW_WARP0;                  // 0..7
        const int aj = aw >> 2;                             // which quarters: aj, aj + 2
        const int row = (warp & 3) * 32 + lane;
        const uint32_t afull_leader = g3_mapa(smem_u32(afull), 0);
        const uint32_t ta = tmem + ((uint32_t)((warp & 3) * 32) << 16) + G3_TM_A0;
        int s = 0, u = 0, gk = 0;
        for (int t = cluster_id; t < num_tiles; t += num_clusters) {
            const G3Tile tl = g3_decode(P, t, MB2, KT, rank);
            for (int kt = tl.kt0; kt < tl.kt1; kt++, gk++) {
                if (aw == 0) G3_STAMP(1, gk, 0);
                g3_wait(&rawfull[s], (uint32_t)(u & 1));
                if (aw == 0) G3_STAMP(1, gk, 1);
                const uint8_t *stg = sRaw + s * G3_RAW_BYTES;
                if (aj == 0) {      // the row factor a7 / 127 of this k-tile, for the epilogue warps
                    const float a7v = *reinterpret_cast<const float *>(stg + G3_RAW_A + G3_RAW_W + G3_RAW_META + row * 4);
                    sRm[(gk & (G3_SN_RING - 1)) * G3_BM + row] = a7v * (1.0f / 127.0f);
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&snfull[gk & (G3_SN_RING - 1)]);
                }
#pragma unroll
                for (int i = 0; i < 2; i++) {
                    const int q = aj + 2 * i;
                    const uint8_t *src = stg + (2 * q) * 2048 + row * 16;      // the quarter's two 32-groups of this row
                    const uint4 w0 = *reinterpret_cast<const uint4 *>(src), w1 = *reinterpret_cast<const uint4 *>(src + 2048);
                    if (gk > 0) {
                        g3_wait(&afree[((gk - 1) & 1) * 4 + q], (uint32_t)(((gk - 1) >> 1) & 1));      // MMA(q) of the previous k-tile: the other issuer's
                        tc_fence_after();
                    }
                    if (aw == 0) G3_STAMP(1, gk, 2 + 3 * i);
                    if (!(PROF && (P.exp & 2)))
#pragma unroll
                    for (int c2 = 0; c2 < 2; c2++) {      // one 32-group = 4 words -> 4 operand cells = 16 TMEM columns per step
                        const uint4 w = c2 ? w1 : w0;
                        const uint4 c0 = g3_cvt_act(w.x), c1 = g3_cvt_act(w.y), c2v = g3_cvt_act(w.z), c3 = g3_cvt_act(w.w);
                        const uint32_t r[16] = {c0.x, c0.y, c0.z, c0.w, c1.x, c1.y, c1.z, c1.w,
                                                c2v.x, c2v.y, c2v.z, c2v.w, c3.x, c3.y, c3.z, c3.w};
                        g3_tmem_st_32x16(ta + q * G3_TM_A_COLS + c2 * 16, r);
                    }
                    g3_tmem_st_wait();
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) {
                        g3_arrive_cluster(afull_leader + (uint32_t)((gk & 1) * 4 + q) * 8);
                        if (i == 1) mbar_arrive(&rawempty[s]);       // every lane has read its raw bytes
                    }
                    if (aw == 0) G3_STAMP(1, gk, 3 + 3 * i);
                }
                if (++s == G3_RAW_STAGES) { s = 0; u++; }
            }
        }
    } else if (warp < G3_EPI_WARP0) {
        asm volatile("setmaxnreg.dec.sync.aligned.u32 32;" ::: "memory");
        // ===================== weight warps: nibbles -> fp16 B operand in shared memory =====================
        // One hand-over per k-tile: a thread converts two (group, channel) units of 4 words each -- the same (group of the
        // quarter, channel) in quarters b_q and b_q + 2 -- into the ring slot of this k-tile (the ring holds two k-tiles, so
        // these warps have a full k-tile of tensor-pipe time for it).
        static_assert(8 * 8 * NTL == 2 * G3_UNPACK_WARPS * 32, "two weight units per weight-warp thread and k-tile");
        const int uw = warp - G3_UNPACK_WARP0;              // 0..11
        const int ut = threadIdx.x - G3_UNPACK_WARP0 * 32;  // 0..383
        const int b_q = ut / NCOL, b_r = ut - b_q * NCOL, b_gg = b_r / (8 * NTL), b_ch = b_r - b_gg * (8 * NTL);
        const int src_off = G3_RAW_A + (b_ch >> 3) * 1024 + (2 * b_q + b_gg) * 128 + (b_ch & 7) * 16;
        const int sc_off = G3_RAW_A + G3_RAW_W + ((int)rank * NTL + (b_ch >> 3)) * 160 + 32 + (b_ch & 7) * 8 + 2 * b_q + b_gg;
        const int dst_off = b_q * G3_F_Q + (b_gg * 4) * LBO_B + b_ch * 16;
        const uint32_t bfull_leader = g3_mapa(smem_u32(bfull), 0);
        // S / zero-point publisher role of the first NCOL threads: channel (half, j, c) of the pair's 16*NTL channels
        const int p_half = ut / (8 * NTL), p_j = (ut % (8 * NTL)) >> 3, p_c = ut & 7;
        const int p_off = G3_RAW_A + G3_RAW_W + (p_half * NTL + p_j) * 160;
        int s = 0, u = 0, gk = 0;
        for (int t = cluster_id; t < num_tiles; t += num_clusters) {
            const G3Tile tl = g3_decode(P, t, MB2, KT, rank);
            for (int kt = tl.kt0; kt < tl.kt1; kt++, gk++) {
                if (uw == 0) G3_STAMP(2, gk, 0);
                g3_wait(&rawfull[s], (uint32_t)(u & 1));
                if (uw == 0) G3_STAMP(2, gk, 1);
                const uint8_t *stg = sRaw + s * G3_RAW_BYTES;
                const uint4 wb0 = *reinterpret_cast<const uint4 *>(stg + src_off);
                const uint4 wb1 = *reinterpret_cast<const uint4 *>(stg + src_off + 512);       // the same unit two quarters (4 groups) later
                const int sc0 = *reinterpret_cast<const int8_t *>(stg + sc_off);
                const int sc1 = *reinterpret_cast<const int8_t *>(stg + sc_off + 4);
                if (uw < SN_WARPS) {      // per-channel super-scales of this k-tile for the epilogue warps
                    const uint8_t *md = stg + p_off;
                    const float sv = __half2float(*reinterpret_cast<const __half *>(md + 2 * p_c));
                    const float zv = __half2float(*reinterpret_cast<const __half *>(md + 16 + 2 * p_c));
                    const uint2 zc = *reinterpret_cast<const uint2 *>(md + 96 + p_c * 8);
                    const int zf = (zv != 0.0f) && ((zc.x | zc.y) != 0u);
                    sSn[(gk & (G3_SN_RING - 1)) * G3_SN_STRIDE + ut] = sv;
                    const int anyz = __any_sync(0xffffffffu, zf);
                    __syncwarp();
                    if (lane == 0) {
                        sZflag[(gk & (G3_SN_RING - 1)) * 8 + uw] = anyz;
                        mbar_arrive(&snfull[gk & (G3_SN_RING - 1)]);
                    }
                }
                const int fs = gk & 1;
                if (gk > 1) g3_wait(&bfree[fs], (uint32_t)(((gk >> 1) - 1) & 1));
                if (uw == 0) G3_STAMP(2, gk, 2);
                uint8_t *dst = sF + fs * G3_F_BYTES + dst_off;
                if (!(PROF && (P.exp & 4)))
#pragma unroll
                for (int i = 0; i < 2; i++) {
                    const uint4 wb = i ? wb1 : wb0;
                    const __half sch = __int2half_rn(i ? sc1 : sc0);
                    const __half2 sc2h = __halves2half2(sch, sch);
                    const __half2 c72h = __hmul2(sc2h, __float2half2_rn(-72.0f));      // exact: 72*sc = 8 * (9*sc), 9*sc < 2048
                    const uint32_t sc2 = *reinterpret_cast<const uint32_t *>(&sc2h), c72 = *reinterpret_cast<const uint32_t *>(&c72h);
                    uint8_t *d2 = dst + i * (2 * G3_F_Q);
                    *reinterpret_cast<uint4 *>(d2) = g3_cvt_w(wb.x, sc2, c72);
                    *reinterpret_cast<uint4 *>(d2 + 1 * LBO_B) = g3_cvt_w(wb.y, sc2, c72);
                    *reinterpret_cast<uint4 *>(d2 + 2 * LBO_B) = g3_cvt_w(wb.z, sc2, c72);
                    *reinterpret_cast<uint4 *>(d2 + 3 * LBO_B) = g3_cvt_w(wb.w, sc2, c72);
                }
                fence_proxy_async_smem();
                __syncwarp();
                if (lane == 0) {
                    g3_arrive_cluster(bfull_leader + (uint32_t)fs * 8);
                    mbar_arrive(&rawempty[s]);       // after the __syncwarp: every lane has read its raw bytes
                }
                if (uw == 0) G3_STAMP(2, gk, 3);
                if (++s == G3_RAW_STAGES) { s = 0; u++; }
            }
        }
    } else {
        // ===================== epilogue =====================
        asm volatile("setmaxnreg.inc.sync.aligned.u32 152;" ::: "memory");
        pdl_wait();                                          // a7 / res come from, and y may be read by, earlier kernels
        const int ew = warp - G3_EPI_WARP0;                 // 0..7
        const int quarter = warp & 3;                       // TMEM lane quarter this warp may access
        const int colhalf = ew >> 2;                        // 0: channels unpacked by CTA 0, 1: by CTA 1
        const int row = quarter * 32 + lane;
        const uint32_t accempty_leader = g3_mapa(smem_u32(accempty), 0);
        const uint32_t t_base = tmem + ((uint32_t)(quarter * 32) << 16) + colhalf * (8 * NTL);
        uint8_t *obuf = sOut + ew * (G3_OUT_BOXES * 1024);  // this warp's 32 x 32 B store boxes
        uint64_t acc2[4 * NTL];                             // 8*NTL fp32 accumulators as f32x2 pairs
        int gk = 0;
        for (int t = cluster_id; t < num_tiles; t += num_clusters) {
            const G3Tile tl = g3_decode(P, t, MB2, KT, rank);
            const int m = tl.mb * G3_BM + row;
            const bool mvalid = m < M;
            const int N = P.N[tl.sg];
            const int n_base = (tl.cbl * (2 * NTL) + colhalf * NTL) * TILE_CH;
#pragma unroll
            for (int i = 0; i < 4 * NTL; i++) acc2[i] = 0ull;
            if (ADD && tl.ks == 0 && (P.bias != nullptr || (P.res != nullptr && !P.res_in_y))) {
#pragma unroll
                for (int ch = 0; ch < NTL; ch++) {
                    const int n = n_base + ch * 8;
                    if (n < N) {
                        float4 v0 = make_float4(0.f, 0.f, 0.f, 0.f), v1 = v0;
                        if (P.bias != nullptr) {
                            v0 = __ldg(reinterpret_cast<const float4 *>(P.bias + n));
                            v1 = __ldg(reinterpret_cast<const float4 *>(P.bias + n + 4));
                        }
                        if (P.res != nullptr && !P.res_in_y && mvalid) {
                            const float4 r0 = __ldcg(reinterpret_cast<const float4 *>(P.res + (size_t)m * N + n));
                            const float4 r1 = __ldcg(reinterpret_cast<const float4 *>(P.res + (size_t)m * N + n + 4));
                            v0.x += r0.x; v0.y += r0.y; v0.z += r0.z; v0.w += r0.w;
                            v1.x += r1.x; v1.y += r1.y; v1.z += r1.z; v1.w += r1.w;
                        }
                        acc2[ch * 4 + 0] = g3_pack(v0.x, v0.y);
                        acc2[ch * 4 + 1] = g3_pack(v0.z, v0.w);
                        acc2[ch * 4 + 2] = g3_pack(v1.x, v1.y);
                        acc2[ch * 4 + 3] = g3_pack(v1.z, v1.w);
                    }
                }
            }
            for (int kt = tl.kt0; kt < tl.kt1; kt++, gk++) {
                const int as = gk & 1, u2 = gk >> 1;
                const int ring = gk & (G3_SN_RING - 1);
                if (ew == 0) G3_STAMP(3, gk, 0);
                g3_wait(&snfull[ring], (uint32_t)((gk / G3_SN_RING) & 1));
                g3_wait(&accfull[as], (uint32_t)(u2 & 1));
                if (ew == 0) G3_STAMP(3, gk, 1);
                tc_fence_after();
                const float rm = sRm[ring * G3_BM + row];       // row factor a7/127 (activation warps); column factor S[n]
                const uint64_t rm2 = g3_pack(rm, rm);
                const float *sn = sSn + ring * G3_SN_STRIDE + colhalf * (8 * NTL);
                int anyz = 0;
#pragma unroll
                for (int i = 0; i < SN_WARPS; i++) anyz |= sZflag[ring * 8 + i];
                const uint32_t t0 = t_base + as * G3_TM_D_STRIDE;
                // 32 columns per TMEM load.  tcgen05.wait::ld waits for ALL loads of the thread, so loads cannot be pipelined
                // behind each other: the read phase costs (load latency + math) per chunk, and it must stay below the k-tile's
                // MMA time or the two accumulator stages stall the tensor pipe (8-column chunks: 12 x ~250 cycles of latency
                // per k-tile WAS the kernel's period).  Three chunks of 32 columns, one staging buffer.
                uint32_t r[32];
#pragma unroll
                for (int c = 0; c < NTL / 4; c++) {
                    tmem_ld_32x32(t0 + c * 32, r);
                    // (S[] and the row factor of this k-tile may still be read after the hand-back below: the weight warps
                    // run at most 4 k-tiles ahead of the last accumulator stage released, the rings hold 8)
                    tmem_ld_wait();
                    if (c + 1 == NTL / 4) {          // all TMEM reads of this stage are done: hand it back
                        tc_fence_before();
                        __syncwarp();
                        if (lane == 0) g3_arrive_cluster(accempty_leader + (uint32_t)as * 8);
                        if (ew == 0) G3_STAMP(3, gk, 2);
                    }
                    if (PROF && (P.exp & 1)) continue;
#pragma unroll
                    for (int j = 0; j < 4; j++) {    // 8 columns per step
                        const ulonglong2 sa = *reinterpret_cast<const ulonglong2 *>(sn + c * 32 + j * 8);
                        const ulonglong2 sb = *reinterpret_cast<const ulonglong2 *>(sn + c * 32 + j * 8 + 4);
                        const uint64_t sv2[4] = {sa.x, sa.y, sb.x, sb.y};
#pragma unroll
                        for (int i = 0; i < 4; i++) {
                            const float f0 = __uint_as_float(r[j * 8 + 2 * i]), f1 = __uint_as_float(r[j * 8 + 2 * i + 1]);   // = sum_g sc*dot, exact integers
                            const uint64_t f2 = g3_mul2(g3_pack(f0, f1), rm2);
                            acc2[c * 16 + j * 4 + i] = g3_fma2(f2, sv2[i], acc2[c * 16 + j * 4 + i]);
                            if (DEBUG) {
                                const int n = n_base + c * 32 + j * 8 + 2 * i;
                                if (dbg_isum && mvalid && n < N) {
                                    dbg_isum[((size_t)m * N + n) * KT + kt] = (int)f0;
                                    dbg_isum[((size_t)m * N + n + 1) * KT + kt] = (int)f1;
                                }
                            }
                        }
                    }
                }
                if (ew == 0) G3_STAMP(3, gk, 3);
                if (anyz && mvalid) {
#pragma unroll
                    for (int i = 0; i < 4 * NTL; i++) {      // unrolled although cold: a dynamic index would put acc2[] in local memory
                        const int n = n_base + 2 * i;
                        if (n < N) {
                            float a, b;
                            g3_unpack(acc2[i], a, b);
                            const float a7_cur = st.a7[(size_t)m * KT + kt];
                            a += g3_zero_point_term(P.w[tl.sg], st.sumq, m, n, kt, KT, a7_cur);
                            b += g3_zero_point_term(P.w[tl.sg], st.sumq, m, n + 1, kt, KT, a7_cur);
                            acc2[i] = g3_pack(a, b);
                        }
                    }
                }
            }
            // ---- the tile's 32 rows x 8*NTL channels of this warp: one 32 x 32 B box per 8 channels through shared memory
            //      and a tiled bulk store (or reduce-add when K is split); rows >= M and channels >= N are clipped by the copy
            if (lane == 0) g3_store_wait_read<0>();         // the boxes are free again (stores of the previous tile)
            __syncwarp();
            const int row0 = tl.mb * G3_BM + quarter * 32;
            const CUtensorMap *my = &P.seg[tl.sg].tm_y;
#pragma unroll
            for (int ch = 0; ch < NTL; ch++) {
                uint8_t *box = obuf + (ch % G3_OUT_BOXES) * 1024;
                if (ch >= G3_OUT_BOXES) {
                    if (lane == 0) g3_store_wait_read<G3_OUT_BOXES - 1>();     // the store that used this box has read it
                    __syncwarp();
                }
                *reinterpret_cast<ulonglong2 *>(box + lane * 32) = make_ulonglong2(acc2[ch * 4 + 0], acc2[ch * 4 + 1]);
                *reinterpret_cast<ulonglong2 *>(box + lane * 32 + 16) = make_ulonglong2(acc2[ch * 4 + 2], acc2[ch * 4 + 3]);
                fence_proxy_async_smem();
                __syncwarp();
                if (lane == 0) {
                    if (n_base + ch * 8 < N && row0 < M) {
                        if (P.ksplit > 1)
                            g3_tma_reduce_add_2d(my, box, n_base + ch * 8, row0);
                        else
                            g3_tma_store_2d(my, box, n_base + ch * 8, row0);
                    }
                    g3_store_commit();
                }
            }
        }
        if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");     // all stores complete before exit
    }

    tc_fence_before();
    g3_cluster_sync();
    if (warp == 1)
        asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512) : "memory");
}

// ---- host side ---------------------------------------------------------------------------------------------------
static int64_t g3_col_blocks(int64_t N, int ntl) { return (N / TILE_CH + 2 * ntl - 1) / (2 * ntl); }

bool gemm3_plan(int64_t M, const int64_t *Ns, int nseg, int64_t K, int num_sms, size_t smem_optin, bool allow_split,
                Gemm3Plan *plan) {
    if (M < 1 || nseg < 1 || nseg > G3_MAX_SEG || (size_t)Gemm3Smem::total > smem_optin || num_sms < 2) return false;
    const int64_t pairs = num_sms / 2, MB2 = (M + 2 * G3_BM - 1) / (2 * G3_BM), KT = K / TILE_K;
    const int ntl = G3_MAXNTL;
    int64_t cbs = 0;
    for (int i = 0; i < nseg; i++) cbs += g3_col_blocks(Ns[i], ntl);
    int64_t tiles = MB2 * cbs;
    // K split in two when fewer than half of the SM pairs would have a tile: the tensor-pipe time of a tile does not
    // shrink with its width (<= 192 columns cost the same), so depth is the only dimension left to spread narrow matrices
    static const bool no_split = getenv("TCQ4_G3_NOSPLIT") != nullptr;
    int ksplit = 1;
    if (allow_split && !no_split && 2 * tiles <= pairs && KT >= 8) ksplit = 2;
    tiles *= ksplit;
    if (tiles > (int64_t)1 << 30) return false;
    plan->ntl = ntl;
    plan->ksplit = ksplit;
    plan->num_tiles = (int)tiles;
    plan->grid = dim3((unsigned)(2 * (tiles < pairs ? tiles : pairs)), 1, 1);
    plan->smem_bytes = Gemm3Smem::total;
    return true;
}

// ---- tensor maps, cached by (base pointer, shape, box): encoding ten maps per launch on the host was pure overhead on
//      the eager (ggml) and host-buffer paths
typedef CUresult (*G3EncodeTiled)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static G3EncodeTiled g3_encoder() {
    static G3EncodeTiled fn = [] {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult q = cudaDriverEntryPointSymbolNotFound;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
            q != cudaDriverEntryPointSuccess)
            p = nullptr;
        return (G3EncodeTiled)p;
    }();
    return fn;
}
struct G3MapKey {
    const void *p;
    int64_t d0, d1, stride;
    uint32_t b0, b1;
    int kind;
    bool operator==(const G3MapKey &o) const {
        return p == o.p && d0 == o.d0 && d1 == o.d1 && stride == o.stride && b0 == o.b0 && b1 == o.b1 && kind == o.kind;
    }
};
struct G3MapHash {
    size_t operator()(const G3MapKey &k) const {
        size_t h = (size_t)k.p * 1099511628211ull;
        h ^= (size_t)k.d0 + 0x9e3779b97f4a7c15ull + (h << 6) + (h >> 2);
        h ^= (size_t)k.d1 + 0x9e3779b97f4a7c15ull + (h << 6) + (h >> 2);
        h ^= (size_t)k.stride + 0x9e3779b97f4a7c15ull + (h << 6) + (h >> 2);
        h ^= ((size_t)k.b0 << 20) ^ ((size_t)k.b1 << 4) ^ (size_t)k.kind;
        return h;
    }
};
static std::mutex g3_map_mu;
static std::unordered_map<G3MapKey, CUtensorMap, G3MapHash> g3_map_cache;

static bool g3_get_map(CUtensorMap *out, const G3MapKey &key) {
    {
        std::lock_guard<std::mutex> lk(g3_map_mu);
        auto it = g3_map_cache.find(key);
        if (it != g3_map_cache.end()) {
            *out = it->second;
            return true;
        }
    }
    G3EncodeTiled enc = g3_encoder();
    if (!enc) return false;
    const cuuint64_t gdim[2] = {(cuuint64_t)key.d0, (cuuint64_t)key.d1};
    const cuuint64_t gstride[1] = {(cuuint64_t)key.stride};
    const cuuint32_t box[2] = {key.b0, key.b1};
    const cuuint32_t estr[2] = {1, 1};
    const CUtensorMapDataType dt = key.kind == 0 ? CU_TENSOR_MAP_DATA_TYPE_UINT64 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32;
    if (enc(out, dt, 2, const_cast<void *>(key.p), gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
            CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
        return false;
    std::lock_guard<std::mutex> lk(g3_map_mu);
    if (g3_map_cache.size() > 8192) g3_map_cache.clear();       // bounded: a model has a few hundred weight tensors
    g3_map_cache.emplace(key, *out);
    return true;
}
// the TCQ4 tensor as [NT rows][KT*148 eight-byte words]
bool weight_tensor_map(CUtensorMap_st *map, const void *w, int64_t NT, int64_t KT, uint32_t box_words, uint32_t box_rows) {
    return g3_get_map(map, G3MapKey{w, KT * (TILE_BYTES / 8), NT, KT * TILE_BYTES, box_words, box_rows, 0});
}
static bool g3_weight_map(CUtensorMap *map, const void *w, int64_t NT, int64_t KT, uint32_t box_words, uint32_t box_rows) {
    return weight_tensor_map(map, w, NT, KT, box_words, box_rows);
}
// the fp32 output [M][ldc] with boxes of 32 rows x 8 channels
static bool g3_y_map(CUtensorMap *map, float *y, int64_t M, int64_t N, int64_t ldc) {
    return g3_get_map(map, G3MapKey{y, N, M, ldc * (int64_t)sizeof(float), 8, 32, 1});
}

template <int NTL>
static cudaError_t launch_gemm3_t(cudaLaunchConfig_t *cfg, G3Params &P, ActStage st, int64_t M, int64_t K, int MB2,
                                  int num_tiles, int32_t *dbg_isum) {
#ifdef TCQ4_DIAG
    static const bool prof = getenv("TCQ4_G3_PROF") != nullptr;
    if (prof && !dbg_isum && !P.res && !P.bias) {
        cudaFuncSetAttribute(tcq4_gemm3_kernel<NTL, false, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, Gemm3Smem::total);
        return cudaLaunchKernelEx(cfg, tcq4_gemm3_kernel<NTL, false, false, true>, P, st, (int)M, (int)K, MB2, num_tiles, (int32_t *)nullptr);
    }
#endif
    if (dbg_isum)
        return cudaLaunchKernelEx(cfg, tcq4_gemm3_kernel<NTL, true, false>, P, st, (int)M, (int)K, MB2, num_tiles, dbg_isum);
    if (P.res != nullptr || P.bias != nullptr)
        return cudaLaunchKernelEx(cfg, tcq4_gemm3_kernel<NTL, false, true>, P, st, (int)M, (int)K, MB2, num_tiles,
                                  (int32_t *)nullptr);
    return cudaLaunchKernelEx(cfg, tcq4_gemm3_kernel<NTL, false, false>, P, st, (int)M, (int)K, MB2, num_tiles,
                              (int32_t *)nullptr);
}

cudaError_t launch_gemm3(const Gemm3Seg *segs, int nseg, ActStage st, int64_t M, int64_t K, const Gemm3Plan &plan,
                         int32_t *dbg_isum, cudaStream_t stream, const float *res, const float *bias) {
    if (nseg < 1 || nseg > G3_MAX_SEG || ((res || bias || dbg_isum) && nseg != 1)) return cudaErrorInvalidValue;
    G3Params P;
    memset(&P, 0, sizeof(P));
    const int64_t KT = K / TILE_K;
    int cb = 0;
    for (int i = 0; i < nseg; i++) {
        const int64_t NT = segs[i].N / TILE_CH;
        if (!g3_weight_map(&P.seg[i].tm_tiles, segs[i].w, NT, KT, 128, (uint32_t)plan.ntl) ||
            !g3_weight_map(&P.seg[i].tm_meta, segs[i].w, NT, KT, 160 / 8, (uint32_t)(2 * plan.ntl)) ||
            !g3_y_map(&P.seg[i].tm_y, segs[i].y, M, segs[i].N, segs[i].ldc))
            return cudaErrorInvalidValue;      // unaligned tensor base or no driver entry point
        P.w[i] = reinterpret_cast<const uint8_t *>(segs[i].w);
        P.N[i] = (int)segs[i].N;
        P.cb0[i] = cb;
        cb += (int)g3_col_blocks(segs[i].N, plan.ntl);
    }
    for (int i = nseg; i <= G3_MAX_SEG; i++) P.cb0[i] = cb;
    P.nseg = nseg;
    P.ksplit = plan.ksplit;
    P.res = res;
    P.bias = bias;
    P.res_in_y = (res != nullptr && res == segs[0].y && plan.ksplit > 1) ? 1 : 0;
    P.exp = getenv("TCQ4_G3_EXP") ? atoi(getenv("TCQ4_G3_EXP")) : 0;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = plan.grid;
    cfg.blockDim = dim3(G3_THREADS);
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
    static const bool no_pdl = getenv("TCQ4_NO_PDL") != nullptr;
    cfg.numAttrs = no_pdl ? 1 : 2;
    const int MB2 = (int)((M + 2 * G3_BM - 1) / (2 * G3_BM));
    switch (plan.ntl) {
        case 12: return launch_gemm3_t<12>(&cfg, P, st, M, K, MB2, plan.num_tiles, dbg_isum);
        default: return cudaErrorInvalidValue;
    }
}

cudaError_t gemm3_read_prof(long long *host_out, size_t bytes) {
#ifdef TCQ4_DIAG
    return cudaMemcpyFromSymbol(host_out, g_g3_prof, bytes < sizeof(g_g3_prof) ? bytes : sizeof(g_g3_prof));
#else
    (void)host_out; (void)bytes;
    return cudaErrorNotSupported;
#endif
}

cudaError_t gemm3_init_attributes(size_t smem_optin) {
    if ((size_t)Gemm3Smem::total > smem_optin) return cudaErrorInvalidValue;
    cudaError_t e;
    if ((e = cudaFuncSetAttribute(tcq4_gemm3_kernel<12, true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  Gemm3Smem::total)) != cudaSuccess)
        return e;
    if ((e = cudaFuncSetAttribute(tcq4_gemm3_kernel<12, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  Gemm3Smem::total)) != cudaSuccess)
        return e;
    return cudaFuncSetAttribute(tcq4_gemm3_kernel<12, false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                Gemm3Smem::total);
}

}  // namespace tcq4
