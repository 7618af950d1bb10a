// tcq4_common.cuh -- shared definitions for the sm_100a TCQ4_K32 W4A4 kernels.
//
// Layout contract: block_tcq4_tile, /root/reference/ggml/src/ggml-common.h:321-341 (1184 B per
// 8 output channels x 256 K), tensor = [N/8][K/256] tiles, n-tile major
// (/root/reference/ggml/src/ggml-quants.c:1550-1573).
#pragma once

#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace tcq4 {

constexpr int TILE_K       = 256;
constexpr int TILE_CH      = 8;
constexpr int TILE_GROUPS  = 8;
constexpr int GROUP_SIZE   = 32;
constexpr int TILE_BYTES   = 1184;
constexpr int TILE_NIB     = 1024;   // bytes of nibbles at the head of a tile
constexpr int TILE_OFF_S   = 1024;   // fp16 S[8]
constexpr int TILE_OFF_Z   = 1040;   // fp16 Z[8]
constexpr int TILE_OFF_SC  = 1056;   // int8 sc[8][8]  (channel major)
constexpr int TILE_OFF_ZC  = 1120;   // int8 zc[8][8]

struct __align__(16) Tile {
    uint8_t tiles[8][128];
    __half  S[8];
    __half  Z[8];
    int8_t  sc[8][8];
    int8_t  zc[8][8];
};
static_assert(sizeof(Tile) == TILE_BYTES, "block_tcq4_tile must be 1184 bytes");

// block_rrs_int4_tc, /root/reference/ggml/src/ggml-cuda/tcq4_k32.cuh:51-56
struct __align__(16) ActBlockTC {
    uint8_t qs[128];
    float   smooth_scale;
    int16_t sum_q[8];
    uint8_t pad[12];
};
static_assert(sizeof(ActBlockTC) == 160, "block_rrs_int4_tc must be 160 bytes");

// ---- GEMM activation staging (internal scratch format, written by the activation kernel) -------
// For every (m-block of 128 rows, k-tile) one 32 KB unit holding s8 codes in the canonical
// no-swizzle K-major UMMA layout:  [16 K-chunks of 16 B][16 row-groups][8 rows][16 B].
//   LBO (K-direction core-matrix stride) = 2048 B, SBO (row-group stride) = 128 B.
// Inside a 32-group the K byte order is sigma-permuted per 8-slice: byte r of slice s holds element
// s*8 + SIGMA[r], SIGMA = {0,2,4,6,1,3,5,7} -- the order the weight unpack (unpack_chunk_split) produces for free.
constexpr int ACT_UNIT_BYTES = 128 * 256;
constexpr int ACT_LBO        = 2048;
constexpr int ACT_SBO        = 128;

// ---- GEMM activation staging, nibble form (tcq4_gemm3.cu) -------------------------------------
// For every (m-block of 128 rows, k-tile) one 16 KB unit [8 groups][128 rows][16 B]: the 16 bytes are the four
// 32-bit words of the row's 32-group, word s = elements s*8 .. s*8+7, element i at bits [4i, 4i+4) as u = q + 8 in
// [1, 15] -- the weight tiles' own nibble order, so the same widening trick serves both operands.  Half the bytes of an
// s8 unit: the L2 -> SM fabric and the depth of the copy ring (bytes in flight = bandwidth x latency) are what bound the
// GEMM's operand supply, not the extra ALU work of widening nibbles.
constexpr int ACT_NIB_UNIT_BYTES = 128 * 128;

struct ActStage {           // pointers into the caller's workspace
    int8_t  *codes;         // [MB][KT][32 KB] s8 units, or [MB][KT][16 KB] nibble units (tcq4_gemm3.cu; same region)
    float   *a7;            // [M][KT]      max_abs / 7
    int16_t *sumq;          // [M][KT][8]   per-group code sums (zero-point term)
    float   *a7t;           // [MB][KT][128] the same scales, contiguous per (128-row block, k-tile): one 512-byte bulk copy
};                          //               (tcq4_gemm3.cu only; null otherwise)

__host__ __device__ inline int64_t round_up(int64_t x, int64_t a) { return (x + a - 1) / a * a; }

#ifdef __CUDACC__
// ---- weight unpack for the tcgen05 GEMMs: one 16-byte chunk (channel c of group g: 32 int4 nibbles) ->
//      32 "hi" bytes + 32 "lo" bytes of the s8 B operand, with the group's int8 sub-scale folded in EXACTLY:
//          sc = 16*sh + sl  (sh = sc >> 4 arithmetic in [-8, 7], sl = sc & 15),   q in [-8, 7]
//          p  = sc*q = 16*(sh*q) + (sl*q),    sh*q in [-56, 64],  sl*q in [-120, 105]   -> both fit s8
//      so  16 * (A . hi) + (A . lo) = sum_k a_k * sc * q_k  in int32, the same identity the epilogue always used.
//      Splitting the SCALE (not the product) keeps every field inside one byte, so a register of four nibble bytes
//      is multiplied at once: with u = q + 8 in [0, 15],  byte = s*u + (128 - 8*s) = s*q + 128 in [8, 233] -- no carry
//      between bytes for any s in [-8, 15] -- and one XOR 0x80 per byte turns it into the two's-complement s8.
//      Per 8-nibble word: 1 shift, 2 mask ops, 4 multiply-adds, 4 XORs (the product-split version needed 17 with
//      four byte permutes).  Even nibbles land in one register and odd ones in the next: K byte order inside an
//      8-slice is {0,2,4,6,1,3,5,7}, which the activation kernel mirrors in its s8 codes.
//      out_hi / out_lo point at this channel's 16-byte row in K-chunk 0; K-chunk 1 is lbo bytes further.
// The arithmetic and the four stores are separate steps so that a caller can do the arithmetic BEFORE it waits for the
// destination stage to become free (tcq4_gemm2.cu).
__device__ __forceinline__ void unpack_chunk_compute(const uint4 wq, int sc, uint32_t (&hi)[8], uint32_t (&lo)[8]) {
    const uint32_t sh = (uint32_t)(sc >> 4), sl = (uint32_t)(sc & 15);
    const uint32_t ch = 0x80808080u - sh * 0x08080808u;        // 128 - 8*sh in every byte
    const uint32_t cl = 0x80808080u - sl * 0x08080808u;
    const uint32_t wv[4] = {wq.x, wq.y, wq.z, wq.w};
#pragma unroll
    for (int s = 0; s < 4; s++) {
        const uint32_t e = (wv[s] ^ 0x08080808u) & 0x0F0F0F0Fu;            // u of elements 0, 2, 4, 6
        const uint32_t o = ((wv[s] >> 4) ^ 0x08080808u) & 0x0F0F0F0Fu;     // u of elements 1, 3, 5, 7
        hi[s * 2 + 0] = (e * sh + ch) ^ 0x80808080u;
        hi[s * 2 + 1] = (o * sh + ch) ^ 0x80808080u;
        lo[s * 2 + 0] = (e * sl + cl) ^ 0x80808080u;
        lo[s * 2 + 1] = (o * sl + cl) ^ 0x80808080u;
    }
}
__device__ __forceinline__ void unpack_chunk_store(const uint32_t (&hi)[8], const uint32_t (&lo)[8], uint8_t *out_hi,
                                                   uint8_t *out_lo, int lbo) {
    *reinterpret_cast<uint4 *>(out_hi) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
    *reinterpret_cast<uint4 *>(out_hi + lbo) = make_uint4(hi[4], hi[5], hi[6], hi[7]);
    *reinterpret_cast<uint4 *>(out_lo) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
    *reinterpret_cast<uint4 *>(out_lo + lbo) = make_uint4(lo[4], lo[5], lo[6], lo[7]);
}
__device__ __forceinline__ void unpack_chunk_split(const uint4 wq, int sc, uint8_t *out_hi, uint8_t *out_lo, int lbo) {
    uint32_t hi[8], lo[8];
    unpack_chunk_compute(wq, sc, hi, lo);
    unpack_chunk_store(hi, lo, out_hi, out_lo, lbo);
}
#endif

// ---------------------------------------------------------------------------------------------
// PTX wrappers (sm_100a)
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_mbar_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t *bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    while (!mbar_try_wait(bar, parity)) { }
}

// 1-D bulk async copy global -> shared (TMA unit, SASS UBLKCP); size and both addresses % 16 == 0
__device__ __forceinline__ void bulk_g2s(void *dst_smem, const void *src_gmem, uint32_t bytes, uint64_t *bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
            smem_u32(dst_smem)),
        "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
        : "memory");
}
// same with an L2 evict-first policy (streamed-once weights)
__device__ __forceinline__ void bulk_g2s_stream(void *dst_smem, const void *src_gmem, uint32_t bytes,
                                                uint64_t *bar, uint64_t policy) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::
            "r"(smem_u32(dst_smem)),
        "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar)), "l"(policy)
        : "memory");
}
__device__ __forceinline__ uint64_t l2_policy_evict_first() {
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
    return p;
}
__device__ __forceinline__ uint64_t l2_policy_evict_last() {
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
    return p;
}

// generic-proxy smem writes -> visible to the async proxy (tcgen05.mma / TMA reads)
__device__ __forceinline__ void fence_proxy_async_smem() {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

// Where a kernel writes its [M][N] result: `n` row-major fp32 matrices with leading dimension ldc, each pointer already
// offset to this call's first column.  n = 1, ldc = N is ggml's dst; n = G with peer-mapped pointers is the N-sharded
// MUL_MAT whose epilogue stores its column block straight into every GPU's full C (the all-gather fused away).
constexpr int MAX_PEERS = 8;
struct YOut {
    float    *p[MAX_PEERS];
    int       n;
    long long ldc;
    float    *mc;          // non-null: ONE multicast (NVLS) address of the symmetric C instead of p[]: every store is
};                         // replicated to all GPUs by the switch (multimem.st), so the outbound traffic is 1x, not (G-1)x
__device__ __forceinline__ void multimem_st_f32(float *addr, float v) {
    asm volatile("multimem.st.weak.global.f32 [%0], %1;" ::"l"(addr), "f"(v) : "memory");
}
__device__ __forceinline__ void multimem_st_v4(float *addr, float4 v) {
    asm volatile("multimem.st.weak.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
                 : "memory");
}

// Programmatic dependent launch: a kernel launched with cudaLaunchAttributeProgrammaticStreamSerialization may start
// once every CTA of its predecessor has executed launch_dependents (or exited); it must not touch memory the
// predecessor reads or writes before griddepcontrol.wait, which returns when the predecessor has completed and
// flushed.  Every kernel of this library executes both (all threads, unconditionally).
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile(
        "{\n\t.reg .pred P;\n\t"
        "elect.sync _|P, 0xffffffff;\n\t"
        "selp.u32 %0, 1, 0, P;\n\t}"
        : "=r"(pred));
    return pred != 0;
}

// ---- tcgen05 / TMEM -----------------------------------------------------------------------
__device__ __forceinline__ void tmem_alloc(uint32_t *dst_smem, uint32_t ncols) {   // whole warp
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)),
                 "r"(ncols)
                 : "memory");
}
__device__ __forceinline__ void tmem_relinquish() {
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {     // whole warp
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() {
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_after() {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}
// D[tmem] (+)= A[smem] * B[smem], int8 inputs, int32 accumulate; one thread issues
__device__ __forceinline__ void umma_i8(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                        uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
        "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// arrive on an mbarrier when all previously issued tcgen05.mma of this thread have completed
__device__ __forceinline__ void umma_commit(uint64_t *bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() {
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
// 32 lanes x 32 columns of 32-bit: thread t of the warp gets columns [c, c+32) of TMEM lane base+t
__device__ __forceinline__ void tmem_ld_32x32(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
          "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
          "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ void tmem_ld_32x16(uint32_t taddr, uint32_t (&r)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr)
        : "memory");
}

__device__ __forceinline__ void tmem_ld_32x8(uint32_t taddr, uint32_t (&r)[8]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
        : "r"(taddr)
        : "memory");
}

// Shared-memory matrix descriptor, K-major, no swizzle (cute::UMMA::SmemDescriptor bit layout):
//  [0,14) start>>4 | [16,30) LBO>>4 | [32,46) SBO>>4 | [46,48) version=1 | [61,64) layout=0
__device__ __forceinline__ uint64_t umma_desc_kmajor_noswz(uint32_t smem_addr, uint32_t lbo, uint32_t sbo) {
    uint64_t d = 0;
    d |= (uint64_t)((smem_addr >> 4) & 0x3FFF);
    d |= (uint64_t)((lbo >> 4) & 0x3FFF) << 16;
    d |= (uint64_t)((sbo >> 4) & 0x3FFF) << 32;
    d |= (uint64_t)1 << 46;
    return d;
}
// Instruction descriptor for kind::i8, S32 accumulate, K-major A and B
//  [4,6) c_format=2(S32) | [7,10) a_format | [10,13) b_format (0=u8, 1=s8) | [17,23) N>>3 | [24,29) M>>4
__host__ __device__ constexpr uint32_t umma_idesc_i8(int M, int N, int a_signed, int b_signed) {
    return (2u << 4) | ((uint32_t)a_signed << 7) | ((uint32_t)b_signed << 10) | ((uint32_t)(N >> 3) << 17) |
           ((uint32_t)(M >> 4) << 24);
}

// ---------------------------------------------------------------------------------------------
// Rotated-Runtime-Smooth activation stage for ONE 256-block held by ONE warp.
// Lane l holds elements i = l*8 + r (r = 0..7), i.e. group g = l/4, 8-slice s = l%4.
// Arithmetic reproduces /root/reference/ggml/src/ggml-cuda/tcq4_k32.cu:781-831 operation for
// operation (same (j, j+h) pairs, lower-minus-upper, *1/16, fmaxf max, IEEE 7/max, rintf, clamp +-7)
// so codes, max_abs and group sums are bit-identical to the scalar oracle.
// ---------------------------------------------------------------------------------------------
struct ActQuant {
    int   q[8];      // codes in [-7, 7]
    float max_abs;   // after the < 1e-10 -> 1.0 rule
    int   gsum;      // sum of the 32 codes of this lane's group (valid in all 4 lanes of the group)
};

__device__ __forceinline__ ActQuant rrs_quantize_block(float (&v)[8], int lane) {
    // stages h = 1, 2, 4: inside the lane
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
    // stages h = 8 .. 128: across lanes, partner = lane ^ (h/8); upper lane computes lower - upper
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const float sgn = (lane & d) ? -1.0f : 1.0f;           // upper lane: lower - upper = o + (-v)
#pragma unroll
        for (int r = 0; r < 8; r++) {
            const float o = __shfl_xor_sync(0xffffffffu, v[r], d);
            v[r] = __fmaf_rn(v[r], sgn, o);                    // exact: v * (+-1) + o rounds once, like the add
        }
    }
    float mx = 0.0f;
#pragma unroll
    for (int r = 0; r < 8; r++) {
        v[r] = v[r] * (1.0f / 16.0f);
        const float a = fabsf(v[r]);
        mx = (r == 0) ? a : fmaxf(mx, a);
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, d));
    if (mx < 1e-10f) mx = 1.0f;
    const float inv = __fdiv_rn(7.0f, mx);
    ActQuant out;
    out.max_abs = mx;
    int s = 0;
#pragma unroll
    for (int r = 0; r < 8; r++) {
        const float scaled = v[r] * inv;
        const int q = (int)fmaxf(-7.0f, fminf(7.0f, rintf(scaled)));
        out.q[r] = q;
        s += q;
    }
    s += __shfl_xor_sync(0xffffffffu, s, 1);
    s += __shfl_xor_sync(0xffffffffu, s, 2);
    out.gsum = s;
    return out;
}

// ---------------------------------------------------------------------------------------------
// Same arithmetic, fewer warp shuffles (SHFL issues at one warp-instruction per clock per SM --
// measured, tests/gpu_pipe_peak.py -- and the 40 shuffles of the version above are what bounds it).
// The block enters in layout A (lane l holds i = l*8 + r): stages h = 1,2,4 in registers, h = 8,16 by
// shuffle (lane bits 0,1).  It is then transposed through a 1 KB per-warp shared-memory scratch into
// layout B (lane l holds i = r*32 + l, i.e. ONE element of each 32-group r), where h = 32,64,128 are
// again in registers.  Same (j, j+h) pairs, same operand order, same stage order -> bit-identical.
// Group sums are one REDUX each.  Output: codes in layout B.
// ---------------------------------------------------------------------------------------------
struct ActQuantT {
    int   q[8];      // q[g] = code of element g*32 + lane
    float max_abs;   // after the < 1e-10 -> 1.0 rule
    int   gsum[8];   // sum of the 32 codes of group g (all lanes)
};

template <bool WANT_GSUM>
__device__ __forceinline__ ActQuantT rrs_quantize_block_t(float (&v)[8], int lane, float *scratch /* 256 floats, this warp's */) {
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
    for (int d = 1; d < 4; d <<= 1) {                      // h = 8, 16
        const float sgn = (lane & d) ? -1.0f : 1.0f;       // upper lane: lower - upper = o + (-v)
#pragma unroll
        for (int r = 0; r < 8; r++) {
            const float o = __shfl_xor_sync(0xffffffffu, v[r], d);
            v[r] = __fmaf_rn(v[r], sgn, o);                // exact: v * (+-1) + o rounds once, like the add
        }
    }
    __syncwarp();
    *reinterpret_cast<float4 *>(scratch + lane * 8)     = make_float4(v[0], v[1], v[2], v[3]);
    *reinterpret_cast<float4 *>(scratch + lane * 8 + 4) = make_float4(v[4], v[5], v[6], v[7]);
    __syncwarp();
#pragma unroll
    for (int r = 0; r < 8; r++) v[r] = scratch[r * 32 + lane];
#pragma unroll
    for (int h = 1; h < 8; h <<= 1) {                      // h = 32, 64, 128 on the register index
#pragma unroll
        for (int r = 0; r < 8; r++) {
            if ((r & h) == 0) {
                const float a = v[r], b = v[r + h];
                v[r]     = a + b;
                v[r + h] = a - b;
            }
        }
    }
    float mx = 0.0f;
#pragma unroll
    for (int r = 0; r < 8; r++) {
        v[r] = v[r] * (1.0f / 16.0f);
        const float a = fabsf(v[r]);
        mx = (r == 0) ? a : fmaxf(mx, a);
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, d));
    if (mx < 1e-10f) mx = 1.0f;
    const float inv = __fdiv_rn(7.0f, mx);
    ActQuantT out;
    out.max_abs = mx;
#pragma unroll
    for (int r = 0; r < 8; r++) {
        const float scaled = v[r] * inv;
        const int q = (int)fmaxf(-7.0f, fminf(7.0f, rintf(scaled)));
        out.q[r] = q;
        out.gsum[r] = WANT_GSUM ? __reduce_add_sync(0xffffffffu, q) : 0;
    }
    return out;
}

// codes of layout B -> layout A through the same scratch: returns the s8 codes of elements lane*8 .. lane*8+7
// (byte r of the 64-bit result = element lane*8 + r)
__device__ __forceinline__ uint2 rrs_codes_layout_a(const ActQuantT &a, int lane, float *scratch) {
    uint8_t *sb = reinterpret_cast<uint8_t *>(scratch);
    __syncwarp();
#pragma unroll
    for (int g = 0; g < 8; g++) sb[g * 32 + lane] = (uint8_t)a.q[g];
    __syncwarp();
    return *reinterpret_cast<const uint2 *>(sb + lane * 8);
}

// load the 8 elements of lane `lane` of block kt of row x_row (optionally through perm)
__device__ __forceinline__ void rrs_load_block(const float *__restrict__ x_row, const int32_t *__restrict__ perm,
                                               int kt, int lane, float (&v)[8]) {
    const int base = kt * TILE_K + lane * 8;
    if (perm == nullptr) {
        const float4 a = __ldg(reinterpret_cast<const float4 *>(x_row + base));
        const float4 b = __ldg(reinterpret_cast<const float4 *>(x_row + base) + 1);
        v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w;
        v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
    } else {
        const int4 p0 = __ldg(reinterpret_cast<const int4 *>(perm + base));
        const int4 p1 = __ldg(reinterpret_cast<const int4 *>(perm + base) + 1);
        v[0] = __ldg(x_row + p0.x); v[1] = __ldg(x_row + p0.y); v[2] = __ldg(x_row + p0.z); v[3] = __ldg(x_row + p0.w);
        v[4] = __ldg(x_row + p1.x); v[5] = __ldg(x_row + p1.y); v[6] = __ldg(x_row + p1.z); v[7] = __ldg(x_row + p1.w);
    }
}

// same, but x through ld.global.cg (L2 only): used by kernels launched with programmatic stream
// serialization, whose CTAs may be resident (with a warm L1) before the producer of x has finished
__device__ __forceinline__ void rrs_load_block_cg(const float *__restrict__ x_row, const int32_t *__restrict__ perm,
                                                  int kt, int lane, float (&v)[8]) {
    const int base = kt * TILE_K + lane * 8;
    if (perm == nullptr) {
        const float4 a = __ldcg(reinterpret_cast<const float4 *>(x_row + base));
        const float4 b = __ldcg(reinterpret_cast<const float4 *>(x_row + base) + 1);
        v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w;
        v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
    } else {
        const int4 p0 = __ldg(reinterpret_cast<const int4 *>(perm + base));
        const int4 p1 = __ldg(reinterpret_cast<const int4 *>(perm + base) + 1);
        v[0] = __ldcg(x_row + p0.x); v[1] = __ldcg(x_row + p0.y); v[2] = __ldcg(x_row + p0.z); v[3] = __ldcg(x_row + p0.w);
        v[4] = __ldcg(x_row + p1.x); v[5] = __ldcg(x_row + p1.y); v[6] = __ldcg(x_row + p1.z); v[7] = __ldcg(x_row + p1.w);
    }
}

}  // namespace tcq4
