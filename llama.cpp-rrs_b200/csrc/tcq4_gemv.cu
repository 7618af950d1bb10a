// tcq4_gemv.cu -- K3: fused decode GEMV / small-batch GEMM for M <= 16 (HBM-bandwidth bound).
//
// Replaces tcq4_rrs_fused_gemm_smallM_kernel (/root/reference/ggml/src/ggml-cuda/tcq4_k32.cu:733-889,
// grid (ceil(N/32), M), every block redoing the activation stage and re-reading the weights per m)
// and tcq4_rrs_fused_gemv_kernel (:296-459).
//
// Shape of the kernel
//  * Persistent, one CTA of 16 warps per SM.  The flattened tile list [N/8][K/256] is exactly memory
//    order (ggml-quants.c:1550-1573); it is cut into contiguous row ranges per cluster and each of those
//    into 32 contiguous warp ranges.  A warp streams its range through a private ring of 1-D bulk async
//    copies (cp.async.bulk -> UBLKCP, 1184 B per tile) and never synchronises with another warp
//    inside the loop.
//  * Programmatic dependent launch: barrier init and the first ring fill touch only the (constant)
//    weights, so they run BEFORE griddepcontrol.wait, and launch_dependents is signalled right after
//    them -- the next MUL_MAT of the stream starts prefetching its weights into the shared memory of the
//    same SMs while this one is still computing.  x is read (ld.global.cg) only after the wait.
//  * The block's int4 x int4 dots run on the legacy warp-level tensor path: the tile format is already
//    the m16n8k32 B-fragment order (ggml-common.h:315-341), so lane L's 32-bit word of a group IS its B
//    fragment; two logic ops turn it into two s8x4 registers carrying 16*q, and one
//    mma.sync.m16n8k32.s8.s8 per 32-group multiplies 8 channels by up to 16 activation rows at once.
//    M <= 8 and M <= 16 therefore cost the same weight traffic and nearly the same instructions as M = 1.
//  * Per-group scale codes enter in the integer domain: the eight per-group dots of a channel fit s16,
//    so four PRMT + four IDP.2A form isum = sum_g sc_g * dot_g exactly (the same int32 the tcgen05 GEMM
//    accumulates in TMEM), and the fp32 step per (row, channel, k-tile) is the GEMM's:
//    acc = fma(float(isum) * (a7/127), S, acc).
//  * CTAs run as clusters of two (one TPC).  A cluster owns whole output rows (n-tiles), so no row is ever
//    cut by a cluster boundary and the kernel needs no global atomics, tickets or fences at its tail; the
//    cluster's tiles are split evenly over its 32 warps, and rows cut by a warp boundary are combined
//    deterministically (k order) in CTA 0's shared memory -- CTA 1 stores its partials there through DSMEM --
//    after one cluster barrier.
#include <cstdlib>

#include "tcq4_common.cuh"
#include "tcq4_kernels.h"

namespace tcq4 {

// Warps per CTA and ring depth are launch parameters (GemvPlan); the planner uses 16 warps x 4 tiles.
constexpr int GV_CLUSTER = 2;
// a cluster has 2*NW workers (warps) and 4*NW partial-row slots (a warp cuts at most two rows)

struct GvSplit {    // even split of T items over W workers, contiguous
    int base, rem;
    __host__ __device__ GvSplit(int T, int W) : base(T / W), rem(T % W) {}
    __host__ __device__ int start(int w) const { return w * base + (w < rem ? w : rem); }
    __host__ __device__ int count(int w) const { return base + (w < rem ? 1 : 0); }
};

__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// address of `p` (this CTA's shared memory) in CTA `rank` of the cluster
__device__ __forceinline__ uint32_t dsmem_addr(const void *p, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(smem_u32(p)), "r"(rank));
    return r;
}
// Asynchronous 4-byte store into a CTA of the cluster that also counts its bytes on an mbarrier of that CTA: the
// receiver learns "all partials have landed" from the barrier's transaction count -- no cluster barrier, no fence
// (barrier.cluster.arrive.release costs MEMBAR.ALL.GPU, ~1 us on the critical path of every decode launch).
__device__ __forceinline__ void dsmem_st_async(uint32_t addr, uint32_t v, uint32_t bar) {
    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.b32 [%0], %1, [%2];" ::"r"(addr), "r"(v),
                 "r"(bar)
                 : "memory");
}

// D(16x8, s32) = A(16x32, s8, row) * B(32x8, s8, col); rows 8..15 of A are a1/a3
__device__ __forceinline__ void mma_s8(int (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0,
                                       uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%10,%10,%10,%10};"
        : "=r"(d[0]), "=r"(d[1]), "=r"(d[2]), "=r"(d[3])
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1), "r"(0));
}

// shared-memory carve-up (bytes), shared by the kernel and the planner
struct GvSmem {
    int off_ring, off_full, off_afrag, off_rm, off_sumq, off_part, off_prow, off_cbar, total;
    __host__ __device__ static int al(int o) { return (o + 15) & ~15; }
    __host__ __device__ GvSmem(int M, int KT, int NW, int depth) {
        const int entries = NW * GV_CLUSTER * 2;
        int o = 0;
        off_ring = o;    o = al(o + NW * depth * TILE_BYTES);            // 75776 (16 x 4) or 113664 (32 x 3)
        off_full = o;    o = al(o + NW * depth * 8);
        off_afrag = o;   o = al(o + M * KT * 256);                       // [KT][8 groups][M][4 k-quarters] uint2 = M*K bytes
        off_rm = o;      o = al(o + 2 * M * KT * 4);                     // [KT][M] a7/127, then [KT][M] a7
        off_sumq = o;    o = al(o + M * KT * 8 * 4);                     // [M][KT][8] group sums (zero-point path)
        off_part = o;    o = al(o + entries * M * 32);                   // [entry][M rows][8 ch] fp32 (used in CTA 0)
        off_prow = o;    o = al(o + entries * 4);                        // [entry] n-tile or -1
        off_cbar = o;    o = al(o + 8);                                  // (CTA 0) "all partial rows have landed"
        total = o;
    }
};

// phase timestamps (ns, %globaltimer) of the DEBUG instantiation: [cta][0..5] = start, prefetch issued,
// dependency resolved, activation stage done, stream loop done, end
__device__ unsigned long long g_gv_prof[160][16];
__device__ unsigned long long g_gv_iter[16][8][3];     // CTA 0: per warp, per tile: wait begin / data ready / done
__device__ __forceinline__ unsigned long long gtime() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
#define GV_ITER(i, k) do { if (DEBUG && blockIdx.x == 0 && lane == 0 && warp < 16 && (i) < 8) g_gv_iter[warp][i][k] = gtime(); } while (0)
// stamp taken only once `dep` (a float) is available
#define GV_STAMP_DEP(i, dep) do { if (DEBUG && threadIdx.x == 0 && blockIdx.x < 160) { unsigned long long t_; \
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_) : "f"(dep)); g_gv_prof[blockIdx.x][i] = t_; } } while (0)
#define GV_STAMP(i) do { if (DEBUG && threadIdx.x == 0 && blockIdx.x < 160) g_gv_prof[blockIdx.x][i] = gtime(); } while (0)

// Zero-point term of one (row, channel, k-tile) (tcq4_k32.cu:850,872-873).  Cold: the fork's quantiser always
// writes Z = 0 (ggml-quants.c:1603).
__device__ __noinline__ float gv_zero_point(unsigned short zh, uint2 zcv, const int *sumq8, float a7) {
    const float Z = __half2float(__ushort_as_half(zh));
    const uint32_t zc[2] = {zcv.x, zcv.y};
    float zs = 0.0f;
    for (int g = 0; g < 8; g++) {
        const float sq = (float)sumq8[g] * a7;
        const int cg = (int)(int8_t)(zc[g >> 2] >> (8 * (g & 3)));
        zs += sq * __fdiv_rn(Z * (float)cg, 127.0f);
    }
    return zs;
}

// One activation block (row m, k-tile kt) quantised by one warp and stored in the layout the inner loops read; `base` is
// the start of the activation area (shared memory inside the GEMV, or the global blob of the pre-pass kernel -- the
// two have the same layout, GvSmem offsets relative to off_afrag).  DP selects the dp4a arrangement of the codes.
__device__ __forceinline__ void gv_store_act_block(const ActQuant &a, uint8_t *base, int rm_off, int sumq_off, int M, int KT,
                                                   int m, int kt, int lane, bool DP) {
    const int q = lane & 3, r = lane >> 2;
    const uint32_t ev = (uint32_t)(a.q[0] & 0xFF) | ((uint32_t)(a.q[2] & 0xFF) << 8) |
                        ((uint32_t)(a.q[4] & 0xFF) << 16) | ((uint32_t)(a.q[6] & 0xFF) << 24);
    const uint32_t od = (uint32_t)(a.q[1] & 0xFF) | ((uint32_t)(a.q[3] & 0xFF) << 8) |
                        ((uint32_t)(a.q[5] & 0xFF) << 16) | ((uint32_t)(a.q[7] & 0xFF) << 24);
    if (DP) {           // dp4a path: per (k-tile, group, token) 32 bytes = {ev of slices 0..3, od of slices 0..3}
        uint32_t *ad = reinterpret_cast<uint32_t *>(base) + ((kt * 8 + r) * M + m) * 8;
        ad[q] = ev;
        ad[4 + q] = od;
    } else {            // IMMA path: [k-tile][group][token][k-quarter] {ev, od}
        reinterpret_cast<uint2 *>(base)[((kt * 8 + r) * M + m) * 4 + q] = make_uint2(ev, od);
    }
    float *rms = reinterpret_cast<float *>(base + rm_off);
    if (lane == 0) {
        const float a7 = __fdiv_rn(a.max_abs, 7.0f);
        rms[M * KT + kt * M + m] = a7;                       // a7s = rms + M*KT
        rms[kt * M + m] = a7 * (1.0f / 127.0f);
    }
    if (q == 0) reinterpret_cast<int *>(base + sumq_off)[(m * KT + kt) * 8 + r] = a.gsum;
}

// Pre-pass for M * K/256 > GV_PREPASS_BLOCKS: inside the GEMV every CTA repeats the activation stage for all rows
// (M = 8, K = 14336: 448 blocks on 16 warps = 28 rounds, 20 us); here the blocks are spread over the whole GPU once and
// the GEMV copies the finished codes and scales (M*K + 40*M*K/256 bytes, L2-resident) into its shared memory.
constexpr int GV_PREPASS_BLOCKS = 16;      // round 2: 32 -> 16 (M = 2, 3 on K = 2560 .. 4096: a decode layer 6 % faster, tests/gpu_decode_layer.py)
constexpr int GV_PREPASS_PERM_BLOCKS = 16;
constexpr int GVP_WARPS = 4;

__global__ void __launch_bounds__(GVP_WARPS * 32)
tcq4_gemv_prepass_kernel(const float *__restrict__ x, const int32_t *__restrict__ perm, uint8_t *__restrict__ blob, int M,
                         int K, int rm_off, int sumq_off, int dp) {
    const int lane = threadIdx.x & 31;
    const int KT = K / TILE_K;
    const int b = blockIdx.x * GVP_WARPS + (threadIdx.x >> 5);
    pdl_launch_dependents();     // the GEMV behind this kernel sets itself up and prefetches its weights meanwhile
    pdl_wait();                  // x, and the blob the previous GEMV may still be reading, belong to earlier kernels
    if (b >= M * KT) return;
    const int m = b / KT, kt = b - m * KT;
    float v[8];
    rrs_load_block_cg(x + (size_t)m * K, perm, kt, lane, v);
    const ActQuant a = rrs_quantize_block(v, lane);
    gv_store_act_block(a, blob, rm_off, sumq_off, M, KT, m, kt, lane, dp != 0);
}

// MT > 0 selects the dp4a inner loop for exactly MT <= 3 tokens (at 4 the IMMA loop is faster: 14.9 vs 17.7 us at
// N=14336, K=4096) (see "dp4a path" below); MT = 0 the IMMA loop for
// up to 8 * MB tokens.
template <int MB, int MT, bool DEBUG, int MODE = 0>
__global__ void __launch_bounds__(512, (MT == 1 && !DEBUG) ? 2 : 1)
tcq4_gemv_kernel(const uint8_t *__restrict__ w, const float *__restrict__ x, const int32_t *__restrict__ perm,
                 const YOut out, int M, int N, int K, int depth, const uint8_t *__restrict__ pre_blob,
                 int32_t *__restrict__ dbg_isum, const GvExpert ex) {
    extern __shared__ __align__(16) uint8_t smem[];
    const int KT = K / TILE_K;
    const int NT = N / TILE_CH;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int q = lane & 3, r = lane >> 2;          // mma fragment coordinates: k-quarter / row (token) or column pair

    const int NW = (int)(blockDim.x >> 5);          // warps of this CTA
    const int workers = NW * GV_CLUSTER, entries = workers * 2;
    // GROUP (MUL_MAT_ID, one token): M = 1 token is computed per channel, but Ma activation rows (one per slot) may sit
    // in shared memory; N is the virtual channel count n_used * N_expert
    // MODE: 0 plain, 1 GROUP (MUL_MAT_ID, below), 2 SEG (fan-out: channel blocks of several weight matrices, same x),
    // 3 plain + bias / residual added at the store
    constexpr bool GROUP = (MODE == 1), SEG = (MODE == 2);
    static_assert(!GROUP || (MB == 1 && MT == 1), "grouped launches use the one-token dp4a loop");
    const int Ma = GROUP ? ex.xslots : M;
    const GvSmem L(Ma, KT, NW, depth);
    uint8_t *ring = smem + L.off_ring;
    uint64_t *full = reinterpret_cast<uint64_t *>(smem + L.off_full);
    uint2 *afrag = reinterpret_cast<uint2 *>(smem + L.off_afrag);      // [KT][8][M][4]
    float *rms = reinterpret_cast<float *>(smem + L.off_rm);           // [KT][M]   a7 / 127
    float *a7s = rms + Ma * KT;                                        // [KT][M]   max_abs / 7
    int *sumq = reinterpret_cast<int *>(smem + L.off_sumq);            // [M][KT][8]
    float *part = reinterpret_cast<float *>(smem + L.off_part);        // [entry][M][8]   (CTA 0's copy is the live one)
    int *prow = reinterpret_cast<int *>(smem + L.off_prow);            // [entry] n-tile or -1
    uint64_t *cbar = reinterpret_cast<uint64_t *>(smem + L.off_cbar);

    // ---- work split: whole rows per cluster, tiles evenly over the cluster's 32 warps ------------------
    const int crank = (int)cluster_ctarank();
    const int cid = blockIdx.x / GV_CLUSTER, ncl = gridDim.x / GV_CLUSTER;
    const GvSplit rs(NT, ncl);
    const int R0 = rs.start(cid), nR = rs.count(cid);
    const int wid = crank * NW + warp;
    const GvSplit wsp(nR * KT, workers);
    const int lt0 = wsp.start(wid), cnt = wsp.count(wid);

    GV_STAMP(0);
    // CTA 0 collects the cut rows of the whole cluster: every warp of the cluster sends both of its entry slots
    // (used or not), so the byte total is fixed and the barrier can be armed before anything else happens
    if (crank == 0 && threadIdx.x == 0) {
        mbar_init(cbar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        mbar_arrive_expect_tx(cbar, (uint32_t)(entries * (M * 32 + 4)));
    }

    // ---- weight prefetch: constant data, legal before the programmatic dependency resolves -----------
    // Each warp owns its ring and its barriers: lane 0 initialises them and issues the first copies without
    // any CTA-wide barrier (the other lanes first touch them after the __syncthreads below).
    uint8_t *myring = ring + warp * depth * TILE_BYTES;
    uint64_t *myfull = full + warp * depth;
    const uint8_t *wsrc = w + ((size_t)R0 * KT + lt0) * TILE_BYTES;
    // weight tile i of this warp.  Plain: tiles are in memory order.  GROUP: the slot's expert comes from the ids
    auto tile_src = [&](int i) -> const uint8_t * {
        if constexpr (SEG) {
            const int g = lt0 + i, rr = g / KT;
            const int row = R0 + rr, k = g - rr * KT;
            int si = 0;
            while (si + 1 < ex.nseg && row >= ex.seg_nt0[si + 1]) si++;
            return ex.seg_w[si] + ((size_t)(row - ex.seg_nt0[si]) * KT + k) * TILE_BYTES;
        } else if constexpr (!GROUP) {
            return wsrc + (size_t)i * TILE_BYTES;
        } else {
            const int g = lt0 + i, rr = g / KT;
            const int row = R0 + rr, k = g - rr * KT;
            const int s = row / ex.nt_expert;
            int e = __ldg(ex.sel + s);
            if ((unsigned)e >= (unsigned)ex.n) {
                if (ex.err != nullptr) *ex.err = 1;
                e = 0;
            }
            return w + (size_t)e * (size_t)ex.stride + ((size_t)(row - s * ex.nt_expert) * KT + k) * TILE_BYTES;
        }
    };
    // y[m][row * 8 + ch] = v.  SEG: the row's own matrix and its own [M][N_i] output
    auto store_out = [&](int m, int row, int ch, float v) {
        if constexpr (SEG) {
            int si = 0;
            while (si + 1 < ex.nseg && row >= ex.seg_nt0[si + 1]) si++;
            ex.seg_y[si][(size_t)m * (size_t)ex.seg_ldc[si] + (row - ex.seg_nt0[si]) * 8 + ch] = v;
        } else {
            if constexpr (MODE == 3) {                                  // tcq4_b200_mul_mat_add
                if (ex.bias) v += __ldg(ex.bias + row * 8 + ch);
                if (ex.res) v += ex.res[(size_t)m * N + row * 8 + ch];
            }
            if (out.mc) multimem_st_f32(out.mc + (size_t)m * out.ldc + row * 8 + ch, v);
            else
                for (int pi = 0; pi < out.n; pi++) out.p[pi][(size_t)m * out.ldc + row * 8 + ch] = v;
        }
    };
    if ((MODE == 0 || MODE == 3) && ex.sel != nullptr) {
        // MUL_MAT_ID: this launch's expert matrix.  The id comes from an earlier kernel of the stream (the router) and is
        // read here, ahead of griddepcontrol.wait: the first slot's launch is made without programmatic serialisation (it
        // starts after the router has completed), the later ones start no earlier than the first one has.
        int e = __ldcg(ex.sel);
        if ((unsigned)e >= (unsigned)ex.n) {
            if (ex.err != nullptr && threadIdx.x == 0 && blockIdx.x == 0) *ex.err = 1;
            e = 0;
        }
        wsrc += (size_t)e * (size_t)ex.stride;
    }
    uint64_t pol = 0;
    if (lane == 0) {
        for (int i = 0; i < depth; i++) mbar_init(&myfull[i], 1);
        fence_mbar_init();
        pol = l2_policy_evict_first();
        if (!GROUP) {
            const int pre = cnt < depth ? cnt : depth;
            for (int i = 0; i < pre; i++) {
                mbar_arrive_expect_tx(&myfull[i], TILE_BYTES);
                bulk_g2s_stream(myring + i * TILE_BYTES, tile_src(i), TILE_BYTES, &myfull[i], pol);
            }
        }
    }
    __syncwarp();
    // first half of a cluster barrier: "CTA 0 is running and its collection barrier is initialised" (execution
    // ordering only -- the init was published by fence.mbarrier_init -- so the relaxed form, which has no fence)
    asm volatile("barrier.cluster.arrive.relaxed.aligned;" ::: "memory");
    GV_STAMP(1);
    // the channel permutation is constant data like the weights: the indices of this warp's first 256-block are fetched before
    // the dependency wait, so that the gather behind it is one round trip instead of two
    int4 pp0 = make_int4(0, 0, 0, 0), pp1 = pp0;
    const bool perm_first = (pre_blob == nullptr) && (perm != nullptr) && (warp < Ma * KT);
    if (perm_first) {
        const int base = (warp % KT) * TILE_K + lane * 8;
        pp0 = __ldg(reinterpret_cast<const int4 *>(perm + base));
        pp1 = __ldg(reinterpret_cast<const int4 *>(perm + base) + 1);
    }
    pdl_launch_dependents();     // the next MUL_MAT may start its own prefetch on this SM (signalling later -- after
                                 // the activation stage or after the loop -- measured 3-20 % slower)
    pdl_wait();                  // x belongs to the previous kernel until here
    GV_STAMP(2);
    if (GROUP && lane == 0) {    // the ids may come from the kernel just before this one: first copies only now
        const int pre = cnt < depth ? cnt : depth;
        for (int i = 0; i < pre; i++) {
            mbar_arrive_expect_tx(&myfull[i], TILE_BYTES);
            bulk_g2s_stream(myring + i * TILE_BYTES, tile_src(i), TILE_BYTES, &myfull[i], pol);
        }
    }

    // ---- activation stage for all M rows -----------------------------------------------------------------
    // (latency-bound here -- at most a few 256-blocks per warp -- so the all-shuffle butterfly, whose dependent
    //  chain is shorter than the shared-memory transposition of K1, is the faster one)
    // The loads of the next block are issued before the butterfly of the current one: a warp with several blocks
    // (K = 9728: 38 blocks per row) pays the L2 round trip once, not once per block.
    if (pre_blob != nullptr) {
        // the pre-pass kernel has done the work: copy codes, scales and group sums (one contiguous area, same layout)
        const int n16 = (L.off_part - L.off_afrag) >> 4;
        const uint4 *src = reinterpret_cast<const uint4 *>(pre_blob);
        uint4 *dst = reinterpret_cast<uint4 *>(smem + L.off_afrag);
        for (int i = threadIdx.x; i < n16; i += blockDim.x) dst[i] = __ldcg(src + i);
    } else {
        float vn[8];
        if (perm_first) {
            const float *xr = x + (size_t)(warp / KT) * K;
            vn[0] = __ldcg(xr + pp0.x); vn[1] = __ldcg(xr + pp0.y); vn[2] = __ldcg(xr + pp0.z); vn[3] = __ldcg(xr + pp0.w);
            vn[4] = __ldcg(xr + pp1.x); vn[5] = __ldcg(xr + pp1.y); vn[6] = __ldcg(xr + pp1.z); vn[7] = __ldcg(xr + pp1.w);
        } else if (warp < Ma * KT) {
            rrs_load_block_cg(x + (size_t)(warp / KT) * K, perm, warp % KT, lane, vn);
        }
        for (int b = warp; b < Ma * KT; b += NW) {
            const int m = b / KT, kt = b - m * KT;
            float v[8];
#pragma unroll
            for (int i = 0; i < 8; i++) v[i] = vn[i];
            if (b + NW < Ma * KT) rrs_load_block_cg(x + (size_t)((b + NW) / KT) * K, perm, (b + NW) % KT, lane, vn);
            if (b == warp) GV_STAMP_DEP(8, v[0] + v[7]);
            const ActQuant a = rrs_quantize_block(v, lane);            // lane = (group r, k-quarter q), elements lane*8..+7
            if (b == warp) GV_STAMP_DEP(9, a.max_abs + (float)(a.q[7] + a.gsum));
            gv_store_act_block(a, smem + L.off_afrag, L.off_rm - L.off_afrag, L.off_sumq - L.off_afrag, Ma, KT, m, kt, lane,
                               MT > 0);
            if (b == warp) GV_STAMP(6);
        }
    }
    __syncthreads();
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
    GV_STAMP(3);

    // Fragment roles (mma.m16n8k32, PTX ISA "Matrix Fragments for mma.m16n8k32"): A (16 x 32, row) = the tile's
    // 8 channels in rows 0..7 -- lane (c = lane/4, s = lane%4) holds a0 = word (c, s) even nibbles, a2 = odd nibbles,
    // rows 8..15 (a1, a3) are zero; B (32 x 8, col) = up to 8 activation rows -- lane (n = lane/4, s) holds the codes
    // of token n, slice s (b0 = even positions, b1 = odd).  D: lane (c, s) gets tokens 2s, 2s+1 of channel c.
    // Lanes whose token does not exist read the last valid row instead (their D columns are never used).
    const int c = r;                                          // channel of this lane's D fragment
    const int tk0 = 2 * q, tk1 = 2 * q + 1;                   // tokens of this lane's D fragment (block mb adds 8)
    const int brow0 = min(r, M - 1), brow1 = min(r + 8, M - 1);
    float acc[MB][2];
#pragma unroll
    for (int i = 0; i < MB; i++) acc[i][0] = acc[i][1] = 0.0f;
    int nseg = 0;
    int nt = R0 + lt0 / KT, kt = lt0 % KT;
    bool from_row_start = (kt == 0);
    const uint32_t part0 = dsmem_addr(part, 0), prow0 = dsmem_addr(prow, 0), cbar0 = dsmem_addr(cbar, 0);    // CTA 0's copies
    const uint32_t ring_u32 = smem_u32(myring), full_u32 = smem_u32(myfull);
    const uint2 *af0 = afrag + brow0 * 4 + q;                  // + ((kt*8 + g) * M) * 4
    const uint2 *af1 = afrag + brow1 * 4 + q;
    const int af_g = M * 4;                                    // uint2 stride between groups

    if constexpr (MT > 0) {
    // ---- dp4a path (M = MT <= 3) ---------------------------------------------------------------------------
    // With one to three tokens the m16n8k32 shape is mostly padding and the loop is bound by instruction issue, so the
    // lanes are re-mapped for wide loads instead: lane l = (gq = l / 8, cc = l % 8) reads words 4l .. 4l+3 of the tile's
    // first and second half with two LDS.128 = ALL 32 nibbles of channel cc in groups gq and gq + 4.  Per word: three
    // logic ops (even / odd nibbles as 16*q in s8 lanes) and two DP4A per token against the token's codes (LDS.128,
    // broadcast within a group).  The two group dots (16*dot fits s16) meet their scale codes in one IDP.2A; two
    // shuffles add the four group pairs: isum = sum_g sc_g * dot_g, the same int32 as the IMMA path and the GEMMs.
    const int cc = lane & 7, gq = lane >> 3;
    const uint32_t scsel = (uint32_t)gq | ((uint32_t)(4 + gq) << 4);      // PRMT: byte 0 = sc[gq], byte 1 = sc[gq + 4]
    const uint4 *afq = reinterpret_cast<const uint4 *>(afrag);           // [(kt*8 + g) * M + m][2] = {ev[4]}, {od[4]}
    float accd[MT];
#pragma unroll
    for (int m = 0; m < MT; m++) accd[m] = 0.0f;
    int slot = 0;
    uint32_t phase = 0;
    const int MS = GROUP ? Ma : MT;                            // activation rows in the shared-memory layout
    int m0 = (GROUP && Ma > 1) ? nt / ex.nt_expert : 0;       // GROUP: this channel block's own activation row
    for (int i = 0; i < cnt; i++) {
        GV_ITER(i, 0);
        mbar_wait(&myfull[slot], phase);
        GV_ITER(i, 1);
        const uint8_t *tp = myring + slot * TILE_BYTES;
        const uint4 w0 = *reinterpret_cast<const uint4 *>(tp + lane * 16);
        const uint4 w1 = *reinterpret_cast<const uint4 *>(tp + 512 + lane * 16);
        const uint2 scv = *reinterpret_cast<const uint2 *>(tp + TILE_OFF_SC + 8 * cc);
        const unsigned short sh = *reinterpret_cast<const unsigned short *>(tp + TILE_OFF_S + 2 * cc);
        const unsigned short zh = *reinterpret_cast<const unsigned short *>(tp + TILE_OFF_Z + 2 * cc);
        const bool anyz = __any_sync(0xffffffffu, (zh & 0x7FFFu) != 0u);
        uint2 zcv = make_uint2(0, 0);
        if (anyz) zcv = *reinterpret_cast<const uint2 *>(tp + TILE_OFF_ZC + 8 * cc);
        __syncwarp();
        if (i + depth < cnt) {
            if (elect_one()) {
                const uint32_t bar = full_u32 + slot * 8, dst = ring_u32 + slot * TILE_BYTES;
                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(TILE_BYTES) : "memory");
                asm volatile(
                    "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::
                        "r"(dst), "l"(tile_src(i + depth)), "r"(TILE_BYTES), "r"(bar), "l"(pol)
                    : "memory");
            }
        }
        const uint32_t ww[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
        uint32_t ae[8], ao[8];
#pragma unroll
        for (int j = 0; j < 8; j++) {
            ae[j] = (ww[j] << 4) & 0xF0F0F0F0u;
            ao[j] = ww[j] & 0xF0F0F0F0u;
        }
        const int scp = (int)__byte_perm(scv.x, scv.y, scsel);
        const float S = __half2float(__ushort_as_half(sh));
#pragma unroll
        for (int m = 0; m < MT; m++) {
            const uint4 *a0p = afq + ((kt * 8 + gq) * MS + m0 + m) * 2;
            const uint4 *a1p = afq + ((kt * 8 + gq + 4) * MS + m0 + m) * 2;
            const uint4 e0 = a0p[0], o0 = a0p[1], e1 = a1p[0], o1 = a1p[1];
            int d0 = __dp4a((int)ae[0], (int)e0.x, 0);
            d0 = __dp4a((int)ao[0], (int)o0.x, d0);
            d0 = __dp4a((int)ae[1], (int)e0.y, d0);
            d0 = __dp4a((int)ao[1], (int)o0.y, d0);
            d0 = __dp4a((int)ae[2], (int)e0.z, d0);
            d0 = __dp4a((int)ao[2], (int)o0.z, d0);
            d0 = __dp4a((int)ae[3], (int)e0.w, d0);
            d0 = __dp4a((int)ao[3], (int)o0.w, d0);
            int d1 = __dp4a((int)ae[4], (int)e1.x, 0);
            d1 = __dp4a((int)ao[4], (int)o1.x, d1);
            d1 = __dp4a((int)ae[5], (int)e1.y, d1);
            d1 = __dp4a((int)ao[5], (int)o1.y, d1);
            d1 = __dp4a((int)ae[6], (int)e1.z, d1);
            d1 = __dp4a((int)ao[6], (int)o1.z, d1);
            d1 = __dp4a((int)ae[7], (int)e1.w, d1);
            d1 = __dp4a((int)ao[7], (int)o1.w, d1);
            int is = __dp2a_lo((int)__byte_perm((uint32_t)d0, (uint32_t)d1, 0x5410), scp, 0);
            is += __shfl_xor_sync(0xffffffffu, is, 8);
            is += __shfl_xor_sync(0xffffffffu, is, 16);
            const int isv = is >> 4;                                 // weights were carried as 16*q: exact
            accd[m] = __fmaf_rn((float)isv * rms[kt * MS + m0 + m], S, accd[m]);
            if (anyz) accd[m] += gv_zero_point(zh, zcv, sumq + ((m0 + m) * KT + kt) * 8, a7s[kt * MS + m0 + m]);
            if (DEBUG && dbg_isum && lane < 8) dbg_isum[((size_t)m * N + nt * 8 + cc) * KT + kt] = isv;
        }
        GV_ITER(i, 2);
        if (++slot == depth) { slot = 0; phase ^= 1u; }

        const bool row_end = (kt == KT - 1);
        if (row_end || i == cnt - 1) {
            if (row_end && from_row_start) {            // the whole row was this warp's: lanes 0..7 hold its 8 channels
                if (lane < 8) {
#pragma unroll
                    for (int m = 0; m < MT; m++) {
                        store_out(m, nt, cc, accd[m]);
                    }
                }
            } else {                                    // a cut row: park the partial in CTA 0's shared memory
                const int e = wid * 2 + nseg;
                if (lane < 8) {
#pragma unroll
                    for (int m = 0; m < MT; m++)
                        dsmem_st_async(part0 + (uint32_t)(((e * MT + m) * 8 + cc) * 4), __float_as_uint(accd[m]), cbar0);
                }
                if (lane == 0) dsmem_st_async(prow0 + (uint32_t)(e * 4), (uint32_t)nt, cbar0);
                nseg++;
            }
#pragma unroll
            for (int m = 0; m < MT; m++) accd[m] = 0.0f;
        }
        if (row_end) {
            kt = 0;
            nt++;
            if (GROUP && Ma > 1) m0 = nt / ex.nt_expert;
            from_row_start = true;
        } else {
            kt++;
        }
    }
    for (; nseg < 2; nseg++) {                          // unused entry slots are sent too: fixed byte total
        const int e = wid * 2 + nseg;
        if (lane < 8) {
#pragma unroll
            for (int m = 0; m < MT; m++) dsmem_st_async(part0 + (uint32_t)(((e * MT + m) * 8 + cc) * 4), 0u, cbar0);
        }
        if (lane == 0) dsmem_st_async(prow0 + (uint32_t)(e * 4), 0xFFFFFFFFu, cbar0);
    }
    } else {
    int slot = 0;
    uint32_t phase = 0;
    for (int i = 0; i < cnt; i++) {
        GV_ITER(i, 0);
        mbar_wait(&myfull[slot], phase);
        GV_ITER(i, 1);

        const uint8_t *tp = myring + slot * TILE_BYTES;
        uint32_t wv[8];
#pragma unroll
        for (int g = 0; g < 8; g++) wv[g] = reinterpret_cast<const uint32_t *>(tp)[g * 32 + lane];
        const uint2 scv = *reinterpret_cast<const uint2 *>(tp + TILE_OFF_SC + 8 * c);      // sc[c][0..7]
        const unsigned short sh = *reinterpret_cast<const unsigned short *>(tp + TILE_OFF_S + 2 * c);
        const unsigned short zh = *reinterpret_cast<const unsigned short *>(tp + TILE_OFF_Z + 2 * c);
        const bool anyz = __any_sync(0xffffffffu, (zh & 0x7FFFu) != 0u);
        uint2 zcv = make_uint2(0, 0);
        if (anyz) zcv = *reinterpret_cast<const uint2 *>(tp + TILE_OFF_ZC + 8 * c);
        __syncwarp();
        if (i + depth < cnt) {                  // refill the slot just drained (warp-uniform operands, one issuer)
            if (elect_one()) {
                const uint32_t bar = full_u32 + slot * 8, dst = ring_u32 + slot * TILE_BYTES;
                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(TILE_BYTES) : "memory");
                asm volatile(
                    "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::
                        "r"(dst), "l"(tile_src(i + depth)), "r"(TILE_BYTES), "r"(bar), "l"(pol)
                    : "memory");
            }
        }

        // one mma per 32-group (per 8 activation rows); group pairs are folded straight into
        // isum = sum_g sc_g * dot_g (exact): the two 16*dot values (each fits s16) are packed by one PRMT and hit
        // two scale bytes with one IDP.2A
        int is[MB][2];
#pragma unroll
        for (int mb = 0; mb < MB; mb++) is[mb][0] = is[mb][1] = 0;
        const int afk = kt * 8 * af_g;
#pragma unroll
        for (int gp = 0; gp < 4; gp++) {
            int dd[MB][2][4];
#pragma unroll
            for (int h = 0; h < 2; h++) {
                const int g = gp * 2 + h;
                const uint32_t a0 = (wv[g] << 4) & 0xF0F0F0F0u;      // nibbles 0,2,4,6 of the word as 16*q (s8)
                const uint32_t a2 = wv[g] & 0xF0F0F0F0u;             // nibbles 1,3,5,7
                const uint2 b = af0[afk + g * af_g];
                mma_s8(dd[0][h], a0, 0u, a2, 0u, b.x, b.y);
                if (MB > 1) {
                    const uint2 b2 = af1[afk + g * af_g];
                    mma_s8(dd[MB - 1][h], a0, 0u, a2, 0u, b2.x, b2.y);
                }
            }
            const uint32_t scw = gp < 2 ? scv.x : scv.y;
#pragma unroll
            for (int mb = 0; mb < MB; mb++) {
#pragma unroll
                for (int j = 0; j < 2; j++) {
                    const int p = (int)__byte_perm((uint32_t)dd[mb][0][j], (uint32_t)dd[mb][1][j], 0x5410);
                    is[mb][j] = (gp & 1) ? __dp2a_hi(p, (int)scw, is[mb][j]) : __dp2a_lo(p, (int)scw, is[mb][j]);
                }
            }
        }
        const float S = __half2float(__ushort_as_half(sh));
#pragma unroll
        for (int mb = 0; mb < MB; mb++) {
#pragma unroll
            for (int j = 0; j < 2; j++) {
                const int m = 8 * mb + tk0 + j;
                const bool valid = m < M;
                const int isv = is[mb][j] >> 4;                      // weights were carried as 16*q: exact
                const float rm = valid ? rms[kt * M + m] : 0.0f;
                acc[mb][j] = __fmaf_rn((float)isv * rm, S, acc[mb][j]);
                if (anyz && valid) acc[mb][j] += gv_zero_point(zh, zcv, sumq + (m * KT + kt) * 8, a7s[kt * M + m]);
                if (DEBUG && dbg_isum && valid) dbg_isum[((size_t)m * N + nt * 8 + c) * KT + kt] = isv;
            }
        }
        GV_ITER(i, 2);
        if (++slot == depth) { slot = 0; phase ^= 1u; }

        // end of this warp's share of row nt?
        const bool row_end = (kt == KT - 1);
        if (row_end || i == cnt - 1) {
            if (row_end && from_row_start) {            // the whole row was this warp's
#pragma unroll
                for (int mb = 0; mb < MB; mb++)
#pragma unroll
                    for (int j = 0; j < 2; j++)
                        if (8 * mb + tk0 + j < M) store_out(8 * mb + tk0 + j, nt, c, acc[mb][j]);
            } else {                                    // a cut row: park the partial in CTA 0's shared memory
                const int e = wid * 2 + nseg;
#pragma unroll
                for (int mb = 0; mb < MB; mb++)
#pragma unroll
                    for (int j = 0; j < 2; j++)
                        if (8 * mb + tk0 + j < M)
                            dsmem_st_async(part0 + (uint32_t)(((e * M + 8 * mb + tk0 + j) * 8 + c) * 4),
                                           __float_as_uint(acc[mb][j]), cbar0);
                if (lane == 0) dsmem_st_async(prow0 + (uint32_t)(e * 4), (uint32_t)nt, cbar0);
                nseg++;
            }
#pragma unroll
            for (int mb = 0; mb < MB; mb++) acc[mb][0] = acc[mb][1] = 0.0f;
        }
        if (row_end) {
            kt = 0;
            nt++;
            from_row_start = true;
        } else {
            kt++;
        }
    }
    (void)tk1;
    // unused entry slots are sent too (row = -1, zeros): the byte total CTA 0 waits for is fixed
    for (; nseg < 2; nseg++) {
        const int e = wid * 2 + nseg;
#pragma unroll
        for (int mb = 0; mb < MB; mb++)
#pragma unroll
            for (int j = 0; j < 2; j++)
                if (8 * mb + tk0 + j < M) dsmem_st_async(part0 + (uint32_t)(((e * M + 8 * mb + tk0 + j) * 8 + c) * 4), 0u, cbar0);
        if (lane == 0) dsmem_st_async(prow0 + (uint32_t)(e * 4), 0xFFFFFFFFu, cbar0);
    }
    }   // IMMA path
    GV_STAMP(4);
    if (crank != 0) return;      // nothing reads this CTA's shared memory: CTA 1 is done
    mbar_wait(cbar, 0u);         // all partials (local and remote) have landed in CTA 0
    GV_STAMP(7);

    // ---- CTA 0 combines the cut rows in k order (entry order = worker order = tile order) ----------------
    if (crank == 0 && M <= 2 && wsp.base >= 2) {
        // Short runs (every warp holds at least two tiles, so a row is cut into two or three pieces): the warp's four
        // entries are handled AT ONCE, one per lane quarter, each summed in entry order -- a quarter of the latency of
        // walking the four entries one after the other (0.8 -> ~0.3 us at the end of every decode launch).
        const int e = warp * 4 + (lane >> 3), c8 = lane & 7;
        const int row = prow[e];
        bool first = row >= 0;
        if (first)
            for (int j = e - 1; j >= 0; j--) {
                const int pr = prow[j];
                if (pr >= 0) { first = (pr != row); break; }
            }
        if (first) {
            for (int m = 0; m < M; m++) {
                float sum = 0.0f;
                for (int j = e; j < entries; j++) {
                    const int pr = prow[j];
                    if (pr < 0) continue;
                    if (pr != row) break;
                    sum += part[(size_t)(j * M + m) * 8 + c8];
                }
                store_out(m, row, c8, sum);
            }
        }
    } else if (crank == 0) {
        for (int sidx = 0; sidx < 4; sidx++) {                  // entries / NW = 4 slots per warp
            const int e = warp * 4 + sidx;
            const int row = prow[e];
            if (row < 0) continue;
            int prev = -1;
            for (int j = e - 1; j >= 0; j--) {
                const int pr = prow[j];
                if (pr >= 0) { prev = pr; break; }
            }
            if (prev == row) continue;                         // an earlier entry owns this row's run
            if (M <= 2) {
                // One or two tokens leave 3/4 of the lanes idle in the loop below and a run of ten or more entries
                // (N=1024, K=2560: every warp holds one tile) costs ~60 cycles per entry.  Here the lanes look at 32
                // entries at once (two ballots give the run), lane quarter jj sums every 4th member and two shuffles
                // add the quarters: a fixed association, so the result stays deterministic.
                const int jj = lane >> 3, c8 = lane & 7;
                float s2[2] = {0.0f, 0.0f};
                int nth = 0;
                for (int base = e;; base += 32) {
                    const int idx = base + lane;
                    const int pr = idx < entries ? prow[idx] : 0x7fffffff;      // past the end: ends the run
                    const unsigned stopm = __ballot_sync(0xffffffffu, pr >= 0 && pr != row);
                    const int lim = stopm ? __ffs((int)stopm) - 1 : 32;
                    unsigned same = __ballot_sync(0xffffffffu, pr == row) & (lim == 32 ? 0xffffffffu : ((1u << lim) - 1u));
                    while (same) {
                        const int i = __ffs((int)same) - 1;
                        same &= same - 1u;
                        if ((nth & 3) == jj) {
                            s2[0] += part[(size_t)((base + i) * M) * 8 + c8];
                            if (M == 2) s2[1] += part[(size_t)((base + i) * M + 1) * 8 + c8];
                        }
                        nth++;
                    }
                    if (lim < 32) break;
                }
                for (int m = 0; m < M; m++) {
                    float v = s2[m];
                    v += __shfl_xor_sync(0xffffffffu, v, 8);
                    v += __shfl_xor_sync(0xffffffffu, v, 16);
                    if (jj == 0) store_out(m, row, c8, v);
                }
                continue;
            }
            // lane -> (token m = lane / 8 (+4 per pass), channel lane % 8): coalesced 32-byte row pieces of y
            for (int m = lane >> 3; m < M; m += 4) {
                float sum = 0.0f;
                for (int j = e; j < entries; j++) {
                    const int pr = prow[j];
                    if (pr < 0) continue;
                    if (pr != row) break;
                    sum += part[(size_t)(j * M + m) * 8 + (lane & 7)];
                }
                store_out(m, row, lane & 7, sum);
            }
        }
    }
    GV_STAMP(5);
}

static int gemv_mb(int64_t M) { return M <= 8 ? 1 : 2; }

bool gemv_plan(int64_t M, int64_t N, int64_t K, int num_sms, size_t smem_optin, GemvPlan *plan) {
    if (M < 1 || M > GEMV_MAX_M) return false;
    const int64_t NT = N / TILE_CH, KT = K / TILE_K;
    if (NT * KT >= (int64_t)1 << 30) return false;                 // 32-bit tile arithmetic in the kernel
    // 16 warps.  (32 warps x 3 tiles -- every tile of the decode shapes requested by the prologue -- measured 20 %
    // SLOWER, 12.1 vs 10.0 us at N=9728, K=2560: the loop is bound by instruction issue, not by latency.)
    // Ring depth = the most tiles any warp will see, at most 3.  Deeper rings measured slower once launches overlap
    // (N=9728, K=2560, M=1, graph of 24 launches: 7.0 us per launch with 2 or 3 slots, 7.4 with 4, 7.8 with 5): the
    // prefetch burst of the next launch shares the SM's return path with the refills of the running one, and
    // two resident CTAs must share the 228 KB.
    const int nw = 16;
    int64_t clusters_ = num_sms / GV_CLUSTER;
    if (NT < clusters_) clusters_ = NT;
    if (clusters_ < 1) clusters_ = 1;
    const int64_t rows_max = (NT + clusters_ - 1) / clusters_;
    const int64_t per_warp = (rows_max * KT + nw * GV_CLUSTER - 1) / (nw * GV_CLUSTER);
    int depth = per_warp < 2 ? 2 : (per_warp > 3 ? 3 : (int)per_warp);
    static const char *denv = getenv("TCQ4_GV_DEPTH");
    if (denv && atoi(denv) >= 2 && atoi(denv) <= 8) depth = atoi(denv);
    while (depth > 2 && (size_t)GvSmem((int)M, (int)KT, nw, depth).total > smem_optin) depth--;
    const GvSmem L((int)M, (int)KT, nw, depth);
    if ((size_t)L.total > smem_optin) return false;
    plan->warps = nw;
    plan->depth = depth;
    static const bool nopre = getenv("TCQ4_GV_NOPRE") != nullptr;
    // M = 1 keeps the stage inside the kernel: consecutive M = 1 launches share an SM (dependent launch), and a kernel
    // between them costs more than the 2-4 rounds it saves (K = 14336: 14.1 us inside, 15.6 us with the pre-pass)
    static const int pre_b = getenv("TCQ4_GV_PRE_BLOCKS") ? atoi(getenv("TCQ4_GV_PRE_BLOCKS")) : GV_PREPASS_BLOCKS;      // experiments
    plan->prepass = !nopre && M >= 2 && M * KT > pre_b;
    // with a channel permutation every block is a gather (two dependent loads per element): there the pre-pass pays from 17
    // blocks, M = 1 included (Llama-3-8B layer with imatrix perms, graph: M = 1 61.2 -> 56.8 us, M = 2 66.1 -> 55.4 us)
    plan->prepass_perm = !nopre && M * KT > GV_PREPASS_PERM_BLOCKS;
    plan->blob_bytes = (size_t)(L.off_part - L.off_afrag);
    int64_t clusters = num_sms / GV_CLUSTER;
    if (NT < clusters) clusters = NT;                              // a cluster owns whole rows
    if (clusters < 1) clusters = 1;
    plan->grid = (int)(clusters * GV_CLUSTER);
    plan->smem_bytes = (size_t)L.total;
    plan->maxslots = 0;
    plan->partial_bytes = 0;                                       // no global scratch: rows never leave a cluster
    plan->ticket_bytes = 0;
    return true;
}

template <int MB, int MT, bool DEBUG>
static cudaError_t launch_gemv_t(const void *w, const float *x, const int32_t *perm, YOut y, int64_t M, int64_t N,
                                 int64_t K, const GemvPlan &plan, const uint8_t *pre_blob, int32_t *dbg,
                                 cudaStream_t stream, const GvExpert *expert) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)plan.grid);
    cfg.blockDim = dim3((unsigned)plan.warps * 32);
    cfg.dynamicSmemBytes = plan.smem_bytes;
    cfg.stream = stream;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = GV_CLUSTER;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    // Programmatic dependent launch (TCQ4_NO_PDL=1 turns it off).  It pays when this launch and the next one fit one SM
    // together (M = 1: 64 registers, ring of <= 5 tiles): the next MUL_MAT's barrier setup and weight prefetch run under
    // this one's loop.  N=9728, K=2560, M=1, 24 launches in a graph: 8.5 us per launch without, 7.0 us with.
    static const bool pdl = getenv("TCQ4_NO_PDL") == nullptr;
    // an expert id is read at kernel entry, before griddepcontrol.wait: the caller says when that is safe
    cfg.numAttrs = (pdl && (expert == nullptr || expert->pdl)) ? 2 : 1;
    const GvExpert ex = expert ? *expert : GvExpert{};
    if constexpr (!DEBUG) {
        if (ex.res != nullptr || ex.bias != nullptr)
            return cudaLaunchKernelEx(&cfg, tcq4_gemv_kernel<MB, MT, false, 3>, (const uint8_t *)w, x, perm, y, (int)M, (int)N,
                                      (int)K, plan.depth, pre_blob, dbg, ex);
        if (ex.nseg > 0)
            return cudaLaunchKernelEx(&cfg, tcq4_gemv_kernel<MB, MT, false, 2>, (const uint8_t *)w, x, perm, y, (int)M, (int)N,
                                      (int)K, plan.depth, pre_blob, dbg, ex);
    }
    return cudaLaunchKernelEx(&cfg, tcq4_gemv_kernel<MB, MT, DEBUG>, (const uint8_t *)w, x, perm, y, (int)M, (int)N, (int)K,
                              plan.depth, pre_blob, dbg, ex);
}

template <int MB, int MT>
static cudaError_t gemv_attr_one(size_t smem_optin) {
    cudaError_t e = cudaFuncSetAttribute(tcq4_gemv_kernel<MB, MT, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)smem_optin);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(tcq4_gemv_kernel<MB, MT, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_optin);
    if (e != cudaSuccess) return e;
    // One shared-memory carve-out for every launch: matmuls of different K ask for different amounts of shared memory,
    // and an SM cannot host the next launch (programmatic dependent launch) while its L1/shared split must change.
    e = cudaFuncSetAttribute(tcq4_gemv_kernel<MB, MT, true>, cudaFuncAttributePreferredSharedMemoryCarveout,
                             (int)cudaSharedmemCarveoutMaxShared);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(tcq4_gemv_kernel<MB, MT, false>, cudaFuncAttributePreferredSharedMemoryCarveout,
                             (int)cudaSharedmemCarveoutMaxShared);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(tcq4_gemv_kernel<MB, MT, false, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_optin);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(tcq4_gemv_kernel<MB, MT, false, 2>, cudaFuncAttributePreferredSharedMemoryCarveout,
                             (int)cudaSharedmemCarveoutMaxShared);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(tcq4_gemv_kernel<MB, MT, false, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_optin);
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(tcq4_gemv_kernel<MB, MT, false, 3>, cudaFuncAttributePreferredSharedMemoryCarveout,
                                (int)cudaSharedmemCarveoutMaxShared);
}

cudaError_t gemv_init_attributes(size_t smem_optin) {
    cudaError_t e = cudaFuncSetAttribute(tcq4_gemv_prepass_kernel, cudaFuncAttributePreferredSharedMemoryCarveout,
                                         (int)cudaSharedmemCarveoutMaxShared);      // same split as the GEMV behind it
    if (e != cudaSuccess) return e;
    e = gemv_attr_one<1, 0>(smem_optin);
    if (e == cudaSuccess) e = gemv_attr_one<2, 0>(smem_optin);
    if (e == cudaSuccess) e = gemv_attr_one<1, 1>(smem_optin);
    if (e == cudaSuccess) e = gemv_attr_one<1, 2>(smem_optin);
    if (e == cudaSuccess) e = gemv_attr_one<1, 3>(smem_optin);
    if (e == cudaSuccess)
        e = cudaFuncSetAttribute(tcq4_gemv_kernel<1, 1, false, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)smem_optin);
    if (e == cudaSuccess)
        e = cudaFuncSetAttribute(tcq4_gemv_kernel<1, 1, false, 1>, cudaFuncAttributePreferredSharedMemoryCarveout,
                                 (int)cudaSharedmemCarveoutMaxShared);
    return e;
}

// MUL_MAT_ID for one token as ONE launch: a GEMV over the virtual matrix of n_used * N channels whose channel blocks
// take their weight tiles from the expert the ids name (GvExpert, tcq4_kernels.h).  The ids are read behind
// griddepcontrol.wait (they may be the previous kernel's output), so only the set-up overlaps the previous launch.
cudaError_t launch_gemv_grouped(const void *w_experts, const float *x, const int32_t *perm, float *y, int64_t n_used,
                                int64_t N, int64_t K, const GemvPlan &plan, const GvExpert &expert, cudaStream_t stream) {
    if (expert.sel == nullptr || expert.xslots < 1 || expert.nt_expert != (int)(N / TILE_CH)) return cudaErrorInvalidValue;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)plan.grid);
    cfg.blockDim = dim3((unsigned)plan.warps * 32);
    cfg.dynamicSmemBytes = plan.smem_bytes;
    cfg.stream = stream;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = GV_CLUSTER;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;      // safe: the ids are read behind griddepcontrol.wait
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    static const bool pdl = getenv("TCQ4_NO_PDL") == nullptr;
    cfg.numAttrs = pdl ? 2 : 1;
    YOut yo;
    for (int i = 0; i < MAX_PEERS; i++) yo.p[i] = y;
    yo.n = 1;
    yo.ldc = n_used * N;
    yo.mc = nullptr;
    return cudaLaunchKernelEx(&cfg, tcq4_gemv_kernel<1, 1, false, 1>, (const uint8_t *)w_experts, x, perm, yo, 1,
                              (int)(n_used * N), (int)K, plan.depth, (const uint8_t *)nullptr, (int32_t *)nullptr, expert);
}

bool gemv_uses_prepass(const GemvPlan &plan, const int32_t *perm) { return plan.prepass || (perm != nullptr && plan.prepass_perm); }

template <int MB, int MT>
static cudaError_t launch_gemv_d(bool debug, const void *w, const float *x, const int32_t *perm, YOut y, int64_t M, int64_t N,
                                 int64_t K, const GemvPlan &plan, uint8_t *blob, int32_t *dbg, cudaStream_t stream,
                                 const GvExpert *expert) {
    const uint8_t *pre = nullptr;
    if (gemv_uses_prepass(plan, perm) && blob != nullptr) {
        // activation stage once for the whole GPU (see tcq4_gemv_prepass_kernel), in the GEMV's own layout
        const GvSmem L((int)M, (int)(K / TILE_K), plan.warps, plan.depth);
        cudaLaunchConfig_t cfg = {};
        const int nblk = (int)(M * (K / TILE_K));
        cfg.gridDim = dim3((unsigned)((nblk + GVP_WARPS - 1) / GVP_WARPS));
        cfg.blockDim = dim3(GVP_WARPS * 32);
        cfg.stream = stream;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = attr;
        static const bool pdl = getenv("TCQ4_NO_PDL") == nullptr;
        cfg.numAttrs = pdl ? 1 : 0;
        cudaError_t e = cudaLaunchKernelEx(&cfg, tcq4_gemv_prepass_kernel, x, perm, blob, (int)M, (int)K,
                                           L.off_rm - L.off_afrag, L.off_sumq - L.off_afrag, MT > 0 ? 1 : 0);
        if (e != cudaSuccess) return e;
        pre = blob;
    }
    // the instrumented instantiation exists for the plain kernel only: a fan-out (nseg) or bias / residual launch always
    // takes its own non-debug instantiation, whatever the diagnostics switches say
    if (expert && (expert->nseg > 0 || expert->res != nullptr || expert->bias != nullptr)) {
        if (dbg != nullptr) return cudaErrorNotSupported;
        debug = false;
    }
    if (debug) return launch_gemv_t<MB, MT, true>(w, x, perm, y, M, N, K, plan, pre, dbg, stream, expert);
    return launch_gemv_t<MB, MT, false>(w, x, perm, y, M, N, K, plan, pre, nullptr, stream, expert);
}

cudaError_t launch_gemv(const void *w_tiles, const float *x, const int32_t *perm, YOut y, int64_t M, int64_t N,
                        int64_t K, const GemvPlan &plan, float *ws_partial, int *ws_ticket, int32_t *dbg_isum,
                        cudaStream_t stream, const GvExpert *expert) {
    uint8_t *blob = reinterpret_cast<uint8_t *>(ws_partial);      // pre-pass output (plan.blob_bytes), may be null
    (void)ws_ticket;
    if (M < 1 || M > GEMV_MAX_M) return cudaErrorInvalidValue;
#ifdef TCQ4_DIAG
    static const bool prof = getenv("TCQ4_GV_PROF") != nullptr;      // phase stamps without the isum dump (diagnostics build)
#else
    const bool prof = false;
#endif
    static const bool nodp = getenv("TCQ4_GV_NODP") != nullptr;      // diagnostics: IMMA loop for every M
    const bool dbg = dbg_isum != nullptr || prof;
    if (!nodp) {
        switch (M) {
            case 1: return launch_gemv_d<1, 1>(dbg, w_tiles, x, perm, y, M, N, K, plan, blob, dbg_isum, stream, expert);
            case 2: return launch_gemv_d<1, 2>(dbg, w_tiles, x, perm, y, M, N, K, plan, blob, dbg_isum, stream, expert);
            case 3: {      // round 2: with the pre-pass in front the IMMA loop is 3-11 % ahead at M = 3 on every shape measured
                static const bool dp3 = getenv("TCQ4_GV_DP3") != nullptr;      // (tests/gpu_g4_vs_gemv.py 3); the dp4a form stays selectable
                if (dp3) return launch_gemv_d<1, 3>(dbg, w_tiles, x, perm, y, M, N, K, plan, blob, dbg_isum, stream, expert);
                break;
            }
            default: break;
        }
    }
    if (gemv_mb(M) == 1) return launch_gemv_d<1, 0>(dbg, w_tiles, x, perm, y, M, N, K, plan, blob, dbg_isum, stream, expert);
    return launch_gemv_d<2, 0>(dbg, w_tiles, x, perm, y, M, N, K, plan, blob, dbg_isum, stream, expert);
}

cudaError_t gemv_read_prof(unsigned long long *host_out, size_t bytes) {
    const size_t a = sizeof(g_gv_prof), b = sizeof(g_gv_iter);
    cudaError_t e = cudaMemcpyFromSymbol(host_out, g_gv_prof, bytes < a ? bytes : a);
    if (e == cudaSuccess && bytes >= a + b)
        e = cudaMemcpyFromSymbol(reinterpret_cast<uint8_t *>(host_out) + a, g_gv_iter, b);
    return e;
}

}  // namespace tcq4
