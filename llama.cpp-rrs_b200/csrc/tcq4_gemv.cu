// tcq4_gemv.cu -- K3: fused decode GEMV/GEMM for M <= 4 (bandwidth-bound).
//
// One persistent CTA per SM.  The flattened tile list [N/8][K/256] is exactly memory order
// (/root/reference/ggml/src/ggml-quants.c:1550-1573), so it is cut into W = grid*8 contiguous byte
// ranges, one per warp; each warp streams its range through a private ring of 1-D bulk async copies
// (cp.async.bulk -> UBLKCP, 1184 B per tile, GV_DEPTH tiles in flight) and never synchronises with
// another warp inside the loop.  While the first tiles are in flight the CTA runs the activation
// stage (FWHT-256 + quantise, tcq4_common.cuh) for all M rows into shared memory -- 148x redundant
// instead of the reference's ceil(N/32)*M x (tcq4_k32.cu:768-831), and off the critical path.
//
// Per tile a lane owns two 16-byte chunks = (group g, channel c) and (g+4, c): 32 nibbles each,
// expanded to s8*16 with two logic ops per word ((w<<4)&0xF0F0F0F0, w&0xF0F0F0F0) and fed to dp4a
// against the s8 codes (stored de-interleaved even/odd per 8-slice to match).  The fp32 epilogue
// reproduces tcq4_k32.cu:841-883 operation for operation, including the 4,2,1 xor tree over
// groups, so whenever one warp owns a whole output row the result is bit-identical to the oracle.
//
// Rows cut by a warp boundary are finished deterministically: every contributing warp writes its
// partial to a workspace slot and takes a ticket; the last one sums the slots in k order.
#include "tcq4_common.cuh"
#include "tcq4_kernels.h"

namespace tcq4 {

constexpr int GV_WARPS = 8;
constexpr int GV_DEPTH = 8;
constexpr int GV_THREADS = GV_WARPS * 32;

struct GvSplit {
    long long T, W, base, rem;
    __host__ __device__ long long start(long long gw) const { return gw * base + (gw < rem ? gw : rem); }
    __host__ __device__ long long count(long long gw) const { return base + (gw < rem ? 1 : 0); }
    __host__ __device__ long long owner(long long t) const {
        const long long cut = rem * (base + 1);
        if (t < cut) return t / (base + 1);
        return rem + (t - cut) / base;     // base >= 1 here: t >= cut can only happen when base > 0
    }
};

template <int MR, bool DEBUG>
__global__ void __launch_bounds__(GV_THREADS, 1)
tcq4_gemv_kernel(const uint8_t *__restrict__ w, const float *__restrict__ x, const int32_t *__restrict__ perm,
                 float *__restrict__ y, int N, int K, float *__restrict__ ws_partial, int *__restrict__ ws_ticket,
                 int maxslots, int32_t *__restrict__ dbg_isum) {
    extern __shared__ __align__(16) uint8_t smem[];
    const int KT = K / TILE_K;
    const int NT = N / TILE_CH;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    // smem carve-up
    uint8_t *ring = smem;                                                  // [GV_WARPS][GV_DEPTH][1184]
    uint64_t *full = reinterpret_cast<uint64_t *>(ring + GV_WARPS * GV_DEPTH * TILE_BYTES);
    uint32_t *codes = reinterpret_cast<uint32_t *>(full + GV_WARPS * GV_DEPTH);   // [MR][K] bytes
    float *a7s = reinterpret_cast<float *>(reinterpret_cast<uint8_t *>(codes) + (size_t)MR * K);   // [MR][KT]
    int *sumq = reinterpret_cast<int *>(a7s + MR * KT);                    // [MR][KT][8]

    GvSplit sp;
    sp.T = (long long)NT * KT;
    sp.W = (long long)gridDim.x * GV_WARPS;
    sp.base = sp.T / sp.W;
    sp.rem = sp.T % sp.W;
    const long long gw = (long long)blockIdx.x * GV_WARPS + warp;
    const long long t0 = sp.start(gw);
    const int cnt = (int)sp.count(gw);

    if (threadIdx.x == 0) {
        for (int i = 0; i < GV_WARPS * GV_DEPTH; i++) mbar_init(&full[i], 1);
        fence_mbar_init();
    }
    __syncthreads();

    uint8_t *myring = ring + warp * GV_DEPTH * TILE_BYTES;
    uint64_t *myfull = full + warp * GV_DEPTH;
    const uint8_t *wsrc = w + (size_t)t0 * TILE_BYTES;
    uint64_t pol = 0;
    if (lane == 0) {
        pol = l2_policy_evict_first();
        const int pre = cnt < GV_DEPTH ? cnt : GV_DEPTH;
        for (int i = 0; i < pre; i++) {
            mbar_arrive_expect_tx(&myfull[i], TILE_BYTES);
            bulk_g2s_stream(myring + i * TILE_BYTES, wsrc + (size_t)i * TILE_BYTES, TILE_BYTES, &myfull[i], pol);
        }
    }

    // activation stage for all MR rows (weights are already in flight)
    for (int b = warp; b < MR * KT; b += GV_WARPS) {
        const int m = b / KT, kt = b - m * KT;
        float v[8];
        rrs_load_block(x + (size_t)m * K, perm, kt, lane, v);
        const ActQuant a = rrs_quantize_block(v, lane);
        const uint32_t ev = (uint32_t)(a.q[0] & 0xFF) | ((uint32_t)(a.q[2] & 0xFF) << 8) |
                            ((uint32_t)(a.q[4] & 0xFF) << 16) | ((uint32_t)(a.q[6] & 0xFF) << 24);
        const uint32_t od = (uint32_t)(a.q[1] & 0xFF) | ((uint32_t)(a.q[3] & 0xFF) << 8) |
                            ((uint32_t)(a.q[5] & 0xFF) << 16) | ((uint32_t)(a.q[7] & 0xFF) << 24);
        reinterpret_cast<uint2 *>(codes)[((size_t)m * K + kt * TILE_K) / 8 + lane] = make_uint2(ev, od);
        if (lane == 0) a7s[b] = __fdiv_rn(a.max_abs, 7.0f);
        if ((lane & 3) == 0) sumq[b * 8 + (lane >> 2)] = a.gsum;
    }
    __syncthreads();

    const int c = lane & 7;
    const int g0 = lane >> 3, g1 = g0 + 4;
    float acc[MR];
#pragma unroll
    for (int m = 0; m < MR; m++) acc[m] = 0.0f;

    for (int i = 0; i < cnt; i++) {
        const long long t = t0 + i;
        const int nt = (int)(t / KT);
        const int kt = (int)(t - (long long)nt * KT);
        const int slot = i % GV_DEPTH;
        mbar_wait(&myfull[slot], (uint32_t)((i / GV_DEPTH) & 1));

        const uint8_t *tp = myring + slot * TILE_BYTES;
        const uint4 wa = *reinterpret_cast<const uint4 *>(tp + lane * 16);
        const uint4 wb = *reinterpret_cast<const uint4 *>(tp + 512 + lane * 16);
        const float S = __half2float(*reinterpret_cast<const __half *>(tp + TILE_OFF_S + 2 * c));
        const float Z = __half2float(*reinterpret_cast<const __half *>(tp + TILE_OFF_Z + 2 * c));
        const int sc0 = *reinterpret_cast<const int8_t *>(tp + TILE_OFF_SC + c * 8 + g0);
        const int sc1 = *reinterpret_cast<const int8_t *>(tp + TILE_OFF_SC + c * 8 + g1);
        const int zc0 = *reinterpret_cast<const int8_t *>(tp + TILE_OFF_ZC + c * 8 + g0);
        const int zc1 = *reinterpret_cast<const int8_t *>(tp + TILE_OFF_ZC + c * 8 + g1);
        __syncwarp();
        if (lane == 0 && i + GV_DEPTH < cnt) {     // refill the slot just drained
            mbar_arrive_expect_tx(&myfull[slot], TILE_BYTES);
            bulk_g2s_stream(myring + slot * TILE_BYTES, wsrc + (size_t)(i + GV_DEPTH) * TILE_BYTES, TILE_BYTES,
                            &myfull[slot], pol);
        }

        // b_scale = S*sc/127, b_zero = Z*zc/127   (tcq4_k32.cu:849-850), IEEE divide
        const float bs0 = __fdiv_rn(S * (float)sc0, 127.0f);
        const float bs1 = __fdiv_rn(S * (float)sc1, 127.0f);
        const float zp0 = Z * (float)zc0, zp1 = Z * (float)zc1;
        const float bz0 = (zp0 == 0.0f) ? zp0 : __fdiv_rn(zp0, 127.0f);
        const float bz1 = (zp1 == 0.0f) ? zp1 : __fdiv_rn(zp1, 127.0f);

        const uint32_t wv0[4] = {wa.x, wa.y, wa.z, wa.w};
        const uint32_t wv1[4] = {wb.x, wb.y, wb.z, wb.w};
        uint32_t lo0[4], hi0[4], lo1[4], hi1[4];
#pragma unroll
        for (int s = 0; s < 4; s++) {
            lo0[s] = (wv0[s] << 4) & 0xF0F0F0F0u;
            hi0[s] = wv0[s] & 0xF0F0F0F0u;
            lo1[s] = (wv1[s] << 4) & 0xF0F0F0F0u;
            hi1[s] = wv1[s] & 0xF0F0F0F0u;
        }

#pragma unroll
        for (int m = 0; m < MR; m++) {
            const uint4 *ap0 = reinterpret_cast<const uint4 *>(reinterpret_cast<const uint8_t *>(codes) +
                                                               (size_t)m * K + kt * TILE_K + g0 * 32);
            const uint4 *ap1 = ap0 + 8;     // +128 bytes = 4 groups
            const uint4 A0 = ap0[0], A1 = ap0[1], B0 = ap1[0], B1 = ap1[1];
            int d0 = 0, d1 = 0;
            d0 = __dp4a((int)lo0[0], (int)A0.x, d0); d0 = __dp4a((int)hi0[0], (int)A0.y, d0);
            d0 = __dp4a((int)lo0[1], (int)A0.z, d0); d0 = __dp4a((int)hi0[1], (int)A0.w, d0);
            d0 = __dp4a((int)lo0[2], (int)A1.x, d0); d0 = __dp4a((int)hi0[2], (int)A1.y, d0);
            d0 = __dp4a((int)lo0[3], (int)A1.z, d0); d0 = __dp4a((int)hi0[3], (int)A1.w, d0);
            d1 = __dp4a((int)lo1[0], (int)B0.x, d1); d1 = __dp4a((int)hi1[0], (int)B0.y, d1);
            d1 = __dp4a((int)lo1[1], (int)B0.z, d1); d1 = __dp4a((int)hi1[1], (int)B0.w, d1);
            d1 = __dp4a((int)lo1[2], (int)B1.x, d1); d1 = __dp4a((int)hi1[2], (int)B1.y, d1);
            d1 = __dp4a((int)lo1[3], (int)B1.z, d1); d1 = __dp4a((int)hi1[3], (int)B1.w, d1);
            d0 >>= 4;   // weights were carried as 16*q
            d1 >>= 4;

            const float a7 = a7s[m * KT + kt];
            const float sq0 = (float)sumq[(m * KT + kt) * 8 + g0];
            const float sq1 = (float)sumq[(m * KT + kt) * 8 + g1];
            // contrib = (float)dot*a7*b_scale + (float)sum_a*a7*b_zero   (tcq4_k32.cu:872-873)
            const float c0 = ((float)d0 * a7) * bs0 + (sq0 * a7) * bz0;
            const float c1 = ((float)d1 * a7) * bs1 + (sq1 * a7) * bz1;
            float v = c0 + c1;                                   // xor 4 : g <-> g+4
            v += __shfl_xor_sync(0xffffffffu, v, 16);            // xor 2
            v += __shfl_xor_sync(0xffffffffu, v, 8);             // xor 1
            acc[m] += v;                                         // tcq4_k32.cu:881-883

            if (DEBUG) {
                int is = sc0 * d0 + sc1 * d1;
                is += __shfl_xor_sync(0xffffffffu, is, 16);
                is += __shfl_xor_sync(0xffffffffu, is, 8);
                if (lane < 8 && dbg_isum) dbg_isum[((size_t)m * N + nt * 8 + c) * KT + kt] = is;
            }
        }

        // end of this warp's share of row nt?
        if (kt == KT - 1 || i == cnt - 1) {
            const long long row_first = (long long)nt * KT;
            const bool whole = (row_first >= t0) && (row_first + KT <= t0 + cnt);
            if (whole) {
                if (lane < 8) {
#pragma unroll
                    for (int m = 0; m < MR; m++) y[(size_t)m * N + nt * 8 + c] = acc[m];
                }
            } else {
                const long long first_owner = sp.owner(row_first);
                const int nslots = (int)(sp.owner(row_first + KT - 1) - first_owner) + 1;
                const int myslot = (int)(gw - first_owner);
                float *prow = ws_partial + (size_t)nt * maxslots * (MR * 8);
                if (lane < 8) {
#pragma unroll
                    for (int m = 0; m < MR; m++) __stcg(prow + ((size_t)myslot * MR + m) * 8 + c, acc[m]);
                }
                __threadfence();
                __syncwarp();
                int last = 0;
                if (lane == 0) last = (atomicAdd(&ws_ticket[nt], 1) == nslots - 1);
                last = __shfl_sync(0xffffffffu, last, 0);
                if (last) {
                    __threadfence();
                    if (lane < 8) {
#pragma unroll
                        for (int m = 0; m < MR; m++) {
                            float s = 0.0f;
                            for (int sl = 0; sl < nslots; sl++) s += __ldcg(prow + ((size_t)sl * MR + m) * 8 + c);
                            y[(size_t)m * N + nt * 8 + c] = s;
                        }
                    }
                    if (lane == 0) ws_ticket[nt] = 0;     // self-cleaning for the next launch
                }
            }
#pragma unroll
            for (int m = 0; m < MR; m++) acc[m] = 0.0f;
        }
    }
}

static size_t gemv_smem_bytes(int64_t M, int64_t K) {
    const int64_t KT = K / TILE_K;
    size_t s = (size_t)GV_WARPS * GV_DEPTH * TILE_BYTES;
    s += (size_t)GV_WARPS * GV_DEPTH * sizeof(uint64_t);
    s += (size_t)M * K;                 // codes
    s += (size_t)M * KT * sizeof(float);
    s += (size_t)M * KT * 8 * sizeof(int);
    return s;
}

bool gemv_plan(int64_t M, int64_t N, int64_t K, int num_sms, size_t smem_optin, GemvPlan *plan) {
    if (M < 1 || M > GEMV_MAX_M) return false;
    const size_t smem = gemv_smem_bytes(M, K);
    if (smem > smem_optin) return false;
    const int64_t NT = N / TILE_CH, KT = K / TILE_K;
    const int64_t T = NT * KT;
    int grid = num_sms;
    const int64_t need = (T + GV_WARPS - 1) / GV_WARPS;
    if (need < grid) grid = (int)(need < 1 ? 1 : need);
    plan->grid = grid;
    plan->smem_bytes = smem;
    const int64_t W = (int64_t)grid * GV_WARPS;
    const int64_t base = T / W;
    // a row of KT tiles can be cut into at most ceil(KT/base)+1 pieces (base >= 1), or KT pieces (base == 0)
    int64_t maxslots = base >= 1 ? (KT + base - 1) / base + 1 : KT;
    if (maxslots > KT) maxslots = KT;
    if (maxslots < 1) maxslots = 1;
    plan->maxslots = (int)maxslots;
    plan->partial_bytes = (size_t)round_up((int64_t)NT * maxslots * M * 8 * sizeof(float), 256);
    plan->ticket_bytes = (size_t)round_up(NT * (int64_t)sizeof(int), 256);
    return true;
}

template <int MR>
static cudaError_t launch_gemv_mr(const void *w, const float *x, const int32_t *perm, float *y, int64_t N, int64_t K,
                                  const GemvPlan &plan, float *ws_partial, int *ws_ticket, int32_t *dbg,
                                  cudaStream_t stream) {
    if (dbg) {
        auto kern = tcq4_gemv_kernel<MR, true>;
        kern<<<plan.grid, GV_THREADS, plan.smem_bytes, stream>>>((const uint8_t *)w, x, perm, y, (int)N, (int)K,
                                                                 ws_partial, ws_ticket, plan.maxslots, dbg);
    } else {
        auto kern = tcq4_gemv_kernel<MR, false>;
        kern<<<plan.grid, GV_THREADS, plan.smem_bytes, stream>>>((const uint8_t *)w, x, perm, y, (int)N, (int)K,
                                                                 ws_partial, ws_ticket, plan.maxslots, nullptr);
    }
    return cudaGetLastError();
}

template <int MR>
static cudaError_t gemv_attr_mr(size_t smem_optin) {
    cudaError_t e = cudaFuncSetAttribute(tcq4_gemv_kernel<MR, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)smem_optin);
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(tcq4_gemv_kernel<MR, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                (int)smem_optin);
}

cudaError_t gemv_init_attributes(size_t smem_optin) {
    cudaError_t e;
    if ((e = gemv_attr_mr<1>(smem_optin)) != cudaSuccess) return e;
    if ((e = gemv_attr_mr<2>(smem_optin)) != cudaSuccess) return e;
    if ((e = gemv_attr_mr<3>(smem_optin)) != cudaSuccess) return e;
    return gemv_attr_mr<4>(smem_optin);
}

cudaError_t launch_gemv(const void *w_tiles, const float *x, const int32_t *perm, float *y, int64_t M, int64_t N,
                        int64_t K, const GemvPlan &plan, float *ws_partial, int *ws_ticket, int32_t *dbg_isum,
                        cudaStream_t stream) {
    switch (M) {
        case 1: return launch_gemv_mr<1>(w_tiles, x, perm, y, N, K, plan, ws_partial, ws_ticket, dbg_isum, stream);
        case 2: return launch_gemv_mr<2>(w_tiles, x, perm, y, N, K, plan, ws_partial, ws_ticket, dbg_isum, stream);
        case 3: return launch_gemv_mr<3>(w_tiles, x, perm, y, N, K, plan, ws_partial, ws_ticket, dbg_isum, stream);
        case 4: return launch_gemv_mr<4>(w_tiles, x, perm, y, N, K, plan, ws_partial, ws_ticket, dbg_isum, stream);
        default: return cudaErrorInvalidValue;
    }
}

}  // namespace tcq4
