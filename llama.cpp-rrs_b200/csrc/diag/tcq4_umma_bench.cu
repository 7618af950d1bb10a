// tcq4_umma_bench.cu -- bare tcgen05.mma issue loop: measures the tensor-pipe peak of this box for
// kind::i8 / kind::f16 / kind::f8f6f4 at a given N and shared-memory layout.  MEASURED_PEAKS.json
// has no INT8 entry, so this is the denominator the GEMM's tensor roofline is
// quoted against, and the experiment that decides the operand layout.  Operand contents are
// irrelevant (zeroed smem); only issue/fetch/execute timing is measured.
#include "../../../include/tcq4_b200_diag.h"
#include <cstdio>
#include <cstdlib>

#include "../tcq4_common.cuh"

namespace tcq4 {

enum { UB_I8 = 0, UB_BF16 = 1, UB_FP8 = 2 };

template <int KIND>
__device__ __forceinline__ void ub_mma(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
    if (KIND == UB_I8) {
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                     "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}" ::"r"(d), "l"(a), "l"(b), "r"(idesc),
                     "r"(acc) : "memory");
    } else if (KIND == UB_BF16) {
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                     "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d), "l"(a), "l"(b), "r"(idesc),
                     "r"(acc) : "memory");
    } else {
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                     "tcgen05.mma.cta_group::1.kind::f8f6f4 [%0], %1, %2, %3, p;\n\t}" ::"r"(d), "l"(a), "l"(b),
                     "r"(idesc), "r"(acc) : "memory");
    }
}

// swizzle: 0 = none (LBO 2048 / SBO 128, the layout of tcq4_gemm.cu), 1 = SWIZZLE_128B (SBO 1024)
template <int KIND>
__global__ void __launch_bounds__(128, 1)
umma_bench_kernel(int N, int swizzle, int iters, int kslices, long long *cycles_out) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint64_t bar;
    __shared__ uint32_t tmem_slot;
    const int warp = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < (64 * 1024) / 4; i += blockDim.x) reinterpret_cast<uint32_t *>(smem)[i] = 0;
    if (threadIdx.x == 0) {
        mbar_init(&bar, 1);
        fence_mbar_init();
    }
    if (warp == 0) {
        tmem_alloc(&tmem_slot, 512);
        tmem_relinquish();
    }
    fence_proxy_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = tmem_slot;
    if (threadIdx.x == 0) {
        uint32_t idesc;
        if (KIND == UB_I8) idesc = umma_idesc_i8(128, N, 1, 1);
        else if (KIND == UB_BF16) idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | (8u << 24);
        else idesc = (1u << 4) | ((uint32_t)(N >> 3) << 17) | (8u << 24);
        const uint32_t a_base = smem_u32(smem), b_base = a_base + 32 * 1024;
        const long long t0 = clock64();
        for (int it = 0; it < iters; it++) {
            for (int ks = 0; ks < kslices; ks++) {
                uint64_t ad, bd;
                if (swizzle == 0) {
                    ad = umma_desc_kmajor_noswz(a_base + ks * 4096, 2048, 128);
                    bd = umma_desc_kmajor_noswz(b_base + ks * (N * 32), N * 16, 128);
                } else {
                    // 128B swizzle, K-major: rows 128 B apart, 8-row groups 1024 B apart, k-slice = +32 B
                    ad = ((uint64_t)(((a_base + ks * 32) >> 4) & 0x3FFF)) | ((uint64_t)1 << 16) |
                         ((uint64_t)(1024 >> 4) << 32) | ((uint64_t)1 << 46) | ((uint64_t)2 << 61);
                    bd = ((uint64_t)(((b_base + ks * 32) >> 4) & 0x3FFF)) | ((uint64_t)1 << 16) |
                         ((uint64_t)(1024 >> 4) << 32) | ((uint64_t)1 << 46) | ((uint64_t)2 << 61);
                }
                ub_mma<KIND>(tmem + (it & 1) * 256, ad, bd, idesc, 1u);
            }
        }
        umma_commit(&bar);
        mbar_wait(&bar, 0);
        const long long t1 = clock64();
        cycles_out[blockIdx.x] = t1 - t0;
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem, 512);
}

}  // namespace tcq4

extern "C" TCQ4_B200_API int tcq4_b200_measure_umma(int kind, int N, int swizzle, int iters, double *tops_out,
                                                    double *cycles_per_mma_out) {
    using namespace tcq4;
    int dev = 0, sms = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return TCQ4_ERR_NO_DEVICE;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int kslices = 4;
    const size_t smem = 64 * 1024 + 1024;
    long long *d_cycles = nullptr;
    if (cudaMalloc(&d_cycles, sizeof(long long) * sms) != cudaSuccess) return TCQ4_ERR_CUDA;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    float ms = 0.f;
    cudaError_t err = cudaSuccess;
    for (int rep = 0; rep < 2; rep++) {     // first rep warms up
        cudaEventRecord(e0);
        if (kind == UB_I8) {
            cudaFuncSetAttribute(umma_bench_kernel<UB_I8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            umma_bench_kernel<UB_I8><<<sms, 128, smem>>>(N, swizzle, iters, kslices, d_cycles);
        } else if (kind == UB_BF16) {
            cudaFuncSetAttribute(umma_bench_kernel<UB_BF16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            umma_bench_kernel<UB_BF16><<<sms, 128, smem>>>(N, swizzle, iters, kslices, d_cycles);
        } else {
            cudaFuncSetAttribute(umma_bench_kernel<UB_FP8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            umma_bench_kernel<UB_FP8><<<sms, 128, smem>>>(N, swizzle, iters, kslices, d_cycles);
        }
        cudaEventRecord(e1);
        err = cudaEventSynchronize(e1);
        if (err != cudaSuccess) break;
        cudaEventElapsedTime(&ms, e0, e1);
    }
    long long h_cycles = 0;
    cudaMemcpy(&h_cycles, d_cycles, sizeof(long long), cudaMemcpyDeviceToHost);
    cudaFree(d_cycles);
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    if (err != cudaSuccess) return TCQ4_ERR_CUDA;
    const double kelems = (kind == UB_BF16) ? 16.0 : 32.0;
    const double ops = 2.0 * 128.0 * N * kelems * (double)iters * kslices * sms;
    if (tops_out) *tops_out = ops / (ms * 1e-3) / 1e12;
    if (cycles_per_mma_out) *cycles_per_mma_out = (double)h_cycles / ((double)iters * kslices);
    return TCQ4_OK;
}

// ---- cta_group::2 variant: a CTA pair issues M = 256 MMAs (each CTA holds 128 rows of A and N/2 rows of B) ----
namespace tcq4 {
// bg_warps > 0 (diagnostic, TCQ4_UMMA2_BG="warps,mode"): warps 4.. of BOTH CTAs stream through a separate 32 KB of shared
// memory while the MMAs run -- mode 1 = STS.128, 2 = LDS.128 -- to see what concurrent shared-memory traffic (the unpack
// stores, TMA fills and scale reads of the real kernel) costs the tensor pipe's operand fetch.
template <int kind>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(512, 1)
umma2_bench_kernel(int N, int iters, int kslices, long long *cycles_out, int bg_warps, int bg_mode) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ __align__(8) uint64_t bar;
    __shared__ uint32_t tmem_slot;
    __shared__ volatile int bg_stop;
    if (threadIdx.x == 0) bg_stop = 0;
    const int warp = threadIdx.x >> 5;
    uint32_t crank;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(crank));
    for (int i = threadIdx.x; i < (64 * 1024) / 4; i += blockDim.x) reinterpret_cast<uint32_t *>(smem)[i] = 0;
    if (threadIdx.x == 0) {
        mbar_init(&bar, 1);
        fence_mbar_init();
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_slot)), "r"(512) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
    fence_proxy_async_smem();
    tc_fence_before();
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
    tc_fence_after();
    const uint32_t tmem = tmem_slot;
    if (warp >= 4) {
        if (warp - 4 < bg_warps) {
            uint4 *area = reinterpret_cast<uint4 *>(smem + 64 * 1024 + 1024) + (threadIdx.x & 31);
            uint4 v = make_uint4(threadIdx.x, 1, 2, 3);
            int k = (warp - 4) * 64;
            while (!bg_stop) {
#pragma unroll
                for (int u = 0; u < 8; u++) {
                    uint4 *pp = area + ((k + u * 4) & 63) * 32;           // 64 x 512 B = 32 KB, conflict-free per instruction
                    if (bg_mode == 1) *pp = v;
                    else {
                        uint4 t;
                        asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(t.x), "=r"(t.y), "=r"(t.z), "=r"(t.w) : "r"(smem_u32(pp)));
                        v.x ^= t.x;
                    }
                }
                k++;
            }
            if (v.x == 0xdeadbeefu) cycles_out[0] = 0;                     // keep the loads alive
        }
    } else if (crank == 0 && threadIdx.x == 0) {
        // kind 0: i8 (K = 32); 1: f16 x f16 -> f32 (K = 16), both operands in shared memory; 2: the same with A in TMEM
        const uint32_t idesc = kind == 0 ? umma_idesc_i8(256, N, 1, 1)
                                         : ((1u << 4) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(256 >> 4) << 24));
        const uint32_t a_base = smem_u32(smem), b_base = a_base + 32 * 1024;
        const int nh = N / 2;     // B rows held by each CTA
        const long long t0 = clock64();
        for (int it = 0; it < iters; it++) {
            for (int ks = 0; ks < kslices; ks++) {
                const uint64_t ad = umma_desc_kmajor_noswz(a_base + ks * 4096, 2048, 128);
                const uint64_t bd = umma_desc_kmajor_noswz(b_base + ks * (nh * 32), nh * 16, 128);
                const uint32_t d = tmem + (it & 1) * 256;
                if (kind == 0)
                    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                                 "tcgen05.mma.cta_group::2.kind::i8 [%0], %1, %2, %3, p;\n\t}" ::"r"(d), "l"(ad), "l"(bd),
                                 "r"(idesc), "r"(1u) : "memory");
                else if (kind == 1)
                    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                                 "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d), "l"(ad), "l"(bd),
                                 "r"(idesc), "r"(1u) : "memory");
                else      // A from TMEM columns 448.. (8 columns per K = 16); D uses columns 0..N-1 / 192..192+N-1
                    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                                 "tcgen05.mma.cta_group::2.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(tmem + (it & 1) * 192),
                                 "r"(tmem + 448 + ks * 8), "l"(bd), "r"(idesc), "r"(1u) : "memory");
            }
        }
        asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::
                         "r"(smem_u32(&bar)), "h"((uint16_t)1) : "memory");
        mbar_wait(&bar, 0);
        const long long t1 = clock64();
        cycles_out[blockIdx.x / 2] = t1 - t0;
        bg_stop = 1;
        // the peer CTA's background warps stop too
        asm volatile("{\n\t.reg .b32 ra;\n\tmapa.shared::cluster.u32 ra, %0, 1;\n\tst.shared::cluster.u32 [ra], %1;\n\t}" ::
                         "r"(smem_u32(const_cast<int *>(&bg_stop))), "r"(1) : "memory");
    }
    tc_fence_before();
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512) : "memory");
}
}  // namespace tcq4

extern "C" TCQ4_B200_API int tcq4_b200_measure_umma2(int N, int iters, double *tops_out, double *cycles_per_mma_out) {
    using namespace tcq4;
    int dev = 0, sms = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return TCQ4_ERR_NO_DEVICE;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    sms &= ~1;
    const int kslices = 4;
    int bg_warps = 0, bg_mode = 0;
    if (const char *bg = getenv("TCQ4_UMMA2_BG")) sscanf(bg, "%d,%d", &bg_warps, &bg_mode);
    if (bg_warps < 0 || bg_warps > 12) bg_warps = 0;
    int kind = 0;      // TCQ4_UMMA2_KIND: 0 = i8, 1 = f16 (SS), 2 = f16 with A in TMEM (N <= 192)
    if (const char *k = getenv("TCQ4_UMMA2_KIND")) kind = atoi(k);
    if (kind < 0 || kind > 2 || (kind == 2 && N > 192)) kind = 0;
    const size_t smem = 64 * 1024 + 1024 + 32 * 1024 + 512;
    long long *d_cycles = nullptr;
    if (cudaMalloc(&d_cycles, sizeof(long long) * sms) != cudaSuccess) return TCQ4_ERR_CUDA;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    float ms = 0.f;
    auto kern = kind == 0 ? umma2_bench_kernel<0> : kind == 1 ? umma2_bench_kernel<1> : umma2_bench_kernel<2>;
    cudaError_t err = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    for (int rep = 0; rep < 2 && err == cudaSuccess; rep++) {
        cudaEventRecord(e0);
        kern<<<sms, bg_warps > 0 ? 128 + 32 * bg_warps : 128, smem>>>(N, iters, kslices, d_cycles, bg_warps, bg_mode);
        cudaEventRecord(e1);
        err = cudaEventSynchronize(e1);
        if (err != cudaSuccess) break;
        cudaEventElapsedTime(&ms, e0, e1);
    }
    long long h_cycles = 0;
    cudaMemcpy(&h_cycles, d_cycles, sizeof(long long), cudaMemcpyDeviceToHost);
    cudaFree(d_cycles);
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    if (err != cudaSuccess) return TCQ4_ERR_CUDA;
    const double ops = 2.0 * 256.0 * N * (kind == 0 ? 32.0 : 16.0) * (double)iters * kslices * (sms / 2);
    if (tops_out) *tops_out = ops / (ms * 1e-3) / 1e12;
    if (cycles_per_mma_out) *cycles_per_mma_out = (double)h_cycles / ((double)iters * kslices);
    return TCQ4_OK;
}

// ---- warp-level pipe microbenchmark: cycles per warp-instruction per SM for the instructions the decode
// kernel could be built on (legacy IMMA int8, IDP.4A, SHFL, REDUX).  16 warps per SM, `unroll` independent
// chains per warp.  kind: 0 = mma.sync.m16n8k32.s8, 1 = dp4a, 2 = shfl.xor, 3 = redux.add, 4 = cvt.rn.f32.s32, 5 = mad.lo.s32
namespace tcq4 {
template <int KIND>
__global__ void __launch_bounds__(512, 1) pipe_bench_kernel(int iters, long long *cycles_out, int *sink) {
    int acc[8][4];
#pragma unroll
    for (int i = 0; i < 8; i++)
#pragma unroll
        for (int j = 0; j < 4; j++) acc[i][j] = threadIdx.x + i + j;
    uint32_t a = threadIdx.x * 2654435761u, b = a ^ 0x9E3779B9u;
    __syncthreads();
    const long long t0 = clock64();
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++) {
            if (KIND == 0) {
                asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                             : "+r"(acc[i][0]), "+r"(acc[i][1]), "+r"(acc[i][2]), "+r"(acc[i][3])
                             : "r"(a), "r"(b), "r"(a), "r"(b), "r"(a), "r"(b));
            } else if (KIND == 1) {
                acc[i][0] = __dp4a((int)a, (int)b, acc[i][0]);
            } else if (KIND == 2) {
                acc[i][0] = __shfl_xor_sync(0xffffffffu, acc[i][0], 1 + (i & 15));
            } else if (KIND == 4) {
                acc[i][0] = __float_as_int((float)acc[i][0]);          // I2FP.F32.S32
            } else if (KIND == 5) {
                acc[i][0] = acc[i][0] * 16 + (int)a;                   // IMAD
            } else {
                acc[i][0] += __reduce_add_sync(0xffffffffu, acc[i][0]);
            }
        }
    }
    __syncthreads();
    const long long t1 = clock64();
    int s = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) s += acc[i][0] + acc[i][1] + acc[i][2] + acc[i][3];
    if (s == 0x7fffffff) *sink = s;
    if (threadIdx.x == 0) cycles_out[blockIdx.x] = t1 - t0;
}
}  // namespace tcq4

extern "C" TCQ4_B200_API int tcq4_b200_measure_pipe(int kind, int iters, double *cycles_per_warp_instr_per_sm) {
    using namespace tcq4;
    int dev = 0, sms = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return TCQ4_ERR_NO_DEVICE;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    long long *d_cycles = nullptr;
    int *d_sink = nullptr;
    if (cudaMalloc(&d_cycles, sizeof(long long) * sms) != cudaSuccess) return TCQ4_ERR_CUDA;
    cudaMalloc(&d_sink, sizeof(int));
    for (int rep = 0; rep < 2; rep++) {
        if (kind == 0) pipe_bench_kernel<0><<<sms, 512>>>(iters, d_cycles, d_sink);
        else if (kind == 1) pipe_bench_kernel<1><<<sms, 512>>>(iters, d_cycles, d_sink);
        else if (kind == 2) pipe_bench_kernel<2><<<sms, 512>>>(iters, d_cycles, d_sink);
        else if (kind == 4) pipe_bench_kernel<4><<<sms, 512>>>(iters, d_cycles, d_sink);
        else if (kind == 5) pipe_bench_kernel<5><<<sms, 512>>>(iters, d_cycles, d_sink);
        else pipe_bench_kernel<3><<<sms, 512>>>(iters, d_cycles, d_sink);
    }
    cudaError_t err = cudaDeviceSynchronize();
    long long h = 0;
    cudaMemcpy(&h, d_cycles, sizeof(long long), cudaMemcpyDeviceToHost);
    cudaFree(d_cycles);
    cudaFree(d_sink);
    if (err != cudaSuccess) return TCQ4_ERR_CUDA;
    if (cycles_per_warp_instr_per_sm) *cycles_per_warp_instr_per_sm = (double)h / ((double)iters * 8 * 16);
    return TCQ4_OK;
}

// ---- TMEM read bandwidth: `warps` warps per SM each loop tcgen05.ld.32x32b.x{8,16,32} over their own lane
// quarter.  The prefill epilogue reads every accumulator once per 256-block, so this rate (not the MMA rate)
// bounds it when the accumulator tile is wide.
namespace tcq4 {
template <int COLS>
__global__ void __launch_bounds__(1024, 1) tmem_ld_bench_kernel(int iters, long long *cycles_out, int *sink) {
    __shared__ uint32_t slot;
    const int warp = threadIdx.x >> 5;
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&slot)), "r"(512)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t base = slot + ((uint32_t)((warp & 3) * 32) << 16);
    uint32_t x = 0;
    __syncthreads();
    const long long t0 = clock64();
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const uint32_t col = (uint32_t)(((warp >> 2) * 4 + j) * COLS) & 511u;
            if (COLS == 8) {
                uint32_t r[8];
                tmem_ld_32x8(base + col, r);
                tmem_ld_wait();
#pragma unroll
                for (int i = 0; i < 8; i++) x ^= r[i];
            } else if (COLS == 16) {
                uint32_t r[16];
                tmem_ld_32x16(base + col, r);
                tmem_ld_wait();
#pragma unroll
                for (int i = 0; i < 16; i++) x ^= r[i];
            } else {
                uint32_t r[32];
                tmem_ld_32x32(base + col, r);
                tmem_ld_wait();
#pragma unroll
                for (int i = 0; i < 32; i++) x ^= r[i];
            }
        }
    }
    __syncthreads();
    const long long t1 = clock64();
    if (x == 0x12345678u) *sink = (int)x;
    if (threadIdx.x == 0) cycles_out[blockIdx.x] = t1 - t0;
    tc_fence_before();
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(slot), "r"(512) : "memory");
}
}  // namespace tcq4

extern "C" TCQ4_B200_API int tcq4_b200_measure_tmem_ld(int cols, int warps, int iters, double *bytes_per_clk_per_sm) {
    using namespace tcq4;
    if ((cols != 8 && cols != 16 && cols != 32) || warps < 1 || warps > 32 || iters < 1) return TCQ4_ERR_ARG;
    int dev = 0, sms = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return TCQ4_ERR_NO_DEVICE;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    long long *d_cycles = nullptr;
    int *d_sink = nullptr;
    if (cudaMalloc(&d_cycles, sizeof(long long) * sms) != cudaSuccess) return TCQ4_ERR_CUDA;
    cudaMalloc(&d_sink, sizeof(int));
    for (int rep = 0; rep < 2; rep++) {
        if (cols == 8) tmem_ld_bench_kernel<8><<<sms, warps * 32>>>(iters, d_cycles, d_sink);
        else if (cols == 16) tmem_ld_bench_kernel<16><<<sms, warps * 32>>>(iters, d_cycles, d_sink);
        else tmem_ld_bench_kernel<32><<<sms, warps * 32>>>(iters, d_cycles, d_sink);
    }
    cudaError_t err = cudaDeviceSynchronize();
    long long h = 0;
    cudaMemcpy(&h, d_cycles, sizeof(long long), cudaMemcpyDeviceToHost);
    cudaFree(d_cycles);
    cudaFree(d_sink);
    if (err != cudaSuccess) return TCQ4_ERR_CUDA;
    if (bytes_per_clk_per_sm) *bytes_per_clk_per_sm = (double)iters * 4 * warps * 32.0 * cols * 4.0 / (double)h;
    return TCQ4_OK;
}

// ---- exactness probe: kind::f16 on small integers ---------------------------------------------------------------------
// A (128 x 256) holds activation codes in [-7, 7] and B (256 x 256) holds sc*q in [-1016, 1016], both exactly
// representable in fp16; every partial sum of a 256-deep dot is an integer below 2^21, so an fp32 accumulator that keeps
// 24 significant bits holds the int32 result EXACTLY.  This kernel checks that the tensor core's accumulation does
// (one CTA, 16 MMAs of 128 x 256 x 16, D read back with tcgen05.ld and compared with an integer dot product).
// pattern 0: random, 1: every product at its maximum (+7 * +1016), 2: alternating signs of maximal products, 3: random
// with A in [-7,7] and B restricted to +-1016 / +-1015 (odd/even mix: low bits alive at the largest magnitudes)
namespace tcq4 {
__device__ __forceinline__ uint32_t f16x_hash(uint32_t x) {
    x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
    return x;
}
__global__ void __launch_bounds__(128, 1) f16_exact_kernel(int pattern, uint32_t seed, int *mismatches, int *max_abs_out) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ __align__(8) uint64_t bar;
    __shared__ uint32_t tmem_slot;
    __half *sA = reinterpret_cast<__half *>(smem);                 // [32 chunks][128 rows][8]
    __half *sB = reinterpret_cast<__half *>(smem + 64 * 1024);     // [32 chunks][256 rows][8]
    const int warp = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < 128 * 256; i += blockDim.x) {
        const int r = i >> 8, k = i & 255;
        int a;
        if (pattern == 1) a = 7;
        else if (pattern == 2) a = (k & 1) ? -7 : 7;
        else a = (int)(f16x_hash(seed ^ (uint32_t)i) % 15u) - 7;
        sA[(k >> 3) * (128 * 8) + r * 8 + (k & 7)] = __int2half_rn(a);
    }
    for (int i = threadIdx.x; i < 256 * 256; i += blockDim.x) {
        const int n = i >> 8, k = i & 255;
        int b;
        const uint32_t h = f16x_hash(seed * 31u + 0x9E3779B9u + (uint32_t)i);
        if (pattern == 1) b = 1016;
        else if (pattern == 2) b = (k & 2) ? -1016 : 1016;
        else if (pattern == 3) b = ((h & 1) ? 1016 : 1015) * ((h & 2) ? 1 : -1);
        else b = (int)(h % 2033u) - 1016;
        sB[(k >> 3) * (256 * 8) + n * 8 + (k & 7)] = __int2half_rn(b);
    }
    if (threadIdx.x == 0) {
        mbar_init(&bar, 1);
        fence_mbar_init();
    }
    if (warp == 0) {
        tmem_alloc(&tmem_slot, 256);
        tmem_relinquish();
    }
    fence_proxy_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = tmem_slot;
    if (threadIdx.x == 0) {
        const uint32_t idesc = (1u << 4) | ((uint32_t)(256 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);   // F32 acc, F16 x F16
        const uint32_t a_base = smem_u32(sA), b_base = smem_u32(sB);
        for (int ks = 0; ks < 16; ks++) {
            const uint64_t ad = umma_desc_kmajor_noswz(a_base + ks * 2 * 2048, 2048, 128);
            const uint64_t bd = umma_desc_kmajor_noswz(b_base + ks * 2 * 4096, 4096, 128);
            ub_mma<UB_BF16>(tmem, ad, bd, idesc, ks > 0 ? 1u : 0u);       // kind::f16 (operand format comes from idesc)
        }
        umma_commit(&bar);
    }
    mbar_wait(&bar, 0);
    tc_fence_after();
    const int r = threadIdx.x;      // TMEM lane = row
    int bad = 0, mx = 0;
    for (int c0 = 0; c0 < 256; c0 += 32) {
        uint32_t v[32];
        tmem_ld_32x32(tmem + ((uint32_t)(warp * 32) << 16) + c0, v);
        tmem_ld_wait();
        for (int j = 0; j < 32; j++) {
            const int n = c0 + j;
            int ref = 0;
            for (int k = 0; k < 256; k++)
                ref += __half2int_rn(sA[(k >> 3) * (128 * 8) + r * 8 + (k & 7)]) *
                       __half2int_rn(sB[(k >> 3) * (256 * 8) + n * 8 + (k & 7)]);
            const float got = __uint_as_float(v[j]);
            if (got != (float)ref || (int)got != ref) bad++;
            mx = max(mx, abs(ref));
        }
    }
    if (bad) atomicAdd(mismatches, bad);
    atomicMax(max_abs_out, mx);
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem, 256);
}
}  // namespace tcq4

extern "C" TCQ4_B200_API int tcq4_b200_measure_f16_exact(int pattern, unsigned seed, int *mismatches, int *max_abs) {
    using namespace tcq4;
    int *d = nullptr;
    if (cudaMalloc(&d, 8) != cudaSuccess) return TCQ4_ERR_CUDA;
    cudaMemset(d, 0, 8);
    const size_t smem = 192 * 1024 + 1024;
    cudaError_t e = cudaFuncSetAttribute(f16_exact_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e == cudaSuccess) {
        f16_exact_kernel<<<1, 128, smem>>>(pattern, seed, d, d + 1);
        e = cudaDeviceSynchronize();
    }
    int h[2] = {-1, -1};
    if (e == cudaSuccess) cudaMemcpy(h, d, 8, cudaMemcpyDeviceToHost);
    cudaFree(d);
    if (e != cudaSuccess) return TCQ4_ERR_CUDA;
    if (mismatches) *mismatches = h[0];
    if (max_abs) *max_abs = h[1];
    return TCQ4_OK;
}
