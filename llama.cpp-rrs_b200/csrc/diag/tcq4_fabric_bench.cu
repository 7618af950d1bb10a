// tcq4_fabric_bench.cu -- L2 -> SM fabric rate as the pair GEMM's operand staging sees it.
//
// The pair GEMM re-fetches a 32 KB activation unit per (CTA, k-tile, column block): at M = 4096 its L2 -> SM traffic
// is 49x the unique bytes (l1tex__m_xbar2l1tex_read_bytes, profiles/).  This microbenchmark is the denominator for that
// traffic: every SM streams units of an L2-resident region through a 4-stage shared-memory ring with cp.async.bulk, with
// or without cluster multicast (does sharing a unit between the CTAs of a cluster lift the delivered rate?).
// Diagnostics only (libtcq4_b200_diag.so).
#include "../../../include/tcq4_b200_diag.h"

#include <cstdio>

#include "../tcq4_common.cuh"

namespace tcq4 {

constexpr int FB_STAGES = 4;
constexpr int FB_MAX_UNIT = 49152;

__device__ __forceinline__ uint32_t fb_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void fb_wait(uint64_t *bar, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "FB_WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1, %2;\n\t"
        "@!p bra FB_WAIT_%=;\n\t}" ::"r"(smem_u32(bar)),
        "r"(parity), "r"(20000u)
        : "memory");
}
__device__ __forceinline__ void fb_bulk_mc(void *dst_smem, const void *src, uint32_t bytes, uint64_t *bar, uint16_t mask) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1], %2, [%3], %4;" ::"r"(
            smem_u32(dst_smem)),
        "l"(src), "r"(bytes), "r"(smem_u32(bar)), "h"(mask)
        : "memory");
}

__global__ void __launch_bounds__(64, 1)
fabric_bench_kernel(const uint8_t *__restrict__ buf, long long n_units, int unit_bytes, int iters, int cluster, int multicast,
                    int pattern, long long *cycles_out) {
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ __align__(8) uint64_t full[FB_STAGES], empty[FB_STAGES];
    const uint32_t rank = cluster > 1 ? fb_ctarank() : 0;
    const int cluster_id = (int)blockIdx.x / cluster;
    if (threadIdx.x == 0) {
        for (int i = 0; i < FB_STAGES; i++) {
            mbar_init(&full[i], 1);
            mbar_init(&empty[i], (uint32_t)(multicast ? cluster : 1));
        }
        fence_mbar_init();
    }
    if (cluster > 1)
        asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
    else
        __syncthreads();
    if (threadIdx.x == 0) {
        const uint16_t mask = (uint16_t)((1u << cluster) - 1u);
        const uint32_t slice = (uint32_t)unit_bytes / (uint32_t)cluster;
        const long long t0 = clock64();
        for (int i = 0; i < iters + FB_STAGES - 1; i++) {
            if (i < iters) {
                const int s = i % FB_STAGES, u = i / FB_STAGES;
                if (u > 0) fb_wait(&empty[s], (uint32_t)((u - 1) & 1));
                const long long unit = pattern == 0 ? ((long long)cluster_id * 977 + i) % n_units : (long long)i % n_units;
                const uint8_t *src = buf + unit * unit_bytes;
                mbar_arrive_expect_tx(&full[s], (uint32_t)unit_bytes);
                if (multicast && cluster > 1)
                    fb_bulk_mc(smem + s * FB_MAX_UNIT + rank * slice, src + rank * slice, slice, &full[s], mask);
                else
                    bulk_g2s(smem + s * FB_MAX_UNIT, src, (uint32_t)unit_bytes, &full[s]);
            }
            const int j = i - (FB_STAGES - 1);
            if (j >= 0) {
                const int s = j % FB_STAGES, u = j / FB_STAGES;
                fb_wait(&full[s], (uint32_t)(u & 1));
                if (multicast && cluster > 1) {
                    // the stage may be overwritten by ANY CTA of the cluster: tell all of them it has been consumed
                    for (int r = 0; r < cluster; r++) {
                        uint32_t remote;
                        asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(smem_u32(&empty[s])), "r"(r));
                        asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(remote) : "memory");
                    }
                } else {
                    mbar_arrive(&empty[s]);
                }
            }
        }
        cycles_out[blockIdx.x] = clock64() - t0;
    }
    if (cluster > 1)
        asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}

}  // namespace tcq4

extern "C" TCQ4_B200_API int tcq4_b200_measure_fabric(int cluster, int multicast, int pattern, int unit_bytes, size_t region_bytes,
                                                      int iters, double *delivered_gbs, int *ctas_out) {
    using namespace tcq4;
    if ((cluster != 1 && cluster != 2 && cluster != 4) || unit_bytes < 1024 || unit_bytes > FB_MAX_UNIT ||
        (unit_bytes % (16 * cluster)) != 0 || iters < FB_STAGES || region_bytes < (size_t)unit_bytes)
        return TCQ4_ERR_ARG;
    int dev = 0, sms = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return TCQ4_ERR_NO_DEVICE;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const size_t smem = (size_t)FB_STAGES * FB_MAX_UNIT;
    if (cudaFuncSetAttribute(fabric_bench_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess)
        return TCQ4_ERR_CUDA;
    const long long n_units = (long long)(region_bytes / (size_t)unit_bytes);
    uint8_t *buf = nullptr;
    long long *d_cycles = nullptr;
    if (cudaMalloc(&buf, (size_t)n_units * unit_bytes) != cudaSuccess) return TCQ4_ERR_CUDA;
    if (cudaMalloc(&d_cycles, sizeof(long long) * 1024) != cudaSuccess) {
        cudaFree(buf);
        return TCQ4_ERR_CUDA;
    }
    cudaMemset(buf, 1, (size_t)n_units * unit_bytes);
    // grid: as many whole clusters as the device can hold at once (cluster size 4 strands SMs in GPCs whose SM count is
    // not a multiple of 4): ask the occupancy API
    cudaLaunchConfig_t cfg = {};
    cfg.blockDim = dim3(64);
    cfg.dynamicSmemBytes = smem;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = (unsigned)cluster;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    int grid = (sms / cluster) * cluster;
    if (cluster > 1) {
        cfg.gridDim = dim3((unsigned)grid);
        int nclusters = 0;
        if (cudaOccupancyMaxActiveClusters(&nclusters, fabric_bench_kernel, &cfg) == cudaSuccess && nclusters > 0 &&
            nclusters * cluster < grid)
            grid = nclusters * cluster;
    }
    cfg.gridDim = dim3((unsigned)grid);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    float ms = 0.f;
    cudaError_t err = cudaSuccess;
    for (int rep = 0; rep < 3 && err == cudaSuccess; rep++) {       // first reps warm L2
        cudaEventRecord(e0);
        err = cudaLaunchKernelEx(&cfg, fabric_bench_kernel, (const uint8_t *)buf, n_units, unit_bytes, iters, cluster, multicast,
                                 pattern, d_cycles);
        cudaEventRecord(e1);
        if (err == cudaSuccess) err = cudaEventSynchronize(e1);
        if (err == cudaSuccess) cudaEventElapsedTime(&ms, e0, e1);
    }
    cudaFree(buf);
    cudaFree(d_cycles);
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    if (err != cudaSuccess) return TCQ4_ERR_CUDA;
    if (delivered_gbs) *delivered_gbs = (double)grid * iters * unit_bytes / (ms * 1e-3) / 1e9;
    if (ctas_out) *ctas_out = grid;       // 148, or fewer when clusters of 4 strand SMs
    return TCQ4_OK;
}
