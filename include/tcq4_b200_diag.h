/*
 * tcq4_b200_diag.h -- diagnostics of the B200-native TCQ4_K32 path: microbenchmarks (tensor-pipe issue peaks, warp-level
 * pipe rates, TMEM read rate, L2 -> SM fabric rate) and the phase / timeline readers of the instrumented kernel
 * instantiations.  NOT part of the product surface (include/tcq4_b200.h): these symbols live in
 * llama.cpp-rrs_b200/libtcq4_b200_diag.so, a superset build of the product sources with -DTCQ4_DIAG plus csrc/diag/.
 * No reference counterpart except the fork's event-timed ggml_cuda_rrs_benchmark (ggml/src/ggml-cuda/rrs.cu:350-419).
 */
#ifndef TCQ4_B200_DIAG_H
#define TCQ4_B200_DIAG_H

#include "tcq4_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

/* Tensor-pipe peak of this device for one tcgen05.mma flavour: a bare single-thread issue loop on
 * every SM (kind: 0 = i8, 1 = bf16 (kind::f16), 2 = e4m3 (kind::f8f6f4); M = 128; N in {64..256};
 * swizzle: 0 = none (the GEMM's operand layout), 1 = 128B).  MEASURED_PEAKS.json has no INT8 figure,
 * so this is the measured denominator for the GEMM's tensor roofline.  Synchronous. */
TCQ4_B200_API int tcq4_b200_measure_umma(int kind, int N, int swizzle, int iters, double *tops_out,
                                         double *cycles_per_mma_out);

/* Same for a CTA pair: tcgen05.mma.cta_group::2.kind::i8, M = 256, N in {64..256}, no-swizzle K-major. */
TCQ4_B200_API int tcq4_b200_measure_umma2(int N, int iters, double *tops_out, double *cycles_per_mma_out);

/* Diagnostic: SM-wide cost (cycles per warp-instruction, 16 resident warps) of the warp-level instructions the
 * decode kernel is built from.  kind: 0 = mma.sync.m16n8k32.s8 (legacy IMMA), 1 = dp4a, 2 = shfl, 3 = redux.add,
 * 4 = cvt.rn.f32.s32 (I2FP), 5 = mad.lo.s32 (IMAD) */
TCQ4_B200_API int tcq4_b200_measure_pipe(int kind, int iters, double *cycles_per_warp_instr_per_sm);
/* TMEM read bandwidth: `warps` warps per SM loop tcgen05.ld.32x32b.x{cols} (cols = 8, 16 or 32) over their lane
 * quarter; returns bytes per clock per SM.  Bounds the prefill epilogue ("TMEM read rate"). */
TCQ4_B200_API int tcq4_b200_measure_tmem_ld(int cols, int warps, int iters, double* bytes_per_clk_per_sm);

/* Diagnostic: per-CTA phase timestamps (ns) of the last GEMV launched through tcq4_b200_mul_mat_debug:
 * out[cta][8] = {start, prefetch issued, dependency resolved, activation stage done, stream done, end}. */
TCQ4_B200_API int tcq4_b200_debug_gemv_phases(unsigned long long *out, size_t bytes);
/* Diagnostics: SM-clock stamps [cta 2][role 6][k-tile 16][event 8] of one cluster of the pair GEMM, recorded when
 * the environment has TCQ4_G2_PROF set (a separate instantiation; the product kernel carries no stamps). */
TCQ4_B200_API int tcq4_b200_debug_gemm2_timeline(long long *out, size_t bytes);
/* Same for the kind::f16 pair GEMM (TCQ4_G3_PROF): [role 4: MMA issuer, activation warp 0, weight warp 0, epilogue warp 0]
 * [k-tile 64][event 8] of the leader CTA of one cluster. */
TCQ4_B200_API int tcq4_b200_debug_gemm3_timeline(long long *out, size_t bytes);

/* Same for the batched-decode GEMM (TCQ4_G4_PROF): [role 4: MMA issuer, widening warp 0, widening warp 15, epilogue warp 0]
 * [k-tile 64][event 8] of the middle CTA.  TCQ4_G4_EXP switches parts of the CUDA-core work off (bit 0 weight widening,
 * bit 1 token widening, bit 2 epilogue arithmetic; results are then wrong on purpose). */
TCQ4_B200_API int tcq4_b200_debug_gemm4_timeline(long long *out, size_t bytes);

/* L2 -> SM fabric rate: every CTA (one per SM, clusters of `cluster` CTAs) streams `unit_bytes` units with cp.async.bulk
 * out of an L2-resident region of `region_bytes` into a 4-stage shared-memory ring, `iters` units per CTA.
 *   pattern 0: clusters read different units at any time;  pattern 1: all clusters read the same unit sequence
 *   multicast 0: every CTA copies the whole unit for itself;  1: CTA r of a cluster copies slice r of the unit and
 *   multicasts it into all CTAs of the cluster (every CTA still RECEIVES unit_bytes per unit)
 * Returns delivered bytes per second summed over all SMs (what the pair GEMM's operand staging can be fed with). */
TCQ4_B200_API int tcq4_b200_measure_fabric(int cluster, int multicast, int pattern, int unit_bytes, size_t region_bytes,
                                           int iters, double *delivered_gbs, int *ctas_out);

/* Exactness probe: tcgen05.mma kind::f16 on integer-valued fp16 operands (A in [-7,7], B in [-1016,1016], K = 256): are
 * the fp32 accumulators in TMEM bit-equal to the int32 dot products?  pattern 0 random, 1 all products maximal, 2
 * alternating signs, 3 odd/even maximal magnitudes.  mismatches = elements of the 128 x 256 tile that differ. */
TCQ4_B200_API int tcq4_b200_measure_f16_exact(int pattern, unsigned seed, int *mismatches, int *max_abs);

#ifdef __cplusplus
}
#endif
#endif /* TCQ4_B200_DIAG_H */
