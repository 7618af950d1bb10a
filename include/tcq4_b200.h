/*
 * tcq4_b200.h -- C ABI of the B200-native TCQ4_K32 W4A4 MUL_MAT hot path.
 *
 * Drop-in boundary for skyne98/llama.cpp-rrs (all file:line under /root/reference):
 * every entry point below replaces one piece of the fork's ggml-cuda plug-in surface for
 * GGML_TYPE_TCQ4_K32 and nothing else.  Plain pointers and sizes only; all data pointers are
 * DEVICE pointers unless the name says host.  No entry point allocates, synchronises or touches
 * host state on the launch path, so every launch is CUDA-graph-capture safe
 * (ggml/src/ggml-cuda/ggml-cuda.cu:3791-3799).
 *
 * There is no CPU fallback: on a machine without an sm_100 device every compute entry point
 * returns TCQ4_ERR_NO_DEVICE.
 */
#ifndef TCQ4_B200_H
#define TCQ4_B200_H

#include <stddef.h>
#include <stdint.h>
#ifndef __cplusplus
#include <stdbool.h>
#endif

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__CUDACC__) || defined(CUDART_VERSION) || defined(__DRIVER_TYPES_H__)
/* cudaStream_t already known */
#else
typedef struct CUstream_st *cudaStream_t;
#endif

#define TCQ4_B200_API __attribute__((visibility("default")))

/* ---- format constants (ggml/src/ggml-common.h:315-341, ggml/src/ggml.c:750-757) ------------- */
#define TCQ4_B200_GGML_TYPE        42     /* GGML_TYPE_TCQ4_K32, ggml/include/ggml.h:432 */
#define TCQ4_B200_TILE_K           256
#define TCQ4_B200_TILE_CHANNELS    8
#define TCQ4_B200_TILE_BYTES       1184
#define TCQ4_B200_ACT_BLOCK_BYTES  160    /* block_rrs_int4_tc, ggml/src/ggml-cuda/tcq4_k32.cuh:51-56 */

/* ---- status codes (the reference aborts via GGML_ASSERT, rrs.cu:256-259; the shim asserts ==0) */
enum {
    TCQ4_OK             = 0,
    TCQ4_ERR_SHAPE      = 1,   /* K % 256 != 0, N % 8 != 0, or non-positive dims (rrs.cu:258-259) */
    TCQ4_ERR_ARG        = 2,   /* null / misaligned pointer */
    TCQ4_ERR_WORKSPACE  = 3,   /* workspace too small (and cannot grow: stream is capturing) */
    TCQ4_ERR_CUDA       = 4,   /* a CUDA runtime call failed; see tcq4_b200_last_error() */
    TCQ4_ERR_NO_DEVICE  = 5    /* no sm_100 device: there is no CPU fallback */
};

/* kernel selection, see tcq4_b200_select_kernel() */
enum {
    TCQ4_KERNEL_AUTO  = 0,
    TCQ4_KERNEL_GEMV  = 1,     /* fused FWHT+quant+streaming dot, small M, HBM-bound          */
    TCQ4_KERNEL_GEMM  = 2,     /* activation kernel + tcgen05.mma kind::i8 GEMM, one CTA per 128 tokens   */
    TCQ4_KERNEL_GEMM2 = 3,     /* same on CTA pairs (cta_group::2, 256 tokens per pair); several destinations       */
    TCQ4_KERNEL_GEMM3 = 4,     /* CTA pairs, tcgen05.mma kind::f16 on exact fp16 integers, nibble operands widened
                                  in the SM, fan-out segments, K split: the prefill kernel (round 2)                */
    TCQ4_KERNEL_GEMM4 = 5      /* batched decode, 9 <= M <= 64: weights widened into TMEM (the MMA's A operand, 128
                                  channels per CTA), tokens as the MMA's N; K split on narrow matrices              */
};

TCQ4_B200_API const char *tcq4_b200_version(void);
TCQ4_B200_API const char *tcq4_b200_last_error(void);

/* ggml_cuda_supports_rrs + the asserts of ggml_cuda_rrs_mul_mat (rrs.cu:237-239, 256-259) and the
 * MUL_MAT case of ggml_backend_cuda_device_supports_op (ggml-cuda.cu:4457-4459).
 * cc = compute capability * 10 (100 for sm_100).  Returns 1 if this library takes the op. */
TCQ4_B200_API int tcq4_b200_supports(int64_t M, int64_t N, int64_t K, int cc);

/* Which kernel ggml_cuda_rrs_mul_mat's dispatch (rrs.cu:317-343) would map to here. */
TCQ4_B200_API int tcq4_b200_select_kernel(int64_t M, int64_t N, int64_t K);

/* Scratch bytes tcq4_b200_mul_mat_ws needs for this shape (quantised activations, split-K partials).
 * The reference's live path needs none (rrs.cu:321); its #else path takes it from ctx.pool()
 * (rrs.cu:326-328) -- a ggml shim does the same with this size. */
TCQ4_B200_API size_t tcq4_b200_workspace_size(int64_t M, int64_t N, int64_t K);

/* THE hot path.  Replaces tcq4_rrs_fused_gemm_smallM (tcq4_k32.cuh:150-156, called from
 * ggml_cuda_rrs_mul_mat, rrs.cu:321):  y[m*N+n] = sum_k W[n][k] * x[m*K+k] in W4A4 arithmetic.
 *   w_tiles : [N/8][K/256] block_tcq4_tile, 16-byte aligned (ggml buffers are 128-byte aligned)
 *   x       : F32 [M][K] contiguous;  y : F32 [M][N] contiguous
 *   d_perm  : int32[K] device channel permutation or NULL (rrs.cu:266-270)
 *   workspace: >= tcq4_b200_workspace_size(M,N,K) bytes, 256-byte aligned, must stay valid until
 *             the launch completes; first use of a workspace must be preceded by
 *             tcq4_b200_workspace_init (zeroes the split-K tickets).
 * Asynchronous on `stream`. */
TCQ4_B200_API int tcq4_b200_mul_mat_ws(const void *w_tiles, const float *x, float *y,
                                       int64_t M, int64_t N, int64_t K, const int32_t *d_perm,
                                       void *workspace, size_t workspace_bytes, int kernel,
                                       cudaStream_t stream);
TCQ4_B200_API int tcq4_b200_workspace_init(void *workspace, size_t workspace_bytes, cudaStream_t stream);

/* Same, with scratch taken from a per-device arena owned by the library.  The arena grows with
 * cudaMalloc only outside stream capture; call tcq4_b200_reserve at load time to pre-size it.  A block the arena has
 * outgrown is retired, not freed (graphs captured earlier may still point into it).  ONE arena per device: every entry
 * without a workspace argument (tcq4_b200_mul_mat, _add, _fanout, _mul_mat_id, _peers, _multicast, the ggml mirror) is
 * asynchronous on `stream` but must not run CONCURRENTLY with another such call on a different stream of the same
 * device -- callers with several streams pass their own workspaces (tcq4_b200_mul_mat_ws); the host-buffer pipeline
 * (tcq4_b200_mul_mat_host_async) owns a workspace of its own. */
TCQ4_B200_API int tcq4_b200_mul_mat(const void *w_tiles, const float *x, float *y,
                                    int64_t M, int64_t N, int64_t K, const int32_t *d_perm,
                                    cudaStream_t stream);
TCQ4_B200_API int tcq4_b200_reserve(int64_t max_M, int64_t max_N, int64_t max_K);

/* Allocation-free variant of tcq4_b200_mul_mat.  NO counterpart in the reference: added
 * for callers that cannot call cudaMalloc (stream capture, tight serving loops, harnesses that forbid it).
 * The scratch (activation codes, scales, group sums) lives in a file-scope `static __device__` array of
 * TCQ4_B200_STATIC_SCRATCH_BYTES inside the library -- every GPU that loads the library holds it until exit -- whose
 * device address is taken once per device with cudaGetSymbolAddress.  The call performs no cudaMalloc / cudaFree and
 * calls no allocating routine; shapes whose tcq4_b200_workspace_size exceeds the array are refused with
 * TCQ4_ERR_WORKSPACE and a message.  All calls on a device share the array (it keeps the previous call's contents; the
 * few bytes that must be zero are cleared on `stream` by each call): calls that overlap on the GPU -- other streams,
 * concurrently replayed graphs -- RACE; serialise them on one stream.  Same kernels, same results as tcq4_b200_mul_mat. */
#define TCQ4_B200_STATIC_SCRATCH_BYTES (8u << 20)
TCQ4_B200_API int tcq4_b200_mul_mat_static(const void *w_tiles, const float *x, float *y,
                                           int64_t M, int64_t N, int64_t K, const int32_t *d_perm,
                                           cudaStream_t stream);

/* Epilogue fusion: y = W x + bias[n] + residual[m][n] -- the ggml_add nodes behind a MUL_MAT (residual
 * stream after attn_output / ffn_down, ggml_add of a bias tensor) folded into its epilogue.  residual: F32 [M][N] or
 * NULL, may alias y; bias: F32 [N] or NULL.  Equal to MUL_MAT followed by the adds up to fp32 summation order. */
TCQ4_B200_API int tcq4_b200_mul_mat_add(const void *w_tiles, const float *x, float *y, int64_t M, int64_t N, int64_t K,
                                        const int32_t *d_perm, const float *residual, const float *bias,
                                        cudaStream_t stream);

/* Fan-out: n <= 8 weight matrices applied to the SAME activations x [M][K] (Q/K/V, gate/up: consecutive
 * MUL_MAT nodes of the ggml graph that share src1; the fusion hook is ggml-cuda.cu:3481-3688).  y[i] = F32 [M][N[i]].
 * The activation stage runs once; M > 64: one segmented launch of the kind::f16 pair GEMM per group of up to four matrices;
 * 9 <= M <= 64: one launch of the batched-decode GEMM per group of up to four; M <= 8 and n <= 4: ONE GEMV launch over all
 * matrices.  Same int32 accumulators as n separate tcq4_b200_mul_mat calls with the same perm; the fp32 sum over a channel's
 * 256-blocks may be split differently (last-bit differences against separate calls). */
TCQ4_B200_API int tcq4_b200_mul_mat_fanout(int n, const void *const *w_tiles, const int64_t *N, float *const *y,
                                           const float *x, int64_t M, int64_t K, const int32_t *d_perm,
                                           cudaStream_t stream);
/* Same with a leading dimension per output (ldc[i] >= N[i], multiple of 4; NULL: contiguous): y[i] may be a column window of a
 * wider row-major matrix (a ggml view; this GPU's column block of an N-sharded C). */
TCQ4_B200_API int tcq4_b200_mul_mat_fanout_ld(int n, const void *const *w_tiles, const int64_t *N, float *const *y,
                                              const int64_t *ldc, const float *x, int64_t M, int64_t K,
                                              const int32_t *d_perm, cudaStream_t stream);

/* Input-side fusion: the element-wise op that produced the activations in the graph, applied while the activation
 * stage loads them -- one fp32 round trip of [M][K] less.
 *   TCQ4_INPUT_RMS_NORM_MUL: x' = (row_scale[m] * x[m][k]) * col_weight[k]   -- RMSNorm * weight in front of Q/K/V and gate/up
 *       (rms_norm_f32<.., do_multiply>, ggml-cuda/norm.cu:75-150); row_scale = rsqrt(mean(x^2) + eps) from tcq4_b200_rms_row_scale
 *   TCQ4_INPUT_SWIGLU:       x' = (x / (1 + expf(-x))) * x2                  -- SwiGLU in front of down (unary.cu:255-267)
 * Bit-identical to tcq4_b200_apply_input followed by tcq4_b200_mul_mat_fanout_ld (the same expression, evaluated in the load);
 * M <= 8 (decode GEMV) takes that unfused route inside the call.  ldc may be NULL (contiguous outputs). */
enum { TCQ4_INPUT_NONE = 0, TCQ4_INPUT_RMS_NORM_MUL = 1, TCQ4_INPUT_SWIGLU = 2 };
TCQ4_B200_API int tcq4_b200_rms_row_scale(const float *x, float *row_scale, int64_t M, int64_t K, float eps, cudaStream_t stream);
TCQ4_B200_API int tcq4_b200_apply_input(int mode, const float *x, const float *x2, const float *row_scale,
                                        const float *col_weight, float *out, int64_t M, int64_t K, cudaStream_t stream);
TCQ4_B200_API int tcq4_b200_mul_mat_fanout_input(int n, const void *const *w_tiles, const int64_t *N, float *const *y,
                                                 const int64_t *ldc, const float *x, int mode, const float *x2,
                                                 const float *row_scale, const float *col_weight, int64_t M, int64_t K,
                                                 const int32_t *d_perm, cudaStream_t stream);

/* MUL_MAT_ID, the mixture-of-experts form of the same operator: replaces ggml_cuda_mul_mat_id
 * (ggml-cuda.cu:2287-2436) for a TCQ4 src0, which the reference admits in supports_op (:4457-4459) and serves through
 * its generic path (ids to the host, rows sorted by expert, one ggml_cuda_mul_mat per expert).
 *   w_experts : n_expert matrices of [N/8][K/256] tiles, expert e at w_experts + e * expert_stride_bytes (ggml nb02)
 *   x   : F32 [n_tokens][x_slots][K] contiguous, x_slots = 1 (one row per token, broadcast to its experts: gate / up)
 *         or n_used (one row per token and slot: down)                       -- ggml src1 [K, ne11, n_tokens]
 *   ids : device int32, ids[t * ids_stride + s] = expert of slot s of token t  -- ggml ids [n_used, n_tokens]
 *   y   : F32 [n_tokens][n_used][N] contiguous:  y[t][s][:] = W[ids[t][s]] * x[t][s % x_slots][:]
 * No host round trip for any n_tokens, capturable in a CUDA graph (warm the scratch arena up once outside the capture).
 * n_tokens <= 8: the GEMV reads the ids on the device, one launch per token; an id outside [0, n_expert) sets *err_flag --
 * device int, may be NULL -- and computes with expert 0.  More tokens: three launches -- a routing kernel bins the rows by
 * expert into chunks of <= 64 rows, the activation stage reads x through the row map, ONE grouped launch of the batched-decode
 * GEMM serves every (chunk, channel block) pair and scatters the rows into y; a bad id sets *err_flag and the row is skipped.
 * TCQ4_MOE_HOSTSORT=1 in the environment selects the reference's scheme instead (ids to the host: synchronises the stream,
 * not capturable, a bad id returns TCQ4_ERR_ARG). */
TCQ4_B200_API int tcq4_b200_mul_mat_id(const void *w_experts, int64_t n_expert, size_t expert_stride_bytes,
                                       const float *x, int64_t x_slots, const int32_t *ids, int64_t ids_stride,
                                       int64_t n_used, int64_t n_tokens, float *y, int64_t N, int64_t K,
                                       const int32_t *d_perm, int *err_flag, cudaStream_t stream);

/* End-to-end entry with HOST activations/outputs (weights stay resident on the device, as they do
 * in a ggml CUDA buffer): H2D copy of x, MUL_MAT, D2H copy of y, all on `stream`, then a stream
 * synchronise.  Mirrors ggml_backend_tensor_set + graph_compute + ggml_backend_tensor_get. */
/* N-sharded MUL_MAT with the all-gather fused into the epilogue (no counterpart in the fork, whose TCQ4 path is
 * single-GPU: ggml-cuda.cu:2195-2200).  Computes the N output columns of this GPU's weight shard and stores them at
 * column col0 of EVERY destination: y_peers[i] is the base of a row-major fp32 [M][ldc] matrix, normally the same
 * symmetric buffer on each GPU of the node, peer-mapped into this process.  ldc % 4 == 0, col0 % 4 == 0, n_peers <= 8.
 * The caller places one cross-GPU barrier between these stores and the consumers of C. */
TCQ4_B200_API int tcq4_b200_mul_mat_peers(const void *w_tiles, const float *x, float *const *y_peers, int n_peers,
                                          int64_t ldc, int64_t col0, int64_t M, int64_t N, int64_t K,
                                          const int32_t *d_perm, cudaStream_t stream);

/* The same exchange through the NVSwitch: y_multicast is the multicast (NVLS) address of the symmetric C; each store
 * of the epilogue is a multimem.st replicated by the switch into every GPU's copy (outbound traffic 1x). */
TCQ4_B200_API int tcq4_b200_mul_mat_multicast(const void *w_tiles, const float *x, float *y_multicast, int64_t ldc,
                                              int64_t col0, int64_t M, int64_t N, int64_t K, const int32_t *d_perm,
                                              cudaStream_t stream);

/* Copy-engine form of the exchange: after a plain/strided MUL_MAT into the LOCAL C (n_peers = 1 above), push this
 * GPU's column window [col0, col0+ncols) of the row-major [M][ldc] matrix into the same window of a peer's matrix
 * (one cudaMemcpy2DAsync; no SM work, overlaps the next matmul when issued on a side stream). */
TCQ4_B200_API int tcq4_b200_push_columns(float *dst_base, const float *src_base, int64_t M, int64_t ldc, int64_t col0,
                                         int64_t ncols, cudaStream_t stream);
/* Cross-GPU barrier for the exchange (no counterpart in the fork): flags_peers[r] = rank r's peer-mapped int32[9] (8 peer
 * slots + this rank's epoch), zeroed once; every rank calls it after the stores / pushes its consumers depend on.  One
 * 32-thread kernel on `stream` (release / acquire at system scope), capturable. */
#define TCQ4_B200_MAX_PEERS 8
TCQ4_B200_API int tcq4_b200_peer_barrier(int *const *flags_peers, int n_peers, int rank, cudaStream_t stream);

TCQ4_B200_API int tcq4_b200_mul_mat_host(const void *w_tiles_dev, const float *x_host, float *y_host,
                                         int64_t M, int64_t N, int64_t K, const int32_t *d_perm,
                                         cudaStream_t stream);

/* Pipelined form of the same: returns once the work is queued.  Consecutive calls overlap the host->device copy
 * of the next matmul, the kernels of the current one and the device->host copy of the previous one (three
 * library-owned streams, double-buffered device staging), so both PCIe directions and the SMs are busy at once.
 * x_host must stay unchanged and y_host must not be read until tcq4_b200_host_sync() returns TCQ4_OK.
 * Pinned host memory is needed for the copies to be asynchronous.  Not capturable (it owns its streams). */
TCQ4_B200_API int tcq4_b200_mul_mat_host_async(const void *w_tiles_dev, const float *x_host, float *y_host, int64_t M,
                                               int64_t N, int64_t K, const int32_t *d_perm);
/* The same for a fan-out group (n <= 8 consecutive MUL_MAT nodes that share src1: Q/K/V, gate/up): x crosses PCIe once for the
 * group, y_host[i] = F32 [M][N[i]] comes back one copy per matrix.  Uses the device arena for scratch (tcq4_b200_reserve
 * is called for the group's shape); same two-deep pipeline and the same tcq4_b200_host_sync. */
TCQ4_B200_API int tcq4_b200_mul_mat_fanout_host_async(int n, const void *const *w_tiles_dev, const int64_t *N,
                                                      float *const *y_host, const float *x_host, int64_t M, int64_t K,
                                                      const int32_t *d_perm);
TCQ4_B200_API int tcq4_b200_host_sync(void);

/* Stand-alone activation stage.  Replaces tcq4_rrs_fwht_quantize_tc / tcq4_rrs_perm_fwht_quantize_tc
 * (tcq4_k32.cuh:79-94, tcq4_k32.cu:181-277): perm gather -> FWHT-256 -> 1/16 -> max-scale ->
 * int4 codes + per-group sums, written as block_rrs_int4_tc [M][K/256] (160 B each). */
TCQ4_B200_API int tcq4_b200_fwht_quantize_tc(const float *x, void *y_blocks, const int32_t *d_perm,
                                             int64_t K, int64_t M, cudaStream_t stream);

/* ---- quantiser side and GET_ROWS (the producer of the format and the only other op that reads it) ------
 * Device weight quantiser, byte-identical to quantize_tcq4_tile (ggml/src/ggml-quants.c:1528-1639; reached through
 * ggml_quantize_chunk, ggml.c:7681) on the same F32 input: per 8 rows and 256 columns FWHT-256, per-group absmax/7
 * scales, fp16 super-scale, int8 sub-scale codes, int4 codes in mma.m16n8k32 B-fragment order, Z = zc = 0.
 *   w      : F32 [nrow][K] row-major (device), 16-byte aligned;  K % 256 == 0
 *   tiles  : out, [ceil(nrow/8)][K/256] block_tcq4_tile = tcq4_b200_quantize_size(nrow, K) bytes; rows past nrow are
 *            quantised as zeros, as the reference pads its last group (ggml-quants.c:1565-1568)
 *   d_perm : int32[K] device permutation or NULL: quantises W'[n][j] = W[n][perm[j]], i.e. apply_channel_perm_to_weights
 *            (src/llama-quant.cpp:26-40) fused into the load.
 * Asynchronous on `stream`, capture-safe. */
TCQ4_B200_API size_t tcq4_b200_quantize_size(int64_t nrow, int64_t K);
TCQ4_B200_API int tcq4_b200_quantize_tiles(const float *w, void *tiles, int64_t nrow, int64_t K, const int32_t *d_perm,
                                           cudaStream_t stream);

/* derive_tcq4_perms_from_imatrix for one tensor (tools/quantize/quantize.cpp:127-165): inside every 256-block the channel
 * indices sorted by imatrix value, descending.  HOST arrays.  std::sort leaves the order of equal values unspecified in
 * the reference; here ties keep ascending index.  A trailing partial block (K % 256 != 0) is sorted on its own, as there. */
TCQ4_B200_API int tcq4_b200_perm_from_imatrix(const float *imatrix_host, int64_t K, int32_t *perm_host);

/* GET_ROWS for a TCQ4 src0 (ggml_cuda_op_get_rows -> get_rows_cuda, ggml/src/ggml-cuda/getrows.cu:348-380: tile
 * dequantisation k_get_rows_tcq4_tile :96-171 followed by the FWHT-256 that returns the rows to the weight domain; CPU
 * counterpart ggml/src/ggml-cpu/ops.cpp:4800-4824), one fused pass:
 *   dst[i][:] = FWHT256( dequant(row rows[i] of the tensor) ),   dst row stride dst_stride floats (>= K, % 4 == 0).
 *   rows: int32[n_idx] device.  An index outside [0, nrows_src) gives a zero row and, when err_flag (device int) is
 *   non-NULL, sets *err_flag = 1 (the reference reads out of bounds).  Channel permutations are NOT undone, as in the fork.
 * Asynchronous on `stream`, capture-safe. */
TCQ4_B200_API int tcq4_b200_get_rows(const void *w_tiles, const int32_t *rows, float *dst, int64_t n_idx,
                                     int64_t nrows_src, int64_t K, int64_t dst_stride, int *err_flag,
                                     cudaStream_t stream);

/* Verification taps (not on the product path): run the chosen kernel and also dump the integer
 * accumulators.  isum: int32 [M][N][K/256] = sum_g sc[n][kt][g]*dot_g (the TMEM accumulator of the
 * GEMM, and the same quantity formed from the per-group dots in the GEMV).  Nullable. */
TCQ4_B200_API int tcq4_b200_mul_mat_debug(const void *w_tiles, const float *x, float *y,
                                          int64_t M, int64_t N, int64_t K, const int32_t *d_perm,
                                          void *workspace, size_t workspace_bytes, int kernel,
                                          int32_t *isum, cudaStream_t stream);

/* ---- channel-permutation registry ------------------------------------------------------------
 * The three symbols libllama resolves by name through ggml_backend_reg_get_proc_address
 * (ggml/include/ggml-cuda.h:48-50, ggml-cuda.cu:4148-4158, 4876-4885; caller
 * src/llama-model.cpp:7176-7197), same names and signatures.  Registration copies h_perm to the
 * CURRENT device (rrs.cu:82-99); the registry is keyed by (device, name) and guarded by a mutex. */
TCQ4_B200_API void ggml_backend_cuda_tcq4_register_perm(const char *tensor_name, const int32_t *h_perm, int K);
TCQ4_B200_API void ggml_backend_cuda_tcq4_clear_perms(void);
TCQ4_B200_API void ggml_backend_cuda_tcq4_set_reorder_enabled(bool enabled);
/* ggml_cuda_rrs_get_perm / has_perm / get_reorder_enabled (rrs.cu:101-127): device pointer for the
 * current device, or NULL when absent or reordering is disabled. */
TCQ4_B200_API const int32_t *tcq4_b200_get_perm(const char *tensor_name);
TCQ4_B200_API int  tcq4_b200_has_perm(const char *tensor_name);
TCQ4_B200_API int  tcq4_b200_get_reorder_enabled(void);

/* ---- N-sharding (no reference counterpart) ------------------------------------------
 * Rows [n0,n1) of an [N][K] TCQ4 tensor are the contiguous byte range
 * [n0*(K/256)*148, n1*(K/256)*148) (tiles are n-tile major, ggml-quants.c:1550-1573).
 * Splits N over `world` ranks at multiples of `align` channels (align % 8 == 0). */
TCQ4_B200_API int tcq4_b200_shard_range(int64_t N, int64_t K, int world, int rank, int64_t align,
                                        int64_t *n0, int64_t *n1, size_t *byte_offset, size_t *byte_len);

/* ---- host-side mirror of the ggml operator interface -------------------------------------------
 * Minimal view of ggml_tensor (ggml/include/ggml.h:650-690): just the fields
 * ggml_cuda_rrs_mul_mat reads (rrs.cu:247-270).  ne/nb follow ggml (ne[0] fastest). */
typedef struct {
    int         type;        /* ggml_type: 0 = F32, 42 = TCQ4_K32 */
    int64_t     ne[4];
    size_t      nb[4];
    void       *data;        /* device pointer */
    const char *name;        /* tensor name, perm lookup key (rrs.cu:268-270) */
} tcq4_tensor_view;

/* MUL_MAT case of ggml_backend_cuda_device_supports_op for a TCQ4 src0 (ggml-cuda.cu:4397-4459). */
TCQ4_B200_API int tcq4_b200_ggml_supports_op(const tcq4_tensor_view *src0, const tcq4_tensor_view *src1,
                                             const tcq4_tensor_view *dst, int cc);
/* ggml_cuda_rrs_mul_mat (rrs.cu:241-344): shape extraction, asserts (returned as status instead
 * of abort), perm lookup by src0->name, launch on `stream` (= ctx.stream()). */
TCQ4_B200_API int tcq4_b200_ggml_mul_mat(const tcq4_tensor_view *src0, const tcq4_tensor_view *src1,
                                         tcq4_tensor_view *dst, void *workspace, size_t workspace_bytes,
                                         cudaStream_t stream);

/* The fan-out through the operator mirror: n consecutive MUL_MAT nodes (src0[i], src1) -> dst[i] with ONE src1.  Perms are
 * looked up by src0[i]->name; when they are not all equal the nodes are run one by one. */
TCQ4_B200_API int tcq4_b200_ggml_mul_mat_fanout(int n, const tcq4_tensor_view *const *src0, const tcq4_tensor_view *src1,
                                                tcq4_tensor_view *const *dst, cudaStream_t stream);

/* The graph-level entry (the node loop of ggml_cuda_graph_evaluate_and_capture, ggml-cuda.cu:3481-3688): the TCQ4 MUL_MAT
 * nodes of a graph in execution order.  Runs of consecutive nodes with one src1 (Q/K/V, gate/up) become one activation
 * stage + one segmented launch (tcq4_b200_ggml_mul_mat_fanout); the rest go node by node.  TCQ4_NO_GRAPH_FUSION=1 in the
 * environment turns the fusion off. */
TCQ4_B200_API int tcq4_b200_ggml_graph_compute(int n_nodes, const tcq4_tensor_view *const *src0,
                                               const tcq4_tensor_view *const *src1, tcq4_tensor_view *const *dst,
                                               cudaStream_t stream);

/* ggml_cuda_mul_mat_id (ggml-cuda.cu:2287-2436) for a TCQ4 src0: src0 [K, N, n_expert], src1 F32 [K, ne11, n_tokens]
 * with ne11 in {1, n_used}, ids I32 [n_used, n_tokens], dst F32 [N, n_used, n_tokens]; perm looked up by src0->name. */
TCQ4_B200_API int tcq4_b200_ggml_mul_mat_id(const tcq4_tensor_view *src0, const tcq4_tensor_view *src1,
                                            const tcq4_tensor_view *ids, tcq4_tensor_view *dst, int *err_flag,
                                            cudaStream_t stream);

/* ggml_cuda_op_get_rows for a TCQ4 src0 (ggml/src/ggml-cuda/getrows.cu:397-430, 348-380): src1 is GGML_TYPE_I32 (26)
 * [ne10, ne11, ne12] with contiguous dim 0, dst F32 [K, ne10, ne11, ne12]; admitted by supports_op at ggml-cuda.cu:4478-4480. */
TCQ4_B200_API int tcq4_b200_ggml_get_rows(const tcq4_tensor_view *src0, const tcq4_tensor_view *src1,
                                          tcq4_tensor_view *dst, int *err_flag, cudaStream_t stream);

/* Quantiser side: the type llama_tensor_get_type picks for a tensor under LLAMA_FTYPE_MOSTLY_TCQ4_K32
 * (src/llama-quant.cpp:204-526): TCQ4_K32 (42) for 2-D weights with K % 256 == 0 and N % 8 == 0, Q4_K (12) for the token
 * embeddings and as the shape fallback, Q6_K (K % 256 == 0, the block size of the ftype's default type, ggml.c:750-753) or
 * Q8_0 for the output matrix.  Returns a ggml_type id, -2 for tensors the reference does not quantise (1-D tensors, norms,
 * the MoE router: they keep their source type), or -1 where the reference throws. */
TCQ4_B200_API int tcq4_b200_quant_type_for(const char *name, int n_dims, int64_t K, int64_t N, int has_output);

/* ---- GGUF files with TCQ4 tensors and permutations (host side) ------------------------------------------
 * The loader's scan for "tcq4.reorder.enabled" / "tcq4.<tensor>.perm" and the registration flow
 * (src/llama-model.cpp:7129-7205), the verbatim upload of tensor bytes (ggml-cuda.cu:627-633), and the writer side of
 * llama_model_quantize for these keys (src/llama-quant.cpp:840-854).  GGUF v2/v3 (ggml/src/gguf.cpp), little-endian. */
typedef struct tcq4_gguf tcq4_gguf;
TCQ4_B200_API int     tcq4_b200_gguf_open(const char *path, tcq4_gguf **out);
TCQ4_B200_API void    tcq4_b200_gguf_close(tcq4_gguf *g);
TCQ4_B200_API int     tcq4_b200_gguf_reorder_enabled(const tcq4_gguf *g);             /* -1: key absent, else 0 / 1 */
TCQ4_B200_API int64_t tcq4_b200_gguf_n_tensors(const tcq4_gguf *g);
TCQ4_B200_API int64_t tcq4_b200_gguf_n_perms(const tcq4_gguf *g);
/* type = ggml_type (42 = TCQ4_K32); ne as ggml (ne[0] = K fastest); file_offset absolute; nbytes exact for F32, F16 and
 * TCQ4_K32 (ne[1] * K/256 * 148), else the extent up to the next tensor.  Returned strings live until close. */
TCQ4_B200_API int     tcq4_b200_gguf_tensor_info(const tcq4_gguf *g, int64_t i, const char **name, int *type, int64_t ne[4],
                                                 uint64_t *file_offset, uint64_t *nbytes);
TCQ4_B200_API int     tcq4_b200_gguf_perm(const tcq4_gguf *g, int64_t i, const char **tensor_name, const int32_t **perm,
                                          int64_t *K);
/* set_reorder_enabled + register_perm for every permutation of the file on the CURRENT device (call once per device: the
 * fork registers with the first GPU backend only, llama-model.cpp:7200) */
TCQ4_B200_API int     tcq4_b200_gguf_register_perms(const tcq4_gguf *g);
TCQ4_B200_API int     tcq4_b200_gguf_read_tensor(const tcq4_gguf *g, int64_t i, void *dst_host, size_t dst_bytes);
/* tensor bytes -> device memory through pinned staging, synchronous on return */
TCQ4_B200_API int     tcq4_b200_gguf_load_tensor(const tcq4_gguf *g, int64_t i, void *dst_device, size_t dst_bytes,
                                                 cudaStream_t stream);
/* Writer: 2-D tensors of type F32 (0), F16 (1) or TCQ4_K32 (42) from host memory, "tcq4.<tensor>.perm" int32 arrays and
 * "tcq4.reorder.enabled" (reorder_enabled < 0: key omitted), general.alignment = 32. */
TCQ4_B200_API int     tcq4_b200_gguf_write(const char *path, int n_tensors, const char *const *names, const int *types,
                                           const int64_t *ne0, const int64_t *ne1, const void *const *data_host, int n_perms,
                                           const char *const *perm_tensors, const int32_t *const *perms,
                                           const int64_t *perm_K, int reorder_enabled);

/* Per-launch timing tap for the roofline leg of bench.py (no reference counterpart; the fork's
 * ggml_cuda_rrs_benchmark, rrs.cu:350-419, times with CUDA events the same way): between begin and
 * end every tcq4_b200_mul_mat* call records CUDA events on its stream around the activation kernel and
 * around the main kernel.  end() synchronises the device and returns, for launches whose main kernel
 * was `kernel` (TCQ4_KERNEL_AUTO = all), the summed main-kernel and activation-kernel milliseconds.
 * Not thread-safe, not usable under stream capture. */
TCQ4_B200_API int tcq4_b200_profile_begin(int max_launches);
TCQ4_B200_API int tcq4_b200_profile_end(int kernel, double *main_ms_sum, double *pre_ms_sum, int *launches);

/* Launch statistics since load (kernels launched by this library), for bench.py's gpu_launches. */
TCQ4_B200_API uint64_t tcq4_b200_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* TCQ4_B200_H */
