// tcq4_kernels.h -- internal launch interface between the C-ABI layer (tcq4_api.cu) and the kernels.
#pragma once

#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

#include "tcq4_common.cuh"

struct CUtensorMap_st;      // <cuda.h>

namespace tcq4 {

// ---- K1: activation stage (tcq4_act.cu) -------------------------------------------------------
// Input-side fusion: the element-wise op that produced the activations in the graph, applied while K1 loads them
// instead of in a kernel of its own (one fp32 round trip of [M][K] less).  Expressions and operation order are the ggml-cuda
// kernels': rms_norm_f32<.., do_multiply> (norm.cu:75-150: scale * x * mul) and unary_gated_op_kernel<op_silu> (unary.cu:255-267,
// unary.cuh:92-94: x / (1 + expf(-x)) * g).
enum { ACT_IN_NONE = 0, ACT_IN_NORM = 1, ACT_IN_SWIGLU = 2 };
struct ActInput {
    int          mode;
    const float *x2;            // SWIGLU: the second operand ("up"), [M][K] like x ("gate")
    const float *row_scale;     // NORM: rsqrt(mean(x^2) + eps) per row, [M]
    const float *col_weight;    // NORM: the norm's weight, [K]
};
// the transform on its own (what the fused load computes, element for element): out [M][K]
cudaError_t launch_apply_input(const ActInput &in, const float *x, float *out, int64_t M, int64_t K, int num_sms, cudaStream_t stream);
// rsqrt(mean(x^2) + eps) per row with the reduction order of rms_norm_f32 (norm.cu:125-137)
cudaError_t launch_rms_row_scale(const float *x, float *row_scale, int64_t M, int64_t K, float eps, cudaStream_t stream);
// the transform the NEXT activation-stage launches of this host thread apply to their input (nullptr: none); set and cleared by
// the fused entry points of tcq4_api.cu around their calls
void set_pending_act_input(const ActInput *in);

cudaError_t launch_act_quant_ref(const float *x, const int32_t *perm, int64_t M, int64_t K, void *out_blocks,
                                 int num_sms, cudaStream_t stream);
cudaError_t launch_act_quant_stage(const float *x, const int32_t *perm, int64_t M, int64_t K, ActStage st,
                                   int num_sms, cudaStream_t stream);
// nibble staging for tcq4_gemm3.cu; zero_y / zero_floats (a multiple of 4, 16-byte aligned): an output matrix the GEMM
// behind this launch accumulates into (K split), cleared by the same launch
cudaError_t launch_act_quant_nib(const float *x, const int32_t *perm, int64_t M, int64_t K, ActStage st, int num_sms,
                                 cudaStream_t stream, float *zero_y = nullptr, int64_t zero_floats = 0);
cudaError_t launch_zero_floats(float *y, int64_t n_floats, int num_sms, cudaStream_t stream);
// the same for up to four strided regions (rows x cols floats, row stride ld; cols and ld multiples of 4, p 16-byte aligned):
// the outputs of a fan-out group whose blocks are cut (tcq4_gemm4.cu), or its arrival counters
struct ZeroList {
    float    *p[4];
    long long rows[4], cols[4], ld[4];
    int       n;
};
cudaError_t launch_act_quant_nib_zl(const float *x, const int32_t *perm, int64_t M, int64_t K, ActStage st, int num_sms,
                                    cudaStream_t stream, const ZeroList *zl);
cudaError_t launch_zero_list(const ZeroList &zl, int num_sms, cudaStream_t stream);
// MUL_MAT_ID (device-routed): max_rows virtual rows, the first *n_chunks * 64 exist, virtual row r reads x[row_map[r]] (-1: skip)
cudaError_t launch_act_quant_nib_rows(const float *x, const int32_t *perm, int64_t max_rows, int64_t K, ActStage st, int num_sms,
                                      cudaStream_t stream, const int32_t *row_map, const int *n_chunks);

// ---- K3: fused decode GEMV (tcq4_gemv.cu) -----------------------------------------------------
constexpr int GEMV_MAX_M = 16;
struct GemvPlan;
bool gemv_uses_prepass(const GemvPlan &plan, const int32_t *perm);
struct GemvPlan {
    int     grid;            // CTAs (persistent, one per SM)
    int     warps;           // warps per CTA (16 or 32)
    int     depth;           // tiles in flight per warp (ring slots)
    bool    prepass;         // activation stage in its own launch (M * K/256 large), result in the workspace blob
    bool    prepass_perm;    // the same when the call has a channel permutation (gathers: pays from fewer blocks, M = 1 included)
    size_t  blob_bytes;      // workspace bytes of that result
    size_t  smem_bytes;
    int     maxslots;        // split-K partial slots per n-tile row
    size_t  partial_bytes;   // workspace: float [NT][maxslots][8 or 16 rows][8]
    size_t  ticket_bytes;    // workspace: int   [NT]
};
bool        gemv_plan(int64_t M, int64_t N, int64_t K, int num_sms, size_t smem_optin, GemvPlan *plan);
cudaError_t gemv_init_attributes(size_t smem_optin);
// MUL_MAT_ID decode: the weight matrix of a launch is chosen ON THE DEVICE, w_tiles + sel[0] * stride, so that the
// router's output never travels to the host (sel == nullptr: plain MUL_MAT).  An id outside [0, n) sets *err and
// computes with expert 0 (the reference asserts: ggml-cuda.cu:2357).
struct GvExpert {
    const int32_t *sel;
    long long stride;      // bytes between expert matrices (ggml nb02)
    int n;
    int *err;
    int pdl;               // host side: this launch may use programmatic serialisation (see tcq4_b200_mul_mat_id)
    // grouped launch (launch_gemv_grouped): ONE launch for all slots of a token.  The kernel sees a virtual matrix of
    // n_used * N channels; channel block nt belongs to slot nt / nt_expert, whose expert is sel[slot] and whose
    // activation row is slot % xslots.
    int xslots;            // activation rows held in shared memory (1 or n_used); 0 = not grouped
    int nt_expert;         // N / 8 of one expert
    // fan-out launch (launch_gemv with nseg > 0): ONE launch for up to GV_MAX_SEG weight matrices over the same activations.
    // The kernel sees sum(N_i) channels; channel block nt in [seg_nt0[i], seg_nt0[i+1]) belongs to matrix i.
    int nseg;
    int seg_nt0[5];
    const uint8_t *seg_w[4];
    float *seg_y[4];       // [M][seg_ldc[i]] each (N_i columns used)
    long long seg_ldc[4];
    // epilogue fusion (tcq4_b200_mul_mat_add, MODE 3 of the kernel): y = W x + bias[n] + res[m][n]; res may alias y
    const float *res;      // [M][N] contiguous or null
    const float *bias;     // [N] or null
};
constexpr int GV_MAX_SEG = 4;
// plan: gemv_plan(x_slots, n_used * N, K); y: [n_used][N] contiguous
cudaError_t launch_gemv_grouped(const void *w_experts, const float *x, const int32_t *perm, float *y, int64_t n_used,
                                int64_t N, int64_t K, const GemvPlan &plan, const GvExpert &expert, cudaStream_t stream);
cudaError_t launch_gemv(const void *w_tiles, const float *x, const int32_t *perm, YOut y, int64_t M, int64_t N,
                        int64_t K, const GemvPlan &plan, float *ws_blob, int *ws_ticket, int32_t *dbg_isum,
                        cudaStream_t stream, const GvExpert *expert = nullptr);

// ---- cross-GPU barrier of the N-shard exchange (tcq4_peer.cu): flags[r] = rank r's peer-mapped int32[MAX_PEERS + 1], zeroed once
cudaError_t launch_peer_barrier(int *const *flags, int n, int rank, cudaStream_t stream);

// ---- MUL_MAT_ID row gather / scatter (tcq4_moe.cu) ------------------------------------------------
cudaError_t launch_copy_rows(const float *src, const int32_t *src_idx, float *dst, const int32_t *dst_idx, int64_t rows,
                             int64_t cols, int num_sms, cudaStream_t stream);
// device-side routing (one CTA): rows binned by expert into chunks of <= 64 rows at 64-aligned virtual rows; see tcq4_moe.cu
cudaError_t launch_moe_route(const int32_t *ids, int64_t ids_stride, int64_t n_used, int64_t n_tokens, int64_t n_expert,
                             int64_t x_slots, int max_chunks, int *n_chunks, int *chunk_expert, int32_t *src_map,
                             int32_t *dst_map, int32_t *pos_tmp, int *err_flag, cudaStream_t stream);

// ---- K4 / K5: device weight quantiser and GET_ROWS (tcq4_quant.cu) ------------------------------
cudaError_t launch_quantize_tiles(const float *w, const int32_t *perm, void *tiles, int64_t nrow, int64_t K, int num_sms,
                                  cudaStream_t stream);
cudaError_t launch_get_rows(const void *tiles, const int32_t *rows, float *dst, int64_t n_idx, int64_t nrows_src,
                            int64_t K, int64_t dst_stride, int *err_flag, int num_sms, cudaStream_t stream);

cudaError_t gemm2_read_prof(long long *host_out, size_t bytes);            // PROF-instantiation timeline stamps
cudaError_t gemv_read_prof(unsigned long long *host_out, size_t bytes);   // DEBUG-instantiation phase stamps

// ---- K2: tcgen05 prefill GEMM (tcq4_gemm.cu) --------------------------------------------------
struct GemmPlan {
    dim3    grid;
    size_t  smem_bytes;
    size_t  stage_codes_bytes;   // workspace: ActStage.codes
    size_t  stage_a7_bytes;
    size_t  stage_sumq_bytes;
};
bool        gemm_plan(int64_t M, int64_t N, int64_t K, int num_sms, size_t smem_optin, GemmPlan *plan);
cudaError_t gemm_init_attributes(size_t smem_optin);
cudaError_t launch_gemm(const void *w_tiles, ActStage st, float *y, int64_t M, int64_t N, int64_t K,
                        const GemmPlan &plan, int32_t *dbg_isum, cudaStream_t stream);

// ---- K2 on CTA pairs, M > 128 (tcq4_gemm2.cu) --------------------------------------------------
struct Gemm2Plan {
    dim3    grid;            // (2 * ceil(M/256), ceil(NT / (2*ntl))), clusters of 2 along x
    size_t  smem_bytes;
    int     ntl;             // n-tiles (8 channels) unpacked per CTA per pair-tile
};
bool        gemm2_plan(int64_t M, int64_t N, int64_t K, int num_sms, size_t smem_optin, Gemm2Plan *plan);
cudaError_t gemm2_init_attributes(size_t smem_optin);
// fuse_x != null: the kernel's own warps run the activation stage (x, perm) into `st` first; fuse_cnt = two zeroed ints
cudaError_t launch_gemm2(const void *w_tiles, ActStage st, YOut y, int64_t M, int64_t N, int64_t K,
                         const Gemm2Plan &plan, int32_t *dbg_isum, cudaStream_t stream, const float *fuse_x = nullptr,
                         const int32_t *fuse_perm = nullptr, int *fuse_cnt = nullptr, const float *res = nullptr,
                         const float *bias = nullptr);

// ---- K2 (round 2): kind::f16 pair GEMM with in-SM nibble conversion, fan-out segments, K split (tcq4_gemm3.cu) ----
struct Gemm3Seg {
    const void *w;       // [N/8][K/256] tiles
    float      *y;       // [M][ldc]
    int64_t     N, ldc;
};
struct Gemm3Plan {
    dim3    grid;
    size_t  smem_bytes;
    int     ntl;         // n-tiles per CTA (12: 192 channels per pair-tile)
    int     ksplit;      // 1, or 2: the two K halves of a tile meet in y through a TMA reduce-add (y must start at zero)
    int     num_tiles;
};
bool        gemm3_plan(int64_t M, const int64_t *Ns, int nseg, int64_t K, int num_sms, size_t smem_optin, bool allow_split,
                       Gemm3Plan *plan);
cudaError_t gemm3_init_attributes(size_t smem_optin);
cudaError_t gemm3_read_prof(long long *host_out, size_t bytes);           // PROF-instantiation timeline stamps (diagnostics build)
cudaError_t launch_gemm3(const Gemm3Seg *segs, int nseg, ActStage st, int64_t M, int64_t K, const Gemm3Plan &plan,
                         int32_t *dbg_isum, cudaStream_t stream, const float *res = nullptr, const float *bias = nullptr);

// cached cuTensorMapEncodeTiled of a TCQ4 tensor seen as [NT rows][KT * 148 eight-byte words] (tcq4_gemm3.cu); box in words x rows
bool weight_tensor_map(::CUtensorMap_st *map, const void *w, int64_t NT, int64_t KT, uint32_t box_words, uint32_t box_rows);

// ---- K2c: batched-decode GEMM, 9 <= M <= 64 (tcq4_gemm4.cu) -----------------------------------------------------
constexpr int GEMM4_MAX_PARTS = 8;
constexpr int GEMM4_MAX_CTAS = 160;        // bound of blocks * parts (workspace sizing is device independent)
constexpr int GEMM4_MAX_SEG = 4;           // matrices of a fan-out group served by one launch
struct Gemm4Seg {
    const void *w;       // tiles of this matrix
    float      *y;       // its output, row stride ldc floats
    int64_t     N, ldc;
};
struct Gemm4Plan {
    int    grid;            // CTAs = contiguous ranges of the blocks * KT line
    size_t smem_bytes;
    int    mt;              // tokens per MMA (32 or 64)
    int    num_blocks;      // 128-channel blocks of all segments
    int    mode;            // 0: every range is whole blocks (plain stores); 1: a block has at most two parts, they meet in y through
                            //    red.global.add.f32 -- y must start at zero; 2: `parts` K ranges per block, partial sums in workspace
                            //    slots, the last part to arrive adds them in order -- the counters must start at zero
    int    parts;
    size_t partial_bytes;   // mode 2: one [mt][128] fp32 slot per CTA
    size_t counter_bytes;   // mode 2: 4 ints per block
};
// no_atomic: never let two parts of a block meet in y through red.add (a fused residual may sit there)
bool        gemm4_plan(int64_t M, const int64_t *Ns, int nseg, int64_t K, int num_sms, size_t smem_optin, Gemm4Plan *plan,
                       bool no_atomic = false);
cudaError_t launch_gemm4_id(const void *w_experts, int64_t n_expert, size_t expert_stride_bytes, ActStage st, float *y, int64_t N,
                            int64_t K, const int *n_chunks, const int *chunk_expert, const int32_t *dst_map, int max_chunks,
                            int num_sms, cudaStream_t stream);
cudaError_t gemm4_init_attributes(size_t smem_optin);
cudaError_t gemm4_read_prof(long long *host_out, size_t bytes);           // timeline stamps (diagnostics build)
// st: MODE_NIB activation stage of the M <= 64 tokens (nibble units + a7t); y row stride ldc floats
// res ([M][N] contiguous, may alias the output) / bias ([N]): fused into the epilogue, one segment only, plan made with no_atomic
cudaError_t launch_gemm4(const Gemm4Seg *segs, int nseg, ActStage st, int64_t M, int64_t K, const Gemm4Plan &plan,
                         float *partials, int *counters, int32_t *dbg_isum, cudaStream_t stream, const float *res = nullptr,
                         const float *bias = nullptr);

}  // namespace tcq4
