/* oracle/ref_baseline.c -- TEST INFRASTRUCTURE ONLY (see tcq4_oracle.h).
 *
 * CPU baseline driver for bench.py: runs the REFERENCE's own ggml CPU backend F32 MUL_MAT
 * (ggml_mul_mat through ggml_graph_compute_with_ctx, /root/reference/ggml/include/ggml-cpu.h:71)
 * on caller-provided buffers.  Linked against oracle/_ref/libggml_ref.so, which is the reference's
 * sources compiled where they lie (oracle/Makefile `make ref`).  This file contains no reference
 * code: it only calls the public ggml C API the way SURVEY 8(d) "CPU baseline" step (4) describes.
 *
 *   W : F32 [N][K]  (Hadamard-domain weights from dequantize_row_tcq4_tile)
 *   x : F32 [M][K]  (FWHT'd, permuted activations)
 *   y : F32 [M][N]
 */
#include <stddef.h>
#include <stdint.h>
#include <string.h>

#include "ggml.h"
#include "ggml-cpu.h"

int refbase_mul_mat_f32(const float *W, const float *x, float *y, int64_t M, int64_t N, int64_t K, int n_threads) {
    /* no_alloc context: tensors point straight at the caller's buffers, only metadata + the
     * graph + ggml's work buffer live in the context */
    size_t work = (size_t)n_threads * (size_t)(K + 64) * sizeof(float) * 4 + (size_t)(M * K) * sizeof(float) + (64u << 20);
    struct ggml_init_params ip;
    ip.mem_size = 8 * ggml_tensor_overhead() + ggml_graph_overhead() + work;
    ip.mem_buffer = NULL;
    ip.no_alloc = 1;
    struct ggml_context *ctx = ggml_init(ip);
    if (!ctx) return 1;
    struct ggml_tensor *tw = ggml_new_tensor_2d(ctx, GGML_TYPE_F32, K, N);
    struct ggml_tensor *tx = ggml_new_tensor_2d(ctx, GGML_TYPE_F32, K, M);
    tw->data = (void *)W;
    tx->data = (void *)x;
    struct ggml_tensor *ty = ggml_mul_mat(ctx, tw, tx);   /* [N, M] : y[m*N + n] */
    ty->data = (void *)y;
    struct ggml_cgraph *gf = ggml_new_graph(ctx);
    ggml_build_forward_expand(gf, ty);
    ggml_set_no_alloc(ctx, 0);                            /* the work buffer is a real allocation */
    enum ggml_status st = ggml_graph_compute_with_ctx(ctx, gf, n_threads);
    ggml_free(ctx);
    return st == GGML_STATUS_SUCCESS ? 0 : 2;
}
