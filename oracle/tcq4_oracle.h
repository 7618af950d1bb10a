/*
 * tcq4_oracle.h -- scalar CPU restatement of the TCQ4_K32 W4A4 MUL_MAT path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product: only
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load this library, and only as the checker / reported CPU baseline.
 * The product path (llama.cpp-rrs_b200/csrc) never links or calls it.
 *
 * Parity pin: the reference holds no golden vectors for this path, so
 * the restatement is pinned against the reference's own CPU code compiled from
 * /root/reference into oracle/_ref/libtcq4_ref.so (quantize_tcq4_tile,
 * dequantize_row_tcq4_tile, ggml_fwht_impl) -- byte-identical tiles / bit-identical
 * floats, see tests/test_oracle_vs_ref.py -- and against the committed golden
 * fixtures under tests/golden/ generated from that library.
 *
 * All paths cited are relative to /root/reference.
 * Build with -ffp-contract=off: every float op below is a single IEEE-754 op.
 */
#ifndef TCQ4_ORACLE_H
#define TCQ4_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ORC_TILE_K        256   /* ggml/src/ggml-common.h:315 */
#define ORC_TILE_CHANNELS 8     /* ggml/src/ggml-common.h:316 */
#define ORC_TILE_GROUPS   8     /* ggml/src/ggml-common.h:317 */
#define ORC_GROUP_SIZE    32    /* ggml/src/ggml-common.h:318 */
#define ORC_TILE_BYTES    1184  /* ggml/src/ggml-common.h:341 */

/* ggml/src/ggml-common.h:321-340 (block_tcq4_tile), byte-identical layout */
typedef struct {
    uint8_t  tiles[8][128];
    uint16_t S[8];      /* fp16 bits */
    uint16_t Z[8];      /* fp16 bits */
    int8_t   sc[8][8];
    int8_t   zc[8][8];
} orc_tile;

/* ggml/src/ggml-cuda/tcq4_k32.cuh:51-56 (block_rrs_int4_tc), 160 bytes */
typedef struct {
    uint8_t qs[128];
    float   smooth_scale;
    int16_t sum_q[8];
    uint8_t pad[12];
} orc_act_block;

uint16_t orc_fp32_to_fp16(float f);           /* ggml/src/ggml-impl.h:397-421 (RNE) */
float    orc_fp16_to_fp32(uint16_t h);        /* ggml/src/ggml-impl.h:374-395 */

/* ggml/src/ggml-cpu/rrs.c:41-123, scalar branch 103-121 */
void orc_fwht(float *data, int n);

/* ggml/src/ggml-quants.c:1381-1400 / 1403-1420 */
void orc_pack_group(const int8_t w[8][32], uint8_t packed[128]);
void orc_unpack_group(const uint8_t packed[128], int8_t w[8][32]);

/* ggml/src/ggml-quants.c:1528-1639 quantize_tcq4_tile; returns bytes written for nrow rows */
size_t orc_quantize_tiles(const float *src, void *dst, int64_t nrow, int64_t n_per_row);

/* ggml/src/ggml-quants.c:1425-1463 dequantize_row_tcq4_tile: y[t*2048 + c*256 + k], k elems total = ntiles*256 */
void orc_dequant_tiles(const void *tiles, float *y, int64_t k);

/* ggml/src/ggml-cuda/getrows.cu:96-171 + 372-378 (CPU: ggml-cpu/ops.cpp:4800-4824): dst[i][K] =
 * FWHT256(dequant(row rows[i])) */
void orc_get_rows(const void *tiles, const int32_t *rows, int64_t n_idx, int64_t K, float *dst);

/* src/llama-quant.cpp:26-40 apply_channel_perm_to_weights: dst[r][c] = src[r][perm[c]] */
void orc_apply_perm_weights(const float *src, float *dst, int64_t nrows, int64_t ncols, const int32_t *perm);

/* tools/quantize/quantize.cpp:127-165 derive_tcq4_perms_from_imatrix (one tensor):
 * per 256-block indices sorted by imatrix value descending (std::sort is not stable; ties are
 * broken here by ascending index, callers must not rely on tie order) */
void orc_perm_from_imatrix(const float *imatrix, int64_t K, int32_t *perm);

/* ggml/src/ggml-cuda/tcq4_k32.cu:771-831 (== 197-276): one activation row.
 *   codes[K] int8 in [-7,7]; max_abs[K/256]; sum_q[K/256][8].  perm may be NULL. */
void orc_quantize_act_row(const float *x_row, const int32_t *perm, int64_t K,
                          int8_t *codes, float *max_abs, int16_t *sum_q);

/* same, emitting the reference's block_rrs_int4_tc array [M][K/256]
 * (ggml/src/ggml-cuda/tcq4_k32.cu:262-276) */
void orc_quantize_act_blocks(const float *x, const int32_t *perm, int64_t M, int64_t K, orc_act_block *out);

/* ggml/src/ggml-cuda/tcq4_k32.cu:833-887: full W4A4 matmul, C[m*N+n].
 *   dots  (nullable): int32 [M][N][KT][8]   per-group integer dots
 *   isum  (nullable): int32 [M][N][KT]      sum_g sc[n][kt][g]*dot  (the tcgen05 accumulator)
 *   codes (nullable): int8  [M][K]; max_abs (nullable) [M][KT]; sum_q (nullable) [M][KT][8] */
void orc_mul_mat(const void *w_tiles, const float *x, const int32_t *perm,
                 int64_t M, int64_t N, int64_t K, float *C,
                 int32_t *dots, int32_t *isum, int8_t *codes, float *max_abs, int16_t *sum_q);

/* O2: C = dequant(W_tiles) . FWHT(perm(x)), all F32, no activation quantisation.
 * This is the arithmetic of the CPU baseline (dequant + FWHT + F32 MUL_MAT). */
void orc_mul_mat_f32(const void *w_tiles, const float *x, const int32_t *perm,
                     int64_t M, int64_t N, int64_t K, float *C);

/* 64-bit FNV-1a over a byte range, for checksum-of-checksums at full sizes */
uint64_t orc_fnv1a64(const void *p, size_t n);

#ifdef __cplusplus
}
#endif
#endif
