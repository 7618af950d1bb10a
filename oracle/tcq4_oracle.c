/*
 * tcq4_oracle.c -- scalar CPU restatement of the TCQ4_K32 W4A4 MUL_MAT path.
 * TEST INFRASTRUCTURE ONLY -- see tcq4_oracle.h for the rules and the parity pin.
 * Compile: gcc -O2 -ffp-contract=off -fno-fast-math -fPIC -shared
 */
#include "tcq4_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------------------------------
 * fp16 <-> fp32 (IEEE binary16, round-to-nearest-even) -- same function as
 * ggml/src/ggml-impl.h:374-421, written directly on the bit fields.
 * ---------------------------------------------------------------------------------------- */
static uint32_t f2u(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
static float u2f(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }

uint16_t orc_fp32_to_fp16(float f) {
    const uint32_t u = f2u(f);
    const uint32_t sign = (u >> 16) & 0x8000u;
    const uint32_t absu = u & 0x7FFFFFFFu;
    if (absu > 0x7F800000u) return (uint16_t)(sign | 0x7E00u);      /* NaN -> canonical qNaN */
    if (absu >= 0x47800000u) {                                       /* |f| >= 65536 (incl. inf) */
        return (uint16_t)(sign | 0x7C00u);
    }
    if (absu >= 0x477FF000u) return (uint16_t)(sign | 0x7C00u);      /* rounds up to inf */
    if (absu < 0x33000001u) return (uint16_t)sign;                   /* <= 2^-25 -> 0 (tie to even) */
    int32_t e = (int32_t)(absu >> 23) - 127;                         /* unbiased exponent */
    uint32_t man = (absu & 0x7FFFFFu) | 0x800000u;                   /* 24-bit significand */
    uint32_t shift, half;
    uint32_t out;
    if (e < -14) {                                                   /* fp16 subnormal */
        shift = (uint32_t)(13 + (-14 - e));                          /* 14..24 */
        uint32_t q = man >> shift;
        uint32_t rem = man & ((1u << shift) - 1u);
        half = 1u << (shift - 1);
        if (rem > half || (rem == half && (q & 1u))) q++;
        out = q;                                                     /* may carry into exponent 1: fine */
    } else {
        uint32_t q = man >> 13;                                      /* 11 bits incl. hidden */
        uint32_t rem = man & 0x1FFFu;
        if (rem > 0x1000u || (rem == 0x1000u && (q & 1u))) q++;
        out = ((uint32_t)(e + 15) << 10) + (q - 0x400u);             /* carry propagates into exponent */
    }
    return (uint16_t)(sign | out);
}

float orc_fp16_to_fp32(uint16_t h) {
    const uint32_t sign = ((uint32_t)h & 0x8000u) << 16;
    const uint32_t e = (h >> 10) & 0x1Fu;
    const uint32_t m = h & 0x3FFu;
    if (e == 0) {
        if (m == 0) return u2f(sign);
        /* subnormal: m * 2^-24 */
        float v = (float)m * 0x1.0p-24f;
        return u2f(sign | f2u(v));
    }
    if (e == 31) return u2f(sign | 0x7F800000u | (m << 13));
    return u2f(sign | ((e + 112u) << 23) | (m << 13));
}

/* ------------------------------------------------------------------------------------------
 * FWHT -- ggml/src/ggml-cpu/rrs.c:41-123.  The scalar branch (103-121) defines the order:
 * stages h = 1, 2, 4, ... n/2; pairs (j, j+h) -> (x+y, x-y); the last stage also multiplies by
 * 1/sqrtf(n).  The AVX branches do the same adds on the same pairs, so results are bit-equal.
 * ---------------------------------------------------------------------------------------- */
void orc_fwht(float *data, int n) {
    if (n <= 1) return;
    const float scale = 1.0f / sqrtf((float)n);
    const int last_h = n >> 1;
    for (int h = 1; h < n; h <<= 1) {
        const int stride = h << 1;
        for (int i = 0; i < n; i += stride) {
            for (int j = i; j < i + h; j++) {
                const float x = data[j];
                const float y = data[j + h];
                if (h == last_h) {
                    data[j]     = (x + y) * scale;
                    data[j + h] = (x - y) * scale;
                } else {
                    data[j]     = x + y;
                    data[j + h] = x - y;
                }
            }
        }
    }
}

/* ------------------------------------------------------------------------------------------
 * Fragment order -- ggml/src/ggml-quants.c:1363-1420.
 * "lane" L = c*4 + s holds the 8 nibbles of channel c, K elements s*8 .. s*8+7, element i at
 * bits [4i, 4i+4) of the little-endian uint32 at byte offset L*4.  Equivalent closed form
 * (SURVEY section 4): element (c,k) -> byte c*16 + (k/8)*4 + (k%8)/2, low nibble iff k even.
 * ---------------------------------------------------------------------------------------- */
void orc_pack_group(const int8_t w[8][32], uint8_t packed[128]) {
    for (int lane = 0; lane < 32; lane++) {
        const int c = lane / 4, s = lane % 4;
        uint32_t reg = 0;
        for (int i = 0; i < 8; i++) reg |= ((uint32_t)(w[c][s * 8 + i] & 0xF)) << (4 * i);
        packed[lane * 4 + 0] = (uint8_t)(reg);
        packed[lane * 4 + 1] = (uint8_t)(reg >> 8);
        packed[lane * 4 + 2] = (uint8_t)(reg >> 16);
        packed[lane * 4 + 3] = (uint8_t)(reg >> 24);
    }
}

void orc_unpack_group(const uint8_t packed[128], int8_t w[8][32]) {
    for (int c = 0; c < 8; c++) {
        for (int k = 0; k < 32; k++) {
            const uint8_t b = packed[c * 16 + (k / 8) * 4 + (k % 8) / 2];
            const int v = (k & 1) ? (b >> 4) : (b & 0xF);
            w[c][k] = (int8_t)(v >= 8 ? v - 16 : v);
        }
    }
}

/* ------------------------------------------------------------------------------------------
 * Weight quantiser -- ggml/src/ggml-quants.c:1528-1639 (quantize_tcq4_tile).
 * ---------------------------------------------------------------------------------------- */
size_t orc_quantize_tiles(const float *src, void *dst, int64_t nrow, int64_t n_per_row) {
    const int64_t KT = n_per_row / ORC_TILE_K;
    const int64_t full = nrow / 8, rem = nrow % 8;
    const int64_t ngroups = full + (rem ? 1 : 0);
    float *scratch = (float *)malloc((size_t)8 * n_per_row * sizeof(float));
    orc_tile *out = (orc_tile *)dst;
    int64_t row = 0, ti = 0;
    for (int64_t rg = 0; rg < ngroups; rg++) {
        const int rows = rg < full ? 8 : (int)rem;
        for (int c = 0; c < 8; c++) {
            float *r = scratch + (size_t)c * n_per_row;
            if (c < rows) {
                memcpy(r, src + (size_t)(row + c) * n_per_row, (size_t)n_per_row * sizeof(float));
                for (int64_t i = 0; i < n_per_row; i += ORC_TILE_K) orc_fwht(r + i, ORC_TILE_K);   /* :1563 */
            } else {
                memset(r, 0, (size_t)n_per_row * sizeof(float));                                   /* :1567 */
            }
        }
        for (int64_t kt = 0; kt < KT; kt++) {
            orc_tile *t = &out[ti++];
            float scales[8][8];
            for (int c = 0; c < 8; c++) {                                                          /* :1576-1588 */
                const float *rd = scratch + (size_t)c * n_per_row + kt * ORC_TILE_K;
                for (int g = 0; g < 8; g++) {
                    float mx = 0.0f;
                    for (int j = 0; j < 32; j++) {
                        const float v = fabsf(rd[g * 32 + j]);
                        if (v > mx) mx = v;
                    }
                    scales[c][g] = mx / 7.0f;
                    if (scales[c][g] < 1e-10f) scales[c][g] = 1.0f;
                }
            }
            for (int c = 0; c < 8; c++) {                                                          /* :1594-1610 */
                float ms = 0.0f;
                for (int g = 0; g < 8; g++) if (scales[c][g] > ms) ms = scales[c][g];
                const float S_f = ms > 0.0f ? ms : 1.0f;
                t->S[c] = orc_fp32_to_fp16(S_f);
                t->Z[c] = orc_fp32_to_fp16(0.0f);
                for (int g = 0; g < 8; g++) {
                    const float sc_f = (scales[c][g] / S_f) * 127.0f;
                    t->sc[c][g] = (int8_t)fmaxf(-127.0f, fminf(127.0f, roundf(sc_f)));
                    t->zc[c][g] = 0;
                }
            }
            for (int g = 0; g < 8; g++) {                                                          /* :1613-1627 */
                int8_t gw[8][32];
                for (int c = 0; c < 8; c++) {
                    const float *rd = scratch + (size_t)c * n_per_row + kt * ORC_TILE_K;
                    const float inv = 1.0f / scales[c][g];
                    for (int j = 0; j < 32; j++) {
                        const float v = rd[g * 32 + j];
                        gw[c][j] = (int8_t)fmaxf(-8.0f, fminf(7.0f, roundf(v * inv)));
                    }
                }
                orc_pack_group(gw, t->tiles[g]);
            }
        }
        row += rows;
    }
    free(scratch);
    return (size_t)nrow * (size_t)KT * 148;                                                        /* :1636-1638 */
}

/* ------------------------------------------------------------------------------------------
 * Tile dequantiser -- ggml/src/ggml-quants.c:1425-1463.  k/256 tiles, 2048 floats each.
 * ---------------------------------------------------------------------------------------- */
void orc_dequant_tiles(const void *tiles, float *y, int64_t k) {
    const orc_tile *tp = (const orc_tile *)tiles;
    const int64_t nt = k / ORC_TILE_K;
    for (int64_t t = 0; t < nt; t++) {
        for (int c = 0; c < 8; c++) {
            const float S = orc_fp16_to_fp32(tp[t].S[c]);
            const float Z = orc_fp16_to_fp32(tp[t].Z[c]);
            for (int g = 0; g < 8; g++) {
                const float scale = S * (float)tp[t].sc[c][g] / 127.0f;
                const float zero  = Z * (float)tp[t].zc[c][g] / 127.0f;
                int8_t gw[8][32];
                orc_unpack_group(tp[t].tiles[g], gw);
                for (int j = 0; j < 32; j++)
                    y[t * 2048 + c * 256 + g * 32 + j] = scale * (float)gw[c][j] + zero;
            }
        }
    }
}

/* ------------------------------------------------------------------------------------------
 * GET_ROWS of a TCQ4 tensor -- ggml/src/ggml-cuda/getrows.cu:96-171 (row i01 = channel i01 % 8 of
 * n-tile i01 / 8; scale = S*sc/127, zero = Z*zc/127, value = scale*q + zero) followed by the
 * FWHT in 256-element chunks that takes the row back to the weight domain (getrows.cu:372-378;
 * CPU: ggml/src/ggml-cpu/ops.cpp:4800-4824, ggml_fwht_impl with step 256).
 * ---------------------------------------------------------------------------------------- */
void orc_get_rows(const void *tiles, const int32_t *rows, int64_t n_idx, int64_t K, float *dst) {
    const orc_tile *tp = (const orc_tile *)tiles;
    const int64_t KT = K / ORC_TILE_K;
    for (int64_t i = 0; i < n_idx; i++) {
        const int64_t r = rows[i];
        const int c = (int)(r % 8);
        float *out = dst + i * K;
        for (int64_t kt = 0; kt < KT; kt++) {
            const orc_tile *t = &tp[(r / 8) * KT + kt];
            const float S = orc_fp16_to_fp32(t->S[c]);
            const float Z = orc_fp16_to_fp32(t->Z[c]);
            for (int g = 0; g < 8; g++) {
                const float scale = S * (float)t->sc[c][g] / 127.0f;
                const float zero  = Z * (float)t->zc[c][g] / 127.0f;
                int8_t gw[8][32];
                orc_unpack_group(t->tiles[g], gw);
                for (int j = 0; j < 32; j++) out[kt * 256 + g * 32 + j] = scale * (float)gw[c][j] + zero;
            }
            orc_fwht(out + kt * 256, 256);
        }
    }
}

/* src/llama-quant.cpp:26-40 */
void orc_apply_perm_weights(const float *src, float *dst, int64_t nrows, int64_t ncols, const int32_t *perm) {
    for (int64_t r = 0; r < nrows; r++)
        for (int64_t c = 0; c < ncols; c++)
            dst[r * ncols + c] = src[r * ncols + perm[c]];
}

/* tools/quantize/quantize.cpp:127-165 */
typedef struct { float v; int32_t i; } orc_iv;
static int orc_iv_cmp(const void *a, const void *b) {
    const orc_iv *x = (const orc_iv *)a, *y = (const orc_iv *)b;
    if (x->v > y->v) return -1;
    if (x->v < y->v) return 1;
    return (x->i > y->i) - (x->i < y->i);
}
void orc_perm_from_imatrix(const float *imatrix, int64_t K, int32_t *perm) {
    orc_iv tmp[256];
    for (int64_t b = 0; b < K / 256; b++) {
        for (int i = 0; i < 256; i++) { tmp[i].v = imatrix[b * 256 + i]; tmp[i].i = (int32_t)(b * 256 + i); }
        qsort(tmp, 256, sizeof(orc_iv), orc_iv_cmp);
        for (int i = 0; i < 256; i++) perm[b * 256 + i] = tmp[i].i;
    }
}

/* ------------------------------------------------------------------------------------------
 * Activation stage -- ggml/src/ggml-cuda/tcq4_k32.cu:771-831.
 *   gather (771-778) -> 8 unnormalised FWHT stages (781-793) -> * 1/16 (795) ->
 *   max |v| by fmaxf reduction (798-821; NaN-ignoring, so the reduction order is immaterial) ->
 *   < 1e-10 -> 1.0 (817) -> q = clamp(rintf(v * (7/max)), -7, 7) (828-829) -> group sums (831).
 * The reference TU is built with -use_fast_math (approximate 7.0f/max_abs); the oracle and the
 * new kernels use the IEEE divide.
 * ---------------------------------------------------------------------------------------- */
void orc_quantize_act_row(const float *x_row, const int32_t *perm, int64_t K,
                          int8_t *codes, float *max_abs, int16_t *sum_q) {
    float v[256];
    for (int64_t kt = 0; kt < K / 256; kt++) {
        for (int i = 0; i < 256; i++) {
            const int64_t src = perm ? (int64_t)perm[kt * 256 + i] : kt * 256 + i;
            v[i] = x_row[src];
        }
        for (int h = 1; h < 256; h <<= 1)
            for (int i = 0; i < 256; i += 2 * h)
                for (int j = i; j < i + h; j++) {
                    const float a = v[j], b = v[j + h];
                    v[j] = a + b;
                    v[j + h] = a - b;
                }
        float mx = 0.0f;
        for (int i = 0; i < 256; i++) {
            v[i] = v[i] * (1.0f / 16.0f);
            const float a = fabsf(v[i]);
            mx = (i == 0) ? a : fmaxf(mx, a);
        }
        if (mx < 1e-10f) mx = 1.0f;
        max_abs[kt] = mx;
        const float inv = 7.0f / mx;
        for (int g = 0; g < 8; g++) {
            int s = 0;
            for (int j = 0; j < 32; j++) {
                const float scaled = v[g * 32 + j] * inv;
                const int8_t q = (int8_t)fmaxf(-7.0f, fminf(7.0f, rintf(scaled)));
                codes[kt * 256 + g * 32 + j] = q;
                s += q;
            }
            sum_q[kt * 8 + g] = (int16_t)s;
        }
    }
}

void orc_quantize_act_blocks(const float *x, const int32_t *perm, int64_t M, int64_t K, orc_act_block *out) {
    const int64_t KT = K / 256;
    int8_t *codes = (int8_t *)malloc((size_t)K);
    float *mx = (float *)malloc((size_t)KT * sizeof(float));
    int16_t *sq = (int16_t *)malloc((size_t)KT * 8 * sizeof(int16_t));
    for (int64_t m = 0; m < M; m++) {
        orc_quantize_act_row(x + m * K, perm, K, codes, mx, sq);
        for (int64_t kt = 0; kt < KT; kt++) {
            orc_act_block *b = &out[m * KT + kt];
            memset(b, 0, sizeof(*b));
            for (int t = 0; t < 128; t++) {                       /* tcq4_k32.cu:266-270 */
                const int8_t q0 = codes[kt * 256 + 2 * t], q1 = codes[kt * 256 + 2 * t + 1];
                b->qs[t] = (uint8_t)(((uint8_t)(q1 & 0xF) << 4) | (uint8_t)(q0 & 0xF));
            }
            b->smooth_scale = mx[kt];
            for (int g = 0; g < 8; g++) b->sum_q[g] = sq[kt * 8 + g];
        }
    }
    free(codes); free(mx); free(sq);
}

/* ------------------------------------------------------------------------------------------
 * Dot + epilogue -- ggml/src/ggml-cuda/tcq4_k32.cu:833-887.
 *   b_scale = S*(float)sc/127 ; b_zero = Z*(float)zc/127 ; a7 = max_abs/7           (845-852)
 *   contrib = (float)dot*a7*b_scale + (float)sum_a*a7*b_zero                          (872-873)
 *   8 group contributions are summed by the xor-shuffle tree 4,2,1                    (876-879)
 *   acc += tree sum, sequentially over k-tiles                                        (881-883)
 * ---------------------------------------------------------------------------------------- */
void orc_mul_mat(const void *w_tiles, const float *x, const int32_t *perm,
                 int64_t M, int64_t N, int64_t K, float *C,
                 int32_t *dots, int32_t *isum, int8_t *codes_out, float *max_abs_out, int16_t *sum_q_out) {
    const int64_t KT = K / 256, NT = N / 8;
    const orc_tile *tiles = (const orc_tile *)w_tiles;
    int8_t *codes = (int8_t *)malloc((size_t)K);
    float *mx = (float *)malloc((size_t)KT * sizeof(float));
    int16_t *sq = (int16_t *)malloc((size_t)KT * 8 * sizeof(int16_t));
    /* unpack all weights once: wq[nt][kt][g][c][32] */
    int8_t *wq = (int8_t *)malloc((size_t)NT * KT * 2048);
    for (int64_t t = 0; t < NT * KT; t++)
        for (int g = 0; g < 8; g++)
            orc_unpack_group(tiles[t].tiles[g], (int8_t (*)[32])(wq + (size_t)t * 2048 + g * 256));

    for (int64_t m = 0; m < M; m++) {
        orc_quantize_act_row(x + m * K, perm, K, codes, mx, sq);
        if (codes_out)   memcpy(codes_out + m * K, codes, (size_t)K);
        if (max_abs_out) memcpy(max_abs_out + m * KT, mx, (size_t)KT * sizeof(float));
        if (sum_q_out)   memcpy(sum_q_out + m * KT * 8, sq, (size_t)KT * 8 * sizeof(int16_t));
        for (int64_t n = 0; n < N; n++) {
            const int64_t nt = n / 8;
            const int c = (int)(n % 8);
            float acc = 0.0f;
            for (int64_t kt = 0; kt < KT; kt++) {
                const orc_tile *t = &tiles[nt * KT + kt];
                const float S = orc_fp16_to_fp32(t->S[c]);
                const float Z = orc_fp16_to_fp32(t->Z[c]);
                const float a7 = mx[kt] / 7.0f;
                float contrib[8];
                int32_t is = 0;
                for (int g = 0; g < 8; g++) {
                    const float b_scale = S * (float)t->sc[c][g] / 127.0f;
                    const float b_zero  = Z * (float)t->zc[c][g] / 127.0f;
                    const int8_t *wg = wq + (size_t)(nt * KT + kt) * 2048 + g * 256 + c * 32;
                    const int8_t *ag = codes + kt * 256 + g * 32;
                    int32_t dot = 0;
                    for (int j = 0; j < 32; j++) dot += (int32_t)ag[j] * (int32_t)wg[j];
                    if (dots) dots[((m * N + n) * KT + kt) * 8 + g] = dot;
                    is += (int32_t)t->sc[c][g] * dot;
                    const float t1 = (float)dot * a7 * b_scale;
                    const float t2 = (float)sq[kt * 8 + g] * a7 * b_zero;
                    contrib[g] = t1 + t2;
                }
                if (isum) isum[(m * N + n) * KT + kt] = is;
                for (int off = 4; off > 0; off >>= 1) {
                    float nx[8];
                    for (int g = 0; g < 8; g++) nx[g] = contrib[g] + contrib[g ^ off];
                    memcpy(contrib, nx, sizeof(nx));
                }
                acc += contrib[0];
            }
            C[m * N + n] = acc;
        }
    }
    free(codes); free(mx); free(sq); free(wq);
}

/* O2: dequant(W) . FWHT(perm(x)) in F32 */
void orc_mul_mat_f32(const void *w_tiles, const float *x, const int32_t *perm,
                     int64_t M, int64_t N, int64_t K, float *C) {
    const int64_t KT = K / 256, NT = N / 8;
    float *wf = (float *)malloc((size_t)NT * KT * 2048 * sizeof(float));
    orc_dequant_tiles(w_tiles, wf, NT * KT * 256);
    float *xr = (float *)malloc((size_t)K * sizeof(float));
    for (int64_t m = 0; m < M; m++) {
        for (int64_t k = 0; k < K; k++) xr[k] = x[m * K + (perm ? perm[k] : k)];
        for (int64_t kt = 0; kt < KT; kt++) orc_fwht(xr + kt * 256, 256);
        for (int64_t n = 0; n < N; n++) {
            const int64_t nt = n / 8;
            const int c = (int)(n % 8);
            float acc = 0.0f;
            for (int64_t kt = 0; kt < KT; kt++) {
                const float *wr = wf + (size_t)(nt * KT + kt) * 2048 + c * 256;
                float s = 0.0f;
                for (int j = 0; j < 256; j++) s += wr[j] * xr[kt * 256 + j];
                acc += s;
            }
            C[m * N + n] = acc;
        }
    }
    free(wf); free(xr);
}

uint64_t orc_fnv1a64(const void *p, size_t n) {
    const uint8_t *b = (const uint8_t *)p;
    uint64_t h = 0xcbf29ce484222325ull;
    for (size_t i = 0; i < n; i++) { h ^= b[i]; h *= 0x100000001b3ull; }
    return h;
}
