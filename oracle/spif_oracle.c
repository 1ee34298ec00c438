/* oracle/spif_oracle.c -- TEST INFRASTRUCTURE ONLY.
 *
 * A plain-C restatement of the algorithm of SparkInfer's activation-sparse FFN path as its CPU backend
 * computes it ("O-cpu"), plus an FP64 restatement of the same math ("O-f64").  It exists to check the
 * CUDA product; only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may load it, and the
 * product (sparkinfer_b200/) never does.
 *
 * PARITY PINNING: the reference ships no golden vectors / KATs for this path, so this file
 * is pinned against outputs of the reference itself: oracle/_ref/libspif_ref.so (the unmodified reference CPU
 * ops compiled from /root/reference by oracle/Makefile) on seeded inputs -- tests/test_oracle_vs_ref.py when
 * _ref is present, and the committed fixtures tests/golden/*.npz (made by tests/golden/make_golden.py from
 * _ref) everywhere else.  For F16 / Q8_0 / Q4_0 weights the restatement reproduces the reference's x86-64-v3
 * (AVX2, 8-lane) reduction trees and matches _ref BIT FOR BIT; for BF16 it matches to ~1e-6 relative (the
 * reference's mul+add there may or may not be contracted by the compiler).
 * Not pinned by any reference code (none exists): spifo_topk_mask and the Q4_0 axpy (the reference aborts,
 * ggml-cpu.c:2225-2226) -- "parity unpinned" for those two.
 *
 * Reference lines restated (relative to /root/reference):
 *   mask convention                 ggml/src/ggml-cuda/mm-sparse.cu:22, ggml/src/ggml-cpu/ggml-cpu.c:224-226,1775
 *   mul_mat_sparse                  ggml/src/ggml-cpu/ggml-cpu.c:1692-1925
 *   axpy_sparse                     ggml/src/ggml-cpu/ggml-cpu.c:1927-2337
 *   vec_dot f16 / bf16              ggml/src/ggml-cpu/vec.cpp:137-336, simd-mappings.h:527-623
 *   vec_dot q8_0.q8_0, q4_0.q8_0    ggml/src/ggml-cpu/arch/x86/quants.c:543-600,1011-1060 (+ :43-49,229-244)
 *   activation quantiser            ggml/src/ggml-cpu/arch/x86/quants.c:290-356
 *   weight quantisers               ggml/src/ggml-quants.c:36-66,199-221 ; dequant :307-325,401-413
 *   predictor / layer recipe        src/llama-graph.cpp:865-894, 896-1142 ; unary-ops.cpp:27-33 ; vec.h:841
 */
#include "spif_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#ifdef _OPENMP
#    include <omp.h>
#endif

#define QK 32 /* QK4_0 == QK8_0 == 32, ggml-common.h */

const char * spifo_describe(void) {
    return "spif_oracle: plain-C restatement of the reference CPU sparse-FFN path (O-cpu, 1-thread order, "
           "AVX2 8-lane reduction trees) + FP64 restatement (O-f64)";
}

/* ------------------------------------------------------------------ scalar formats */

static inline uint32_t f2u(float f) {
    uint32_t u;
    memcpy(&u, &f, 4);
    return u;
}
static inline float u2f(uint32_t u) {
    float f;
    memcpy(&f, &u, 4);
    return f;
}

/* IEEE binary16 <-> binary32, round-to-nearest-even; value-identical to F16C and to ggml-impl.h:378-425 */
static float h2f(uint16_t h) {
    uint32_t sign = (uint32_t) (h & 0x8000u) << 16;
    uint32_t e    = (h >> 10) & 0x1Fu;
    uint32_t man  = h & 0x3FFu;
    if (e == 0) {
        if (man == 0) {
            return u2f(sign);
        }
        /* subnormal: man * 2^-24 */
        float v = (float) man * 0x1.0p-24f;
        return u2f(sign | f2u(v));
    }
    if (e == 31) {
        return u2f(sign | 0x7F800000u | (man << 13));
    }
    return u2f(sign | ((e + 112u) << 23) | (man << 13));
}

static uint16_t f2h(float f) {
    uint32_t u    = f2u(f);
    uint16_t sign = (uint16_t) ((u >> 16) & 0x8000u);
    uint32_t a    = u & 0x7FFFFFFFu;
    if (a > 0x7F800000u) {
        return sign | 0x7E00u; /* NaN */
    }
    if (a >= 0x477FF000u) { /* >= 65520 rounds to inf */
        return sign | 0x7C00u;
    }
    if (a < 0x33000001u) { /* <= 2^-25: rounds to zero (2^-25 exactly ties to even = 0) */
        return sign;
    }
    int32_t  e   = (int32_t) (a >> 23) - 127;
    uint32_t man = (a & 0x7FFFFFu) | 0x800000u; /* 24-bit significand */
    int      shift;
    uint32_t base;
    if (e < -14) { /* subnormal half: value = man * 2^(e-23), unit = 2^-24 */
        shift = (-14 - e) + 13;
        base  = 0;
    } else {
        shift = 13;
        base  = (uint32_t) (e + 15) << 10;
        man &= 0x7FFFFFu;
    }
    uint32_t q    = man >> shift;
    uint32_t rem  = man & ((1u << shift) - 1u);
    uint32_t half = 1u << (shift - 1);
    if (rem > half || (rem == half && (q & 1u))) {
        q++;
    }
    return sign | (uint16_t) (base + q); /* carry into the exponent is the correct rounding */
}

static inline float bf2f(uint16_t b) {
    return u2f((uint32_t) b << 16);
}
static uint16_t f2bf(float f) { /* RNE, quiet NaN: ggml-impl.h:550-563 */
    uint32_t u = f2u(f);
    if ((u & 0x7FFFFFFFu) > 0x7F800000u) {
        return (uint16_t) ((u >> 16) | 64u);
    }
    return (uint16_t) ((u + (0x7FFFu + ((u >> 16) & 1u))) >> 16);
}

void spifo_fp32_to_fp16_row(const float * x, uint16_t * y, int64_t n) {
    for (int64_t i = 0; i < n; ++i) {
        y[i] = f2h(x[i]);
    }
}
void spifo_fp16_to_fp32_row(const uint16_t * x, float * y, int64_t n) {
    for (int64_t i = 0; i < n; ++i) {
        y[i] = h2f(x[i]);
    }
}
void spifo_fp32_to_bf16_row(const float * x, uint16_t * y, int64_t n) {
    for (int64_t i = 0; i < n; ++i) {
        y[i] = f2bf(x[i]);
    }
}

/* ------------------------------------------------------------------ block formats (ggml-common.h:170-174,218-222) */

#pragma pack(push, 1)
typedef struct {
    uint16_t d;
    uint8_t  qs[QK / 2];
} blk_q4_0; /* 18 bytes */
typedef struct {
    uint16_t d;
    int8_t   qs[QK];
} blk_q8_0; /* 34 bytes */
#pragma pack(pop)

size_t spifo_row_size(int type, int64_t n) {
    switch (type) {
        case SPIFO_F32:
            return (size_t) n * 4;
        case SPIFO_F16:
        case SPIFO_BF16:
            return (size_t) n * 2;
        case SPIFO_Q8_0:
            return (size_t) (n / QK) * sizeof(blk_q8_0);
        case SPIFO_Q4_0:
            return (size_t) (n / QK) * sizeof(blk_q4_0);
        default:
            return 0;
    }
}

/* quantize_row_q8_0_ref, ggml-quants.c:199-221 (weights) */
static void quant_q8_0_ref(const float * x, blk_q8_0 * y, int64_t k) {
    for (int64_t i = 0; i < k / QK; ++i) {
        float amax = 0.0f;
        for (int j = 0; j < QK; ++j) {
            float v = fabsf(x[i * QK + j]);
            amax    = amax > v ? amax : v;
        }
        const float d  = amax / 127;
        const float id = d ? 1.0f / d : 0.0f;
        y[i].d         = f2h(d);
        for (int j = 0; j < QK; ++j) {
            y[i].qs[j] = (int8_t) roundf(x[i * QK + j] * id);
        }
    }
}

/* quantize_row_q4_0_ref, ggml-quants.c:36-66 (weights) */
static void quant_q4_0_ref(const float * x, blk_q4_0 * y, int64_t k) {
    for (int64_t i = 0; i < k / QK; ++i) {
        float amax = 0.0f, max = 0.0f;
        for (int j = 0; j < QK; ++j) {
            const float v = x[i * QK + j];
            if (amax < fabsf(v)) {
                amax = fabsf(v);
                max  = v;
            }
        }
        const float d  = max / -8;
        const float id = d ? 1.0f / d : 0.0f;
        y[i].d         = f2h(d);
        for (int j = 0; j < QK / 2; ++j) {
            const float x0  = x[i * QK + 0 + j] * id;
            const float x1  = x[i * QK + QK / 2 + j] * id;
            int         q0  = (int8_t) (x0 + 8.5f);
            int         q1  = (int8_t) (x1 + 8.5f);
            const uint8_t a = (uint8_t) (q0 < 15 ? q0 : 15);
            const uint8_t b = (uint8_t) (q1 < 15 ? q1 : 15);
            y[i].qs[j]      = (uint8_t) (a | (b << 4));
        }
    }
}

/* CPU-backend activation quantiser (from_float of Q8_0), AVX2 branch of arch/x86/quants.c:290-356:
 * id = 127/amax (not 1/d) and round-half-to-even (_mm256_round_ps NEAREST), unlike the _ref quantiser */
void spifo_quantize_row_q8_0_act(const float * x, void * vy, int64_t n) {
    blk_q8_0 * y = (blk_q8_0 *) vy;
    for (int64_t i = 0; i < n / QK; ++i) {
        float amax = 0.0f;
        for (int j = 0; j < QK; ++j) {
            float v = fabsf(x[i * QK + j]);
            amax    = amax > v ? amax : v;
        }
        const float d  = amax / 127.f;
        const float id = (amax != 0.0f) ? 127.f / amax : 0.0f;
        y[i].d         = f2h(d);
        for (int j = 0; j < QK; ++j) {
            y[i].qs[j] = (int8_t) (int) nearbyintf(x[i * QK + j] * id); /* default FE_TONEAREST = ties-to-even */
        }
    }
}

int spifo_quantize_rows(int type, const float * src, void * dst, int64_t nrows, int64_t n) {
    const size_t rs = spifo_row_size(type, n);
    for (int64_t r = 0; r < nrows; ++r) {
        const float * s = src + r * n;
        char *        d = (char *) dst + r * rs;
        switch (type) {
            case SPIFO_F16:
                spifo_fp32_to_fp16_row(s, (uint16_t *) d, n);
                break;
            case SPIFO_BF16:
                spifo_fp32_to_bf16_row(s, (uint16_t *) d, n);
                break;
            case SPIFO_Q8_0:
                quant_q8_0_ref(s, (blk_q8_0 *) d, n);
                break;
            case SPIFO_Q4_0:
                quant_q4_0_ref(s, (blk_q4_0 *) d, n);
                break;
            case SPIFO_F32:
                memcpy(d, s, (size_t) n * 4);
                break;
            default:
                return 1;
        }
    }
    return 0;
}

/* dequantize_row_q4_0 / q8_0, ggml-quants.c:307-325,401-413 */
int spifo_dequantize_row(int type, const void * src, float * dst, int64_t n) {
    switch (type) {
        case SPIFO_F32:
            memcpy(dst, src, (size_t) n * 4);
            return 0;
        case SPIFO_F16:
            spifo_fp16_to_fp32_row((const uint16_t *) src, dst, n);
            return 0;
        case SPIFO_BF16:
            for (int64_t i = 0; i < n; ++i) {
                dst[i] = bf2f(((const uint16_t *) src)[i]);
            }
            return 0;
        case SPIFO_Q8_0:
            {
                const blk_q8_0 * b = (const blk_q8_0 *) src;
                for (int64_t i = 0; i < n / QK; ++i) {
                    const float d = h2f(b[i].d);
                    for (int j = 0; j < QK; ++j) {
                        dst[i * QK + j] = b[i].qs[j] * d;
                    }
                }
                return 0;
            }
        case SPIFO_Q4_0:
            {
                const blk_q4_0 * b = (const blk_q4_0 *) src;
                for (int64_t i = 0; i < n / QK; ++i) {
                    const float d = h2f(b[i].d);
                    for (int j = 0; j < QK / 2; ++j) {
                        dst[i * QK + j]          = ((b[i].qs[j] & 0x0F) - 8) * d;
                        dst[i * QK + j + QK / 2] = ((b[i].qs[j] >> 4) - 8) * d;
                    }
                }
                return 0;
            }
        default:
            return 1;
    }
}

/* ------------------------------------------------------------------ dot products, AVX2 lane structure */

/* 8-lane tree of GGML_F32x8_REDUCE with 4 accumulators, simd-mappings.h:542-560 */
static float reduce_4x8(float acc[4][8]) {
    float a[8], t0[4];
    for (int l = 0; l < 8; ++l) {
        float s02 = acc[0][l] + acc[2][l];
        float s13 = acc[1][l] + acc[3][l];
        a[l]      = s02 + s13;
    }
    for (int l = 0; l < 4; ++l) {
        t0[l] = a[l] + a[l + 4];
    }
    return (t0[0] + t0[1]) + (t0[2] + t0[3]);
}

/* ggml_vec_dot_f16, vec.cpp:207-336 generic-SIMD branch; x = weight row, y = activations rounded to F16 */
static float dot_f16(int n, const uint16_t * x, const uint16_t * y) {
    float     acc[4][8];
    const int np = n & ~31;
    memset(acc, 0, sizeof(acc));
    for (int i = 0; i < np; i += 32) {
        for (int j = 0; j < 4; ++j) {
            for (int l = 0; l < 8; ++l) {
                acc[j][l] = fmaf(h2f(x[i + j * 8 + l]), h2f(y[i + j * 8 + l]), acc[j][l]);
            }
        }
    }
    double sumf = (double) reduce_4x8(acc);
    for (int i = np; i < n; ++i) {
        sumf += (double) (h2f(x[i]) * h2f(y[i]));
    }
    return (float) sumf;
}

/* ggml_vec_dot_bf16, vec.cpp:137-205 AVX2 branch (mul then add; tree (c1+c3)+(c2+c4), then hsum) */
static float dot_bf16(int n, const uint16_t * x, const uint16_t * y) {
    float c[4][8];
    int   i = 0;
    memset(c, 0, sizeof(c));
    for (; i + 32 <= n; i += 32) {
        for (int j = 0; j < 4; ++j) {
            for (int l = 0; l < 8; ++l) {
                float p = bf2f(x[i + j * 8 + l]) * bf2f(y[i + j * 8 + l]);
                c[j][l] = p + c[j][l];
            }
        }
    }
    float a[8], g[4];
    for (int l = 0; l < 8; ++l) {
        a[l] = (c[0][l] + c[2][l]) + (c[1][l] + c[3][l]);
    }
    for (int l = 0; l < 4; ++l) {
        g[l] = a[l + 4] + a[l];
    }
    g[0] += g[2];
    g[1] += g[3];
    double sumf = (double) (g[0] + g[1]);
    for (; i < n; ++i) {
        sumf += (double) (bf2f(x[i]) * bf2f(y[i]));
    }
    return (float) sumf;
}

/* hsum_float_8, arch/x86/quants.c:43-49 */
static float hsum8(const float a[8]) {
    float r[4];
    for (int l = 0; l < 4; ++l) {
        r[l] = a[l + 4] + a[l];
    }
    r[0] += r[2];
    r[1] += r[3];
    return r[0] + r[1];
}

/* ggml_vec_dot_q8_0_q8_0 AVX2, arch/x86/quants.c:1011-1045: lane L of a block = sum of its 4 int products */
static float dot_q8_0_q8_0(int n, const blk_q8_0 * x, const blk_q8_0 * y) {
    float acc[8] = { 0 };
    for (int ib = 0; ib < n / QK; ++ib) {
        const float d = h2f(x[ib].d) * h2f(y[ib].d);
        for (int l = 0; l < 8; ++l) {
            int s = 0;
            for (int j = 0; j < 4; ++j) {
                s += (int) x[ib].qs[4 * l + j] * (int) y[ib].qs[4 * l + j];
            }
            acc[l] = fmaf(d, (float) s, acc[l]);
        }
    }
    return hsum8(acc);
}

/* ggml_vec_dot_q4_0_q8_0 AVX2, arch/x86/quants.c:543-583: bytes 0..15 = low nibbles, 16..31 = high nibbles */
static float dot_q4_0_q8_0(int n, const blk_q4_0 * x, const blk_q8_0 * y) {
    float acc[8] = { 0 };
    for (int ib = 0; ib < n / QK; ++ib) {
        const float d = h2f(x[ib].d) * h2f(y[ib].d);
        int         q[QK];
        for (int j = 0; j < QK / 2; ++j) {
            q[j]          = (x[ib].qs[j] & 0x0F) - 8;
            q[j + QK / 2] = (x[ib].qs[j] >> 4) - 8;
        }
        for (int l = 0; l < 8; ++l) {
            int s = 0;
            for (int j = 0; j < 4; ++j) {
                s += q[4 * l + j] * (int) y[ib].qs[4 * l + j];
            }
            acc[l] = fmaf(d, (float) s, acc[l]);
        }
    }
    return hsum8(acc);
}

/* vec_dot_type of a weight type: F16->F16, BF16->BF16, Q8_0/Q4_0->Q8_0 (ggml-cpu.c:241-330 type_traits_cpu) */
static size_t act_row_size(int type, int n) {
    return (type == SPIFO_F16 || type == SPIFO_BF16) ? (size_t) n * 2 : (type == SPIFO_F32 ? (size_t) n * 4
                                                                                            : spifo_row_size(SPIFO_Q8_0, n));
}
static void convert_act(int type, const float * x, void * out, int n) {
    switch (type) {
        case SPIFO_F16:
            spifo_fp32_to_fp16_row(x, (uint16_t *) out, n);
            break;
        case SPIFO_BF16:
            spifo_fp32_to_bf16_row(x, (uint16_t *) out, n);
            break;
        case SPIFO_F32:
            memcpy(out, x, (size_t) n * 4);
            break;
        default:
            spifo_quantize_row_q8_0_act(x, out, n);
            break;
    }
}
static float dot_f32(int n, const float * x, const float * y) { /* not on the reference's sparse path; plain */
    double s = 0;
    for (int i = 0; i < n; ++i) {
        s += (double) x[i] * y[i];
    }
    return (float) s;
}
static float vec_dot(int type, int n, const void * w, const void * a) {
    switch (type) {
        case SPIFO_F16:
            return dot_f16(n, (const uint16_t *) w, (const uint16_t *) a);
        case SPIFO_BF16:
            return dot_bf16(n, (const uint16_t *) w, (const uint16_t *) a);
        case SPIFO_Q8_0:
            return dot_q8_0_q8_0(n, (const blk_q8_0 *) w, (const blk_q8_0 *) a);
        case SPIFO_Q4_0:
            return dot_q4_0_q8_0(n, (const blk_q4_0 *) w, (const blk_q8_0 *) a);
        case SPIFO_F32:
            return dot_f32(n, (const float *) w, (const float *) a);
        default:
            return NAN;
    }
}
static int type_ok(int type) {
    return type == SPIFO_F16 || type == SPIFO_BF16 || type == SPIFO_Q8_0 || type == SPIFO_Q4_0;
}

/* ------------------------------------------------------------------ mask */

int spifo_mask_compact(const float * p, int n_ff, float thr, int32_t * idx_out) {
    int c = 0;
    for (int i = 0; i < n_ff; ++i) {
        if (p[i] < thr) { /* NaN < thr is false => NaN counts as active, like mm-sparse.cu:22 */
            continue;
        }
        if (idx_out) {
            idx_out[c] = i;
        }
        ++c;
    }
    return c;
}

int spifo_topk_mask(const float * score, int n_ff, int k, float * out) {
    if (k > n_ff) {
        k = n_ff;
    }
    for (int i = 0; i < n_ff; ++i) {
        out[i] = 0.0f;
    }
    /* selection by repeated threshold counting would be O(n log) -- sizes are small, do an index sort */
    int32_t * ord = (int32_t *) malloc(sizeof(int32_t) * (size_t) (n_ff > 0 ? n_ff : 1));
    for (int i = 0; i < n_ff; ++i) {
        ord[i] = i;
    }
    /* insertion-free: simple O(n*k) partial selection keeps ties stable (lower index first) */
    for (int s = 0; s < k; ++s) {
        int best = s;
        for (int j = s + 1; j < n_ff; ++j) {
            /* total order on the bit patterns (so -0 < +0 and NaNs are ordered), same key as the CUDA radix select */
            uint32_t a = f2u(score[ord[j]]), b = f2u(score[ord[best]]);
            a = (a & 0x80000000u) ? ~a : (a | 0x80000000u);
            b = (b & 0x80000000u) ? ~b : (b | 0x80000000u);
            if (a > b || (a == b && ord[j] < ord[best])) {
                best = j;
            }
        }
        int32_t t = ord[s];
        ord[s]    = ord[best];
        ord[best] = t;
        out[ord[s]] = 1.0f;
    }
    free(ord);
    return k;
}

/* ------------------------------------------------------------------ the two sparse ops */

int spifo_mul_mat_sparse(int type, const void * w, int m, const int32_t * neuron_idx, const int32_t * neuron_mask,
                         const float * x, const float * sparse_idx, float * dst, int n_embd, int n_ff, int n_tok,
                         float thr) {
    if (!type_ok(type) || n_embd % QK != 0) {
        return 1;
    }
    const size_t rs = spifo_row_size(type, n_embd);
    const size_t as = act_row_size(type, n_embd);
    void *       a  = malloc(as ? as : 1);
    memset(dst, 0, sizeof(float) * (size_t) n_ff * n_tok); /* ggml-cpu.c:1800 */
    for (int t = 0; t < n_tok; ++t) {
        convert_act(type, x + (size_t) t * n_embd, a, n_embd); /* ggml-cpu.c:1832-1856 */
        const float * p = sparse_idx + (size_t) t * n_ff;
        float *       d = dst + (size_t) t * n_ff;
        for (int row = 0; row < m; ++row) {
            const int neu = neuron_idx ? neuron_idx[row] : row;
            if (neu < 0 || neu >= n_ff) {
                continue;
            }
            if ((neuron_mask && neuron_mask[neu] == 1) || p[neu] < thr) { /* ggml-cpu.c:1775 */
                continue;
            }
            d[neu] = vec_dot(type, n_embd, (const char *) w + rs * (size_t) row, a);
        }
    }
    free(a);
    return 0;
}

int spifo_axpy_sparse(int type, const void * wt, int m, const int32_t * neuron_idx, const int32_t * neuron_mask,
                      const float * hid, const float * sparse_idx, float * dst, int n_embd, int n_ff, int n_tok,
                      float thr) {
    if (!type_ok(type) || n_embd % QK != 0) {
        return 1;
    }
    const size_t rs = spifo_row_size(type, n_embd);
    for (int t = 0; t < n_tok; ++t) {
        const float * p   = sparse_idx + (size_t) t * n_ff;
        const float * h   = hid + (size_t) t * n_ff;
        float *       buf = dst + (size_t) t * n_embd;
        memset(buf, 0, sizeof(float) * (size_t) n_embd);
        for (int row = 0; row < m; ++row) {
            const int neu = neuron_idx ? neuron_idx[row] : row;
            if (neu < 0 || neu >= n_ff) {
                continue;
            }
            /* alpha is rounded to the weight's float type for F16/BF16, kept F32 for quants: ggml-cpu.c:2196-2218 */
            float alpha = h[neu];
            if (type == SPIFO_F16) {
                alpha = h2f(f2h(alpha));
            } else if (type == SPIFO_BF16) {
                alpha = bf2f(f2bf(alpha));
            }
            if ((neuron_mask && neuron_mask[neu] == 1) || p[neu] < thr || alpha == 0.0f) {
                continue;
            }
            const char * wr = (const char *) wt + rs * (size_t) row;
            if (type == SPIFO_F16) { /* ggml_axpy_avx_f16_alphaf32: fma(w, alpha, y) */
                const uint16_t * v = (const uint16_t *) wr;
                for (int i = 0; i < n_embd; ++i) {
                    buf[i] = fmaf(h2f(v[i]), alpha, buf[i]);
                }
            } else if (type == SPIFO_BF16) {
                const uint16_t * v = (const uint16_t *) wr;
                for (int i = 0; i < n_embd; ++i) {
                    buf[i] = fmaf(bf2f(v[i]), alpha, buf[i]);
                }
            } else if (type == SPIFO_Q8_0) { /* ggml_axpy_avx_q8_0_alphaf32: scale = d*alpha; fma(q, scale, y) */
                const blk_q8_0 * b = (const blk_q8_0 *) wr;
                for (int ib = 0; ib < n_embd / QK; ++ib) {
                    const float sc = h2f(b[ib].d) * alpha;
                    for (int j = 0; j < QK; ++j) {
                        buf[ib * QK + j] = fmaf((float) b[ib].qs[j], sc, buf[ib * QK + j]);
                    }
                }
            } else { /* Q4_0: no reference code (aborts ggml-cpu.c:2225-2226); restated by analogy with Q8_0 */
                const blk_q4_0 * b = (const blk_q4_0 *) wr;
                for (int ib = 0; ib < n_embd / QK; ++ib) {
                    const float sc = h2f(b[ib].d) * alpha;
                    for (int j = 0; j < QK / 2; ++j) {
                        const int lo              = (b[ib].qs[j] & 0x0F) - 8;
                        const int hi              = (b[ib].qs[j] >> 4) - 8;
                        buf[ib * QK + j]          = fmaf((float) lo, sc, buf[ib * QK + j]);
                        buf[ib * QK + j + QK / 2] = fmaf((float) hi, sc, buf[ib * QK + j + QK / 2]);
                    }
                }
            }
        }
    }
    return 0;
}

/* ------------------------------------------------------------------ predictor, activation, whole layer */

int spifo_predictor(int ptype, const void * wpu, const float * bpu, const void * wpd, const float * bpd,
                    const float * x, float * out, float * logits_out, int n_embd, int r, int n_ff, int n_tok) {
    if (!(ptype == SPIFO_F16 || ptype == SPIFO_BF16 || ptype == SPIFO_F32) || n_embd % QK != 0 || r % QK != 0) {
        return 1;
    }
    const size_t rs_u = spifo_row_size(ptype, n_embd), rs_d = spifo_row_size(ptype, r);
    void *       ax = malloc(act_row_size(ptype, n_embd));
    void *       ah = malloc(act_row_size(ptype, r));
    float *      h  = (float *) malloc(sizeof(float) * (size_t) r);
    for (int t = 0; t < n_tok; ++t) {
        convert_act(ptype, x + (size_t) t * n_embd, ax, n_embd);
        for (int j = 0; j < r; ++j) {
            float v = vec_dot(ptype, n_embd, (const char *) wpu + rs_u * (size_t) j, ax);
            if (bpu) {
                v = v + bpu[j];
            }
            h[j] = v > 0.f ? v : 0.f;
        }
        convert_act(ptype, h, ah, r);
        for (int i = 0; i < n_ff; ++i) {
            float z = vec_dot(ptype, r, (const char *) wpd + rs_d * (size_t) i, ah);
            if (bpd) {
                z = z + bpd[i];
            }
            if (logits_out) {
                logits_out[(size_t) t * n_ff + i] = z;
            }
            out[(size_t) t * n_ff + i] = 1.f / (1.f + expf(-z));
        }
    }
    free(ax);
    free(ah);
    free(h);
    return 0;
}

void spifo_activation(int arch, const float * up, const float * gate, const float * b_up, const float * b_gate,
                      float * hid, int n_ff, int n_tok) {
    for (int t = 0; t < n_tok; ++t) {
        for (int i = 0; i < n_ff; ++i) {
            const size_t k = (size_t) t * n_ff + i;
            float        u = up[k];
            if (b_up) {
                u = u + b_up[i]; /* NB added to inactive rows too, llama-graph.cpp:1049-1059 */
            }
            float g = 0.f;
            if (arch != SPIFO_ARCH_RELU_UP) {
                g = gate[k];
                if (b_gate) {
                    g = g + b_gate[i];
                }
            }
            switch (arch) {
                case SPIFO_ARCH_PROSPARSE:
                    hid[k] = ((g > 0.01f) ? g : 0.0f) * u;
                    break;
                case SPIFO_ARCH_RELU_GATED:
                    hid[k] = ((g > 0.f) ? g : 0.f) * ((u > 0.f) ? u : 0.f);
                    break;
                case SPIFO_ARCH_SILU_GATED: /* no reference code (llama-graph.cpp:1092-1093 aborts): parity unpinned */
                    hid[k] = (g / (1.0f + expf(-g))) * u;
                    break;
                default:
                    hid[k] = (u > 0.f) ? u : 0.f;
                    break;
            }
        }
    }
}

int spifo_sparse_ffn(int arch, int type, const void * w_up, const void * w_gate, const void * w_down_t,
                     const float * b_up, const float * b_gate, const float * b_down, const float * x,
                     const float * sparse_idx, float * y, float * o_up, float * o_gate, float * o_hid, int n_embd,
                     int n_ff, int n_tok, float thr) {
    const size_t nh  = (size_t) n_ff * n_tok;
    float *      up  = (float *) malloc(sizeof(float) * nh);
    float *      gt  = (float *) malloc(sizeof(float) * nh);
    float *      hid = (float *) malloc(sizeof(float) * nh);
    int          rc  = spifo_mul_mat_sparse(type, w_up, n_ff, NULL, NULL, x, sparse_idx, up, n_embd, n_ff, n_tok, thr);
    if (!rc && arch != SPIFO_ARCH_RELU_UP) {
        rc = spifo_mul_mat_sparse(type, w_gate, n_ff, NULL, NULL, x, sparse_idx, gt, n_embd, n_ff, n_tok, thr);
    }
    if (!rc) {
        spifo_activation(arch, up, gt, b_up, b_gate, hid, n_ff, n_tok);
        rc = spifo_axpy_sparse(type, w_down_t, n_ff, NULL, NULL, hid, sparse_idx, y, n_embd, n_ff, n_tok, thr);
    }
    if (!rc && b_down) {
        for (int t = 0; t < n_tok; ++t) {
            for (int i = 0; i < n_embd; ++i) {
                y[(size_t) t * n_embd + i] += b_down[i];
            }
        }
    }
    if (!rc) {
        if (o_up) {
            memcpy(o_up, up, sizeof(float) * nh);
        }
        if (o_gate && arch != SPIFO_ARCH_RELU_UP) {
            memcpy(o_gate, gt, sizeof(float) * nh);
        }
        if (o_hid) {
            memcpy(o_hid, hid, sizeof(float) * nh);
        }
    }
    free(up);
    free(gt);
    free(hid);
    return rc;
}

int spifo_sparse_ffn_f64(int arch, int type, const void * w_up, const void * w_gate, const void * w_down_t,
                         const float * b_up, const float * b_gate, const float * b_down, const float * x,
                         const float * sparse_idx, double * y, double * o_hid, int n_embd, int n_ff, int n_tok,
                         float thr) {
    if (!type_ok(type)) {
        return 1;
    }
    const size_t rs  = spifo_row_size(type, n_embd);
    float *      row = (float *) malloc(sizeof(float) * (size_t) n_embd);
    for (int t = 0; t < n_tok; ++t) {
        const float * xt = x + (size_t) t * n_embd;
        const float * p  = sparse_idx + (size_t) t * n_ff;
        double *      yt = y + (size_t) t * n_embd;
        for (int i = 0; i < n_embd; ++i) {
            yt[i] = b_down ? (double) b_down[i] : 0.0;
        }
        for (int n = 0; n < n_ff; ++n) {
            const int act = !(p[n] < thr);
            double    u = 0.0, g = 0.0;
            if (act) {
                spifo_dequantize_row(type, (const char *) w_up + rs * (size_t) n, row, n_embd);
                for (int i = 0; i < n_embd; ++i) {
                    u += (double) row[i] * (double) xt[i];
                }
                if (arch != SPIFO_ARCH_RELU_UP) {
                    spifo_dequantize_row(type, (const char *) w_gate + rs * (size_t) n, row, n_embd);
                    for (int i = 0; i < n_embd; ++i) {
                        g += (double) row[i] * (double) xt[i];
                    }
                }
            }
            if (b_up) {
                u += b_up[n];
            }
            if (b_gate && arch != SPIFO_ARCH_RELU_UP) {
                g += b_gate[n];
            }
            double h;
            switch (arch) {
                case SPIFO_ARCH_PROSPARSE:
                    h = ((g > (double) 0.01f) ? g : 0.0) * u;
                    break;
                case SPIFO_ARCH_RELU_GATED:
                    h = (g > 0 ? g : 0) * (u > 0 ? u : 0);
                    break;
                case SPIFO_ARCH_SILU_GATED:
                    h = (g / (1.0 + exp(-g))) * u;
                    break;
                default:
                    h = u > 0 ? u : 0;
                    break;
            }
            if (o_hid) {
                o_hid[(size_t) t * n_ff + n] = h;
            }
            if (!act || h == 0.0) {
                continue;
            }
            spifo_dequantize_row(type, (const char *) w_down_t + rs * (size_t) n, row, n_embd);
            for (int i = 0; i < n_embd; ++i) {
                yt[i] += h * (double) row[i];
            }
        }
    }
    free(row);
    return 0;
}

/* ------------------------------------------------------------------ timed multi-thread port (fallback CPU baseline) */

static double now_us(void) {
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec * 1e6 + ts.tv_nsec * 1e-3;
}
static int cmp_d(const void * a, const void * b) {
    double x = *(const double *) a, y = *(const double *) b;
    return x < y ? -1 : x > y;
}

/* one layer, OpenMP over neurons; per-thread y buffers merged in thread order (deterministic for a fixed nth) */
static void layer_mt(int arch, int type, const void * wu, const void * wg, const void * wd, const float * x,
                     const float * p, float * y, float * hid, int n_embd, int n_ff, int nth, float thr) {
    const size_t rs = spifo_row_size(type, n_embd);
    void *       a  = malloc(act_row_size(type, n_embd));
    convert_act(type, x, a, n_embd);
    float * ybuf = (float *) calloc((size_t) nth * n_embd, sizeof(float));
#pragma omp parallel num_threads(nth)
    {
#ifdef _OPENMP
        const int ith = omp_get_thread_num();
#else
        const int ith = 0;
#endif
        float * yb = ybuf + (size_t) ith * n_embd;
#pragma omp for schedule(static)
        for (int n = 0; n < n_ff; ++n) {
            hid[n] = 0.f;
            if (p[n] < thr) {
                continue;
            }
            float u = vec_dot(type, n_embd, (const char *) wu + rs * (size_t) n, a), g = 0.f, h;
            if (arch != SPIFO_ARCH_RELU_UP) {
                g = vec_dot(type, n_embd, (const char *) wg + rs * (size_t) n, a);
            }
            h = arch == SPIFO_ARCH_PROSPARSE ? ((g > 0.01f) ? g : 0.f) * u :
                arch == SPIFO_ARCH_RELU_GATED ? ((g > 0.f) ? g : 0.f) * ((u > 0.f) ? u : 0.f) :
                arch == SPIFO_ARCH_SILU_GATED ? (g / (1.0f + expf(-g))) * u :
                                                ((u > 0.f) ? u : 0.f);
            hid[n] = h;
            float alpha = type == SPIFO_F16 ? h2f(f2h(h)) : type == SPIFO_BF16 ? bf2f(f2bf(h)) : h;
            if (alpha == 0.f) {
                continue;
            }
            const char * wr = (const char *) wd + rs * (size_t) n;
            if (type == SPIFO_F16) {
                const uint16_t * v = (const uint16_t *) wr;
                for (int i = 0; i < n_embd; ++i) {
                    yb[i] = fmaf(h2f(v[i]), alpha, yb[i]);
                }
            } else if (type == SPIFO_BF16) {
                const uint16_t * v = (const uint16_t *) wr;
                for (int i = 0; i < n_embd; ++i) {
                    yb[i] = fmaf(bf2f(v[i]), alpha, yb[i]);
                }
            } else if (type == SPIFO_Q8_0) {
                const blk_q8_0 * b = (const blk_q8_0 *) wr;
                for (int ib = 0; ib < n_embd / QK; ++ib) {
                    const float sc = h2f(b[ib].d) * alpha;
                    for (int j = 0; j < QK; ++j) {
                        yb[ib * QK + j] = fmaf((float) b[ib].qs[j], sc, yb[ib * QK + j]);
                    }
                }
            } else {
                const blk_q4_0 * b = (const blk_q4_0 *) wr;
                for (int ib = 0; ib < n_embd / QK; ++ib) {
                    const float sc = h2f(b[ib].d) * alpha;
                    for (int j = 0; j < QK / 2; ++j) {
                        yb[ib * QK + j] = fmaf((float) ((b[ib].qs[j] & 0x0F) - 8), sc, yb[ib * QK + j]);
                        yb[ib * QK + j + QK / 2] =
                            fmaf((float) ((b[ib].qs[j] >> 4) - 8), sc, yb[ib * QK + j + QK / 2]);
                    }
                }
            }
        }
    }
    for (int i = 0; i < n_embd; ++i) {
        float s = 0.f;
        for (int t = 0; t < nth; ++t) {
            s += ybuf[(size_t) t * n_embd + i];
        }
        y[i] = s;
    }
    free(ybuf);
    free(a);
}

static void predictor_mt(int ptype, const void * wpu, const void * wpd, const float * bpd, const float * x,
                         float * out, int n_embd, int r, int n_ff, int nth) {
    const size_t rs_u = spifo_row_size(ptype, n_embd), rs_d = spifo_row_size(ptype, r);
    void *       ax = malloc(act_row_size(ptype, n_embd));
    void *       ah = malloc(act_row_size(ptype, r));
    float *      h  = (float *) malloc(sizeof(float) * (size_t) r);
    convert_act(ptype, x, ax, n_embd);
#pragma omp parallel for num_threads(nth) schedule(static)
    for (int j = 0; j < r; ++j) {
        float v = vec_dot(ptype, n_embd, (const char *) wpu + rs_u * (size_t) j, ax);
        h[j]    = v > 0.f ? v : 0.f;
    }
    convert_act(ptype, h, ah, r);
#pragma omp parallel for num_threads(nth) schedule(static)
    for (int i = 0; i < n_ff; ++i) {
        float z = vec_dot(ptype, r, (const char *) wpd + rs_d * (size_t) i, ah);
        if (bpd) {
            z = z + bpd[i];
        }
        out[i] = 1.f / (1.f + expf(-z));
    }
    free(ax);
    free(ah);
    free(h);
}

double spifo_ffn_stack_time(int arch, int wtype, int ptype, int n_layers, const void * const * w_up,
                            const void * const * w_gate, const void * const * w_down_t, const void * const * wpu,
                            const void * const * wpd, const float * const * bpd, const float * x, float * y,
                            int n_embd, int n_ff, int r, int n_tok, int n_threads, int warmup, int iters) {
    if (!type_ok(wtype) || n_tok < 1 || iters < 1) {
        return -1.0;
    }
    float *  p   = (float *) malloc(sizeof(float) * (size_t) n_ff * 2);
    float *  hid = (float *) malloc(sizeof(float) * (size_t) n_ff);
    float *  yl  = (float *) malloc(sizeof(float) * (size_t) n_embd);
    double * us  = (double *) malloc(sizeof(double) * (size_t) iters);
    for (int it = 0; it < warmup + iters; ++it) {
        const double t0 = now_us();
        for (int t = 0; t < n_tok; ++t) {
            float * cur = p, *nxt = p + n_ff;
            for (int l = 0; l < n_layers; ++l) {
                const float * xl = x + ((size_t) l * n_tok + t) * n_embd;
                if (l == 0) {
                    predictor_mt(ptype, wpu[0], wpd[0], bpd ? bpd[0] : NULL, xl, cur, n_embd, r, n_ff, n_threads);
                }
                if (l + 1 < n_layers) {
                    predictor_mt(ptype, wpu[l + 1], wpd[l + 1], bpd ? bpd[l + 1] : NULL, xl, nxt, n_embd, r, n_ff,
                                 n_threads);
                }
                layer_mt(arch, wtype, w_up[l], w_gate ? w_gate[l] : NULL, w_down_t[l], xl, cur, yl, hid, n_embd,
                         n_ff, n_threads, 0.5f);
                if (y) {
                    memcpy(y + ((size_t) l * n_tok + t) * n_embd, yl, sizeof(float) * (size_t) n_embd);
                }
                float * tmp = cur;
                cur         = nxt;
                nxt         = tmp;
            }
        }
        const double t1 = now_us();
        if (it >= warmup) {
            us[it - warmup] = t1 - t0;
        }
    }
    qsort(us, (size_t) iters, sizeof(double), cmp_d);
    const double med = us[iters / 2];
    free(p);
    free(hid);
    free(yl);
    free(us);
    return med;
}
