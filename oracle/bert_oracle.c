/*
 * oracle/bert_oracle.c -- CPU restatement of the reference's bert_eval path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the product (build/libbert.so) links,
 * loads or calls this file; only tests/, __graft_entry__.smoke() and the
 * cpu_baseline / --impl reference legs of bench.py may.  It is an independent,
 * single-file C restatement of what the reference computes for one sentence,
 * written op by op from the reference sources cited at each function, with the
 * same rounding points as the reference's x86 AVX2/FMA/F16C build
 * (CMakeLists.txt:164-177):
 *     - f16 look-up-table GELU and softmax-exp           ggml.c:1433-1441,10547-10549,2193-2201
 *     - F16 weights: activations rounded to f16          ggml.c:9493-9509,315-332
 *     - Q4_0/Q4_1 weights: activations quantised to Q8_0/Q8_1 per 32 and
 *       integer dot products inside an fp32 scale        ggml-quants.c:702-747,950-985,2500-2523,2808-2841
 *     - 32-lane (4x8) fp32 FMA partial sums and their reduction tree
 *                                                         ggml.c:757-790,1163-1234
 *     - LayerNorm / sum / softmax denominators in double ggml.c:8976-9004,10536-10558
 *
 * Parity pin: tests/test_oracle.py runs this file against
 * oracle/_ref/libbert_ref.so (the reference's own sources compiled here, see
 * oracle/build_ref.sh) on the same GGUF files and token ids; the reference
 * ships no golden embedding vectors (examples/test_embedding.cpp only prints).
 * The outputs of oracle/_ref are also committed under tests/golden/.
 */
#include <math.h>
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <unistd.h>

#define ORACLE_API __attribute__((visibility("default")))

/* ------------------------------------------------------------------ tiny pthread parallel-for
 * Work is split by OUTPUT element, every dot product is computed whole by one thread, so results do not
 * depend on the thread count (same property as the reference, ggml.c:9558-9596). */
static int g_threads = 0;
ORACLE_API void oracle_set_threads(int n) { g_threads = n; }
typedef void (*range_fn)(int64_t lo, int64_t hi, void *arg);
typedef struct { range_fn fn; void *arg; int64_t lo, hi; } par_job;
static void *par_tramp(void *p) { par_job *j = (par_job *)p; j->fn(j->lo, j->hi, j->arg); return NULL; }
static void par_for(int64_t n, range_fn fn, void *arg) {
    int nt = g_threads;
    if (nt <= 0) { const char *e = getenv("ORACLE_THREADS"); nt = e ? atoi(e) : (int)sysconf(_SC_NPROCESSORS_ONLN); }
    if (nt > 64) nt = 64;
    if (nt > n) nt = (int)n;
    if (nt <= 1) { fn(0, n, arg); return; }
    pthread_t th[64]; par_job jb[64];
    int64_t per = (n + nt - 1) / nt;
    for (int t = 0; t < nt; t++) {
        jb[t].fn = fn; jb[t].arg = arg; jb[t].lo = t * per; jb[t].hi = (t + 1) * per < n ? (t + 1) * per : n;
        if (t) pthread_create(&th[t], NULL, par_tramp, &jb[t]);
    }
    par_tramp(&jb[0]);
    for (int t = 1; t < nt; t++) pthread_join(th[t], NULL);
}

enum { T_F32 = 0, T_F16 = 1, T_Q4_0 = 2, T_Q4_1 = 3 }; /* ggml.h:310-313 */

typedef struct {
    const void *data; /* raw GGUF bytes, row-major, ne0 contiguous */
    int32_t type;
    int32_t _pad;
    int64_t ne0, ne1;
} oracle_tensor;

typedef struct {
    int32_t n_vocab, n_max_tokens, n_embd, n_intermediate, n_head, n_layer;
    float eps;
    int32_t _pad;
    /* canonical order: word, type, pos, ln_e_w, ln_e_b, then per layer
       q_w q_b k_w k_b v_w v_b o_w o_b ln_att_w ln_att_b ff_i_w ff_i_b ff_o_w ff_o_b ln_out_w ln_out_b
       (embedding.cpp_b200/gguf_io.py: tensor_names) */
    const oracle_tensor *t;
} oracle_model;

/* ------------------------------------------------------------------ fp16
 * IEEE binary16 <-> binary32, round-to-nearest-even, what F16C's
 * _cvtss_sh(x,0) / _cvtsh_ss do (ggml-impl.h:85-101). */
static inline float h2f(uint16_t h) {
    uint32_t s = (uint32_t)(h & 0x8000u) << 16, e = (h >> 10) & 0x1Fu, m = h & 0x3FFu, u;
    if (e == 0) {
        if (m == 0) u = s;
        else { int sh = 0; while (!(m & 0x400u)) { m <<= 1; sh++; } m &= 0x3FFu; u = s | ((uint32_t)(113 - sh) << 23) | (m << 13); }
    } else if (e == 31) u = s | 0x7F800000u | (m << 13);
    else u = s | ((e + 112u) << 23) | (m << 13);
    float f; memcpy(&f, &u, 4); return f;
}
static inline uint16_t f2h(float f) {
    uint32_t u; memcpy(&u, &f, 4);
    uint32_t s = (u >> 16) & 0x8000u, a = u & 0x7FFFFFFFu;
    if (a >= 0x7F800000u) return (uint16_t)(s | 0x7C00u | ((a > 0x7F800000u) ? (0x200u | ((a >> 13) & 0x3FFu)) : 0));
    if (a >= 0x477FF000u) return (uint16_t)(s | 0x7C00u);                /* rounds to inf */
    if (a < 0x33000001u) return (uint16_t)s;                             /* rounds to zero (<= 2^-25) */
    int32_t e = (int32_t)(a >> 23) - 127;
    uint32_t m = (a & 0x7FFFFFu) | 0x800000u;
    int shift = (e < -14) ? (13 + (-14 - e)) : 13;                        /* subnormal: extra shift */
    uint32_t q = m >> shift, rem = m & ((1u << shift) - 1u), half = 1u << (shift - 1);
    if (rem > half || (rem == half && (q & 1u))) q++;
    uint32_t r = (e < -14) ? q : (((uint32_t)(e + 15) << 10) + (q - 0x400u));
    return (uint16_t)(s | r);
}
ORACLE_API float oracle_h2f(uint16_t h) { return h2f(h); }
ORACLE_API uint16_t oracle_f2h(float f) { return f2h(f); }

/* ------------------------------------------------------------------ f16 tables (ggml.c:2193-2201) */
static uint16_t tab_gelu[1 << 16], tab_exp[1 << 16];
static int tab_ready = 0;
static inline float gelu_f32(float x) { /* ggml.c:1418-1424 */
    return 0.5f * x * (1.0f + tanhf(0.79788456080286535587989211986876f * x * (1.0f + 0.044715f * x * x)));
}
static void init_tables(void) {
    if (tab_ready) return;
    for (int i = 0; i < (1 << 16); i++) {
        float f = h2f((uint16_t)i);
        tab_gelu[i] = f2h(gelu_f32(f));
        tab_exp[i] = f2h(expf(f));
    }
    tab_ready = 1;
}
ORACLE_API void oracle_init(void) { init_tables(); }
/* y = GELU16(x): ggml_vec_gelu_f32 with GGML_GELU_FP16, ggml.c:1433-1441 */
ORACLE_API void oracle_gelu(int n, const float *x, float *y) {
    init_tables();
    for (int i = 0; i < n; i++) y[i] = h2f(tab_gelu[f2h(x[i])]);
}

/* ------------------------------------------------------------------ dot products */
/* reduction tree GGML_F32x8_REDUCE (ggml.c:771-789): acc[4][8] -> scalar */
static inline float reduce32(float acc[4][8]) {
    float t[8];
    for (int l = 0; l < 8; l++) { float a = acc[0][l] + acc[2][l], b = acc[1][l] + acc[3][l]; t[l] = a + b; }
    float u0 = t[0] + t[4], u1 = t[1] + t[5], u2 = t[2] + t[6], u3 = t[3] + t[7];
    return (u0 + u1) + (u2 + u3);
}
/* ggml_vec_dot_f32, ggml.c:1163-1198 (GGML_F32_STEP 32, 4 accumulators of 8 lanes, FMA) */
static float dot_f32(int n, const float *x, const float *y) {
    float acc[4][8] = {{0}};
    int np = n & ~31;
    for (int i = 0; i < np; i += 32)
        for (int j = 0; j < 4; j++)
            for (int l = 0; l < 8; l++) acc[j][l] = fmaf(x[i + 8 * j + l], y[i + 8 * j + l], acc[j][l]);
    float s = reduce32(acc);
    for (int i = np; i < n; i++) { float p = x[i] * y[i]; s += p; }
    return s;
}
/* ggml_vec_dot_f16, ggml.c:1200-1234: both sides widened to f32, same lanes, leftovers in double */
static float dot_f16(int n, const uint16_t *x, const uint16_t *y) {
    float acc[4][8] = {{0}};
    int np = n & ~31;
    for (int i = 0; i < np; i += 32)
        for (int j = 0; j < 4; j++)
            for (int l = 0; l < 8; l++) acc[j][l] = fmaf(h2f(x[i + 8 * j + l]), h2f(y[i + 8 * j + l]), acc[j][l]);
    double s = (double)reduce32(acc);
    for (int i = np; i < n; i++) s += (double)(h2f(x[i]) * h2f(y[i]));
    return (float)s;
}
/* hsum_float_8, ggml-quants.c:67-73 */
static inline float hsum8(const float a[8]) {
    float r0 = a[4] + a[0], r1 = a[5] + a[1], r2 = a[6] + a[2], r3 = a[7] + a[3];
    r0 += r2; r1 += r3;
    return r0 + r1;
}

typedef struct { uint16_t d; int8_t qs[32]; } q8_0_blk;        /* block_q8_0, ggml-quants.h:40-46 */
typedef struct { float d, s; int8_t qs[32]; } q8_1_blk;        /* block_q8_1, ggml-quants.h:49-55 */

static inline float rne(float v) { return nearbyintf(v); }      /* _mm256_round_ps(_MM_ROUND_NEAREST) */

/* quantize_row_q8_0, AVX2 branch, ggml-quants.c:702-747 */
static void quant_q8_0(const float *x, q8_0_blk *y, int k) {
    for (int b = 0; b < k / 32; b++, x += 32) {
        float amax = 0.0f;
        for (int j = 0; j < 32; j++) { float a = fabsf(x[j]); if (a > amax) amax = a; }
        float d = amax / 127.f;
        y[b].d = f2h(d);
        float id = (amax != 0.0f) ? 127.f / amax : 0.0f;
        for (int j = 0; j < 32; j++) y[b].qs[j] = (int8_t)(int)rne(x[j] * id);
    }
}
/* quantize_row_q8_1, AVX2 branch, ggml-quants.c:944-985: d kept in f32, s = d * sum(q) */
static void quant_q8_1(const float *x, q8_1_blk *y, int k) {
    for (int b = 0; b < k / 32; b++, x += 32) {
        float amax = 0.0f;
        for (int j = 0; j < 32; j++) { float a = fabsf(x[j]); if (a > amax) amax = a; }
        float d = amax / 127.f;
        y[b].d = d;
        float id = (amax != 0.0f) ? 127.f / amax : 0.0f;
        int sum = 0;
        for (int j = 0; j < 32; j++) { int q = (int)rne(x[j] * id); y[b].qs[j] = (int8_t)q; sum += q; }
        y[b].s = d * (float)sum;
    }
}
/* ggml_vec_dot_q4_0_q8_0, AVX2 branch, ggml-quants.c:2500-2523.
   block_q4_0 = { f16 d; u8 qs[16] }: element j in the low nibble of qs[j], element j+16 in the high nibble. */
static float dot_q4_0_q8_0(int n, const uint8_t *wx, const q8_0_blk *y) {
    float acc[8] = {0};
    for (int b = 0; b < n / 32; b++) {
        const uint8_t *blk = wx + 18 * b;
        uint16_t dh; memcpy(&dh, blk, 2);
        float d = h2f(dh) * h2f(y[b].d);
        for (int l = 0; l < 8; l++) {
            int s = 0;
            for (int j = 4 * l; j < 4 * l + 4; j++) {
                int q = (j < 16) ? (blk[2 + j] & 0x0F) : (blk[2 + j - 16] >> 4);
                s += (q - 8) * (int)y[b].qs[j];
            }
            acc[l] = fmaf(d, (float)s, acc[l]);
        }
    }
    return hsum8(acc);
}
/* ggml_vec_dot_q4_1_q8_1, AVX2 branch, ggml-quants.c:2808-2841.  block_q4_1 = { f16 d; f16 m; u8 qs[16] } */
static float dot_q4_1_q8_1(int n, const uint8_t *wx, const q8_1_blk *y) {
    float acc[8] = {0};
    float summs = 0;
    for (int b = 0; b < n / 32; b++) {
        const uint8_t *blk = wx + 20 * b;
        uint16_t dh, mh; memcpy(&dh, blk, 2); memcpy(&mh, blk + 2, 2);
        float d0 = h2f(dh), d1 = y[b].d;
        { float p = h2f(mh) * y[b].s; summs += p; }
        float d0d1 = d0 * d1;
        for (int l = 0; l < 8; l++) {
            int s = 0;
            for (int j = 4 * l; j < 4 * l + 4; j++) {
                int q = (j < 16) ? (blk[4 + j] & 0x0F) : (blk[4 + j - 16] >> 4);
                s += q * (int)y[b].qs[j];
            }
            acc[l] = fmaf(d0d1, (float)s, acc[l]);
        }
    }
    return hsum8(acc) + summs;
}

/* ------------------------------------------------------------------ rows of a weight tensor as f32
 * ggml_compute_forward_get_rows_{q,f16,f32}, ggml.c:10108-10224; decode ggml-quants.c:1074-1113 */
static void row_to_f32(const oracle_tensor *t, int64_t r, float *out) {
    int64_t k = t->ne0;
    if (t->type == T_F32) memcpy(out, (const float *)t->data + r * k, (size_t)k * 4);
    else if (t->type == T_F16) { const uint16_t *p = (const uint16_t *)t->data + r * k; for (int64_t i = 0; i < k; i++) out[i] = h2f(p[i]); }
    else if (t->type == T_Q4_0) {
        const uint8_t *p = (const uint8_t *)t->data + r * (k / 32) * 18;
        for (int64_t b = 0; b < k / 32; b++, p += 18) {
            uint16_t dh; memcpy(&dh, p, 2); float d = h2f(dh);
            for (int j = 0; j < 16; j++) {
                out[b * 32 + j] = (float)((p[2 + j] & 0x0F) - 8) * d;
                out[b * 32 + j + 16] = (float)((p[2 + j] >> 4) - 8) * d;
            }
        }
    } else {
        const uint8_t *p = (const uint8_t *)t->data + r * (k / 32) * 20;
        for (int64_t b = 0; b < k / 32; b++, p += 20) {
            uint16_t dh, mh; memcpy(&dh, p, 2); memcpy(&mh, p + 2, 2); float d = h2f(dh), m = h2f(mh);
            for (int j = 0; j < 16; j++) {
                out[b * 32 + j] = (float)(p[4 + j] & 0x0F) * d + m;
                out[b * 32 + j + 16] = (float)(p[4 + j] >> 4) * d + m;
            }
        }
    }
}
ORACLE_API void oracle_dequant_row(const oracle_tensor *t, int64_t r, float *out) { row_to_f32(t, r, out); }

/* ------------------------------------------------------------------ Lin(x; W, b) = b + mul_mat(W, x)
 * ggml_compute_forward_mul_mat, ggml.c:9388-9597 (INIT converts src1 rows to the weight's vec_dot_type,
 * COMPUTE does one whole vec_dot per output element); bias add bert.cpp:909-910.
 * x [n][K] f32, W [N][K] in its file type, y [n][N] f32. */
typedef struct { const oracle_tensor *W; const float *bv; int n; const float *x; float *y; const void *conv; size_t rs; } lin_job;
/* Same arithmetic as dot_f16 / dot_q4_0_q8_0 / dot_q4_1_q8_1 above (kept as the readable statement and
   used by the unit tests); here the weight row is decoded once per output row instead of once per dot. */
static float dot_f16_pre(int n, const float *xw, const float *ya) {  /* both already widened f16 values */
    float acc[4][8] = {{0}};
    int np = n & ~31;
    for (int i = 0; i < np; i += 32)
        for (int j = 0; j < 4; j++)
            for (int l = 0; l < 8; l++) acc[j][l] = fmaf(xw[i + 8 * j + l], ya[i + 8 * j + l], acc[j][l]);
    double s = (double)reduce32(acc);
    for (int i = np; i < n; i++) s += (double)(xw[i] * ya[i]);
    return (float)s;
}
static void lin_range(int64_t lo, int64_t hi, void *arg) {
    const lin_job *J = (const lin_job *)arg;
    const oracle_tensor *W = J->W;
    int64_t K = W->ne0, N = W->ne1;
    int nb = (int)(K / 32);
    float *wf = malloc(sizeof(float) * (size_t)K);          /* F16: widened row */
    int8_t *wq = malloc((size_t)K);                          /* Q4: nibbles (already -8 for Q4_0) */
    float *wd = malloc(sizeof(float) * 2 * (size_t)(nb + 1)), *wm = wd + nb + 1;
    for (int64_t o = lo; o < hi; o++) {
        if (W->type == T_F16) { const uint16_t *p = (const uint16_t *)W->data + o * K; for (int64_t k = 0; k < K; k++) wf[k] = h2f(p[k]); }
        else if (W->type == T_Q4_0 || W->type == T_Q4_1) {
            int bb = (W->type == T_Q4_0) ? 18 : 20, off = bb - 16, sub = (W->type == T_Q4_0) ? 8 : 0;
            const uint8_t *p = (const uint8_t *)W->data + o * nb * bb;
            for (int b = 0; b < nb; b++, p += bb) {
                uint16_t h; memcpy(&h, p, 2); wd[b] = h2f(h);
                if (bb == 20) { memcpy(&h, p + 2, 2); wm[b] = h2f(h); }
                for (int j = 0; j < 16; j++) { wq[b * 32 + j] = (int8_t)((p[off + j] & 0x0F) - sub); wq[b * 32 + j + 16] = (int8_t)((p[off + j] >> 4) - sub); }
            }
        }
        for (int i = 0; i < J->n; i++) {
            float v;
            if (W->type == T_F32) v = dot_f32((int)K, (const float *)W->data + o * K, J->x + (size_t)i * K);
            else if (W->type == T_F16) v = dot_f16_pre((int)K, wf, (const float *)J->conv + (size_t)i * K);
            else if (W->type == T_Q4_0) {
                const q8_0_blk *y = (const q8_0_blk *)((const char *)J->conv + J->rs * i);
                float acc[8] = {0};
                for (int b = 0; b < nb; b++) {
                    float d = wd[b] * h2f(y[b].d);
                    for (int l = 0; l < 8; l++) {
                        int s = 0;
                        for (int j = 4 * l; j < 4 * l + 4; j++) s += (int)wq[b * 32 + j] * (int)y[b].qs[j];
                        acc[l] = fmaf(d, (float)s, acc[l]);
                    }
                }
                v = hsum8(acc);
            } else {
                const q8_1_blk *y = (const q8_1_blk *)((const char *)J->conv + J->rs * i);
                float acc[8] = {0}, summs = 0;
                for (int b = 0; b < nb; b++) {
                    { float p = wm[b] * y[b].s; summs += p; }
                    float d0d1 = wd[b] * y[b].d;
                    for (int l = 0; l < 8; l++) {
                        int s = 0;
                        for (int j = 4 * l; j < 4 * l + 4; j++) s += (int)wq[b * 32 + j] * (int)y[b].qs[j];
                        acc[l] = fmaf(d0d1, (float)s, acc[l]);
                    }
                }
                v = hsum8(acc) + summs;
            }
            J->y[(size_t)i * N + o] = J->bv ? J->bv[o] + v : v;
        }
    }
    free(wf); free(wq); free(wd);
}
ORACLE_API void oracle_linear(const oracle_tensor *W, const oracle_tensor *bias, int n, const float *x, float *y) {
    int64_t K = W->ne0, N = W->ne1;
    const float *bv = bias ? (const float *)bias->data : NULL;
    void *conv = NULL;
    size_t rs = 0;
    if (W->type == T_F16) { rs = (size_t)K * 4; conv = malloc(rs * n);   /* ggml_fp32_to_fp16_row, kept widened */
        for (int i = 0; i < n; i++) for (int64_t k = 0; k < K; k++) ((float *)conv)[(size_t)i * K + k] = h2f(f2h(x[(size_t)i * K + k]));
    } else if (W->type == T_Q4_0) { rs = (size_t)(K / 32) * sizeof(q8_0_blk); conv = malloc(rs * n);
        for (int i = 0; i < n; i++) quant_q8_0(x + (size_t)i * K, (q8_0_blk *)((char *)conv + rs * i), (int)K);
    } else if (W->type == T_Q4_1) { rs = (size_t)(K / 32) * sizeof(q8_1_blk); conv = malloc(rs * n);
        for (int i = 0; i < n; i++) quant_q8_1(x + (size_t)i * K, (q8_1_blk *)((char *)conv + rs * i), (int)K);
    }
    lin_job job = { W, bv, n, x, y, conv, rs };
    par_for(N, lin_range, &job);
    free(conv);
}

/* ------------------------------------------------------------------ LayerNorm
 * ggml_compute_forward_norm_f32 (ggml.c:8976-9004) then w*x (ggml_mul) then +b (ggml_add), bert.cpp:889-898 */
ORACLE_API void oracle_layernorm(int n, int d, float eps, const float *x, const float *w, const float *b, float *y) {
    for (int i = 0; i < n; i++) {
        const float *xr = x + (size_t)i * d; float *yr = y + (size_t)i * d;
        double sum = 0.0;
        for (int k = 0; k < d; k++) sum += (double)xr[k];
        float mean = (float)(sum / d);
        double sum2 = 0.0;
        for (int k = 0; k < d; k++) { float v = xr[k] - mean; yr[k] = v; sum2 += (double)(v * v); }
        float variance = (float)(sum2 / d);
        const float scale = 1.0f / sqrtf(variance + eps);
        for (int k = 0; k < d; k++) { float v = yr[k] * scale; v = w[k] * v; yr[k] = v + b[k]; }
    }
}

/* ------------------------------------------------------------------ embeddings (bert.cpp:868-898)
 * x = LN((type[0] + word[tok]) + pos[i]) ; token type is always 0 (bert.cpp:871-872) */
ORACLE_API void oracle_embed(const oracle_model *m, const int32_t *tok, int n, float *x) {
    int d = m->n_embd;
    float *tw = malloc(sizeof(float) * d * 3), *tt = tw + d, *tp = tt + d;
    row_to_f32(&m->t[1], 0, tt);
    for (int i = 0; i < n; i++) {
        row_to_f32(&m->t[0], tok[i], tw);
        row_to_f32(&m->t[2], i, tp);
        for (int k = 0; k < d; k++) { float a = tt[k] + tw[k]; x[(size_t)i * d + k] = tp[k] + a; }
    }
    free(tw);
    oracle_layernorm(n, d, m->eps, x, (const float *)m->t[3].data, (const float *)m->t[4].data, x);
}

/* ------------------------------------------------------------------ attention (bert.cpp:928-943)
 * q,k,v [n][d] f32 (heads interleaved along d).  Per head: S = K.Q (dot over dh), S *= 1/sqrt(dh)
 * (separate f32 multiply, ggml.c:9884), row softmax with the f16 exp table and a double denominator
 * (ggml.c:10536-10558), O = P.V (dot over n).  No mask: one sentence per graph. */
typedef struct { int n, d, H; const float *q, *k, *v; float *o; } att_job;
static void att_range(int64_t lo, int64_t hi, void *arg) {
    const att_job *J = (const att_job *)arg;
    int n = J->n, d = J->d, H = J->H;
    const float *q = J->q, *k = J->k, *v = J->v; float *o = J->o;
    int dh = d / H;
    const float scale = 1.0f / sqrt((float)dh); /* bert.cpp:934: double sqrt, rounded to f32 */
    for (int h = (int)lo; h < (int)hi; h++) {
        float *S = malloc(sizeof(float) * n), *vt = malloc(sizeof(float) * (size_t)n * dh);
        for (int e = 0; e < dh; e++) for (int j = 0; j < n; j++) vt[(size_t)e * n + j] = v[(size_t)j * d + h * dh + e];
        for (int i = 0; i < n; i++) {
            float mx = -INFINITY;
            for (int j = 0; j < n; j++) {
                float s = dot_f32(dh, k + (size_t)j * d + h * dh, q + (size_t)i * d + h * dh);
                s = s * scale;
                S[j] = s;
                if (s > mx) mx = s;
            }
            double sum = 0.0;
            for (int j = 0; j < n; j++) { float p = h2f(tab_exp[f2h(S[j] - mx)]); sum += (double)p; S[j] = p; }
            const float inv = (float)(1.0 / sum);
            for (int j = 0; j < n; j++) S[j] *= inv;
            for (int e = 0; e < dh; e++) o[(size_t)i * d + h * dh + e] = dot_f32(n, vt + (size_t)e * n, S);
        }
        free(S); free(vt);
    }
}
ORACLE_API void oracle_attention(int n, int d, int H, const float *q, const float *k, const float *v, float *o) {
    init_tables();
    att_job job = { n, d, H, q, k, v, o };
    par_for(H, att_range, &job);
}

/* ------------------------------------------------------------------ pooling + normalise (bert.cpp:995-1006)
 * e = dot(x[:,k], 1/n) (f32 vec_dot over tokens), len = sqrtf((float)sum_double(e^2)), out = e * (1/len) */
ORACLE_API void oracle_pool(int n, int d, const float *x, float *out) {
    float *col = malloc(sizeof(float) * n * 2), *w = col + n;
    const float inv_n = 1.0f / n;
    for (int i = 0; i < n; i++) w[i] = inv_n;
    for (int k = 0; k < d; k++) {
        for (int i = 0; i < n; i++) col[i] = x[(size_t)i * d + k];
        out[k] = dot_f32(n, col, w);
    }
    free(col);
    double s = 0.0;
    for (int k = 0; k < d; k++) { float sq = out[k] * out[k]; s += (double)sq; }
    float len = sqrtf((float)s);
    float sc = 1.0f / len;
    for (int k = 0; k < d; k++) out[k] *= sc;
}

/* ------------------------------------------------------------------ whole forward pass (bert_build, bert.cpp:845-1012)
 * layer_out (optional): [n_layer+1][n][d] f32 -- output of the embedding LN and of every encoder layer. */
ORACLE_API int oracle_forward(const oracle_model *m, const int32_t *tok, int n, float *out, float *layer_out) {
    init_tables();
    int d = m->n_embd, f = m->n_intermediate;
    if (n < 1 || n > m->n_max_tokens) return -1;
    size_t nd = (size_t)n * d;
    float *x = malloc(sizeof(float) * (nd * 7 + (size_t)n * f));
    if (!x) return -2;
    float *q = x + nd, *k = q + nd, *v = k + nd, *a = v + nd, *y = a + nd, *z = y + nd, *hh = z + nd;
    oracle_embed(m, tok, n, x);
    if (layer_out) memcpy(layer_out, x, nd * 4);
    for (int l = 0; l < m->n_layer; l++) {
        const oracle_tensor *t = m->t + 5 + 16 * l;
        oracle_linear(&t[0], &t[1], n, x, q);
        oracle_linear(&t[2], &t[3], n, x, k);
        oracle_linear(&t[4], &t[5], n, x, v);
        oracle_attention(n, d, m->n_head, q, k, v, a);
        oracle_linear(&t[6], &t[7], n, a, y);                       /* bert.cpp:945-947 */
        for (size_t i = 0; i < nd; i++) y[i] = y[i] + x[i];         /* bert.cpp:951 */
        oracle_layernorm(n, d, m->eps, y, (const float *)t[8].data, (const float *)t[9].data, y);
        oracle_linear(&t[10], &t[11], n, y, hh);                    /* bert.cpp:967-970 */
        oracle_gelu(n * f, hh, hh);                                 /* bert.cpp:971 */
        oracle_linear(&t[12], &t[13], n, hh, z);                    /* bert.cpp:974-977 */
        for (size_t i = 0; i < nd; i++) z[i] = y[i] + z[i];         /* bert.cpp:981 */
        oracle_layernorm(n, d, m->eps, z, (const float *)t[14].data, (const float *)t[15].data, x);
        if (layer_out) memcpy(layer_out + (size_t)(l + 1) * nd, x, nd * 4);
    }
    oracle_pool(n, d, x, out);
    free(x);
    return 0;
}

/* ------------------------------------------------------------------ single dot products (unit tests) */
ORACLE_API float oracle_dot_f16(int n, const uint16_t *w, const float *x) {
    uint16_t *xh = malloc(2 * (size_t)n); for (int i = 0; i < n; i++) xh[i] = f2h(x[i]);
    float r = dot_f16(n, w, xh); free(xh); return r;
}
ORACLE_API float oracle_dot_q4_0(int n, const uint8_t *w, const float *x) {
    q8_0_blk *b = malloc(sizeof(q8_0_blk) * (n / 32)); quant_q8_0(x, b, n);
    float r = dot_q4_0_q8_0(n, w, b); free(b); return r;
}
ORACLE_API float oracle_dot_q4_1(int n, const uint8_t *w, const float *x) {
    q8_1_blk *b = malloc(sizeof(q8_1_blk) * (n / 32)); quant_q8_1(x, b, n);
    float r = dot_q4_1_q8_1(n, w, b); free(b); return r;
}

/* ------------------------------------------------------------------ codecs exposed for kernel unit tests */
ORACLE_API void oracle_quant_q8_0_roundtrip(const float *x, float *y, int k) { /* y = f16(d) * q : what a Q4_0 GEMM sees */
    q8_0_blk *b = malloc(sizeof(q8_0_blk) * (k / 32));
    quant_q8_0(x, b, k);
    for (int i = 0; i < k; i++) y[i] = h2f(b[i / 32].d) * (float)b[i / 32].qs[i % 32];
    free(b);
}
ORACLE_API void oracle_quant_q8_1_roundtrip(const float *x, float *y, int k) {
    q8_1_blk *b = malloc(sizeof(q8_1_blk) * (k / 32));
    quant_q8_1(x, b, k);
    for (int i = 0; i < k; i++) y[i] = b[i / 32].d * (float)b[i / 32].qs[i % 32];
    free(b);
}
