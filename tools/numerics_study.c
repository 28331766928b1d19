/*
 * tools/numerics_study.c -- CPU study of WHICH rounding points of a GPU pipeline cost parity on Q4 models.
 *
 * TEST / DESIGN TOOLING (same status as oracle/: never linked into build/libbert.so).  It includes the oracle's C file
 * for its bit-exact building blocks and adds one alternative forward pass whose deviations from the reference are
 * switched on one by one (bit mask), so that their effect on the final embedding can be measured on the CPU, without
 * GPU time:
 *     1   GEMM activation operand  a~ = d_a*q_a rounded to f16            (what a kind::f16 MMA can be fed)
 *     2   GEMM weight operand      w~ = d_w*(q_w-8) | d_w*q_w+m_w rounded to f16
 *     4   Q, K, V stored as f16
 *     8   residual stream read back as f16
 *     16  pre-LayerNorm sum stored as f16 (unfused GEMM -> add_ln path)
 *     32  a~ and w~ carried as f16 hi + f16 lo pairs (overrides 1 and 2: exact to 2^-22)
 * Mask 0 = exact operands, different summation order only (double accumulation): the floor any GPU pipeline has.
 * The Q8 rounding of the next GEMM's input (ggml-quants.c:702-747) amplifies every upstream perturbation: a value that
 * moves across a rounding boundary changes by a whole Q8 step (amax/127), so 2^-12 relative noise upstream becomes
 * ~1e-2 relative noise per token after 12-24 layers (SURVEY.md finding 4).
 *
 *     gcc -O2 -shared -fPIC -o /tmp/libstudy.so tools/numerics_study.c -lm -lpthread     (driver: tools/numerics_study.py)
 */
#include "../oracle/bert_oracle.c"

static inline float r16(float v) { return h2f(f2h(v)); }

/* y[n][N] = x[n][K] . W^T + b with the reference's activation quantisation, operands perturbed per `mask`, double accumulation */
typedef struct { const oracle_tensor *W; const float *bv; int n; const float *av; float *y; int mask; } emu_job;
static void emu_range(int64_t lo, int64_t hi, void *arg) {
    const emu_job *J = (const emu_job *)arg;
    const oracle_tensor *W = J->W;
    int64_t K = W->ne0, N = W->ne1;
    float *wf = malloc(sizeof(float) * (size_t)K);
    for (int64_t o = lo; o < hi; o++) {
        row_to_f32(W, o, wf);                                   /* exact fp32 decode: (q-8)*d, or fp32(q*d+m) */
        if ((J->mask & 2) && !(J->mask & 32)) for (int64_t k = 0; k < K; k++) wf[k] = r16(wf[k]);
        for (int i = 0; i < J->n; i++) {
            const float *a = J->av + (size_t)i * K;
            double s = 0.0;
            for (int64_t k = 0; k < K; k++) s += (double)a[k] * (double)wf[k];
            J->y[(size_t)i * N + o] = J->bv ? J->bv[o] + (float)s : (float)s;
        }
    }
    free(wf);
}
static void emu_linear(const oracle_tensor *W, const oracle_tensor *bias, int n, const float *x, float *y, int mask) {
    int64_t K = W->ne0, N = W->ne1;
    float *av = malloc(sizeof(float) * (size_t)n * K);
    if (W->type == T_Q4_0) for (int i = 0; i < n; i++) oracle_quant_q8_0_roundtrip(x + (size_t)i * K, av + (size_t)i * K, (int)K);
    else if (W->type == T_Q4_1) for (int i = 0; i < n; i++) oracle_quant_q8_1_roundtrip(x + (size_t)i * K, av + (size_t)i * K, (int)K);
    else if (W->type == T_F16) for (size_t i = 0; i < (size_t)n * K; i++) av[i] = r16(x[i]);
    else memcpy(av, x, sizeof(float) * (size_t)n * K);
    if ((mask & 1) && !(mask & 32)) for (size_t i = 0; i < (size_t)n * K; i++) av[i] = r16(av[i]);
    emu_job job = { W, bias ? (const float *)bias->data : NULL, n, av, y, mask };
    par_for(N, emu_range, &job);
    free(av);
}

ORACLE_API int study_forward(const oracle_model *m, const int32_t *tok, int n, float *out, float *layer_out, int mask) {
    init_tables();
    int d = m->n_embd, f = m->n_intermediate;
    if (n < 1 || n > m->n_max_tokens) return -1;
    size_t nd = (size_t)n * d;
    float *x = malloc(sizeof(float) * (nd * 8 + (size_t)n * f));
    float *q = x + nd, *k = q + nd, *v = k + nd, *a = v + nd, *y = a + nd, *z = y + nd, *r = z + nd, *hh = r + nd;
    oracle_embed(m, tok, n, x);
    if (layer_out) memcpy(layer_out, x, nd * 4);
    for (int l = 0; l < m->n_layer; l++) {
        const oracle_tensor *t = m->t + 5 + 16 * l;
        emu_linear(&t[0], &t[1], n, x, q, mask);
        emu_linear(&t[2], &t[3], n, x, k, mask);
        emu_linear(&t[4], &t[5], n, x, v, mask);
        if (mask & 4) for (size_t i = 0; i < nd; i++) { q[i] = r16(q[i]); k[i] = r16(k[i]); v[i] = r16(v[i]); }
        oracle_attention(n, d, m->n_head, q, k, v, a);
        emu_linear(&t[6], &t[7], n, a, y, mask);
        for (size_t i = 0; i < nd; i++) { float res = (mask & 8) ? r16(x[i]) : x[i]; y[i] = y[i] + res; if (mask & 16) y[i] = r16(y[i]); }
        oracle_layernorm(n, d, m->eps, y, (const float *)t[8].data, (const float *)t[9].data, y);
        emu_linear(&t[10], &t[11], n, y, hh, mask);
        oracle_gelu(n * f, hh, hh);
        emu_linear(&t[12], &t[13], n, hh, z, mask);
        for (size_t i = 0; i < nd; i++) { float res = (mask & 8) ? r16(y[i]) : y[i]; z[i] = res + z[i]; if (mask & 16) z[i] = r16(z[i]); }
        oracle_layernorm(n, d, m->eps, z, (const float *)t[14].data, (const float *)t[15].data, x);
        if (layer_out) memcpy(layer_out + (size_t)(l + 1) * nd, x, nd * 4);
    }
    oracle_pool(n, d, x, out);
    free(x);
    return 0;
}
