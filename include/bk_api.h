/*
 * include/bk_api.h -- the thin C ABI between the C++ host layer (api.cpp, host.cpp) and the
 * sm_100a CUDA engine (engine.cu + kernels).  Plain pointers and sizes only.
 *
 * One bk_engine = one GPU: device weights, activation workspaces, one stream.  The bert.h layer
 * (include/bert.h) owns one engine per GPU and shards a batch of independent sentences over them
 * (reference loop bert.cpp:1065 has no cross-sentence state).
 *
 * What each call replaces in the reference is cited per function.  There is no CPU fallback: every
 * function fails (non-zero return, message via bk_last_error()) when no CUDA device is usable.
 */
#ifndef BK_API_H
#define BK_API_H

#include <stddef.h>
#include <stdint.h>

#define BK_API __attribute__((visibility("default")))

#ifdef __cplusplus
extern "C" {
#endif

/* ggml_type ids as stored in GGUF (ggml.h:310-313) */
enum { BK_F32 = 0, BK_F16 = 1, BK_Q4_0 = 2, BK_Q4_1 = 3 };

/* bert_hparams, bert.cpp:20-31 */
typedef struct bk_hparams {
    int32_t n_vocab, n_max_tokens, n_embd, n_intermediate, n_head, n_layer;
    float   eps;
    int32_t wtype; /* file type = type of the 2-D weights (BK_F32..BK_Q4_1), bert.cpp:206-252 */
} bk_hparams;

/* kernel classes, for per-kernel timing (bk_engine_kernel_ms) */
enum {
    BK_K_PACK = 0, BK_K_EMBED_LN, BK_K_GEMM_QKV, BK_K_ATTN, BK_K_GEMM_O, BK_K_ADD_LN,
    BK_K_GEMM_FF1, BK_K_GEMM_FF2, BK_K_POOL, BK_K_COUNT
};

typedef struct bk_engine bk_engine;

BK_API int         bk_device_count(void);          /* <0 on CUDA error, 0 = no device */
BK_API const char *bk_last_error(void);            /* thread-local message of the last failure */

/* Replaces the compute arena + ggml context of bert_ctx (bert.cpp:80-137, 821-842): workspaces
 * for up to cap_tokens packed tokens / cap_seqs sentences per launch are allocated once. */
BK_API bk_engine *bk_engine_create(int device, const bk_hparams *hp, int32_t cap_tokens, int32_t cap_seqs);
BK_API void       bk_engine_destroy(bk_engine *e);

/* Replaces llm_load_tensors/load_all_data (bert.cpp:580-658): `slot` indexes the canonical
 * tensor order of bert.cpp:623-652 (0 word, 1 type, 2 pos, 3 ln_e.w, 4 ln_e.b, then 16 per layer:
 * q.w q.b k.w k.b v.w v.b o.w o.b ln_att.w ln_att.b ff_i.w ff_i.b ff_o.w ff_o.b ln_out.w ln_out.b).
 * `data` are the raw GGUF bytes of the tensor (host); ne0 = contiguous dim. */
BK_API int bk_engine_set_tensor(bk_engine *e, int32_t slot, int32_t ggml_type, const void *data,
                                int64_t ne0, int64_t ne1);
/* after all tensors: builds the fused QKV weight/bias, UMMA operand copies and TMA descriptors */
BK_API int bk_engine_finalize(bk_engine *e);

/* Replaces one pass of the bert_eval_batch loop body (bert.cpp:1065-1099) for n_seq sentences:
 * tokens = the sentences' ids concatenated (host), lens[n_seq] their lengths (host, each
 * 1..n_max_tokens, sum <= cap_tokens), out = n_seq rows of n_embd floats (host).
 * H2D copy, all kernels and the D2H copy run on the engine's stream; returns after it drained. */
BK_API int bk_engine_forward(bk_engine *e, const int32_t *tokens, const int32_t *lens, int32_t n_seq, float *out);

/* The same pass split in two so that consecutive batches overlap: submit stages the ids (pinned copy + H2D),
 * enqueues all kernels and the D2H copy of the embeddings on the engine's stream and returns at once; collect
 * waits for that batch and copies n_seq rows to `out`.  Two slots (0, 1): while the host collects and scatters
 * slot s, the batch submitted to slot 1-s is already running. */
BK_API int bk_engine_submit(bk_engine *e, int32_t slot, const int32_t *tokens, const int32_t *lens, int32_t n_seq);
BK_API int bk_engine_collect(bk_engine *e, int32_t slot, float *out);
/* submit straight from the caller's ragged arrays (bert_eval_batch's `batch_tokens` / `n_tokens`, bert.h:79-85): sent_tokens[i] points at
 * sentence i's ids, sent_lens[i] is its length, in the CALLER'S order -- sorting by length, packing and positions happen on the device
 * (rowwise.cu: batch_sort / batch_pack; the reference loops over sentences instead, bert.cpp:1065).  The n_seq embeddings are copied,
 * in the caller's order, to host_out (pinned memory from bk_host_alloc; NULL = the slot's own buffer, fetched by bk_engine_collect(out)).
 * bk_engine_collect(e, slot, NULL) then only waits for the batch. */
BK_API int bk_engine_submit_ragged(bk_engine *e, int32_t slot, const int32_t *const *sent_tokens, const int32_t *sent_lens, int32_t n_seq,
                                   float *host_out);
/* page-locked host memory usable by every engine's stream as an asynchronous copy target */
BK_API void *bk_host_alloc(size_t bytes);
BK_API void  bk_host_free(void *p);
/* the small-batch (latency) path replays captured CUDA graphs: launches served by a replay, shapes captured (on their second
 * sighting), launches done directly (first sighting) since the engine was created */
BK_API int bk_engine_graph_stats(bk_engine *e, int64_t *replayed, int64_t *captured, int64_t *direct);
/* the same counters summed over the engines of a bert.h context (`ctx` = what bert_load_from_file returned) */
struct bert_ctx;
BK_API int bk_bert_graph_stats(struct bert_ctx *ctx, int64_t *replayed, int64_t *captured, int64_t *direct);

/* Same work with the inputs already resident in HBM (bench.py `value` leg): stage a whole job once
 * (any number of sentences, in the caller's order; it is cut into micro-batches of <= cap_tokens /
 * cap_seqs), then run the kernels over all of it `iters` times, timed with CUDA events on the
 * engine's stream.  ms_total = elapsed ms for all iters; per-kernel-class ms via bk_engine_kernel_ms(). */
BK_API int bk_engine_stage(bk_engine *e, const int32_t *tokens, const int32_t *lens, int32_t n_seq);
BK_API int bk_engine_run_staged(bk_engine *e, int32_t iters, int32_t time_kernels, float *ms_total);
BK_API int bk_engine_read_output(bk_engine *e, float *out, int32_t n_seq);
/* ms[BK_K_COUNT], launches[BK_K_COUNT]: accumulated over the last bk_engine_run_staged(time_kernels=1) */
BK_API int bk_engine_kernel_ms(bk_engine *e, float *ms, int32_t *launches);
/* intermediate activations of the last forward, widened to f32 (tests): which = 0 embeddings+LN
 * output, 1..n_layer = output of that encoder layer; out[T][n_embd] */
BK_API int bk_engine_read_hidden(bk_engine *e, int32_t which, float *out);
BK_API int bk_engine_set_keep_hidden(bk_engine *e, int32_t on);
BK_API int64_t bk_engine_launch_count(bk_engine *e); /* kernels launched since creation */
/* engines created under BERT_B200_GUARD=1 carry a 4 KB canary on both sides of every activation / index buffer: the number of
 * canary bytes overwritten so far (0 = no kernel wrote out of bounds), -1 on error.  Stand-in for compute-sanitizer memcheck. */
BK_API int64_t bk_engine_check_guards(bk_engine *e);
/* cudaProfilerStart (on=1) / cudaProfilerStop (on=0): lets `ncu --profile-from-start off` see only the timed region */
BK_API int bk_profiler_range(int32_t on);

/* ---- single-kernel entry points (tests/, profiles): host buffers in, host buffers out --------
 * f16 buffers are raw IEEE binary16 bit patterns (uint16_t). */
enum { BK_EPI_BIAS = 0, BK_EPI_BIAS_RESID = 1, BK_EPI_BIAS_GELU = 2, BK_EPI_BIAS_GELU_Q8_0 = 3, BK_EPI_BIAS_GELU_Q8_1 = 4 };
/* out[M][N] = epi(A[M][K] . W[N][K]^T + bias[N] (+ resid[M][N]))   -- ggml_mul_mat + bias add,
 * ggml.c:9388-9597, bert.cpp:909-910.  wtype selects how W is stored (BK_F16 raw halves,
 * BK_Q4_0/BK_Q4_1 GGUF blocks along K, BK_F32 floats).  epi may carry a forced tile width in bits 8..:
 * epi | (BN << 8) | (CG << 20), BN in {64,128,192,256}, CG in {1,2} CTAs per tile (0 = automatic choice);
 * bit 28 set with Q4 weights: W stays packed in device memory and is decoded in shared memory by the GEMM (per K stage);
 * bit 29: the weight-stationary kernel (gemm_ws.cu: each CTA pair decodes its column tile once; K <= 1024, N % 128 == 0). */
BK_API int bk_test_gemm(int device, int32_t M, int32_t N, int32_t K, const uint16_t *A, const void *W,
                        int32_t wtype, const float *bias, const uint16_t *resid, int32_t epi,
                        uint16_t *out, float *ms);
/* the fused projection + residual + LayerNorm kernel (gemm_ln.cu): y = LN(A.W^T + bias + resid)*gamma + beta, f16 in /
 * out, W f16 [N][K]; yq = the Q8-rounded copy when qmode 1=Q8_0 / 2=Q8_1.  N must be 384, 768 or 1024.
 * Replaces ggml_mul_mat + add + add + ggml_norm + mul + add, bert.cpp:945-963 and 974-992. */
BK_API int bk_test_gemm_ln(int device, int32_t M, int32_t N, int32_t K, const uint16_t *A, const uint16_t *W,
                           const float *bias, const uint16_t *resid, const float *gamma, const float *beta, float eps,
                           int32_t qmode, uint16_t *y, uint16_t *yq, float *ms);
/* the GEMM epilogue's GELU alone: y[i] = f16(gelu(f16(x[i]))) -- ggml_vec_gelu_f32's f16 table, ggml.c:1418-1441 */
BK_API int bk_test_gelu(int device, int32_t n, const float *x, float *y);
/* LN(x)*gamma+beta over rows of `pre` (f16) -> y f16 (+ yq = Q8-rounded copy when qmode 1=Q8_0, 2=Q8_1);
 * ggml_norm + mul + add, ggml.c:8976-9004, bert.cpp:958-963 */
BK_API int bk_test_add_ln(int device, int32_t T, int32_t d, float eps, const uint16_t *pre, const float *gamma,
                          const float *beta, int32_t qmode, uint16_t *y, uint16_t *yq);
/* softmax(Q.K^T/sqrt(dh)).V per sentence and head on packed qkv[T][3d] (f16); bert.cpp:928-943 */
BK_API int bk_test_attention(int device, int32_t n_seq, const int32_t *lens, int32_t d, int32_t n_head,
                             const uint16_t *qkv, int32_t qmode, uint16_t *out);
/* the same for short sentences sharing 128-row tiles (attention_pack.cuh): lens = n64 sentences of 33..64 tokens, then n32 of 17..32,
 * then n16 of 1..16 (the order the on-device batcher produces); qkv / out as above */
BK_API int bk_test_attention_packed(int device, int32_t n64, int32_t n32, int32_t n16, const int32_t *lens, int32_t d, int32_t n_head,
                                    const uint16_t *qkv, int32_t qmode, uint16_t *out);
/* mean over tokens + L2 normalise; bert.cpp:995-1006 */
BK_API int bk_test_pool(int device, int32_t n_seq, const int32_t *lens, int32_t d, const uint16_t *x, float *out);
/* the on-device batcher alone: lens[n_seq] and the ids concatenated in the caller's order -> order[n_seq] (sorted position ->
 * caller index; stable, longest first), cu[n_seq+1], packed tok[T] and pos[T] */
BK_API int bk_test_batcher(int device, int32_t n_seq, const int32_t *lens, const int32_t *tokens, int32_t *order, int32_t *cu,
                           int32_t *tok, int32_t *pos);
/* The on-device batcher for callers that cannot allocate device memory (CUDA graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Same kernels and results as bk_test_batcher / the engine's batcher, but every pointer is a DEVICE pointer owned by the caller
 * (d_lens[n_seq], d_tokens[n_tokens] in; d_order[n_seq], d_cu[n_seq+1], d_tok[n_tokens], d_pos[n_tokens] out), the two launches go
 * to `stream` (a cudaStream_t; NULL = the default stream) without a synchronisation, and the kernels' scratch (two int32 per
 * sentence) is a fixed 512 KiB __device__ array inside the library instead of a cudaMalloc'd workspace: the call allocates and
 * frees nothing.  n_tokens = sum of the lengths (each clamped to 1..512 as everywhere else).
 * Limit: n_seq <= BK_STATIC_BATCHER_MAX_SEQ, larger batches are rejected (BK_STATIC_E_TOO_LARGE, message on stderr).
 * All calls on a device share the scratch: calls whose kernels overlap on the GPU (different streams) race; calls on one stream
 * are safe.  The first call on each device resolves the array's address (cudaGetSymbolAddress) -- make it outside stream capture.
 * Returns 0, or BK_STATIC_E_ARG / BK_STATIC_E_TOO_LARGE / BK_STATIC_E_CUDA. */
#define BK_STATIC_BATCHER_MAX_SEQ 65536
#define BK_STATIC_E_ARG (-1)
#define BK_STATIC_E_TOO_LARGE (-2)
#define BK_STATIC_E_CUDA (-3)
BK_API int bk_batcher_static_scratch(int device, int32_t n_seq, int32_t n_tokens, const int32_t *d_lens, const int32_t *d_tokens,
                                     int32_t *d_order, int32_t *d_cu, int32_t *d_tok, int32_t *d_pos, void *stream);
/* embeddings gather + LayerNorm alone: x = LN((type[0] + word[tok]) + pos[p]) * gamma + beta, tables as raw GGUF bytes of `wtype`
 * (word [n_vocab][d], type [2][d], posw [n_pos][d]); x f16 [T][d], xq its Q8-rounded copy (qmode 1 / 2).  bert.cpp:868-898,
 * ggml_get_rows ggml.c:10108-10224 */
BK_API int bk_test_embed_ln(int device, int32_t T, int32_t d, int32_t n_vocab, int32_t n_pos, int32_t wtype, const void *word, const void *type,
                            const void *posw, const float *gamma, const float *beta, float eps, const int32_t *tokens, const int32_t *pos,
                            int32_t qmode, uint16_t *x, uint16_t *xq);

/* ---- host-side tokenizer alone (no GPU needed; tests/test_tokenizer.py) -------------------------
 * Replaces bert_tokenizer::load / encode (tokenizer.cpp:30-53 -> Rust tokenizers crate):
 * json = contents of the GGUF key blob.tokenizer.json; encode() = ids without [CLS]/[SEP]
 * (add_special_tokens = false, huggingface_tokenizer.cc:28-35).  Returns the number of ids
 * (written up to max_ids), or -1. */
typedef struct bk_tokenizer bk_tokenizer;
BK_API bk_tokenizer *bk_tokenizer_create(const char *json, int64_t json_len);
BK_API void          bk_tokenizer_destroy(bk_tokenizer *t);
BK_API int32_t       bk_tokenizer_encode(const bk_tokenizer *t, const char *text, int32_t *ids, int32_t max_ids);

#ifdef __cplusplus
}
#endif
#endif /* BK_API_H */
