/*
 * include/bert.h -- the drop-in boundary of build/libbert.so.
 *
 * Same symbols, argument meaning and ABI as the reference's public header
 * (/root/reference/bert.h:1-108): a program or ctypes client written against the
 * reference (examples/main.cpp, examples/sample_dylib.py:17-57,
 * benchmarks/run_mteb.py:39-89) links/loads this library unchanged.  Behind the
 * symbols the reference's per-sentence ggml CPU graph (bert.cpp:845-1108) is
 * replaced by the sm_100a engine under embedding.cpp_b200/csrc/.
 *
 * Each declaration cites the reference interface it replaces.
 */
#ifndef BERT_H
#define BERT_H

#include <stdbool.h>
#include <stddef.h>
#include <stdint.h>

#define BERT_API __attribute__((visibility("default")))

#ifdef __cplusplus
extern "C" {
#endif

/* bert.h:19-24 (declared by the reference, unused by it as well) */
enum llama_log_level {
    LLAMA_LOG_LEVEL_ERROR = 2,
    LLAMA_LOG_LEVEL_WARN  = 3,
    LLAMA_LOG_LEVEL_INFO  = 4
};

/* bert.h:26-33.  Default member initialisers only exist for C++ callers, as in the reference. */
struct bert_params {
#ifdef __cplusplus
    int32_t n_threads  = 6;
    int32_t port       = 8080;
    const char *model  = "models/all-MiniLM-L6-v2/ggml-model-q4_0.bin";
    const char *prompt = "test prompt";
#else
    int32_t n_threads;
    int32_t port;
    const char *model;
    const char *prompt;
#endif
};

#ifdef __cplusplus
/* bert.h:35, bert.cpp:697-733: -t/--threads, -p/--prompt, --port, -m/--model, -h */
BERT_API bool bert_params_parse(int argc, char **argv, bert_params &params);
#endif

struct bert_ctx;
typedef int32_t bert_vocab_id; /* bert.h:39 */

/* bert.h:41, bert.cpp:783-819.  Parses the GGUF file, checks tensor names/shapes as
 * bert.cpp:623-652 does, uploads the packed weight bytes to every GPU the context uses
 * (env BERT_B200_DEVICES="0,1,.." or BERT_B200_NGPU=n; default: all visible).
 * Returns NULL (message on stderr) on any failure, including "no CUDA device". */
BERT_API struct bert_ctx *bert_load_from_file(const char *fname);
/* bert.h:42, bert.cpp:1014-1018 */
BERT_API void bert_free(struct bert_ctx *ctx);

/* bert.h:46-50, bert.cpp:1110-1117: tokenise + eval one text */
BERT_API void bert_encode(struct bert_ctx *ctx, int32_t n_threads, const char *texts, float *embeddings);

/* bert.h:54-60, bert.cpp:1119-1198: tokenise n_inputs texts, embed them all.  The reference
 * forces n_batch_size to 1 (bert.cpp:1128); here it is ignored too: texts are tokenised on host threads in chunks of 16384
 * (the next chunk while the current one is on the GPUs), each chunk is cut into token-budget micro-batches that the device sorts by
 * length and packs. */
BERT_API void bert_encode_batch(struct bert_ctx *ctx, int32_t n_threads, int32_t n_batch_size,
                                int32_t n_inputs, const char **texts, float **embeddings);

/* bert.h:64-69, bert.cpp:738-781: [CLS] ids.. [SEP], truncated so that the last id is SEP */
BERT_API void bert_tokenize(struct bert_ctx *ctx, const char *text, bert_vocab_id *tokens,
                            int32_t *n_tokens, int32_t n_max_tokens);

/* bert.h:71-76, bert.cpp:1020-1028: one sentence; embeddings == NULL is tolerated (the
 * reference's memory-probe mode, bert.cpp:1039) and does nothing. n_threads is ignored. */
BERT_API void bert_eval(struct bert_ctx *ctx, int32_t n_threads, bert_vocab_id *tokens,
                        int32_t n_tokens, float *embeddings);

/* bert.h:78-85, bert.cpp:1030-1108: ragged host arrays in, bert_n_embd() floats per sentence
 * out (mean-pooled over all tokens, L2-normalised; graph node n-1, see SURVEY.md finding 1).
 * Any n_tokens[b] > bert_n_max_tokens() or < 1: message on stderr, nothing is computed
 * and no output is written (bert.cpp:1071-1075). */
BERT_API void bert_eval_batch(struct bert_ctx *ctx, int32_t n_threads, int32_t n_batch_size,
                              bert_vocab_id **batch_tokens, int32_t *n_tokens, float **batch_embeddings);

/* bert.h:87-88, bert.cpp:661-669 */
BERT_API int32_t bert_n_embd(struct bert_ctx *ctx);
BERT_API int32_t bert_n_max_tokens(struct bert_ctx *ctx);
/* bert.h:90, bert.cpp:671-675 (NULL instead of a C++ exception for an id out of range) */
BERT_API const char *bert_vocab_id_to_token(struct bert_ctx *ctx, bert_vocab_id id);

/* bert.h:92, bert.cpp:1313-1599: F32/F16 GGUF -> Q4_0 (ftype 2) / Q4_1 (ftype 3) GGUF.
 * Host-side file tool (not a GPU kernel); same block codec as ggml-quants.c:428-510. */
BERT_API bool bert_model_quantize(const char *fname_inp, const char *fname_out, int ftype);

#ifdef __cplusplus
}
#endif

/* bert.h:98-106 (declared by the reference, unused) */
typedef struct llama_model_quantize_params {
    int nthread;
    bool allow_requantize;
    bool quantize_output_tensor;
    bool only_copy;
} llama_model_quantize_params;

#endif /* BERT_H */
