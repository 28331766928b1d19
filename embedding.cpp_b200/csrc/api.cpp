// api.cpp -- the bert.h exports of libbert.so (include/bert.h) on top of the CUDA engine (bk_api.h).
//
// Host-side replacement of the reference's model layer: bert_loader / bert_load_from_file
// (bert.cpp:143-659, 783-819), bert_eval / bert_eval_batch (bert.cpp:1020-1108), bert_encode_batch
// (bert.cpp:1119-1198), bert_tokenize (bert.cpp:738-781).  Where the reference loops over sentences and
// builds one CPU graph each (bert.cpp:1065), this layer cuts the call into token-budget micro-batches
// and deals those to persistent workers, one per engine / GPU (sentences are independent: no collective,
// host-side gather only); each micro-batch is sorted by length and packed on the device.
#include <algorithm>
#include <atomic>
#include <condition_variable>
#include <functional>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <mutex>
#include <numeric>
#include <string>
#include <thread>
#include <vector>

#include "../../include/bert.h"
#include "../../include/bk_api.h"
#include "gguf.h"
#include "tokenizer_wp.h"

using namespace bk;

// One persistent worker thread per engine (created at load): bert_eval_batch posts a job, the workers run it, the caller waits.
// Spawning std::threads per call (the first version) cost more than a 128-sentence shard's GPU time at 8 GPUs.
class WorkerPool {
  public:
    explicit WorkerPool(int n) : n_(n) {
        for (int i = 1; i < n_; i++) th_.emplace_back([this, i] { loop(i); });  // worker 0 is the calling thread
    }
    ~WorkerPool() {
        { std::lock_guard<std::mutex> l(mu_); stop_ = true; }
        cv_.notify_all();
        for (auto &t : th_) t.join();
    }
    // runs fn(worker) on workers 0..n_active-1 (worker 0 = caller) and returns when all are done
    void run(int n_active, const std::function<void(int)> &fn) {
        if (n_active > n_) n_active = n_;
        if (n_active <= 1) { fn(0); return; }
        {
            std::lock_guard<std::mutex> l(mu_);
            fn_ = &fn; active_ = n_active; pending_ = n_active - 1; gen_++;
        }
        cv_.notify_all();
        fn(0);
        std::unique_lock<std::mutex> l(mu_);
        done_.wait(l, [this] { return pending_ == 0; });
        fn_ = nullptr;
    }

  private:
    void loop(int id) {
        uint64_t seen = 0;
        for (;;) {
            const std::function<void(int)> *fn = nullptr;
            {
                std::unique_lock<std::mutex> l(mu_);
                cv_.wait(l, [&] { return stop_ || gen_ != seen; });
                if (stop_) return;
                seen = gen_;
                if (id < active_) fn = fn_;
            }
            if (!fn) continue;
            (*fn)(id);
            std::lock_guard<std::mutex> l(mu_);
            if (--pending_ == 0) done_.notify_one();
        }
    }
    int n_;
    std::vector<std::thread> th_;
    std::mutex mu_;
    std::condition_variable cv_, done_;
    const std::function<void(int)> *fn_ = nullptr;
    int active_ = 0, pending_ = 0;
    uint64_t gen_ = 0;
    bool stop_ = false;
};

struct bert_ctx {
    GgufFile file;
    bk_hparams hp{};
    std::vector<std::string> vocab;
    int32_t cls_id = -1, sep_id = -1, pad_id = -1, unk_id = 0;  // defaults bert.cpp:119-124
    WordPieceTokenizer tok;
    bool tok_ok = false;
    std::vector<bk_engine *> engines;  // n_gpus distinct devices, engines.size() / n_gpus engines on each
    int n_gpus = 1;
    int cap_tokens = 0, cap_seqs = 0;
    std::unique_ptr<WorkerPool> pool;
    // the job's embeddings land here (pinned; every engine's device->host copy targets its own rows) and are handed to the caller's
    // rows only after every micro-batch succeeded: a failing call leaves the caller's outputs untouched (bert.h:83)
    float *job_out = nullptr;
    size_t job_out_rows = 0;
    std::mutex mu;  // one in-flight call per context (the reference is not re-entrant either, bert.cpp:1056-1062)
};

namespace {

int env_int(const char *name, int dflt) {
    const char *v = getenv(name);
    return (v && *v) ? atoi(v) : dflt;
}

// canonical tensor order of bert.cpp:623-652 == slot order of bk_engine_set_tensor
std::vector<std::string> tensor_names(int n_layer) {
    std::vector<std::string> n = {"embeddings.word_embeddings.weight", "embeddings.token_type_embeddings.weight",
                                  "embeddings.position_embeddings.weight", "embeddings.LayerNorm.weight", "embeddings.LayerNorm.bias"};
    static const char *per[16] = {"attention.self.query.weight", "attention.self.query.bias", "attention.self.key.weight",
                                  "attention.self.key.bias", "attention.self.value.weight", "attention.self.value.bias",
                                  "attention.output.dense.weight", "attention.output.dense.bias", "attention.output.LayerNorm.weight",
                                  "attention.output.LayerNorm.bias", "intermediate.dense.weight", "intermediate.dense.bias",
                                  "output.dense.weight", "output.dense.bias", "output.LayerNorm.weight", "output.LayerNorm.bias"};
    for (int l = 0; l < n_layer; l++)
        for (int k = 0; k < 16; k++) n.push_back("encoder.layer." + std::to_string(l) + "." + per[k]);
    return n;
}

// BERT_B200_ENGINES_PER_GPU (default 2) engines per listed GPU, each with its own stream and activation buffers: two
// micro-batches of a job are then in flight per GPU, so the copies of one overlap the kernels of the other and the
// SMs a kernel leaves empty (cluster placement, last wave) run the other micro-batch.  Order g0 g1 .. g0 g1 ..: the
// first workers of a small job land on distinct GPUs.
std::vector<int> replicate_devices(const std::vector<int> &devs) {
    int per = env_int("BERT_B200_ENGINES_PER_GPU", 2);
    if (per < 1) per = 1;
    if (per > 4) per = 4;
    std::vector<int> out;
    for (int r = 0; r < per; r++) out.insert(out.end(), devs.begin(), devs.end());
    return out;
}

std::vector<int> pick_devices() {
    std::vector<int> devs;
    const int n = bk_device_count();
    if (n <= 0) return devs;
    if (const char *s = getenv("BERT_B200_DEVICES")) {
        const char *p = s;
        while (*p) {
            char *end;
            long v = strtol(p, &end, 10);
            if (end == p) break;
            if (v >= 0 && v < n) devs.push_back((int)v);
            p = (*end == ',') ? end + 1 : end;
        }
        if (!devs.empty()) return devs;
    }
    int use = env_int("BERT_B200_NGPU", n);
    if (use < 1) use = 1;
    if (use > n) use = n;
    for (int i = 0; i < use; i++) devs.push_back(i);
    return devs;
}

struct MicroBatch { int begin, end, tokens; };  // sentence range of the job, in the caller's order

}  // namespace

extern "C" {

BERT_API struct bert_ctx *bert_load_from_file(const char *fname) {
    std::unique_ptr<bert_ctx> ctx(new bert_ctx());
    std::string err;
    if (!ctx->file.open(fname, err)) { fprintf(stderr, "%s: failed to load '%s': %s\n", __func__, fname, err.c_str()); return nullptr; }
    const GgufFile &g = ctx->file;
    uint32_t v;
    bk_hparams &hp = ctx->hp;
    // llm_load_hparams, bert.cpp:496-513
    const GgufValue *toks = g.find("tokenizer.ggml.tokens");
    if (!toks || toks->type != GT_ARR || toks->elem_type != GT_STR) { fprintf(stderr, "%s: key not found in model: tokenizer.ggml.tokens\n", __func__); return nullptr; }
    ctx->vocab = toks->strs;
    hp.n_vocab = (int32_t)ctx->vocab.size();
    struct { const char *key; int32_t *dst; } req[] = {{"bert.context_length", &hp.n_max_tokens}, {"bert.embedding_length", &hp.n_embd},
                                                       {"bert.feed_forward_length", &hp.n_intermediate}, {"bert.attention.head_count", &hp.n_head},
                                                       {"bert.block_count", &hp.n_layer}};
    for (auto &r : req) {
        if (!g.get_u32(r.key, v)) { fprintf(stderr, "%s: key not found in model: %s\n", __func__, r.key); return nullptr; }
        *r.dst = (int32_t)v;
    }
    if (!g.get_f32("bert.attention.layer_norm_epsilon", hp.eps)) { fprintf(stderr, "%s: key not found in model: bert.attention.layer_norm_epsilon\n", __func__); return nullptr; }
    if (hp.n_layer < 1 || hp.n_embd < 1 || hp.n_head < 1 || hp.n_max_tokens < 1) { fprintf(stderr, "%s: invalid hyper-parameters\n", __func__); return nullptr; }
    if (hp.n_max_tokens > 512) {  // fail at load, not at the first long sentence: the attention kernels hold at most 512 keys per sentence
        fprintf(stderr, "%s: bert.context_length %d is not supported (maximum 512)\n", __func__, hp.n_max_tokens);
        return nullptr;
    }
    // llm_load_tokenizer, bert.cpp:515-578 (special ids optional)
    if (g.get_u32("tokenizer.ggml.cls_token_id", v)) ctx->cls_id = (int32_t)v;
    if (g.get_u32("tokenizer.ggml.seperator_token_id", v)) ctx->sep_id = (int32_t)v;  // (sic) gguf.h:297
    if (g.get_u32("tokenizer.ggml.padding_token_id", v)) ctx->pad_id = (int32_t)v;
    if (g.get_u32("tokenizer.ggml.unknown_token_id", v)) ctx->unk_id = (int32_t)v;
    // blob.tokenizer.json is a required key in the reference (bert.cpp:576), and a tokenizer that cannot be built must not turn every
    // text into [CLS][SEP] silently.  BERT_B200_ALLOW_NO_TOKENIZER=1 opts in to an ids-only context (bert_eval / bert_eval_batch work,
    // bert_tokenize / bert_encode* report an error and leave their outputs untouched).
    {
        const bool allow = env_int("BERT_B200_ALLOW_NO_TOKENIZER", 0) != 0;
        const GgufValue *blob = g.find("blob.tokenizer.json");
        std::string terr = "key not found in model: blob.tokenizer.json";
        if (blob) ctx->tok_ok = ctx->tok.load(blob->s, terr);
        if (!ctx->tok_ok) {
            fprintf(stderr, "%s: %s tokenizer not usable (%s)%s\n", __func__, allow ? "warning:" : "error:", terr.c_str(),
                    allow ? "; bert_tokenize / bert_encode disabled" : " -- set BERT_B200_ALLOW_NO_TOKENIZER=1 to load for bert_eval only");
            if (!allow) return nullptr;
        }
    }
    // file type = most frequent tensor type among 2-D tensors (bert.cpp:206-252 counts all tensors; the
    // 2-D ones decide which GEMM path is used and are what the converter/quantiser change)
    const std::vector<std::string> names = tensor_names(hp.n_layer);
    int cnt[4] = {0, 0, 0, 0};
    for (const auto &nm : names) {
        const GgufTensor *t = g.tensor(nm);
        if (!t) { fprintf(stderr, "%s: tensor '%s' not found\n", __func__, nm.c_str()); return nullptr; }
        if (t->type > 3 || t->nbytes == 0) { fprintf(stderr, "%s: tensor '%s' has unsupported type %u (supported: F32, F16, Q4_0, Q4_1)\n", __func__, nm.c_str(), t->type); return nullptr; }
        if (t->n_dims > 2 || t->ne[2] != 1 || t->ne[3] != 1) { fprintf(stderr, "%s: tensor '%s' has more than 2 dimensions\n", __func__, nm.c_str()); return nullptr; }
        if (t->n_dims == 2) cnt[t->type]++;
    }
    hp.wtype = (int32_t)(std::max_element(cnt, cnt + 4) - cnt);

    const std::vector<int> gpus = pick_devices();
    const std::vector<int> devs = replicate_devices(gpus);
    ctx->n_gpus = std::max<int>(1, (int)gpus.size());
    if (devs.empty()) { fprintf(stderr, "%s: no usable CUDA device (%s); this library has no CPU path\n", __func__, bk_last_error()); return nullptr; }
    ctx->cap_tokens = env_int("BERT_B200_MB_TOKENS", 65536);
    if (ctx->cap_tokens < hp.n_max_tokens) ctx->cap_tokens = hp.n_max_tokens;
    ctx->cap_seqs = std::min(ctx->cap_tokens, env_int("BERT_B200_MB_SEQS", 8192));
    for (int dev : devs) {
        bk_engine *e = bk_engine_create(dev, &hp, ctx->cap_tokens, ctx->cap_seqs);
        if (!e) { fprintf(stderr, "%s: engine on device %d: %s\n", __func__, dev, bk_last_error()); bert_free(ctx.release()); return nullptr; }
        ctx->engines.push_back(e);
        for (size_t i = 0; i < names.size(); i++) {
            const GgufTensor *t = g.tensor(names[i]);
            if (bk_engine_set_tensor(e, (int)i, (int)t->type, g.tensor_data(*t), (int64_t)t->ne[0], t->n_dims > 1 ? (int64_t)t->ne[1] : 1) != 0) {
                fprintf(stderr, "%s: tensor '%s': %s\n", __func__, names[i].c_str(), bk_last_error());
                bert_free(ctx.release());
                return nullptr;
            }
        }
        if (bk_engine_finalize(e) != 0) { fprintf(stderr, "%s: %s\n", __func__, bk_last_error()); bert_free(ctx.release()); return nullptr; }
    }
    ctx->pool.reset(new WorkerPool((int)ctx->engines.size()));
    if (env_int("BERT_B200_VERBOSE", 0))
        fprintf(stderr, "%s: %s: d=%d f=%d heads=%d layers=%d vocab=%d wtype=%d on %d GPU(s), %zu engine(s), micro-batch %d tokens\n", __func__, fname,
                hp.n_embd, hp.n_intermediate, hp.n_head, hp.n_layer, hp.n_vocab, hp.wtype, ctx->n_gpus, ctx->engines.size(), ctx->cap_tokens);
    return ctx.release();
}

BERT_API void bert_free(struct bert_ctx *ctx) {
    if (!ctx) return;
    ctx->pool.reset();
    for (bk_engine *e : ctx->engines) bk_engine_destroy(e);
    bk_host_free(ctx->job_out);
    delete ctx;
}

BERT_API int32_t bert_n_embd(struct bert_ctx *ctx) { return ctx->hp.n_embd; }
BERT_API int32_t bert_n_max_tokens(struct bert_ctx *ctx) { return ctx->hp.n_max_tokens; }
BERT_API const char *bert_vocab_id_to_token(struct bert_ctx *ctx, bert_vocab_id id) {
    if (id < 0 || id >= (int32_t)ctx->vocab.size()) return nullptr;
    return ctx->vocab[id].c_str();
}

// sentences [b0, b0 + n) of the caller's arrays: micro-batches cut in the caller's order by token budget, dealt to the engines; the
// embeddings land in ctx->job_out (row i = sentence b0 + i).  Sorting by length, packing and the return to the caller's order happen on
// the device (on-device batcher, rowwise.cu).  Returns false if any engine failed.
static bool run_job(bert_ctx *ctx, int b0, int n, bert_vocab_id **batch_tokens, const int32_t *n_tokens) {
    const int N = ctx->hp.n_max_tokens, d = ctx->hp.n_embd;
    long long total = 0;
    for (int b = 0; b < n; b++) total += n_tokens[b0 + b];
    const int n_eng = (int)ctx->engines.size();
    long long budget = ctx->cap_tokens;
    // several GPUs: one share per GPU (a GPU's second engine only takes work when its share exceeds a micro-batch)
    if (ctx->n_gpus > 1) budget = std::min<long long>(budget, std::max<long long>((total + ctx->n_gpus - 1) / ctx->n_gpus, N));
    std::vector<MicroBatch> mbs;
    for (int i = 0; i < n;) {
        int j = i, t = 0;
        while (j < n && j - i < ctx->cap_seqs && t + n_tokens[b0 + j] <= budget) { t += n_tokens[b0 + j]; j++; }
        mbs.push_back({i, j, t});
        i = j;
    }
    std::atomic<int> next{0};
    std::atomic<int> failed{0};
    const std::function<void(int)> worker = [&](int ei) {
        // two micro-batches in flight per engine: while batch k runs on the GPU, batch k+1 is staged, uploaded and enqueued
        bk_engine *e = ctx->engines[ei];
        bool pending[2] = {false, false};
        auto wait_slot = [&](int slot) -> bool {
            if (!pending[slot]) return true;
            pending[slot] = false;
            if (bk_engine_collect(e, slot, nullptr) != 0) {
                fprintf(stderr, "bert_eval_batch: %s\n", bk_last_error());
                failed.store(1);
                return false;
            }
            return true;
        };
        for (int k = 0;; k++) {
            const int slot = k & 1;
            const int m = next.fetch_add(1);
            if (m >= (int)mbs.size() || failed.load()) break;
            if (!wait_slot(slot)) break;  // the slot's previous batch (k-2) must have left its staging buffers
            const MicroBatch &mb = mbs[m];
            if (bk_engine_submit_ragged(e, slot, batch_tokens + b0 + mb.begin, n_tokens + b0 + mb.begin, mb.end - mb.begin,
                                        ctx->job_out + (size_t)mb.begin * d) != 0) {
                fprintf(stderr, "bert_eval_batch: %s\n", bk_last_error());
                failed.store(1);
                break;
            }
            pending[slot] = true;
        }
        wait_slot(0);
        wait_slot(1);
    };
    ctx->pool->run(std::min<int>(n_eng, (int)mbs.size()), worker);
    return failed.load() == 0;
}

BERT_API void bert_eval_batch(struct bert_ctx *ctx, int32_t n_threads, int32_t n_batch_size, bert_vocab_id **batch_tokens,
                              int32_t *n_tokens, float **batch_embeddings) {
    (void)n_threads;
    if (!ctx || n_batch_size <= 0) return;
    if (!batch_embeddings) return;  // memory-probe mode of the reference (bert.cpp:1039): nothing to measure here
    const int N = ctx->hp.n_max_tokens, d = ctx->hp.n_embd;
    for (int b = 0; b < n_batch_size; b++) {
        if (n_tokens[b] > N) { fprintf(stderr, "Too many tokens, maximum is %d\n", N); return; }  // bert.cpp:1071-1075
        if (n_tokens[b] < 1) { fprintf(stderr, "Empty input %d: at least one token is required\n", b); return; }
    }
    std::lock_guard<std::mutex> lock(ctx->mu);
    // jobs of at most kJobRows sentences (bounds the pinned result buffer); the all-or-nothing guarantee is per job
    const int kJobRows = 1 << 18;
    const size_t want = (size_t)std::min(n_batch_size, kJobRows);
    if (ctx->job_out_rows < want) {
        bk_host_free(ctx->job_out);
        ctx->job_out_rows = 0;
        size_t rows = 1024;
        while (rows < want) rows *= 2;
        ctx->job_out = (float *)bk_host_alloc(rows * (size_t)d * sizeof(float));
        if (!ctx->job_out) { fprintf(stderr, "bert_eval_batch: %s\n", bk_last_error()); return; }
        ctx->job_out_rows = rows;
    }
    for (int b0 = 0; b0 < n_batch_size; b0 += kJobRows) {
        const int n = std::min(kJobRows, n_batch_size - b0);
        if (!run_job(ctx, b0, n, batch_tokens, n_tokens)) return;  // outputs of this job untouched
        // hand the rows to the caller (the only host pass over the embeddings); large jobs: spread over the workers
        const std::function<void(int)> scatter = [&](int w) {
            const int nw = std::min<int>((int)ctx->engines.size(), std::max(1, n / 2048));
            for (int i = (int)((long long)n * w / nw), hi = (int)((long long)n * (w + 1) / nw); i < hi; i++)
                memcpy(batch_embeddings[b0 + i], ctx->job_out + (size_t)i * d, (size_t)d * sizeof(float));
        };
        ctx->pool->run(std::min<int>((int)ctx->engines.size(), std::max(1, n / 2048)), scatter);
    }
}

BERT_API void bert_eval(struct bert_ctx *ctx, int32_t n_threads, bert_vocab_id *tokens, int32_t n_tokens, float *embeddings) {
    bert_eval_batch(ctx, n_threads, 1, &tokens, &n_tokens, embeddings ? &embeddings : nullptr);  // bert.cpp:1020-1028
}

BERT_API void bert_tokenize(struct bert_ctx *ctx, const char *text, bert_vocab_id *tokens, int32_t *n_tokens, int32_t n_max_tokens) {
    // bert.cpp:738-781: [CLS] + ids (stop at the first PAD id or when full) + [SEP]; when full the last slot is SEP
    *n_tokens = 0;
    if (n_max_tokens < 2) return;
    if (!ctx->tok_ok) { fprintf(stderr, "bert_tokenize: this context was loaded without a usable tokenizer\n"); return; }
    const std::vector<int32_t> ids = ctx->tok.encode(text ? text : "");
    int t = 0;
    tokens[t++] = ctx->cls_id;
    for (int32_t id : ids) {
        if (id == ctx->pad_id) break;
        if (t >= n_max_tokens) break;
        tokens[t++] = id;
    }
    if (t >= n_max_tokens) tokens[n_max_tokens - 1] = ctx->sep_id;
    else tokens[t++] = ctx->sep_id;
    *n_tokens = t;
}

BERT_API void bert_encode_batch(struct bert_ctx *ctx, int32_t n_threads, int32_t n_batch_size, int32_t n_inputs, const char **texts,
                                float **embeddings) {
    (void)n_batch_size;  // the reference overrides it to 1 (bert.cpp:1128); here the call is cut into micro-batches by token count
    if (!ctx || n_inputs <= 0) return;
    if (!ctx->tok_ok) { fprintf(stderr, "bert_encode_batch: this context was loaded without a usable tokenizer; outputs untouched\n"); return; }
    const int N = ctx->hp.n_max_tokens;
    // Chunks of kChunk texts, two id buffers: chunk c + 1 is tokenised (spread over host threads) while chunk c is on the GPUs.
    // (The reference tokenises and evaluates text by text, bert.cpp:1138-1190.)  A failing chunk leaves its own outputs untouched.
    const int kChunk = 16384;
    const int chunk_cap = std::min(n_inputs, kChunk);
    struct Chunk {
        std::unique_ptr<int32_t[]> ids;
        std::vector<int32_t *> ptrs;
        std::vector<int32_t> lens;
    } ch[2];
    for (auto &c : ch) {
        c.ids.reset(new int32_t[(size_t)chunk_cap * N]);  // not zero-filled: only the tokens written are ever read
        c.ptrs.resize(chunk_cap);
        c.lens.resize(chunk_cap);
        if (n_inputs <= kChunk) break;  // one chunk: one buffer
    }
    const int hw = (int)std::max(1u, std::thread::hardware_concurrency());
    auto tokenize = [&](Chunk &c, int first, int n) {
        const int nt = std::max(1, std::min({hw, n / 64, 32}));  // tokenisation is independent per text
        auto range = [&](int lo, int hi) {
            for (int i = lo; i < hi; i++) {
                c.ptrs[i] = c.ids.get() + (size_t)i * N;
                bert_tokenize(ctx, texts[first + i], c.ptrs[i], &c.lens[i], N);
            }
        };
        if (nt <= 1) { range(0, n); return; }
        std::vector<std::thread> th;
        const int per = (n + nt - 1) / nt;
        for (int t = 0; t < nt; t++) th.emplace_back(range, std::min(n, t * per), std::min(n, (t + 1) * per));
        for (auto &t : th) t.join();
    };
    tokenize(ch[0], 0, std::min(n_inputs, kChunk));
    for (int first = 0, k = 0; first < n_inputs; first += kChunk, k++) {
        const int n = std::min(kChunk, n_inputs - first);
        Chunk &cur = ch[k & 1];
        std::thread ahead;
        if (first + kChunk < n_inputs)
            ahead = std::thread([&, first, k] { tokenize(ch[(k + 1) & 1], first + kChunk, std::min(kChunk, n_inputs - first - kChunk)); });
        bert_eval_batch(ctx, n_threads, n, cur.ptrs.data(), cur.lens.data(), embeddings + first);
        if (ahead.joinable()) ahead.join();
    }
}

BERT_API void bert_encode(struct bert_ctx *ctx, int32_t n_threads, const char *texts, float *embeddings) {
    bert_encode_batch(ctx, n_threads, 1, 1, &texts, &embeddings);  // bert.cpp:1110-1117
}

}  // extern "C"

struct bk_tokenizer { WordPieceTokenizer tok; };
extern "C" {
BK_API int bk_bert_graph_stats(struct bert_ctx *ctx, int64_t *replayed, int64_t *captured, int64_t *direct) {
    if (!ctx) return -1;
    int64_t r = 0, c = 0, d = 0;
    for (bk_engine *e : ctx->engines) {
        int64_t a = 0, b = 0, x = 0;
        bk_engine_graph_stats(e, &a, &b, &x);
        r += a; c += b; d += x;
    }
    if (replayed) *replayed = r;
    if (captured) *captured = c;
    if (direct) *direct = d;
    return 0;
}
BK_API bk_tokenizer *bk_tokenizer_create(const char *json, int64_t json_len) {
    std::unique_ptr<bk_tokenizer> t(new bk_tokenizer());
    std::string err;
    if (!t->tok.load(std::string(json, (size_t)json_len), err)) { fprintf(stderr, "bk_tokenizer_create: %s\n", err.c_str()); return nullptr; }
    return t.release();
}
BK_API void bk_tokenizer_destroy(bk_tokenizer *t) { delete t; }
BK_API int32_t bk_tokenizer_encode(const bk_tokenizer *t, const char *text, int32_t *ids, int32_t max_ids) {
    if (!t || !text) return -1;
    const std::vector<int32_t> v = t->tok.encode(text);
    for (int32_t i = 0; i < (int32_t)v.size() && i < max_ids; i++) ids[i] = v[i];
    return (int32_t)v.size();
}
}

// bert.cpp:681-733
static void bert_print_usage(char **argv, const bert_params &params) {
    fprintf(stderr, "usage: %s [options]\n\noptions:\n", argv[0]);
    fprintf(stderr, "  -h, --help            show this help message and exit\n");
    fprintf(stderr, "  -t N, --threads N     number of threads to use during computation (default: %d)\n", params.n_threads);
    fprintf(stderr, "  -p PROMPT, --prompt PROMPT\n                        prompt to start generation with (default: random)\n");
    fprintf(stderr, "  --port p     port to bind in server mode (default: %d)\n", params.port);
    fprintf(stderr, "  -m FNAME, --model FNAME\n                        model path (default: %s)\n\n", params.model);
}
extern "C" BERT_API bool bert_params_parse(int argc, char **argv, bert_params &params) {
    for (int i = 1; i < argc; i++) {
        const std::string arg = argv[i];
        const bool has_next = i + 1 < argc;
        if ((arg == "-t" || arg == "--threads") && has_next) params.n_threads = atoi(argv[++i]);
        else if ((arg == "-p" || arg == "--prompt") && has_next) params.prompt = argv[++i];
        else if (arg == "--port" && has_next) params.port = atoi(argv[++i]);
        else if ((arg == "-m" || arg == "--model") && has_next) params.model = argv[++i];
        else if (arg == "-h" || arg == "--help") { bert_print_usage(argv, params); exit(0); }
        else { fprintf(stderr, "error: unknown argument: %s\n", arg.c_str()); bert_print_usage(argv, params); exit(0); }
    }
    return true;
}
