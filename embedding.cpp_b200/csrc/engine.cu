// engine.cu -- one GPU's worth of the encoder: device weights, activation workspaces, the launch
// sequence of one forward pass over a packed, variable-length batch, and the C ABI of include/bk_api.h.
//
// Replaces bert_build + ggml_graph_compute_with_ctx (bert.cpp:845-1012, 1077-1099; ggml.c:15961-16049):
// the reference builds and walks a ~45-node-per-layer graph per sentence on CPU threads; here the plan
// is static -- 2 (on-device batcher) + 1 + 7*n_layer + 1 kernel launches per batch on one CUDA stream.
//
// HBM layout (T = packed tokens of the launch, d = n_embd, f = n_intermediate; all activations f16):
//   x   [T][d]   residual stream / LN output        xq  [T][d]  its Q8-rounded copy (Q4 models only)
//   qkv [T][3d]  Q | K | V                          att [T][d]  attention output (Q8-rounded for Q4)
//   pre [T][d]   GEMM epilogue output = Lin + bias + residual, input of add_ln
//   y   [T][d]   post-attention LN output           yq  [T][d]  (Q4 models only)
//   h   [T][f]   GELU(FF1) (Q8-rounded for Q4)
// Weights: 2-D layer matrices as f16 [N][K] (K contiguous) -- the UMMA B operand; Q/K/V fused to one
// [3d][d] matrix; embedding tables stay in their GGUF storage type and are decoded by the gather.
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include <cuda_profiler_api.h>

#include "kernels.h"
#include "../../include/bk_api.h"

namespace bk {

static thread_local std::string g_err;
static int fail(const char *fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_err = buf;
    return -1;
}
#define CK(call)                                                                                  \
    do {                                                                                          \
        cudaError_t _e = (call);                                                                  \
        if (_e != cudaSuccess) return fail("%s:%d %s: %s", __FILE__, __LINE__, #call, cudaGetErrorString(_e)); \
    } while (0)

struct HostTensor {  // device copy of one GGUF tensor in its storage type
    void *dev = nullptr;
    int type = -1;
    int64_t ne0 = 0, ne1 = 0;
    size_t bytes = 0;
};

struct Layer {
    __half *wqkv = nullptr, *wo = nullptr, *w1 = nullptr, *w2 = nullptr;
    uint8_t *pqkv = nullptr, *po = nullptr, *p1 = nullptr, *p2 = nullptr;  // Q4 models: the same matrices still packed (re-tiled GGUF blocks)
    float *bqkv = nullptr;
    const float *bo = nullptr, *b1 = nullptr, *b2 = nullptr, *ln1g = nullptr, *ln1b = nullptr, *ln2g = nullptr, *ln2b = nullptr;
    WeightMaps tm_wqkv, tm_wo, tm_w1, tm_w2;
};

static size_t tensor_bytes(int type, int64_t ne0, int64_t ne1) {
    switch (type) {
        case BK_F32: return (size_t)ne0 * ne1 * 4;
        case BK_F16: return (size_t)ne0 * ne1 * 2;
        case BK_Q4_0: return (size_t)(ne0 / 32) * 18 * ne1;
        case BK_Q4_1: return (size_t)(ne0 / 32) * 20 * ne1;
    }
    return 0;
}

}  // namespace bk

using namespace bk;

struct bk_engine {
    int device = 0;
    bk_hparams hp{};
    int cap_tokens = 0, cap_seqs = 0;
    int num_sms = 148;
    int qmode = 0;  // 0: F32/F16 weights; 1: Q4_0 (activations Q8_0-rounded); 2: Q4_1 (Q8_1)
    cudaStream_t st = nullptr;
    bool finalized = false;

    std::vector<HostTensor> tens;
    std::vector<Layer> layers;

    // workspaces
    // two input/output slots: batch i+1 is staged, uploaded and enqueued while batch i's results are still being
    // read back and scattered by the host (bk_engine_submit / bk_engine_collect)
    int32_t *d_tokens2[2] = {nullptr, nullptr}, *d_lens2[2] = {nullptr, nullptr};
    float *d_out2[2] = {nullptr, nullptr};
    int32_t *h_tokens2[2] = {nullptr, nullptr}, *h_lens2[2] = {nullptr, nullptr};
    float *h_out2[2] = {nullptr, nullptr};
    cudaEvent_t slot_done[2] = {nullptr, nullptr};
    int slot_nseq[2] = {0, 0};
    int32_t *d_cu = nullptr, *d_pos = nullptr;
    // on-device batcher (rowwise.cu): packed, length-sorted ids; sorted position -> caller index; scratch
    int32_t *d_tok = nullptr, *d_order = nullptr, *d_src_off = nullptr, *d_caller_off = nullptr;
    float *slot_host_out[2] = {nullptr, nullptr};  // where slot's embeddings were copied to (caller's pinned buffer or h_out2)
    __half *x = nullptr, *xq = nullptr, *qkv = nullptr, *att = nullptr, *pre = nullptr, *y = nullptr, *yq = nullptr, *h = nullptr;
    CUtensorMap tm_x, tm_att, tm_y, tm_h;  // A-operand maps (x/y point at the Q8-rounded copy for Q4 models)
    CUtensorMap to_qkv, to_pre, to_h;      // GEMM output maps (TMA stores), box {64, 128}
    CUtensorMap tr_x, tr_y;                // residual maps (unrounded x / y), box {64, 128}
    CUtensorMap to_xq, to_yq;              // Q4 models: the Q8-rounded copies as outputs of the fused GEMM + LayerNorm kernel
    CUtensorMap ta_qkv, ta_qkv32, ta_qkv16;  // attention's views of qkv: boxes {dh, 64 | 32 | 16 rows} (the narrow ones: packed short sentences)
    // pinned staging
    // staged batch
    int st_T = 0, st_nseq = 0, st_maxlen = 0;  // last forward (hidden-state readback)
    // staged set (bench `value` leg): a whole job resident in HBM, cut into micro-batches of <= cap_tokens
    // split: how many sentences of the slice fall into each length class of the attention kernels (the on-device batcher sorts by
    // length, so the classes are consecutive runs of the packed batch) -> up to three attention launches per layer
    struct Slice { int tok_off, seq_off, T, n_seq, max_len; AttSplit split; };
    std::vector<Slice> slices;
    int32_t *sg_tokens = nullptr, *sg_lens = nullptr;
    float *sg_out = nullptr;
    size_t sg_cap_tokens = 0, sg_cap_seqs = 0;
    int sg_nseq = 0;
    // hidden-state capture (tests)
    bool keep_hidden = false;
    __half *hidden = nullptr;  // [n_layer+1][cap_tokens][d]
    // CUDA graphs of the launch plan for small batches (latency path): key = (T, n_seq, max_len)
    struct GraphKey {
        int T, n_seq, max_len, slot;
        AttSplit split;
        bool operator==(const GraphKey &o) const {
            return T == o.T && n_seq == o.n_seq && max_len == o.max_len && slot == o.slot && split.n_long == o.split.n_long &&
                   split.n_mid == o.split.n_mid && split.max_mid == o.split.max_mid && split.n64 == o.split.n64 && split.n32 == o.split.n32 &&
                   split.n16 == o.split.n16;
        }
    };
    struct GraphEntry { GraphKey key; cudaGraphExec_t exec; uint64_t last_use; };
    std::vector<GraphEntry> graphs;
    std::vector<GraphKey> seen_once;  // shapes launched directly once; the second sighting captures (ring, newest last)
    uint64_t graph_clock = 0;
    int graph_max_tokens = 4096;  // BERT_B200_GRAPH_TOKENS; 0 disables
    int graph_cache = 1024;       // BERT_B200_GRAPH_CACHE: captured shapes kept (least recently used is evicted)
    int64_t graph_hits = 0, graph_captures = 0, graph_direct = 0;
    // Q4 models: batches of up to this many tokens (the latency regime) run the GEMMs from the packed weights (decoded in shared memory,
    // only Q4 bytes cross HBM: weight traffic dominates there); larger ones use the f16 operand copies, where
    // weights are < 2% of the bytes moved and the decode would compete with the epilogue for issue slots.
    int q4_smem_tokens = 1024;    // BERT_B200_Q4_SMEM_TOKENS; -1 = always, 0 = never
    // BERT_B200_Q4_WS=1 (Q4 models): QKV and FFN-up (K = n_embd <= 1024) run weight-stationary from the packed blocks at every batch
    // size above the latency regime (gemm_ws.cu: every CTA pair decodes its column tile once) and have NO f16 operand copy --
    // the packed GGUF blocks are the only copy of those weights in HBM.  Costs 10-25 % on those two GEMMs (the resident slice
    // takes the shared memory that hides latency), so the default keeps f16 operand copies for batches above q4_smem_tokens.
    // Attention-output and FFN-down run in the fused projection + LayerNorm kernel from f16 copies either way.
    bool q4_ws = false;
    // batches of more than this many tokens run attention-output and FFN-down as ONE kernel each with the residual add and
    // the LayerNorm in the GEMM epilogue (gemm_ln.cu: a cluster per 256-row block); smaller ones have too few row blocks to
    // fill the clusters and use GEMM(+bias+residual) followed by add_ln.  BERT_B200_FUSED_LN_TOKENS; -1 = never
    int fused_ln_tokens = 4096;
    // BERT_B200_GUARD=1 (tests; compute-sanitizer is closed on the GPU pool): every activation / index buffer is allocated with a
    // 4 KB canary on both sides, bk_engine_check_guards() counts canary bytes a kernel has overwritten
    struct Guarded { uint8_t *base; size_t bytes; };
    std::vector<Guarded> guarded;
    bool guard_on = false;
    // timing
    std::vector<cudaEvent_t> ev;
    std::vector<int> ev_kind;
    bool timing = false;
    float k_ms[BK_K_COUNT] = {0};
    int k_launches[BK_K_COUNT] = {0};
    int64_t launch_count = 0;
};

constexpr size_t kGuardBytes = 4096;
constexpr int kGuardByte = 0xA5;
// activation / index buffers go through here: plain cudaMalloc, or (guard mode) canaries on both sides
template <class T>
static cudaError_t act_alloc(bk_engine *e, T **p, size_t bytes) {
    if (!e->guard_on) return cudaMalloc(p, bytes);
    uint8_t *base = nullptr;
    const size_t padded = (bytes + 255) & ~(size_t)255;
    cudaError_t err = cudaMalloc(&base, padded + 2 * kGuardBytes);
    if (err != cudaSuccess) return err;
    err = cudaMemset(base, kGuardByte, padded + 2 * kGuardBytes);
    if (err == cudaSuccess) err = cudaDeviceSynchronize();  // the engine's stream does not synchronise with the default stream
    if (err != cudaSuccess) return err;
    e->guarded.push_back({base, bytes});
    *p = reinterpret_cast<T *>(base + kGuardBytes);
    return cudaSuccess;
}
static void act_free(bk_engine *e, void *p) {
    if (!p) return;
    if (e->guard_on) {
        for (auto &g : e->guarded)
            if (g.base + kGuardBytes == p) { cudaFree(g.base); g.base = nullptr; return; }
    }
    cudaFree(p);
}

extern "C" {

BK_API const char *bk_last_error(void) { return g_err.c_str(); }

BK_API int bk_device_count(void) {
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) { fail("cudaGetDeviceCount: %s", cudaGetErrorString(e)); return e == cudaErrorNoDevice ? 0 : -1; }
    return n;
}

static int engine_alloc(bk_engine *e) {
    const bk_hparams &hp = e->hp;
    const size_t T = (size_t)e->cap_tokens, d = hp.n_embd, f = hp.n_intermediate;
    for (int sl = 0; sl < 2; sl++) {
        CK(act_alloc(e, &e->d_tokens2[sl], T * 4));
        CK(act_alloc(e, &e->d_lens2[sl], (size_t)e->cap_seqs * 4));
        CK(act_alloc(e, &e->d_out2[sl], (size_t)e->cap_seqs * d * 4));
        CK(cudaMallocHost(&e->h_tokens2[sl], T * 4));
        CK(cudaMallocHost(&e->h_lens2[sl], (size_t)e->cap_seqs * 4));
        CK(cudaMallocHost(&e->h_out2[sl], (size_t)e->cap_seqs * d * 4));
        CK(cudaEventCreateWithFlags(&e->slot_done[sl], cudaEventDisableTiming));
    }
    CK(act_alloc(e, &e->d_pos, T * 4));
    CK(act_alloc(e, &e->d_tok, T * 4));
    CK(act_alloc(e, &e->d_cu, ((size_t)e->cap_seqs + 1) * 4));
    CK(act_alloc(e, &e->d_order, (size_t)e->cap_seqs * 4));
    CK(act_alloc(e, &e->d_src_off, (size_t)e->cap_seqs * 4));
    CK(act_alloc(e, &e->d_caller_off, (size_t)e->cap_seqs * 4));
    CK(act_alloc(e, &e->x, T * d * 2));
    CK(act_alloc(e, &e->qkv, T * 3 * d * 2));
    CK(act_alloc(e, &e->att, T * d * 2));
    CK(act_alloc(e, &e->pre, T * d * 2));
    CK(act_alloc(e, &e->y, T * d * 2));
    CK(act_alloc(e, &e->h, T * f * 2));
    if (e->qmode) {
        CK(act_alloc(e, &e->xq, T * d * 2));
        CK(act_alloc(e, &e->yq, T * d * 2));
    }
    // the GEMM A operand may read (never store) rows past T inside the last 128-row tile: keep them finite
    CK(cudaMemsetAsync(e->x, 0, T * d * 2, e->st));
    CK(cudaMemsetAsync(e->qkv, 0, T * 3 * d * 2, e->st));  // attention's MMAs read up to 15 K/V rows past a sentence: finite
    CK(cudaMemsetAsync(e->att, 0, T * d * 2, e->st));
    CK(cudaMemsetAsync(e->y, 0, T * d * 2, e->st));
    CK(cudaMemsetAsync(e->h, 0, T * f * 2, e->st));
    if (e->qmode) {
        CK(cudaMemsetAsync(e->xq, 0, T * d * 2, e->st));
        CK(cudaMemsetAsync(e->yq, 0, T * d * 2, e->st));
    }
    return 0;
}

BK_API bk_engine *bk_engine_create(int device, const bk_hparams *hp, int32_t cap_tokens, int32_t cap_seqs) {
    int n = bk_device_count();
    if (n <= 0) { fail("no CUDA device available (this engine has no CPU path)"); return nullptr; }
    if (device < 0 || device >= n) { fail("device %d out of range (have %d)", device, n); return nullptr; }
    if (hp->n_embd % 32 || hp->n_intermediate % 32 || hp->n_embd % hp->n_head) { fail("unsupported shape: n_embd/n_intermediate must be multiples of 32"); return nullptr; }
    const int dh = hp->n_embd / hp->n_head;
    if (dh != 32 && dh != 64) { fail("unsupported head size %d (supported: 32, 64)", dh); return nullptr; }
    if (hp->n_embd > 2048) { fail("unsupported n_embd %d (> 2048)", hp->n_embd); return nullptr; }
    if (hp->wtype < BK_F32 || hp->wtype > BK_Q4_1) { fail("unsupported weight type %d", hp->wtype); return nullptr; }
    // the attention kernels hold one sentence's keys on chip: 512 tokens at most (every BERT checkpoint's context length)
    if (hp->n_max_tokens < 1 || hp->n_max_tokens > 512) { fail("unsupported context length %d (supported: 1..512)", hp->n_max_tokens); return nullptr; }
    if (cap_tokens < hp->n_max_tokens) cap_tokens = hp->n_max_tokens;
    cap_tokens = (cap_tokens + 255) & ~255;  // whole CTA-pair tiles (256 rows): stores past T stay inside the buffers
    if (cap_seqs < 1) cap_seqs = 1;
    if (cudaSetDevice(device) != cudaSuccess) { fail("cudaSetDevice(%d) failed", device); return nullptr; }
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) { fail("cudaGetDeviceProperties failed"); return nullptr; }
    if (prop.major != 10) { fail("device %d is sm_%d%d; this engine is built for sm_100a (B200) only", device, prop.major, prop.minor); return nullptr; }
    bk_engine *e = new bk_engine();
    e->device = device;
    e->hp = *hp;
    e->cap_tokens = cap_tokens;
    e->cap_seqs = cap_seqs;
    e->num_sms = prop.multiProcessorCount;
    e->qmode = hp->wtype == BK_Q4_0 ? 1 : hp->wtype == BK_Q4_1 ? 2 : 0;
    if (const char *v = getenv("BERT_B200_GUARD")) e->guard_on = atoi(v) != 0;
    if (const char *v = getenv("BERT_B200_GRAPH_TOKENS")) e->graph_max_tokens = atoi(v);
    if (const char *v = getenv("BERT_B200_GRAPH_CACHE")) e->graph_cache = atoi(v) > 0 ? atoi(v) : 1;
    if (const char *v = getenv("BERT_B200_Q4_SMEM_TOKENS")) e->q4_smem_tokens = atoi(v);
    if (const char *v = getenv("BERT_B200_FUSED_LN_TOKENS")) e->fused_ln_tokens = atoi(v);
    {
        const char *v = getenv("BERT_B200_Q4_WS");
        e->q4_ws = e->qmode != 0 && v && atoi(v) != 0 && gemm_ws_supported(3 * hp->n_embd, hp->n_embd, e->num_sms) &&
                   gemm_ws_supported(hp->n_intermediate, hp->n_embd, e->num_sms);
    }
    e->tens.resize(5 + 16 * (size_t)hp->n_layer);
    e->layers.resize(hp->n_layer);
    if (cudaStreamCreateWithFlags(&e->st, cudaStreamNonBlocking) != cudaSuccess) { fail("cudaStreamCreate failed"); delete e; return nullptr; }
    if (engine_alloc(e) != 0) { bk_engine_destroy(e); return nullptr; }
    return e;
}

BK_API void bk_engine_destroy(bk_engine *e) {
    if (!e) return;
    cudaSetDevice(e->device);
    if (e->st) cudaStreamSynchronize(e->st);
    for (auto &t : e->tens) if (t.dev) cudaFree(t.dev);
    for (auto &l : e->layers) {
        cudaFree(l.wqkv); cudaFree(l.wo); cudaFree(l.w1); cudaFree(l.w2); cudaFree(l.bqkv);
        cudaFree(l.pqkv); cudaFree(l.po); cudaFree(l.p1); cudaFree(l.p2);
    }
    void *bufs[] = {e->d_tokens2[0], e->d_tokens2[1], e->d_lens2[0], e->d_lens2[1], e->d_out2[0], e->d_out2[1], e->d_cu, e->d_pos, e->d_tok, e->d_order,
                    e->d_src_off, e->d_caller_off, e->x, e->xq, e->qkv,
                    e->att, e->pre, e->y, e->yq, e->h, e->hidden, e->sg_tokens, e->sg_lens, e->sg_out};
    void *staged[] = {e->hidden, e->sg_tokens, e->sg_lens, e->sg_out};  // plain allocations
    for (void *b : bufs) {
        bool plain = false;
        for (void *q : staged) plain = plain || q == b;
        if (plain) cudaFree(b); else act_free(e, b);
    }
    for (int sl = 0; sl < 2; sl++) {
        if (e->h_tokens2[sl]) cudaFreeHost(e->h_tokens2[sl]);
        if (e->h_lens2[sl]) cudaFreeHost(e->h_lens2[sl]);
        if (e->h_out2[sl]) cudaFreeHost(e->h_out2[sl]);
        if (e->slot_done[sl]) cudaEventDestroy(e->slot_done[sl]);
    }
    for (auto ev : e->ev) cudaEventDestroy(ev);
    for (auto &g : e->graphs) cudaGraphExecDestroy(g.exec);
    if (e->st) cudaStreamDestroy(e->st);
    delete e;
}

BK_API int bk_engine_set_tensor(bk_engine *e, int32_t slot, int32_t type, const void *data, int64_t ne0, int64_t ne1) {
    if (!e || slot < 0 || slot >= (int)e->tens.size()) return fail("bad tensor slot %d", slot);
    if (type < BK_F32 || type > BK_Q4_1) return fail("tensor slot %d: unsupported ggml type %d", slot, type);
    if ((type == BK_Q4_0 || type == BK_Q4_1) && ne0 % 32) return fail("tensor slot %d: ne0 %lld not a multiple of 32", slot, (long long)ne0);
    CK(cudaSetDevice(e->device));
    HostTensor &t = e->tens[slot];
    if (t.dev) { cudaFree(t.dev); t.dev = nullptr; }
    t.type = type; t.ne0 = ne0; t.ne1 = ne1;
    t.bytes = tensor_bytes(type, ne0, ne1);
    CK(cudaMalloc(&t.dev, t.bytes));
    CK(cudaMemcpy(t.dev, data, t.bytes, cudaMemcpyHostToDevice));
    return 0;
}

// slot helpers (order of bert.cpp:623-652)
static inline int L_SLOT(int layer, int k) { return 5 + 16 * layer + k; }
enum { S_QW = 0, S_QB, S_KW, S_KB, S_VW, S_VB, S_OW, S_OB, S_LN1G, S_LN1B, S_W1, S_B1, S_W2, S_B2, S_LN2G, S_LN2B };

static int expect(const bk_engine *e, int slot, int64_t ne0, int64_t ne1, bool must_f32, const char *what) {
    const HostTensor &t = e->tens[slot];
    if (!t.dev) return fail("tensor '%s' (slot %d) missing", what, slot);
    if (t.ne0 != ne0 || t.ne1 != ne1) return fail("tensor '%s' has wrong shape: got {%lld,%lld}, expected {%lld,%lld}", what, (long long)t.ne0, (long long)t.ne1, (long long)ne0, (long long)ne1);
    if (must_f32 && t.type != BK_F32) return fail("tensor '%s' must be F32", what);
    return 0;
}

BK_API int bk_engine_finalize(bk_engine *e) {
    if (!e) return fail("null engine");
    CK(cudaSetDevice(e->device));
    const bk_hparams &hp = e->hp;
    const int64_t d = hp.n_embd, f = hp.n_intermediate;
    if (expect(e, 0, d, hp.n_vocab, false, "embeddings.word_embeddings.weight")) return -1;
    if (expect(e, 1, d, 2, false, "embeddings.token_type_embeddings.weight")) return -1;
    if (expect(e, 2, d, hp.n_max_tokens, false, "embeddings.position_embeddings.weight")) return -1;
    if (expect(e, 3, d, 1, true, "embeddings.LayerNorm.weight")) return -1;
    if (expect(e, 4, d, 1, true, "embeddings.LayerNorm.bias")) return -1;
    if (e->tens[0].type != e->tens[1].type || e->tens[0].type != e->tens[2].type) return fail("embedding tables have mixed types");
    for (int l = 0; l < hp.n_layer; l++) {
        Layer &L = e->layers[l];
        const int wslots[4] = {S_QW, S_KW, S_VW, S_OW};
        for (int k : wslots) if (expect(e, L_SLOT(l, k), d, d, false, "attention weight")) return -1;
        const int bslots[8] = {S_QB, S_KB, S_VB, S_OB, S_LN1G, S_LN1B, S_LN2G, S_LN2B};
        for (int k : bslots) if (expect(e, L_SLOT(l, k), d, 1, true, "bias / LayerNorm vector")) return -1;
        if (expect(e, L_SLOT(l, S_W1), d, f, false, "intermediate.dense.weight")) return -1;
        if (expect(e, L_SLOT(l, S_B1), f, 1, true, "intermediate.dense.bias")) return -1;
        if (expect(e, L_SLOT(l, S_W2), f, d, false, "output.dense.weight")) return -1;
        if (expect(e, L_SLOT(l, S_B2), d, 1, true, "output.dense.bias")) return -1;

        const bool ws = e->q4_ws;  // QKV and FFN-up have no f16 operand copy: they always run from the packed blocks
        if (!ws) CK(cudaMalloc(&L.wqkv, (size_t)3 * d * d * 2));
        CK(cudaMalloc(&L.wo, (size_t)d * d * 2));
        if (!ws) CK(cudaMalloc(&L.w1, (size_t)f * d * 2));
        CK(cudaMalloc(&L.w2, (size_t)d * f * 2));
        CK(cudaMalloc(&L.bqkv, (size_t)3 * d * 4));
        const int qkvw[3] = {S_QW, S_KW, S_VW}, qkvb[3] = {S_QB, S_KB, S_VB};
        for (int i = 0; i < 3; i++) {
            const HostTensor &w = e->tens[L_SLOT(l, qkvw[i])];
            if (!ws) CK(launch_weight_to_f16(w.dev, w.type, d, d, L.wqkv + (size_t)i * d * d, e->st));
            CK(cudaMemcpyAsync(L.bqkv + (size_t)i * d, e->tens[L_SLOT(l, qkvb[i])].dev, (size_t)d * 4, cudaMemcpyDeviceToDevice, e->st));
        }
        if (e->qmode) {
            const int wt = hp.wtype;
            const size_t bb2 = wt == BK_Q4_0 ? 36 : 40;
            CK(cudaMalloc(&L.pqkv, (size_t)3 * d * (d / 64) * bb2));
            CK(cudaMalloc(&L.po, (size_t)d * (d / 64) * bb2));
            CK(cudaMalloc(&L.p1, (size_t)f * (d / 64) * bb2));
            CK(cudaMalloc(&L.p2, (size_t)d * (f / 64) * bb2));
            for (int i = 0; i < 3; i++) {
                const HostTensor &w = e->tens[L_SLOT(l, qkvw[i])];
                if (w.type != wt) return fail("layer %d: mixed weight types", l);
                CK(launch_q4_retile(w.dev, wt, (int)d, (int)d, L.pqkv, 3 * (int)d, i * (int)d, e->st));
            }
            const int ws[3] = {S_OW, S_W1, S_W2};
            uint8_t *dst[3] = {L.po, L.p1, L.p2};
            const int rows[3] = {(int)d, (int)f, (int)d}, ks[3] = {(int)d, (int)d, (int)f};
            for (int i = 0; i < 3; i++) {
                const HostTensor &w = e->tens[L_SLOT(l, ws[i])];
                if (w.type != wt) return fail("layer %d: mixed weight types", l);
                CK(launch_q4_retile(w.dev, wt, rows[i], ks[i], dst[i], rows[i], 0, e->st));
            }
        }
        { const HostTensor &w = e->tens[L_SLOT(l, S_OW)]; CK(launch_weight_to_f16(w.dev, w.type, d, d, L.wo, e->st)); }
        if (!ws) { const HostTensor &w = e->tens[L_SLOT(l, S_W1)]; CK(launch_weight_to_f16(w.dev, w.type, f, d, L.w1, e->st)); }
        { const HostTensor &w = e->tens[L_SLOT(l, S_W2)]; CK(launch_weight_to_f16(w.dev, w.type, d, f, L.w2, e->st)); }
        L.bo = (const float *)e->tens[L_SLOT(l, S_OB)].dev;
        L.b1 = (const float *)e->tens[L_SLOT(l, S_B1)].dev;
        L.b2 = (const float *)e->tens[L_SLOT(l, S_B2)].dev;
        L.ln1g = (const float *)e->tens[L_SLOT(l, S_LN1G)].dev;
        L.ln1b = (const float *)e->tens[L_SLOT(l, S_LN1B)].dev;
        L.ln2g = (const float *)e->tens[L_SLOT(l, S_LN2G)].dev;
        L.ln2b = (const float *)e->tens[L_SLOT(l, S_LN2B)].dev;
        bool ok = make_weight_maps(&L.tm_wo, L.wo, (int)d, (int)d) && make_weight_maps(&L.tm_w2, L.w2, (int)d, (int)f);
        if (!ws) ok = ok && make_weight_maps(&L.tm_wqkv, L.wqkv, 3 * (int)d, (int)d) && make_weight_maps(&L.tm_w1, L.w1, (int)f, (int)d);
        else {  // shape only: the packed kernels take no weight tensor map (an A-operand map stands in as the unused argument)
            for (int i = 0; i < 6; i++) { L.tm_wqkv.tm[i] = L.tm_wo.tm[i]; L.tm_w1.tm[i] = L.tm_wo.tm[i]; }
            L.tm_wqkv.N = 3 * (int)d; L.tm_wqkv.K = (int)d; L.tm_w1.N = (int)f; L.tm_w1.K = (int)d;
        }
        if (!ok) return fail("cuTensorMapEncodeTiled failed for layer %d weights", l);
    }
    CK(cudaStreamSynchronize(e->st));
    // the packed 2-D layer weights are no longer needed on the device (only their f16 operand copies are)
    for (int l = 0; l < hp.n_layer; l++) {
        const int wslots[6] = {S_QW, S_KW, S_VW, S_OW, S_W1, S_W2};
        for (int k : wslots) { HostTensor &t = e->tens[L_SLOT(l, k)]; cudaFree(t.dev); t.dev = nullptr; }
    }
    const uint64_t T = e->cap_tokens;
    bool ok = make_tmap_f16_2d(&e->tm_x, e->qmode ? e->xq : e->x, d, T, d * 2, GEMM_BLOCK_K, GEMM_BLOCK_M) &&
              make_tmap_f16_2d(&e->tm_att, e->att, d, T, d * 2, GEMM_BLOCK_K, GEMM_BLOCK_M) &&
              make_tmap_f16_2d(&e->tm_y, e->qmode ? e->yq : e->y, d, T, d * 2, GEMM_BLOCK_K, GEMM_BLOCK_M) &&
              make_tmap_f16_2d(&e->tm_h, e->h, f, T, f * 2, GEMM_BLOCK_K, GEMM_BLOCK_M) &&
              make_tmap_f16_2d(&e->to_qkv, e->qkv, 3 * d, T, 3 * d * 2, GEMM_OUT_BOX_N, GEMM_BLOCK_M) &&
              make_tmap_f16_2d(&e->to_pre, e->pre, d, T, d * 2, GEMM_OUT_BOX_N, GEMM_BLOCK_M) &&
              make_tmap_f16_2d(&e->to_h, e->h, f, T, f * 2, GEMM_OUT_BOX_N, GEMM_BLOCK_M) &&
              make_tmap_f16_2d(&e->tr_x, e->x, d, T, d * 2, GEMM_OUT_BOX_N, GEMM_BLOCK_M) &&
              make_tmap_f16_2d(&e->tr_y, e->y, d, T, d * 2, GEMM_OUT_BOX_N, GEMM_BLOCK_M) &&
              (!e->qmode || (make_tmap_f16_2d(&e->to_xq, e->xq, d, T, d * 2, GEMM_OUT_BOX_N, GEMM_BLOCK_M) &&
                             make_tmap_f16_2d(&e->to_yq, e->yq, d, T, d * 2, GEMM_OUT_BOX_N, GEMM_BLOCK_M))) &&
              make_tmap_attention(&e->ta_qkv, e->qkv, (int)d, hp.n_head, T) && make_tmap_attention(&e->ta_qkv32, e->qkv, (int)d, hp.n_head, T, 32) &&
              make_tmap_attention(&e->ta_qkv16, e->qkv, (int)d, hp.n_head, T, 16);
    if (!ok) return fail("cuTensorMapEncodeTiled failed for activation buffers");
    e->finalized = true;
    return 0;
}

}  // extern "C"

// =============================================================================== forward pass
namespace {

struct KTimer {  // brackets one kernel class with two events when timing is on
    bk_engine *e;
    int kind;
    KTimer(bk_engine *e_, int kind_) : e(e_), kind(kind_) {
        if (e->timing) push();
    }
    ~KTimer() {
        if (e->timing) push();
    }
    void push() {
        cudaEvent_t ev;
        cudaEventCreate(&ev);
        cudaEventRecord(ev, e->st);
        e->ev.push_back(ev);
        e->ev_kind.push_back(kind);
    }
};

int run_gemm(bk_engine *e, int kind, const CUtensorMap &tmA, const WeightMaps &W, const uint8_t *packed, int M, int N, int K, const float *bias,
             const CUtensorMap *tmRes, const CUtensorMap &tmOut, int epi) {
    KTimer t(e, kind);
    GemmParams p;
    p.M = M; p.N = N; p.K = K; p.bias = bias; p.epi = epi;
    const bool ws_kind = e->q4_ws && (kind == BK_K_GEMM_QKV || kind == BK_K_GEMM_FF1);  // no f16 copy exists for these
    if (e->qmode && packed && K % 64 == 0 && (e->q4_smem_tokens < 0 || M <= e->q4_smem_tokens || ws_kind)) { p.wq = packed; p.wq_type = e->hp.wtype; }
    if (ws_kind && !(e->q4_smem_tokens < 0 || M <= e->q4_smem_tokens)) {
        CK(launch_gemm_ws(tmA, tmOut, p, e->num_sms, e->st));  // weight-stationary: decode once per CTA pair
        e->launch_count++;
        return 0;
    }
    CK(launch_gemm(tmA, W, tmOut, tmRes, p, e->num_sms, e->st));
    e->launch_count++;
    return 0;
}

// all kernels of one batch already resident on the device (tokens, lens in d_tokens/d_lens)
int forward_device(bk_engine *e, const int32_t *d_tokens, const int32_t *d_lens, int T, int n_seq, int max_len, const AttSplit &sp,
                   float *d_out) {
    const bk_hparams &hp = e->hp;
    const int d = hp.n_embd, f = hp.n_intermediate;
    const int q = e->qmode;
    {   // on-device batcher: sort the sentences by length, pack them, positions; d_tokens / d_lens are in the caller's order
        KTimer t(e, BK_K_PACK);
        BatcherParams b{d_tokens, d_lens, n_seq, T, e->d_tok, e->d_pos, e->d_cu, e->d_order, e->d_src_off, e->d_caller_off};
        CK(launch_batcher(b, e->st));
        e->launch_count += 2;
    }
    {
        KTimer t(e, BK_K_EMBED_LN);
        EmbedParams p;
        p.word = e->tens[0].dev; p.type = e->tens[1].dev; p.posw = e->tens[2].dev; p.wtype = e->tens[0].type;
        p.gamma = (const float *)e->tens[3].dev; p.beta = (const float *)e->tens[4].dev;
        p.tokens = e->d_tok; p.pos = e->d_pos; p.T = T; p.d = d; p.n_vocab = hp.n_vocab; p.eps = hp.eps;
        p.x = e->x; p.xq = e->xq; p.qmode = q;
        CK(launch_embed_ln(p, e->st));
        e->launch_count++;
    }
    if (e->keep_hidden) CK(cudaMemcpyAsync(e->hidden, e->x, (size_t)T * d * 2, cudaMemcpyDeviceToDevice, e->st));
    for (int l = 0; l < hp.n_layer; l++) {
        Layer &L = e->layers[l];
        if (run_gemm(e, BK_K_GEMM_QKV, e->tm_x, L.tm_wqkv, L.pqkv, T, 3 * d, d, L.bqkv, nullptr, e->to_qkv, BK_EPI_BIAS)) return -1;
        {
            KTimer t(e, BK_K_ATTN);
            // sentences > 128 tokens: streaming kernel; 65..128 (and short ones when there are too few to pack): looping CTAs, one
            // (sentence, head) per tile; <= 64 tokens: several sentences per tile (attention_pack.cuh)
            const int n_short = sp.n64 + sp.n32 + sp.n16;
            if (sp.n_long) {
                CK(launch_attention(e->ta_qkv, e->d_cu, sp.n_long, max_len, d, hp.n_head, e->att, q, e->st));
                e->launch_count++;
            }
            if (sp.n_mid) {
                CK(launch_attention(e->ta_qkv, e->d_cu + sp.n_long, sp.n_mid, sp.max_mid, d, hp.n_head, e->att, q, e->st));
                e->launch_count++;
            }
            if (n_short) {
                CK(launch_attention_packed(e->ta_qkv, e->ta_qkv32, e->ta_qkv16, e->d_cu + sp.n_long + sp.n_mid, sp.n64, sp.n32, sp.n16, d, hp.n_head,
                                           e->att, q, e->st));
                e->launch_count++;
            }
        }
        const bool fused_ln = e->fused_ln_tokens >= 0 && T > e->fused_ln_tokens && gemm_ln_supported(d, d) && gemm_ln_supported(d, f);
        if (fused_ln) {  // y = LN(att.Wo^T + bo + x) in one kernel (tr_y doubles as the output map of y: same box geometry)
            KTimer t(e, BK_K_GEMM_O);
            CK(launch_gemm_ln(e->tm_att, L.tm_wo, e->tr_x, e->tr_y, q ? &e->to_yq : nullptr, L.bo, L.ln1g, L.ln1b, hp.eps, q, T, d, d, e->num_sms, e->st));
            e->launch_count++;
        } else {
            if (run_gemm(e, BK_K_GEMM_O, e->tm_att, L.tm_wo, L.po, T, d, d, L.bo, &e->tr_x, e->to_pre, BK_EPI_BIAS_RESID)) return -1;
            KTimer t(e, BK_K_ADD_LN);
            CK(launch_add_ln(e->pre, L.ln1g, L.ln1b, T, d, hp.eps, e->y, e->yq, q, e->st));
            e->launch_count++;
        }
        const int epi1 = q == 0 ? BK_EPI_BIAS_GELU : q == 1 ? BK_EPI_BIAS_GELU_Q8_0 : BK_EPI_BIAS_GELU_Q8_1;
        if (run_gemm(e, BK_K_GEMM_FF1, e->tm_y, L.tm_w1, L.p1, T, f, d, L.b1, nullptr, e->to_h, epi1)) return -1;
        if (fused_ln) {  // x = LN(h.W2^T + b2 + y)
            KTimer t(e, BK_K_GEMM_FF2);
            CK(launch_gemm_ln(e->tm_h, L.tm_w2, e->tr_y, e->tr_x, q ? &e->to_xq : nullptr, L.b2, L.ln2g, L.ln2b, hp.eps, q, T, d, f, e->num_sms, e->st));
            e->launch_count++;
        } else {
            if (run_gemm(e, BK_K_GEMM_FF2, e->tm_h, L.tm_w2, L.p2, T, d, f, L.b2, &e->tr_y, e->to_pre, BK_EPI_BIAS_RESID)) return -1;
            KTimer t(e, BK_K_ADD_LN);
            CK(launch_add_ln(e->pre, L.ln2g, L.ln2b, T, d, hp.eps, e->x, e->xq, q, e->st));
            e->launch_count++;
        }
        if (e->keep_hidden)
            CK(cudaMemcpyAsync(e->hidden + (size_t)(l + 1) * e->cap_tokens * d, e->x, (size_t)T * d * 2, cudaMemcpyDeviceToDevice, e->st));
    }
    {
        KTimer t(e, BK_K_POOL);
        CK(launch_pool_norm(e->x, e->d_cu, e->d_order, n_seq, d, d_out, e->st));  // row order[s]: the caller's order
        e->launch_count++;
    }
    return 0;
}

// Small batches (the latency path): the launch sequence is captured per shape key and replayed as a CUDA graph (kernel arguments --
// sizes, tensor maps, engine-owned device pointers -- are identical for equal keys).  A shape is launched DIRECTLY the first time it
// is seen and captured only when it comes back (capture + instantiation of 4 + 7L nodes costs several direct passes; traffic with
// ever-new shapes must not pay it per call).  Direct small-batch launches use programmatic dependent launch like the captured plan.
// Returns 1 = ran as a graph, 0 = caller must launch directly, < 0 error.
int att_launches(const AttSplit &sp);
int forward_graph(bk_engine *e, int slot, int T, int n_seq, int max_len, const AttSplit &sp) {
    e->graph_clock++;
    const bk_engine::GraphKey key{T, n_seq, max_len, slot, sp};
    for (auto &g : e->graphs)
        if (g.key == key) {
            g.last_use = e->graph_clock;
            CK(cudaGraphLaunch(g.exec, e->st));
            e->launch_count += 4 + (6 + att_launches(sp)) * e->hp.n_layer;
            e->graph_hits++;
            return 1;
        }
    bool second = false;
    for (size_t i = 0; i < e->seen_once.size(); i++)
        if (e->seen_once[i] == key) { e->seen_once.erase(e->seen_once.begin() + (long)i); second = true; break; }
    if (!second) {
        if (e->seen_once.size() >= 4096) e->seen_once.erase(e->seen_once.begin());
        e->seen_once.push_back(key);
        e->graph_direct++;
        return 0;
    }
    cudaGraph_t graph = nullptr;
    CK(cudaStreamBeginCapture(e->st, cudaStreamCaptureModeThreadLocal));
    const int rc = forward_device(e, e->d_tokens2[slot], e->d_lens2[slot], T, n_seq, max_len, sp, e->d_out2[slot]);
    const cudaError_t ce = cudaStreamEndCapture(e->st, &graph);
    if (rc) { if (graph) cudaGraphDestroy(graph); return rc; }
    if (ce != cudaSuccess) return fail("cudaStreamEndCapture: %s", cudaGetErrorString(ce));
    cudaGraphExec_t exec = nullptr;
    const cudaError_t ie = cudaGraphInstantiate(&exec, graph, 0);
    cudaGraphDestroy(graph);
    if (ie != cudaSuccess) return fail("cudaGraphInstantiate: %s", cudaGetErrorString(ie));
    if ((int)e->graphs.size() >= e->graph_cache) {
        size_t lru = 0;
        for (size_t i = 1; i < e->graphs.size(); i++) if (e->graphs[i].last_use < e->graphs[lru].last_use) lru = i;
        cudaGraphExecDestroy(e->graphs[lru].exec);
        e->graphs.erase(e->graphs.begin() + (long)lru);
    }
    e->graphs.push_back({key, exec, e->graph_clock});
    e->graph_captures++;
    CK(cudaGraphLaunch(exec, e->st));
    return 1;
}

// The batcher sorts by length on the device, so the length classes of the attention kernels are consecutive runs of the packed
// batch whatever the caller's order; the host only counts them (no sort): > 128 tokens | 65..128 | 33..64 | 17..32 | 1..16.
// Fewer than kPackMin short sentences are not worth a launch of their own: they join the middle class (one per tile).
// BERT_B200_ATT_PACK=0: never pack.
constexpr int kPackMin = 8;
void split_by_length(const int32_t *lens, int n_seq, AttSplit *sp) {
    static const bool pack = [] { const char *v = getenv("BERT_B200_ATT_PACK"); return !v || atoi(v) != 0; }();
    AttSplit a{0, 0, 0, 0, 0, 0};
    int max_short = 0;
    for (int i = 0; i < n_seq; i++) {
        const int L = lens[i];
        if (L > 128) a.n_long++;
        else if (L > 64) { a.n_mid++; if (L > a.max_mid) a.max_mid = L; }
        else {
            if (L > 32) a.n64++; else if (L > 16) a.n32++; else a.n16++;
            if (L > max_short) max_short = L;
        }
    }
    const int n_short = a.n64 + a.n32 + a.n16;
    if (n_short && (!pack || n_short < kPackMin)) {
        a.n_mid += n_short;
        if (max_short > a.max_mid) a.max_mid = max_short;
        a.n64 = a.n32 = a.n16 = 0;
    }
    *sp = a;
}
int att_launches(const AttSplit &sp) { return (sp.n_long ? 1 : 0) + (sp.n_mid ? 1 : 0) + ((sp.n64 + sp.n32 + sp.n16) ? 1 : 0); }

int check_batch(bk_engine *e, const int32_t *lens, int n_seq, int *T_out, int *maxlen_out) {
    if (!e || !e->finalized) return fail("engine not finalized");
    if (n_seq < 1 || n_seq > e->cap_seqs) return fail("n_seq %d outside 1..%d", n_seq, e->cap_seqs);
    long long T = 0;
    int mx = 0;
    for (int i = 0; i < n_seq; i++) {
        if (lens[i] < 1 || lens[i] > e->hp.n_max_tokens) return fail("sentence %d has %d tokens (allowed 1..%d)", i, lens[i], e->hp.n_max_tokens);
        T += lens[i];
        if (lens[i] > mx) mx = lens[i];
    }
    if (T > e->cap_tokens) return fail("batch has %lld tokens, capacity %d", T, e->cap_tokens);
    *T_out = (int)T;
    *maxlen_out = mx;
    return 0;
}

}  // namespace

extern "C" {

BK_API int bk_engine_stage(bk_engine *e, const int32_t *tokens, const int32_t *lens, int32_t n_seq) {
    if (!e || !e->finalized) return fail("engine not finalized");
    if (n_seq < 1) return fail("n_seq %d < 1", n_seq);
    CK(cudaSetDevice(e->device));
    // cut the job into micro-batches in the caller's order; each one is sorted by length and packed on the device (rowwise.cu)
    e->slices.clear();
    long long total = 0;
    for (int i = 0; i < n_seq;) {
        bk_engine::Slice sl{(int)total, i, 0, 0, 0, AttSplit{0, 0, 0, 0, 0, 0}};
        while (i < n_seq && sl.n_seq < e->cap_seqs) {
            if (lens[i] < 1 || lens[i] > e->hp.n_max_tokens) return fail("sentence %d has %d tokens (allowed 1..%d)", i, lens[i], e->hp.n_max_tokens);
            if (sl.T + lens[i] > e->cap_tokens) break;
            sl.T += lens[i]; sl.n_seq++;
            if (lens[i] > sl.max_len) sl.max_len = lens[i];
            i++;
        }
        total += sl.T;
        split_by_length(lens + sl.seq_off, sl.n_seq, &sl.split);
        e->slices.push_back(sl);
    }
    if ((size_t)total > e->sg_cap_tokens || (size_t)n_seq > e->sg_cap_seqs) {
        cudaFree(e->sg_tokens); cudaFree(e->sg_lens); cudaFree(e->sg_out);
        e->sg_tokens = e->sg_lens = nullptr; e->sg_out = nullptr;
        CK(cudaMalloc(&e->sg_tokens, (size_t)total * 4));
        CK(cudaMalloc(&e->sg_lens, (size_t)n_seq * 4));
        CK(cudaMalloc(&e->sg_out, (size_t)n_seq * e->hp.n_embd * 4));
        e->sg_cap_tokens = (size_t)total; e->sg_cap_seqs = (size_t)n_seq;
    }
    CK(cudaMemcpyAsync(e->sg_tokens, tokens, (size_t)total * 4, cudaMemcpyHostToDevice, e->st));
    CK(cudaMemcpyAsync(e->sg_lens, lens, (size_t)n_seq * 4, cudaMemcpyHostToDevice, e->st));
    CK(cudaStreamSynchronize(e->st));
    e->sg_nseq = n_seq;
    return 0;
}

// BERT_B200_PDL_ALL=1: programmatic dependent launch between the kernels of large batches too (default: the latency path only)
static bool pdl_all() {
    static const bool on = [] { const char *v = getenv("BERT_B200_PDL_ALL"); return v && atoi(v) != 0; }();
    return on;
}

// ids and lengths of the batch are in the slot's pinned staging buffers: upload, all kernels, download into host_out (pinned, or
// the slot's own buffer when null)
static int enqueue_slot(bk_engine *e, int slot, int T, int n_seq, int mx, float *host_out) {
    const int32_t *lens = e->h_lens2[slot];
    CK(cudaMemcpyAsync(e->d_tokens2[slot], e->h_tokens2[slot], (size_t)T * 4, cudaMemcpyHostToDevice, e->st));
    CK(cudaMemcpyAsync(e->d_lens2[slot], e->h_lens2[slot], (size_t)n_seq * 4, cudaMemcpyHostToDevice, e->st));
    e->st_T = T; e->st_nseq = n_seq; e->st_maxlen = mx;
    AttSplit sp;
    split_by_length(lens, n_seq, &sp);
    const bool small = T <= e->graph_max_tokens && !e->keep_hidden && !e->timing;
    // the latency path: kernel boundaries are programmatic dependencies, so every kernel's set-up (and its launch) runs under the
    // tail of the one before it (common.cuh: pdl_wait) -- in the captured plan and in direct launches.  BERT_B200_PDL=0: plain edges.
    static const bool use_pdl = [] { const char *v = getenv("BERT_B200_PDL"); return !v || atoi(v) != 0; }();
    int ran = 0;
    if (small) {
        pdl_flag() = use_pdl;
        ran = forward_graph(e, slot, T, n_seq, mx, sp);
        if (ran == 0) ran = forward_device(e, e->d_tokens2[slot], e->d_lens2[slot], T, n_seq, mx, sp, e->d_out2[slot]) ? -1 : 1;
        pdl_flag() = false;
        if (ran < 0) return -1;
    } else {
        pdl_flag() = use_pdl && pdl_all();
        const int rc = forward_device(e, e->d_tokens2[slot], e->d_lens2[slot], T, n_seq, mx, sp, e->d_out2[slot]);
        pdl_flag() = false;
        if (rc) return -1;
    }
    float *dst = host_out ? host_out : e->h_out2[slot];
    CK(cudaMemcpyAsync(dst, e->d_out2[slot], (size_t)n_seq * e->hp.n_embd * 4, cudaMemcpyDeviceToHost, e->st));
    CK(cudaEventRecord(e->slot_done[slot], e->st));
    e->slot_nseq[slot] = n_seq;
    e->slot_host_out[slot] = dst;
    return 0;
}

BK_API int bk_engine_submit(bk_engine *e, int32_t slot, const int32_t *tokens, const int32_t *lens, int32_t n_seq) {
    int T, mx;
    if (slot < 0 || slot > 1) return fail("slot %d outside 0..1", slot);
    if (check_batch(e, lens, n_seq, &T, &mx)) return -1;
    CK(cudaSetDevice(e->device));
    memcpy(e->h_tokens2[slot], tokens, (size_t)T * 4);
    memcpy(e->h_lens2[slot], lens, (size_t)n_seq * 4);
    return enqueue_slot(e, slot, T, n_seq, mx, nullptr);
}

BK_API int bk_engine_submit_ragged(bk_engine *e, int32_t slot, const int32_t *const *sent_tokens, const int32_t *sent_lens, int32_t n_seq,
                                   float *host_out) {
    int T, mx;
    if (slot < 0 || slot > 1) return fail("slot %d outside 0..1", slot);
    if (check_batch(e, sent_lens, n_seq, &T, &mx)) return -1;
    CK(cudaSetDevice(e->device));
    int32_t *dst = e->h_tokens2[slot];
    for (int i = 0; i < n_seq; i++) {  // the only host pass over the ids: caller's rows -> pinned staging, in the caller's order
        memcpy(dst, sent_tokens[i], (size_t)sent_lens[i] * 4);
        dst += sent_lens[i];
    }
    memcpy(e->h_lens2[slot], sent_lens, (size_t)n_seq * 4);
    return enqueue_slot(e, slot, T, n_seq, mx, host_out);
}

BK_API int bk_engine_collect(bk_engine *e, int32_t slot, float *out) {
    if (!e || slot < 0 || slot > 1 || e->slot_nseq[slot] <= 0) return fail("nothing submitted in slot %d", slot);
    CK(cudaSetDevice(e->device));
    CK(cudaEventSynchronize(e->slot_done[slot]));
    if (out && out != e->slot_host_out[slot]) memcpy(out, e->slot_host_out[slot], (size_t)e->slot_nseq[slot] * e->hp.n_embd * 4);
    e->slot_nseq[slot] = 0;
    return 0;
}

BK_API void *bk_host_alloc(size_t bytes) {
    void *p = nullptr;
    if (cudaHostAlloc(&p, bytes ? bytes : 16, cudaHostAllocPortable) != cudaSuccess) { fail("cudaHostAlloc(%zu) failed", bytes); cudaGetLastError(); return nullptr; }
    return p;
}
BK_API void bk_host_free(void *p) { if (p) cudaFreeHost(p); }

BK_API int bk_engine_graph_stats(bk_engine *e, int64_t *replayed, int64_t *captured, int64_t *direct) {
    if (!e) return fail("null engine");
    if (replayed) *replayed = e->graph_hits;
    if (captured) *captured = e->graph_captures;
    if (direct) *direct = e->graph_direct;
    return 0;
}

BK_API int bk_engine_forward(bk_engine *e, const int32_t *tokens, const int32_t *lens, int32_t n_seq, float *out) {
    if (bk_engine_submit(e, 0, tokens, lens, n_seq)) return -1;
    return bk_engine_collect(e, 0, out);
}

BK_API int bk_engine_run_staged(bk_engine *e, int32_t iters, int32_t time_kernels, float *ms_total) {
    if (!e || !e->finalized || e->slices.empty()) return fail("nothing staged");
    CK(cudaSetDevice(e->device));
    for (auto ev : e->ev) cudaEventDestroy(ev);
    e->ev.clear(); e->ev_kind.clear();
    for (int k = 0; k < BK_K_COUNT; k++) { e->k_ms[k] = 0; e->k_launches[k] = 0; }
    cudaEvent_t a, b;
    CK(cudaEventCreate(&a));
    CK(cudaEventCreate(&b));
    e->timing = time_kernels != 0;
    CK(cudaEventRecord(a, e->st));
    int rc = 0;
    pdl_flag() = pdl_all() && !e->timing;
    for (int i = 0; i < iters && rc == 0; i++)
        for (const auto &sl : e->slices) {
            e->st_T = sl.T; e->st_nseq = sl.n_seq; e->st_maxlen = sl.max_len;
            rc = forward_device(e, e->sg_tokens + sl.tok_off, e->sg_lens + sl.seq_off, sl.T, sl.n_seq, sl.max_len, sl.split,
                                e->sg_out + (size_t)sl.seq_off * e->hp.n_embd);
            if (rc) break;
        }
    e->timing = false;
    pdl_flag() = false;
    if (rc) { cudaEventDestroy(a); cudaEventDestroy(b); return rc; }
    CK(cudaEventRecord(b, e->st));
    CK(cudaStreamSynchronize(e->st));
    float ms = 0;
    CK(cudaEventElapsedTime(&ms, a, b));
    if (ms_total) *ms_total = ms;
    cudaEventDestroy(a); cudaEventDestroy(b);
    for (size_t i = 0; i + 1 < e->ev.size(); i += 2) {
        float t = 0;
        cudaEventElapsedTime(&t, e->ev[i], e->ev[i + 1]);
        e->k_ms[e->ev_kind[i]] += t;
        e->k_launches[e->ev_kind[i]] += (e->ev_kind[i] == BK_K_PACK) ? 2 : 1;
    }
    for (auto ev : e->ev) cudaEventDestroy(ev);
    e->ev.clear(); e->ev_kind.clear();
    return 0;
}

BK_API int bk_engine_read_output(bk_engine *e, float *out, int32_t n_seq) {
    if (!e || n_seq < 1 || n_seq > e->sg_nseq) return fail("bad n_seq");
    CK(cudaSetDevice(e->device));
    CK(cudaMemcpyAsync(out, e->sg_out, (size_t)n_seq * e->hp.n_embd * 4, cudaMemcpyDeviceToHost, e->st));
    CK(cudaStreamSynchronize(e->st));
    return 0;
}

BK_API int bk_engine_kernel_ms(bk_engine *e, float *ms, int32_t *launches) {
    if (!e) return fail("null engine");
    for (int k = 0; k < BK_K_COUNT; k++) { if (ms) ms[k] = e->k_ms[k]; if (launches) launches[k] = e->k_launches[k]; }
    return 0;
}

BK_API int bk_engine_set_keep_hidden(bk_engine *e, int32_t on) {
    if (!e) return fail("null engine");
    CK(cudaSetDevice(e->device));
    if (on && !e->hidden) CK(cudaMalloc(&e->hidden, (size_t)(e->hp.n_layer + 1) * e->cap_tokens * e->hp.n_embd * 2));
    e->keep_hidden = on != 0;
    return 0;
}

BK_API int bk_engine_read_hidden(bk_engine *e, int32_t which, float *out) {
    if (!e || !e->hidden || which < 0 || which > e->hp.n_layer) return fail("hidden state %d not captured", which);
    CK(cudaSetDevice(e->device));
    const size_t n = (size_t)e->st_T * e->hp.n_embd;
    std::vector<__half> tmp(n);
    CK(cudaMemcpy(tmp.data(), e->hidden + (size_t)which * e->cap_tokens * e->hp.n_embd, n * 2, cudaMemcpyDeviceToHost));
    for (size_t i = 0; i < n; i++) out[i] = __half2float(tmp[i]);
    return 0;
}

BK_API int64_t bk_engine_launch_count(bk_engine *e) { return e ? e->launch_count : 0; }

BK_API int64_t bk_engine_check_guards(bk_engine *e) {
    if (!e) return -1;
    if (!e->guard_on) { fail("engine was not created with BERT_B200_GUARD=1"); return -1; }
    if (cudaSetDevice(e->device) != cudaSuccess || cudaStreamSynchronize(e->st) != cudaSuccess) { fail("guard check: stream error"); return -1; }
    int64_t bad = 0;
    std::vector<uint8_t> h(kGuardBytes);
    for (const auto &g : e->guarded) {
        if (!g.base) continue;
        const size_t padded = (g.bytes + 255) & ~(size_t)255;
        const uint8_t *parts[2] = {g.base, g.base + kGuardBytes + padded};
        for (const uint8_t *part : parts) {
            if (cudaMemcpy(h.data(), part, kGuardBytes, cudaMemcpyDeviceToHost) != cudaSuccess) { fail("guard check: copy failed"); return -1; }
            for (uint8_t b : h) bad += b != kGuardByte;
        }
        // the slack between the requested size and its 256-byte rounding belongs to the canary too
        if (padded > g.bytes) {
            std::vector<uint8_t> t(padded - g.bytes);
            if (cudaMemcpy(t.data(), g.base + kGuardBytes + g.bytes, t.size(), cudaMemcpyDeviceToHost) != cudaSuccess) { fail("guard check: copy failed"); return -1; }
            for (uint8_t b : t) bad += b != kGuardByte;
        }
    }
    return bad;
}

BK_API int bk_profiler_range(int32_t on) {
    CK(on ? cudaProfilerStart() : cudaProfilerStop());
    return 0;
}

}  // extern "C"
