// test_entry.cu -- single-kernel C-ABI entry points (include/bk_api.h, bk_test_*): host buffers in,
// one kernel on the device, host buffers out.  Used by tests/ (parity per kernel against oracle/) and
// for isolated ncu captures; the product path (engine.cu) launches the same kernels directly.
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

#include "kernels.h"
#include "../../include/bk_api.h"

using namespace bk;

namespace {
thread_local std::string t_err;
struct DevBuf {
    void *p = nullptr;
    ~DevBuf() { if (p) cudaFree(p); }
    cudaError_t alloc(size_t n) { return cudaMalloc(&p, n ? n : 16); }
    template <class T> T *as() { return reinterpret_cast<T *>(p); }
};
#define TCK(call)                                                                                       \
    do {                                                                                                \
        cudaError_t _e = (call);                                                                        \
        if (_e != cudaSuccess) { fprintf(stderr, "bk_test: %s:%d %s: %s\n", __FILE__, __LINE__, #call, cudaGetErrorString(_e)); return -1; } \
    } while (0)

int upload_lens(const int32_t *lens, int n_seq, DevBuf &d_lens, DevBuf &d_cu, DevBuf &d_pos, int *T_out, int *max_out, cudaStream_t st) {
    int T = 0, mx = 0;
    for (int i = 0; i < n_seq; i++) { T += lens[i]; if (lens[i] > mx) mx = lens[i]; }
    TCK(d_lens.alloc((size_t)n_seq * 4));
    TCK(d_cu.alloc(((size_t)n_seq + 1) * 4));
    TCK(d_pos.alloc((size_t)T * 4));
    TCK(cudaMemcpy(d_lens.p, lens, (size_t)n_seq * 4, cudaMemcpyHostToDevice));
    std::vector<int32_t> cu(n_seq + 1, 0);  // the kernels under test take the sentences in the given order
    for (int i = 0; i < n_seq; i++) cu[i + 1] = cu[i] + lens[i];
    TCK(cudaMemcpy(d_cu.p, cu.data(), ((size_t)n_seq + 1) * 4, cudaMemcpyHostToDevice));
    (void)st;
    *T_out = T; *max_out = mx;
    return 0;
}
}  // namespace

extern "C" {

BK_API int bk_test_gemm(int device, int32_t M, int32_t N, int32_t K, const uint16_t *A, const void *W, int32_t wtype,
                        const float *bias, const uint16_t *resid, int32_t epi, uint16_t *out, float *ms) {
    TCK(cudaSetDevice(device));
    cudaDeviceProp prop;
    TCK(cudaGetDeviceProperties(&prop, device));
    size_t wbytes = wtype == BK_F32 ? (size_t)N * K * 4 : wtype == BK_F16 ? (size_t)N * K * 2 : (size_t)N * (K / 32) * (wtype == BK_Q4_0 ? 18 : 20);
    DevBuf dA, dWraw, dW, dB, dR, dO;
    TCK(dA.alloc((size_t)M * K * 2)); TCK(dWraw.alloc(wbytes)); TCK(dW.alloc((size_t)N * K * 2));
    TCK(dB.alloc((size_t)N * 4)); TCK(dR.alloc((size_t)M * N * 2)); TCK(dO.alloc((size_t)M * N * 2));
    TCK(cudaMemcpy(dA.p, A, (size_t)M * K * 2, cudaMemcpyHostToDevice));
    TCK(cudaMemcpy(dWraw.p, W, wbytes, cudaMemcpyHostToDevice));
    TCK(cudaMemcpy(dB.p, bias, (size_t)N * 4, cudaMemcpyHostToDevice));
    if (resid) TCK(cudaMemcpy(dR.p, resid, (size_t)M * N * 2, cudaMemcpyHostToDevice));
    TCK(cudaMemset(dO.p, 0xFF, (size_t)M * N * 2));
    TCK(launch_weight_to_f16(dWraw.p, wtype, N, K, dW.as<__half>(), 0));
    CUtensorMap tmA, tmO, tmR;
    WeightMaps wm;
    if (!make_tmap_f16_2d(&tmA, dA.p, K, M, (uint64_t)K * 2, GEMM_BLOCK_K, GEMM_BLOCK_M) || !make_weight_maps(&wm, dW.p, N, K) ||
        !make_tmap_f16_2d(&tmO, dO.p, N, M, (uint64_t)N * 2, GEMM_OUT_BOX_N, GEMM_BLOCK_M) ||
        !make_tmap_f16_2d(&tmR, dR.p, N, M, (uint64_t)N * 2, GEMM_OUT_BOX_N, GEMM_BLOCK_M)) {
        fprintf(stderr, "bk_test_gemm: cuTensorMapEncodeTiled failed\n");
        return -1;
    }
    GemmParams p;
    p.M = M; p.N = N; p.K = K; p.bias = dB.as<float>(); p.epi = epi;
    const int bn = (epi >> 8) & 0xFFF, cg = (epi >> 20) & 0xF;  // tests may force the tile: epi | (BN << 8) | (CG << 20)
    const bool ws = (epi >> 29) & 1;                              // ... the weight-stationary kernel (gemm_ws.cu; Q4 weights, K <= 1024)
    const bool in_smem = ((epi >> 28) & 1) || ws;                 // ... and, for Q4 weights, the in-smem decode path
    p.epi = epi & 0xFF;
    DevBuf dWt;
    if (in_smem) {
        if (wtype != BK_Q4_0 && wtype != BK_Q4_1) { fprintf(stderr, "bk_test_gemm: in-smem decode needs Q4 weights\n"); return -1; }
        TCK(dWt.alloc(wbytes));
        TCK(launch_q4_retile(dWraw.p, wtype, N, K, dWt.p, N, 0, 0));
        p.wq = dWt.as<uint8_t>(); p.wq_type = wtype;
    }
    cudaEvent_t a, b;
    TCK(cudaEventCreate(&a)); TCK(cudaEventCreate(&b));
    auto run = [&]() -> cudaError_t {
        if (ws) return launch_gemm_ws(tmA, tmO, p, prop.multiProcessorCount, 0, bn);
        return launch_gemm(tmA, wm, tmO, resid ? &tmR : nullptr, p, prop.multiProcessorCount, 0, bn, cg);
    };
    TCK(run());  // warm-up (also first-use setup)
    TCK(cudaDeviceSynchronize());
    const int reps = ms ? 5 : 1;
    TCK(cudaEventRecord(a, 0));
    for (int i = 0; i < reps; i++) TCK(run());
    TCK(cudaEventRecord(b, 0));
    TCK(cudaDeviceSynchronize());
    float t = 0;
    TCK(cudaEventElapsedTime(&t, a, b));
    if (ms) *ms = t / reps;
    cudaEventDestroy(a); cudaEventDestroy(b);
    TCK(cudaMemcpy(out, dO.p, (size_t)M * N * 2, cudaMemcpyDeviceToHost));
    return 0;
}

BK_API int bk_test_gemm_ln(int device, int32_t M, int32_t N, int32_t K, const uint16_t *A, const uint16_t *W, const float *bias,
                           const uint16_t *resid, const float *gamma, const float *beta, float eps, int32_t qmode,
                           uint16_t *y, uint16_t *yq, float *ms) {
    TCK(cudaSetDevice(device));
    cudaDeviceProp prop;
    TCK(cudaGetDeviceProperties(&prop, device));
    if (!gemm_ln_supported(N, K)) { fprintf(stderr, "bk_test_gemm_ln: unsupported N=%d K=%d\n", N, K); return -1; }
    const size_t Mp = ((size_t)M + 255) / 256 * 256;  // whole cluster tiles: rows past M are written (finite garbage)
    DevBuf dA, dW, dB, dG, dE, dR, dY, dQ;
    TCK(dA.alloc(Mp * K * 2)); TCK(dW.alloc((size_t)N * K * 2)); TCK(dB.alloc((size_t)N * 4)); TCK(dG.alloc((size_t)N * 4));
    TCK(dE.alloc((size_t)N * 4)); TCK(dR.alloc(Mp * N * 2)); TCK(dY.alloc(Mp * N * 2)); TCK(dQ.alloc(Mp * N * 2));
    TCK(cudaMemset(dA.p, 0, Mp * K * 2)); TCK(cudaMemset(dR.p, 0, Mp * N * 2));
    TCK(cudaMemcpy(dA.p, A, (size_t)M * K * 2, cudaMemcpyHostToDevice));
    TCK(cudaMemcpy(dW.p, W, (size_t)N * K * 2, cudaMemcpyHostToDevice));
    TCK(cudaMemcpy(dB.p, bias, (size_t)N * 4, cudaMemcpyHostToDevice));
    TCK(cudaMemcpy(dG.p, gamma, (size_t)N * 4, cudaMemcpyHostToDevice));
    TCK(cudaMemcpy(dE.p, beta, (size_t)N * 4, cudaMemcpyHostToDevice));
    TCK(cudaMemcpy(dR.p, resid, (size_t)M * N * 2, cudaMemcpyHostToDevice));
    TCK(cudaMemset(dY.p, 0xFF, Mp * N * 2)); TCK(cudaMemset(dQ.p, 0xFF, Mp * N * 2));
    CUtensorMap tmA, tmR, tmY, tmQ;
    WeightMaps wm;
    if (!make_tmap_f16_2d(&tmA, dA.p, K, Mp, (uint64_t)K * 2, GEMM_BLOCK_K, GEMM_BLOCK_M) || !make_weight_maps(&wm, dW.p, N, K) ||
        !make_tmap_f16_2d(&tmR, dR.p, N, Mp, (uint64_t)N * 2, GEMM_OUT_BOX_N, GEMM_BLOCK_M) ||
        !make_tmap_f16_2d(&tmY, dY.p, N, Mp, (uint64_t)N * 2, GEMM_OUT_BOX_N, GEMM_BLOCK_M) ||
        !make_tmap_f16_2d(&tmQ, dQ.p, N, Mp, (uint64_t)N * 2, GEMM_OUT_BOX_N, GEMM_BLOCK_M)) {
        fprintf(stderr, "bk_test_gemm_ln: cuTensorMapEncodeTiled failed\n");
        return -1;
    }
    cudaEvent_t a, b;
    TCK(cudaEventCreate(&a)); TCK(cudaEventCreate(&b));
    const int reps = ms ? 5 : 1;
    for (int w = 0; w < 2; w++) {  // w = 0: warm-up (first-use setup)
        TCK(cudaEventRecord(a, 0));
        for (int i = 0; i < (w ? reps : 1); i++)
            TCK(launch_gemm_ln(tmA, wm, tmR, tmY, &tmQ, dB.as<float>(), dG.as<float>(), dE.as<float>(), eps, qmode, M, N, K,
                               prop.multiProcessorCount, 0));
        TCK(cudaEventRecord(b, 0));
        TCK(cudaDeviceSynchronize());
    }
    float t = 0;
    TCK(cudaEventElapsedTime(&t, a, b));
    if (ms) *ms = t / reps;
    cudaEventDestroy(a); cudaEventDestroy(b);
    if (getenv("BERT_B200_LN_TRACE")) {  // phase durations (cycles) of the first tiles of block 0
        DevBuf dT;
        TCK(dT.alloc(16 * 8 * 8));
        TCK(cudaMemset(dT.p, 0, 16 * 8 * 8));
        gemm_ln_set_trace(dT.as<long long>());
        TCK(launch_gemm_ln(tmA, wm, tmR, tmY, &tmQ, dB.as<float>(), dG.as<float>(), dE.as<float>(), eps, qmode, M, N, K, prop.multiProcessorCount, 0));
        TCK(cudaDeviceSynchronize());
        gemm_ln_set_trace(nullptr);
        long long h[16 * 8];
        TCK(cudaMemcpy(h, dT.p, sizeof h, cudaMemcpyDeviceToHost));
        for (int i = 0; i < 16 && h[i * 8]; i++)
            fprintf(stderr, "tile %2d: wait_acc %6lld pass1 %6lld stats %6lld pass2 %6lld flush %6lld | tile %6lld\n", i, h[i * 8 + 1] - h[i * 8],
                    h[i * 8 + 2] - h[i * 8 + 1], h[i * 8 + 3] - h[i * 8 + 2], h[i * 8 + 4] - h[i * 8 + 3], h[i * 8 + 5] - h[i * 8 + 4],
                    h[i * 8 + 5] - h[i * 8]);
    }
    TCK(cudaMemcpy(y, dY.p, (size_t)M * N * 2, cudaMemcpyDeviceToHost));
    if (qmode && yq) TCK(cudaMemcpy(yq, dQ.p, (size_t)M * N * 2, cudaMemcpyDeviceToHost));
    return 0;
}

BK_API int bk_test_gelu(int device, int32_t n, const float *x, float *y) {
    TCK(cudaSetDevice(device));
    DevBuf dX, dY;
    TCK(dX.alloc((size_t)n * 4)); TCK(dY.alloc((size_t)n * 4));
    TCK(cudaMemcpy(dX.p, x, (size_t)n * 4, cudaMemcpyHostToDevice));
    TCK(launch_gelu_test(dX.as<float>(), dY.as<float>(), n, 0));
    TCK(cudaDeviceSynchronize());
    TCK(cudaMemcpy(y, dY.p, (size_t)n * 4, cudaMemcpyDeviceToHost));
    return 0;
}

BK_API int bk_test_add_ln(int device, int32_t T, int32_t d, float eps, const uint16_t *pre, const float *gamma,
                          const float *beta, int32_t qmode, uint16_t *y, uint16_t *yq) {
    TCK(cudaSetDevice(device));
    DevBuf dP, dG, dB, dY, dQ;
    TCK(dP.alloc((size_t)T * d * 2)); TCK(dG.alloc((size_t)d * 4)); TCK(dB.alloc((size_t)d * 4));
    TCK(dY.alloc((size_t)T * d * 2)); TCK(dQ.alloc((size_t)T * d * 2));
    TCK(cudaMemcpy(dP.p, pre, (size_t)T * d * 2, cudaMemcpyHostToDevice));
    TCK(cudaMemcpy(dG.p, gamma, (size_t)d * 4, cudaMemcpyHostToDevice));
    TCK(cudaMemcpy(dB.p, beta, (size_t)d * 4, cudaMemcpyHostToDevice));
    TCK(launch_add_ln(dP.as<__half>(), dG.as<float>(), dB.as<float>(), T, d, eps, dY.as<__half>(), dQ.as<__half>(), qmode, 0));
    TCK(cudaDeviceSynchronize());
    TCK(cudaMemcpy(y, dY.p, (size_t)T * d * 2, cudaMemcpyDeviceToHost));
    if (qmode && yq) TCK(cudaMemcpy(yq, dQ.p, (size_t)T * d * 2, cudaMemcpyDeviceToHost));
    return 0;
}

BK_API int bk_test_attention(int device, int32_t n_seq, const int32_t *lens, int32_t d, int32_t n_head,
                             const uint16_t *qkv, int32_t qmode, uint16_t *out) {
    TCK(cudaSetDevice(device));
    DevBuf dL, dC, dPos, dQ, dO;
    int T, mx;
    if (upload_lens(lens, n_seq, dL, dC, dPos, &T, &mx, 0)) return -1;
    TCK(dQ.alloc((size_t)T * 3 * d * 2)); TCK(dO.alloc((size_t)T * d * 2));
    TCK(cudaMemcpy(dQ.p, qkv, (size_t)T * 3 * d * 2, cudaMemcpyHostToDevice));
    CUtensorMap tmQ;
    if (!make_tmap_attention(&tmQ, dQ.p, d, n_head, (uint64_t)T)) { fprintf(stderr, "bk_test_attention: cuTensorMapEncodeTiled failed\n"); return -1; }
    TCK(launch_attention(tmQ, dC.as<int32_t>(), n_seq, mx, d, n_head, dO.as<__half>(), qmode, 0));
    TCK(cudaDeviceSynchronize());
    TCK(cudaMemcpy(out, dO.p, (size_t)T * d * 2, cudaMemcpyDeviceToHost));
    return 0;
}

BK_API int bk_test_attention_packed(int device, int32_t n64, int32_t n32, int32_t n16, const int32_t *lens, int32_t d, int32_t n_head,
                                    const uint16_t *qkv, int32_t qmode, uint16_t *out) {
    TCK(cudaSetDevice(device));
    const int n_seq = n64 + n32 + n16;
    for (int i = 0; i < n_seq; i++) {  // the classes must be consecutive runs: 33..64 | 17..32 | 1..16 tokens
        const int lo = i < n64 ? 33 : i < n64 + n32 ? 17 : 1, hi = i < n64 ? 64 : i < n64 + n32 ? 32 : 16;
        if (lens[i] < lo || lens[i] > hi) { fprintf(stderr, "bk_test_attention_packed: sentence %d has %d tokens, class %d..%d\n", i, lens[i], lo, hi); return -1; }
    }
    DevBuf dL, dC, dPos, dQ, dO;
    int T, mx;
    if (upload_lens(lens, n_seq, dL, dC, dPos, &T, &mx, 0)) return -1;
    TCK(dQ.alloc(((size_t)T + 64) * 3 * d * 2)); TCK(dO.alloc((size_t)T * d * 2));   // the last slot's box may reach 63 rows past T
    TCK(cudaMemset(dQ.p, 0, ((size_t)T + 64) * 3 * d * 2));
    TCK(cudaMemcpy(dQ.p, qkv, (size_t)T * 3 * d * 2, cudaMemcpyHostToDevice));
    CUtensorMap t64, t32, t16;
    if (!make_tmap_attention(&t64, dQ.p, d, n_head, (uint64_t)T + 64, 64) || !make_tmap_attention(&t32, dQ.p, d, n_head, (uint64_t)T + 64, 32) ||
        !make_tmap_attention(&t16, dQ.p, d, n_head, (uint64_t)T + 64, 16)) { fprintf(stderr, "bk_test_attention_packed: cuTensorMapEncodeTiled failed\n"); return -1; }
    TCK(launch_attention_packed(t64, t32, t16, dC.as<int32_t>(), n64, n32, n16, d, n_head, dO.as<__half>(), qmode, 0));
    TCK(cudaDeviceSynchronize());
    TCK(cudaMemcpy(out, dO.p, (size_t)T * d * 2, cudaMemcpyDeviceToHost));
    return 0;
}

BK_API int bk_test_pool(int device, int32_t n_seq, const int32_t *lens, int32_t d, const uint16_t *x, float *out) {
    TCK(cudaSetDevice(device));
    DevBuf dL, dC, dPos, dX, dO;
    int T, mx;
    if (upload_lens(lens, n_seq, dL, dC, dPos, &T, &mx, 0)) return -1;
    TCK(dX.alloc((size_t)T * d * 2)); TCK(dO.alloc((size_t)n_seq * d * 4));
    TCK(cudaMemcpy(dX.p, x, (size_t)T * d * 2, cudaMemcpyHostToDevice));
    TCK(launch_pool_norm(dX.as<__half>(), dC.as<int32_t>(), nullptr, n_seq, d, dO.as<float>(), 0));
    TCK(cudaDeviceSynchronize());
    TCK(cudaMemcpy(out, dO.p, (size_t)n_seq * d * 4, cudaMemcpyDeviceToHost));
    return 0;
}

BK_API int bk_test_batcher(int device, int32_t n_seq, const int32_t *lens, const int32_t *tokens, int32_t *order, int32_t *cu,
                           int32_t *tok, int32_t *pos) {
    TCK(cudaSetDevice(device));
    int T = 0;
    for (int i = 0; i < n_seq; i++) T += lens[i];
    DevBuf dL, dTin, dTok, dPos, dCu, dOrd, dSrc, dOff;
    TCK(dL.alloc((size_t)n_seq * 4)); TCK(dTin.alloc((size_t)T * 4)); TCK(dTok.alloc((size_t)T * 4)); TCK(dPos.alloc((size_t)T * 4));
    TCK(dCu.alloc(((size_t)n_seq + 1) * 4)); TCK(dOrd.alloc((size_t)n_seq * 4)); TCK(dSrc.alloc((size_t)n_seq * 4)); TCK(dOff.alloc((size_t)n_seq * 4));
    TCK(cudaMemcpy(dL.p, lens, (size_t)n_seq * 4, cudaMemcpyHostToDevice));
    TCK(cudaMemcpy(dTin.p, tokens, (size_t)T * 4, cudaMemcpyHostToDevice));
    BatcherParams b{dTin.as<int32_t>(), dL.as<int32_t>(), n_seq, T, dTok.as<int32_t>(), dPos.as<int32_t>(), dCu.as<int32_t>(),
                    dOrd.as<int32_t>(), dSrc.as<int32_t>(), dOff.as<int32_t>()};
    TCK(launch_batcher(b, 0));
    TCK(cudaDeviceSynchronize());
    TCK(cudaMemcpy(order, dOrd.p, (size_t)n_seq * 4, cudaMemcpyDeviceToHost));
    TCK(cudaMemcpy(cu, dCu.p, ((size_t)n_seq + 1) * 4, cudaMemcpyDeviceToHost));
    TCK(cudaMemcpy(tok, dTok.p, (size_t)T * 4, cudaMemcpyDeviceToHost));
    TCK(cudaMemcpy(pos, dPos.p, (size_t)T * 4, cudaMemcpyDeviceToHost));
    return 0;
}

BK_API int bk_test_embed_ln(int device, int32_t T, int32_t d, int32_t n_vocab, int32_t n_pos, int32_t wtype, const void *word, const void *type,
                            const void *posw, const float *gamma, const float *beta, float eps, const int32_t *tokens, const int32_t *pos,
                            int32_t qmode, uint16_t *x, uint16_t *xq) {
    TCK(cudaSetDevice(device));
    auto row_bytes = [&](int64_t rows) -> size_t {
        return wtype == BK_F32 ? (size_t)rows * d * 4 : wtype == BK_F16 ? (size_t)rows * d * 2 : (size_t)rows * (d / 32) * (wtype == BK_Q4_0 ? 18 : 20);
    };
    DevBuf dW, dT, dP, dG, dB, dTok, dPos, dX, dXq;
    TCK(dW.alloc(row_bytes(n_vocab))); TCK(dT.alloc(row_bytes(2))); TCK(dP.alloc(row_bytes(n_pos)));
    TCK(dG.alloc((size_t)d * 4)); TCK(dB.alloc((size_t)d * 4)); TCK(dTok.alloc((size_t)T * 4)); TCK(dPos.alloc((size_t)T * 4));
    TCK(dX.alloc((size_t)T * d * 2)); TCK(dXq.alloc((size_t)T * d * 2));
    TCK(cudaMemcpy(dW.p, word, row_bytes(n_vocab), cudaMemcpyHostToDevice));
    TCK(cudaMemcpy(dT.p, type, row_bytes(2), cudaMemcpyHostToDevice));
    TCK(cudaMemcpy(dP.p, posw, row_bytes(n_pos), cudaMemcpyHostToDevice));
    TCK(cudaMemcpy(dG.p, gamma, (size_t)d * 4, cudaMemcpyHostToDevice));
    TCK(cudaMemcpy(dB.p, beta, (size_t)d * 4, cudaMemcpyHostToDevice));
    TCK(cudaMemcpy(dTok.p, tokens, (size_t)T * 4, cudaMemcpyHostToDevice));
    TCK(cudaMemcpy(dPos.p, pos, (size_t)T * 4, cudaMemcpyHostToDevice));
    EmbedParams p;
    p.word = dW.p; p.type = dT.p; p.posw = dP.p; p.wtype = wtype; p.gamma = dG.as<float>(); p.beta = dB.as<float>();
    p.tokens = dTok.as<int32_t>(); p.pos = dPos.as<int32_t>(); p.T = T; p.d = d; p.n_vocab = n_vocab; p.eps = eps;
    p.x = dX.as<__half>(); p.xq = dXq.as<__half>(); p.qmode = qmode;
    TCK(launch_embed_ln(p, 0));
    TCK(cudaDeviceSynchronize());
    TCK(cudaMemcpy(x, dX.p, (size_t)T * d * 2, cudaMemcpyDeviceToHost));
    if (qmode && xq) TCK(cudaMemcpy(xq, dXq.p, (size_t)T * d * 2, cudaMemcpyDeviceToHost));
    return 0;
}

}  // extern "C"
