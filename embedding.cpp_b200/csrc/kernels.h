// kernels.h -- host-side launch interface of the sm_100a kernels (internal to libbert.so).
#pragma once
#include <cuda.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace bk {

// ---- gemm_tcgen05.cu ------------------------------------------------------------------------
// out[M][N] = epi(A[M][K] . W[N][K]^T + bias)   A, W: f16, K contiguous; fp32 accumulation in TMEM.
struct GemmParams {
    int M, N, K;
    const float *bias;    // [N]
    const __half *resid;  // [M][ldr] (EPI_BIAS_RESID) or nullptr
    int ldr;
    __half *out;          // [M][ldo]
    int ldo;
    int epi;              // BK_EPI_*
};
// box rows of the two TMA descriptors the kernel expects for a given N (tile width selection)
int gemm_block_n(int N);
constexpr int GEMM_BLOCK_M = 128;
constexpr int GEMM_BLOCK_K = 64;
// tmA: 2-D map over A {K, M} box {64, 128}; tmB: map over W {K, N} box {64, gemm_block_n(N)}; both SWIZZLE_128B
cudaError_t launch_gemm(const CUtensorMap &tmA, const CUtensorMap &tmB, const GemmParams &p, int num_sms, cudaStream_t st);
// 2-D f16 row-major tensor map (inner = contiguous dim) with 128-byte swizzle
bool make_tmap_f16_2d(CUtensorMap *tm, const void *base, uint64_t inner, uint64_t rows, uint64_t row_stride_bytes,
                      uint32_t box_inner, uint32_t box_rows);

// ---- rowwise.cu -----------------------------------------------------------------------------
// cu_seqlens (exclusive prefix sum of lens, n_seq+1 entries) and per-token position inside its sentence
cudaError_t launch_pack(const int32_t *lens, int n_seq, int32_t *cu_seqlens, int32_t *pos, int T, cudaStream_t st);
// x = LN(word[tok] + type[0] + pos[p]) * gamma + beta ; tables in their GGUF storage type
struct EmbedParams {
    const void *word, *type, *posw;  // tables, raw GGUF bytes on device
    int wtype;                       // BK_F32 / F16 / Q4_0 / Q4_1 (same for the three tables)
    const float *gamma, *beta;
    const int32_t *tokens, *pos;
    int T, d, n_vocab;
    float eps;
    __half *x;   // [T][d]
    __half *xq;  // Q8-rounded copy (qmode != 0) or nullptr
    int qmode;   // 0 none, 1 Q8_0, 2 Q8_1
};
cudaError_t launch_embed_ln(const EmbedParams &p, cudaStream_t st);
cudaError_t launch_add_ln(const __half *pre, const float *gamma, const float *beta, int T, int d, float eps,
                          __half *y, __half *yq, int qmode, cudaStream_t st);
cudaError_t launch_pool_norm(const __half *x, const int32_t *cu_seqlens, int n_seq, int d, float *out, cudaStream_t st);

// ---- attention.cu ---------------------------------------------------------------------------
// qkv [T][3d] f16 (Q | K | V along the row), out [T][d] f16; one sentence = rows cu[s]..cu[s+1]
cudaError_t launch_attention(const __half *qkv, const int32_t *cu_seqlens, int n_seq, int max_len, int d, int n_head,
                             __half *out, int qmode, cudaStream_t st);

// ---- dequant.cu -----------------------------------------------------------------------------
// weight matrix [rows][K] in GGUF storage type -> f16 [rows][K] (UMMA operand copy, done once at load)
cudaError_t launch_weight_to_f16(const void *src, int wtype, int64_t rows, int64_t K, __half *dst, cudaStream_t st);

}  // namespace bk
