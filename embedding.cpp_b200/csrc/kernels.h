// kernels.h -- host-side launch interface of the sm_100a kernels (internal to libbert.so).
#pragma once
#include <cuda.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace bk {

// Programmatic dependent launch (common.cuh): on for the launches of the calling thread while the engine captures the
// small-batch launch plan.  Every kernel of that plan is launched through launch_config(), which adds the attribute.
inline bool &pdl_flag() { static thread_local bool on = false; return on; }
struct LaunchConfig {
    cudaLaunchConfig_t cfg;
    cudaLaunchAttribute attr[2];
    LaunchConfig(dim3 grid, dim3 block, size_t smem, cudaStream_t st, int cluster = 0) : cfg{} {
        cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
        unsigned n = 0;
        if (cluster > 0) {
            attr[n].id = cudaLaunchAttributeClusterDimension;
            attr[n].val.clusterDim.x = (unsigned)cluster; attr[n].val.clusterDim.y = 1; attr[n].val.clusterDim.z = 1;
            n++;
        }
        if (pdl_flag()) {
            attr[n].id = cudaLaunchAttributeProgrammaticStreamSerialization;
            attr[n].val.programmaticStreamSerializationAllowed = 1;
            n++;
        }
        cfg.attrs = attr;
        cfg.numAttrs = n;
    }
};

// ---- gemm_tcgen05.cu ------------------------------------------------------------------------
// out[M][N] = epi(A[M][K] . W[N][K]^T + bias (+ resid))   A, W, out, resid: f16, fp32 accumulation in TMEM.
struct GemmParams {
    int M, N, K;
    const float *bias;  // [N]
    int epi;            // BK_EPI_*
    // non-null: W is read as re-tiled Q4 blocks (launch_q4_retile) and decoded in shared memory; wq_type = BK_Q4_0 / BK_Q4_1
    const uint8_t *wq = nullptr;
    int wq_type = 0;
};
// GGUF Q4_0 / Q4_1 matrix [n_rows][K] (blocks along K) -> rows n_off.. of the K-slice-major tiling [K/64][N_total][2 blocks]
// that the in-smem decode path streams; same bytes, same size
cudaError_t launch_q4_retile(const void *src, int wtype, int n_rows, int K, void *dst, int N_total, int n_off, cudaStream_t st);
constexpr int GEMM_BLOCK_M = 128;
constexpr int GEMM_BLOCK_K = 64;
constexpr int GEMM_OUT_BOX_N = 64;  // columns of one output / residual TMA box (128 B of f16, 128-byte swizzle)
// TMA descriptors of one weight matrix W[N][K] (K contiguous), one per box height 32/64/96/128/192/256 rows
// (= tile width BN for one-CTA tiles, BN/2 for CTA-pair tiles)
struct WeightMaps {
    CUtensorMap tm[6];
    int N = 0, K = 0;
};
bool make_weight_maps(WeightMaps *w, const void *base, int N, int K);
// tile width BN and CTAs per tile (1, or 2 = CTA pair on a 256 x BN tile) that minimise wave-quantisation loss
// for an M x N problem on num_sms persistent CTAs
void gemm_pick_tile(int M, int N, int num_sms, int *bn, int *cg);
// tmA: map over A {K, M} box {64, 128}; tmOut / tmRes: maps over out / resid {N, M} box {64, 128}; all SWIZZLE_128B.
// tmRes may be null unless epi == BK_EPI_BIAS_RESID.  Rows >= the maps' row count are never written.
cudaError_t launch_gemm(const CUtensorMap &tmA, const WeightMaps &W, const CUtensorMap &tmOut, const CUtensorMap *tmRes,
                        const GemmParams &p, int num_sms, cudaStream_t st, int force_bn = 0, int force_cg = 0);
// 2-D f16 row-major tensor map (inner = contiguous dim) with 128-byte (or 64-byte) swizzle
bool make_tmap_f16_2d(CUtensorMap *tm, const void *base, uint64_t inner, uint64_t rows, uint64_t row_stride_bytes,
                      uint32_t box_inner, uint32_t box_rows, int swizzle_bytes = 128);
// y[i] = the epilogue's GELU of x[i] (tests)
cudaError_t launch_gelu_test(const float *x, float *y, int n, cudaStream_t st);

// ---- gemm_ws.cu -----------------------------------------------------------------------------
// weight-stationary Q4 GEMM: p.wq = the re-tiled packed blocks (launch_q4_retile), the ONLY copy of the weight; every CTA pair decodes
// its column tile (256 / 192 / 128 wide: what its resident slice leaves room for) once into shared memory and streams all row tiles
// past it.  K % 64 == 0, the slice of the narrowest tile that divides N must fit (K <= 1024 at 128 columns), no residual.
bool gemm_ws_supported(int N, int K, int num_sms);
cudaError_t launch_gemm_ws(const CUtensorMap &tmA, const CUtensorMap &tmOut, const GemmParams &p, int num_sms, cudaStream_t st, int force_bn = 0);

// ---- gemm_ln.cu -----------------------------------------------------------------------------
// y = LN(A . W^T + bias + resid) * gamma + beta (and its Q8-rounded copy yq when qmode != 0) in one kernel:
// clusters of CTA pairs hold whole 256 x N row blocks in TMEM.  N must be 384, 768 or 1024.
// tmA: {K, M} box {64, 128}; tmRes / tmY / tmYQ: {N, M} box {64, 128}; all SWIZZLE_128B.
bool gemm_ln_supported(int N, int K);
cudaError_t launch_gemm_ln(const CUtensorMap &tmA, const WeightMaps &W, const CUtensorMap &tmRes, const CUtensorMap &tmY,
                           const CUtensorMap *tmYQ, const float *bias, const float *gamma, const float *beta, float eps, int qmode,
                           int M, int N, int K, int num_sms, cudaStream_t st);

// tuning aid: 16 tiles x 8 clock64 stamps of block 0's first epilogue thread are written to dev_buf (null = off)
void gemm_ln_set_trace(long long *dev_buf);

// ---- rowwise.cu -----------------------------------------------------------------------------
// the on-device batcher: sentences in the caller's order (ids concatenated in tok_in, lens) -> sorted by length (longest first) and
// packed: tok / pos [T], cu [n_seq+1], order[s] = caller index of sorted sentence s; src_off / caller_off: scratch [n_seq]
struct BatcherParams {
    const int32_t *tok_in, *lens;
    int n_seq, T;
    int32_t *tok, *pos, *cu, *order, *src_off, *caller_off;
};
cudaError_t launch_batcher(const BatcherParams &b, cudaStream_t st);
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
// sentence s (rows cu[s]..cu[s+1] of x) -> out row order[s] (order == nullptr: row s)
cudaError_t launch_pool_norm(const __half *x, const int32_t *cu_seqlens, const int32_t *order, int n_seq, int d, float *out, cudaStream_t st);

// ---- attention_tc.cu ------------------------------------------------------------------------
// qkv [T][3d] f16 (Q | K | V along the row) read through tmQKV (make_tmap_attention), out [T][d] f16;
// one sentence = rows cu[s]..cu[s+1], at most 512 tokens; head size 32 or 64
bool make_tmap_attention(CUtensorMap *tm, const void *qkv, int d, int n_head, uint64_t rows, int box_rows = 64);
// length classes of one packed, length-sorted batch (consecutive runs, longest first): n_long sentences of more than 128 tokens,
// n_mid of up to max_mid <= 128 tokens that get a tile each, then n64 / n32 / n16 short ones (33..64 / 17..32 / 1..16 tokens) that
// share tiles
struct AttSplit { int n_long, n_mid, max_mid, n64, n32, n16; };
// short sentences, several per 128-row tile (attention_pack.cuh): cu_seqlens points at the first of them; tm64 / tm32 / tm16 =
// make_tmap_attention with boxes of 64 / 32 / 16 rows.  Bit-identical to launch_attention on the same sentences.
cudaError_t launch_attention_packed(const CUtensorMap &tm64, const CUtensorMap &tm32, const CUtensorMap &tm16, const int32_t *cu_seqlens,
                                    int n64, int n32, int n16, int d, int n_head, __half *out, int qmode, cudaStream_t st);
cudaError_t launch_attention(const CUtensorMap &tmQKV, const int32_t *cu_seqlens, int n_seq, int max_len, int d, int n_head,
                             __half *out, int qmode, cudaStream_t st);

// ---- dequant.cu -----------------------------------------------------------------------------
// weight matrix [rows][K] in GGUF storage type -> f16 [rows][K] (UMMA operand copy, done once at load)
cudaError_t launch_weight_to_f16(const void *src, int wtype, int64_t rows, int64_t K, __half *dst, cudaStream_t st);

}  // namespace bk
