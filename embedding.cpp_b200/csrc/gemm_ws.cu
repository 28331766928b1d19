// gemm_ws.cu -- weight-stationary projection GEMM for Q4_0 / Q4_1 models: the packed GGUF blocks are the ONLY copy of the
// weight in HBM, each CTA decodes its slice ONCE into shared memory and then streams every M tile past it.
//
//   out[M][N] = epi( A[M][K] . dequant(Wq)[N][K]^T + bias[N] )        K <= 1024 (the QKV and FFN-up projections: K = n_embd)
//
// Replaces ggml_mul_mat on a Q4 weight + the bias / GELU nodes (ggml.c:9388-9597, bert.cpp:907-926, 967-971; block layout
// ggml-quants.h:10-55, decode ggml-quants.c:1074-1113) for batches above the latency regime, where round 1 ran the GEMM from f16
// operand copies expanded at load: the per-stage decode of gemm_tcgen05.cu (WQ = 1 / 2) re-decodes every W tile for every M tile and
// competes with the epilogue for issue slots (690 vs 1316 TFLOP/s at M = 32 768), this kernel decodes each weight once per CTA.
//
// Shape: CTA pairs (cta_group::2) on 256 x BN tiles, BN = 256 / 192 / 128.  Pair u owns ONE column tile n = u % (N / BN) for its
// whole life and walks over the row tiles m = u / (N/BN), + pairs / (N/BN), ...  Each CTA of the pair keeps its half of the decoded
// tile -- BN/2 weight rows x K halves, as K/64 UMMA operand blocks (BN/2 rows x 128 B, SWIZZLE_128B layout) -- next to a ring that
// carries ONLY the A operand (16 KB per stage, as many stages as fit, >= 3) and two output staging boxes.  The tile width is what
// the resident slice leaves room for: 144 KB (BN = 192) at K = 768, 128 KB (BN = 128) at K = 1024, 96 KB (BN = 256) at K = 384.
// What it costs (tools/gemm_ws_probe.py, M = 65 536, stand-alone): BERT-base QKV 1000-1027 TFLOP/s against 1326 from f16 copies on
// 256 x 256 tiles, FFN-up (+GELU, Q8) 1005 against 1131, BERT-large QKV 928 against 1386 -- and 750-805 for the per-stage decode.
// The resident slice takes the shared memory that otherwise hides latency: the A ring holds 96 KB (6 stages x 256 MMA cycles
// = 1.5 k cycles of cover at K = 768; 3 stages at BN = 192, 4 at K = 1024) where the f16 kernel has 160 KB (5 x 512 cycles), and a
// narrow tile reads 1.5x the operand bytes per FLOP from shared memory.  Hence an opt-in (BERT_B200_Q4_WS=1: half the weight memory
// of a Q4 model, no f16 copy of QKV / FFN-up), not the default.
//   prologue  16 epilogue warps: packed blocks straight from global memory (re-tiled [K/64][N][2 blocks], launch_q4_retile) ->
//             q4_decode_row -> swizzled shared memory; fence.proxy.async; cluster barrier.  Weights do not depend on the previous
//             kernel, so under programmatic dependent launch the whole prologue runs before griddepcontrol.wait.
//   warp 0    TMA producer (A only)            warp 1  tcgen05.mma issuer (leader CTA), B descriptors index the resident blocks
//   warps 4-19 epilogue: two teams of 8 warps, team t owns the 64-column boxes t, t+2 (one staging box per team);
//             bias | bias + GELU (+ Q8 rounding)
// Bit-identical to the f16-copy path and to the per-stage decode path: same decoder, same MMA K order, same epilogue arithmetic
// (gemm_parts.cuh).
#include <atomic>
#include <cstdlib>

#include "common.cuh"
#include "gemm_parts.cuh"
#include "kernels.h"
#include "../../include/bk_api.h"

namespace bk {

constexpr int kWsThreads = 640;
constexpr int kWsEpiWarp0 = 4;
constexpr int kWsABytes = GEMM_BLOCK_M * GEMM_BLOCK_K * 2;    // 16 KB: one A stage
constexpr int kWsBoxBytes = GEMM_BLOCK_M * GEMM_OUT_BOX_N * 2;  // 16 KB
constexpr int kWsMaxStages = 8;
constexpr int kWsBarBytes = 512;
constexpr int kWsFixedBytes = 1024 /*align*/ + kWsBarBytes + 2 * kWsBoxBytes;
constexpr int kWsSmemLimit = 227 * 1024;
static_assert((2 * kWsMaxStages + 4 + 2) * 8 + 8 <= kWsBarBytes, "barrier area");

// stages of the A ring that fit next to num_kb resident K blocks of a BN-wide tile (0: does not fit)
static inline int ws_stages(int bn, int num_kb) {
    const int bres = num_kb * (bn / 2) * GEMM_BLOCK_K * 2;
    int st = (kWsSmemLimit - kWsFixedBytes - bres) / kWsABytes;
    if (st > kWsMaxStages) st = kWsMaxStages;
    return st >= 3 ? st : 0;
}

template <int WQ, int BN>
__global__ void __launch_bounds__(kWsThreads, 1)
gemm_ws_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmOut, const GemmParams p, const int stages) {
    pdl_launch_dependents();
    constexpr int PK_ROWB = WQ == 1 ? 36 : 40;      // packed bytes of 64 weights of one row
    constexpr int B_ROWS = BN / 2;                   // weight rows resident per CTA
    constexpr int B_BLOCK = B_ROWS * GEMM_BLOCK_K * 2;  // one resident K block of W (a multiple of 1024: whole swizzle atoms)
    constexpr int NBOXL = BN / GEMM_OUT_BOX_N;       // logical 64-column boxes of a tile (2, 3, 4); physical staging boxes: one per team
    constexpr uint32_t TMEM_COLS = 2 * BN <= 256 ? 256 : 512;
    extern __shared__ uint8_t smem_raw[];
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    const int num_kb = p.K / GEMM_BLOCK_K;           // K % 64 == 0
    uint8_t *ring = smem;                                   // stages x A tile
    uint8_t *bres = ring + stages * kWsABytes;              // num_kb resident W blocks
    uint8_t *obuf = bres + num_kb * B_BLOCK;                // 2 output boxes
    uint64_t *full_bar = reinterpret_cast<uint64_t *>(obuf + 2 * kWsBoxBytes);
    uint64_t *empty_bar = full_bar + kWsMaxStages;
    uint64_t *tfull_bar = empty_bar + kWsMaxStages;  // [2] accumulator ready
    uint64_t *tempty_bar = tfull_bar + 2;            // [2] accumulator drained
    uint64_t *sfree_bar = tempty_bar + 2;            // [2] (per team) the store of the team's box has finished reading it
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(sfree_bar + 2);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int rank = (int)cluster_ctarank();
    const int unit = (int)(blockIdx.x >> 1), units = (int)(gridDim.x >> 1);
    const int n_tiles = p.N / BN;                    // the launcher guarantees N % BN == 0 and units % n_tiles == 0
    const int m_tiles = (p.M + 2 * GEMM_BLOCK_M - 1) / (2 * GEMM_BLOCK_M);
    const int n_blk = unit % n_tiles;
    const int m_first = unit / n_tiles, m_step = units / n_tiles;
    const int n0 = n_blk * BN;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmA);
        tma_prefetch_desc(&tmOut);
        for (int i = 0; i < stages; i++) { mbar_init(&full_bar[i], 2); mbar_init(&empty_bar[i], 1); }
        for (int i = 0; i < 2; i++) { mbar_init(&tfull_bar[i], 1); mbar_init(&tempty_bar[i], 2 * 16); mbar_init(&sfree_bar[i], 1); }
        mbar_fence_init();
    }
    if (warp == 1) tmem_alloc<TMEM_COLS, 2>(tmem_slot);
    if (warp >= kWsEpiWarp0) {
        // ---- prologue: this CTA's BN/2 weight rows x K, decoded once.  item = (K block, row); 512 threads
        const int tid = threadIdx.x - kWsEpiWarp0 * 32;
        const int nrow0 = n0 + rank * B_ROWS;
        for (int item = tid; item < num_kb * B_ROWS; item += 512) {
            const int kb = item / B_ROWS, r = item % B_ROWS;
            const uint8_t *src = p.wq + ((size_t)kb * p.N + nrow0 + r) * PK_ROWB;  // 4-byte aligned (36 / 40 bytes per row)
            uint4 o[8];
            q4_decode_row<WQ>(reinterpret_cast<const uint32_t *>(src), o);
            uint8_t *dst = bres + kb * B_BLOCK + r * 128;
            const uint32_t sw = (uint32_t)(r & 7);
#pragma unroll
            for (int j = 0; j < 8; j++) *reinterpret_cast<uint4 *>(dst + ((((uint32_t)j) ^ sw) << 4)) = o[j];
        }
        fence_proxy_async_smem();  // generic-proxy writes -> visible to the MMAs' operand fetch (async proxy), both CTAs
    }
    tc_fence_before();
    cluster_sync_all();  // barriers initialised, TMEM allocated, both halves of the weight tile resident
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();          // everything above ran under the previous kernel's tail; from here on activations are touched

    if (warp == 0) {
        // ===================================================================== TMA producer (A only)
        int s = 0; uint32_t ph = 0;
        for (int m_blk = m_first; m_blk < m_tiles; m_blk += m_step) {
            const int m0 = m_blk * 2 * GEMM_BLOCK_M + rank * GEMM_BLOCK_M;
            for (int kb = 0; kb < num_kb; kb++) {
                mbar_wait(&empty_bar[s], ph ^ 1);
                const uint32_t lead_full = smem_u32(&full_bar[s]) & kPeerBitMask;  // both CTAs report to the leader's barrier
                mbar_arrive_expect_tx_cluster_e(lead_full, kWsABytes);
                tma_load_2d_pair_e(ring + s * kWsABytes, &tmA, lead_full, kb * GEMM_BLOCK_K, m0);
                if (++s == stages) { s = 0; ph ^= 1; }
            }
        }
    } else if (warp == 1 && rank == 0) {
        // ======================================================================= MMA issuer (leader CTA)
        constexpr uint32_t idesc = umma_idesc_f16(2 * GEMM_BLOCK_M, BN);
        const uint32_t bres_addr = smem_u32(bres);
        int s = 0; uint32_t ph = 0;
        int as = 0; uint32_t aph = 0;
        for (int m_blk = m_first; m_blk < m_tiles; m_blk += m_step) {
            mbar_wait(&tempty_bar[as], aph ^ 1);  // epilogue has drained this accumulator stage
            tc_fence_after();
            const uint32_t tmem_d = tmem_base + (uint32_t)(as * BN);
            for (int kb = 0; kb < num_kb; kb++) {
                mbar_wait(&full_bar[s], ph);
                tc_fence_after();
                const uint64_t adesc = umma_desc_sw128(smem_u32(ring + s * kWsABytes));
                const uint64_t bdesc = umma_desc_sw128(bres_addr + (uint32_t)(kb * B_BLOCK));
#pragma unroll
                for (int k = 0; k < GEMM_BLOCK_K / 16; k++)
                    umma_f16_e<2>(tmem_d, adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), idesc, (kb | k) != 0);
                umma_commit_e<2>(&empty_bar[s]);                        // A stage reusable in both CTAs when these MMAs retire
                if (kb == num_kb - 1) umma_commit_e<2>(&tfull_bar[as]);  // accumulator complete
                if (++s == stages) { s = 0; ph ^= 1; }
            }
            if (++as == 2) { as = 0; aph ^= 1; }
        }
    } else if (warp >= kWsEpiWarp0) {
        // ========================================================================= epilogue (no residual on these projections)
        const int ew = warp - kWsEpiWarp0;
        const int team = ew >> 3;         // logical boxes team, team + 2; physical staging box = team
        const int quad = ew & 3;          // == warp % 4: TMEM lanes 32*quad .. 32*quad+31
        const int half = (ew >> 2) & 1;   // which 32 columns of the 64-column box
        const int r = quad * 32 + lane;   // row inside this CTA's 128 rows
        const bool issuer = (ew & 7) == 0 && lane == 0;
        const uint32_t row_off = (uint32_t)r * 128u;
        const uint32_t sw = (uint32_t)(r & 7);
        uint8_t *box = obuf + team * kWsBoxBytes;
        int as = 0; uint32_t aph = 0;
        uint32_t bph = 0;
        const uint32_t tempty_addr0 = smem_u32(&tempty_bar[0]) & kPeerBitMask;  // accumulator hand-back goes to the leader's MMA warp
        for (int m_blk = m_first; m_blk < m_tiles; m_blk += m_step) {
            const int m0 = m_blk * 2 * GEMM_BLOCK_M + rank * GEMM_BLOCK_M;
            mbar_wait(&tfull_bar[as], aph);
            tc_fence_after();
            if (team >= NBOXL) {  // (not reached: NBOXL >= 2)
                tc_fence_before();
                if (lane == 0) mbar_arrive_cluster(tempty_addr0 + 8u * as);
            }
#pragma unroll 1
            for (int b = team; b < NBOXL; b += 2) {
                uint32_t acc[32];
                tmem_ld_32x32(tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(as * BN + b * GEMM_OUT_BOX_N + half * 32), acc);
                tmem_ld_wait();
                if (b + 2 >= NBOXL) {  // last read of this stage by this warp: hand the accumulator back
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive_cluster(tempty_addr0 + 8u * as);
                }
                mbar_wait(&sfree_bar[team], bph ^ 1);  // the previous store of this team's box has read it
                bph ^= 1;
                const int n = n0 + b * GEMM_OUT_BOX_N + half * 32;
                float2 v[16];
                epi_add_bias(acc, p.bias + n, v);
                if (p.epi >= BK_EPI_BIAS_GELU) epi_gelu_q8(v, p.epi);
                epi_stage_f16(v, box + row_off, half, sw);
                fence_proxy_async_smem();              // my generic-proxy writes -> visible to the TMA store
                named_bar_sync(1 + team, 256);
                if (issuer) {
                    tma_store_2d(&tmOut, box, n0 + b * GEMM_OUT_BOX_N, m0);
                    bulk_commit_group();
                    bulk_wait_group_read<0>();
                    mbar_arrive(&sfree_bar[team]);
                }
            }
            if (++as == 2) { as = 0; aph ^= 1; }
        }
        if (issuer) bulk_wait_group<0>();  // all stores complete before the CTA (and its smem) goes away
    }

    tc_fence_before();
    cluster_sync_all();  // the peer may still be read by the leader's MMAs / signalled until here
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc<TMEM_COLS, 2>(tmem_base);
    }
}

// tile width for an N x K weight on num_sms / 2 CTA pairs: the widest tile whose resident slice fits, weighted by the SMs the
// column-tile count leaves busy (pairs per column tile is an integer) and by the measured relative speed of the tile width
static int ws_pick_bn(int N, int K, int num_sms) {
    static const int bns[3] = {256, 192, 128};
    static const double eff[3] = {1.00, 0.80, 0.78};  // stand-alone TFLOP/s relative to 256 columns (tools/gemm_ws_probe.py)
    const int pairs = num_sms / 2;
    int best = 0;
    double best_score = 0;
    for (int i = 0; i < 3; i++) {
        const int bn = bns[i];
        if (N % bn != 0 || !ws_stages(bn, K / GEMM_BLOCK_K)) continue;
        const int n_tiles = N / bn, per_col = pairs / n_tiles;
        if (per_col < 1) continue;
        const double score = eff[i] * (double)(per_col * n_tiles) / pairs;
        if (score > best_score) { best_score = score; best = bn; }
    }
    return best;
}

bool gemm_ws_supported(int N, int K, int num_sms) {
    if (K % GEMM_BLOCK_K != 0 || K < GEMM_BLOCK_K) return false;
    return ws_pick_bn(N, K, num_sms) != 0;
}

template <int WQ, int BN>
static cudaError_t launch_ws(const CUtensorMap &tmA, const CUtensorMap &tmOut, const GemmParams &p, int num_sms, cudaStream_t st) {
    static std::atomic<bool> attr_set[64];  // per kernel instance and device (the attribute is per device)
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64 || !attr_set[dev]) {
        cudaError_t err = cudaFuncSetAttribute(gemm_ws_kernel<WQ, BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, kWsSmemLimit);
        if (err != cudaSuccess) return err;
        if (dev >= 0 && dev < 64) attr_set[dev] = true;
    }
    const int num_kb = p.K / GEMM_BLOCK_K;
    const int stages = ws_stages(BN, num_kb);
    if (!stages) return cudaErrorInvalidValue;
    const size_t smem = (size_t)stages * kWsABytes + (size_t)num_kb * (BN / 2) * GEMM_BLOCK_K * 2 + kWsFixedBytes;
    const int n_tiles = p.N / BN, m_tiles = (p.M + 2 * GEMM_BLOCK_M - 1) / (2 * GEMM_BLOCK_M);
    int per_col = (num_sms / 2) / n_tiles;            // pairs per column tile
    if (per_col > m_tiles) per_col = m_tiles;         // small batches: no pair without a row tile
    const int grid = 2 * per_col * n_tiles;
    LaunchConfig lc(dim3((unsigned)grid), dim3(kWsThreads), smem, st, 2);
    return cudaLaunchKernelEx(&lc.cfg, gemm_ws_kernel<WQ, BN>, tmA, tmOut, p, stages);
}

cudaError_t launch_gemm_ws(const CUtensorMap &tmA, const CUtensorMap &tmOut, const GemmParams &p, int num_sms, cudaStream_t st, int force_bn) {
    if (!p.wq || (p.wq_type != BK_Q4_0 && p.wq_type != BK_Q4_1) || p.K % GEMM_BLOCK_K != 0) return cudaErrorInvalidValue;
    if (p.epi == BK_EPI_BIAS_RESID || p.M < 1) return cudaErrorInvalidValue;
    static const int env_bn = [] { const char *v = getenv("BERT_B200_WS_BN"); return v ? atoi(v) : 0; }();
    int bn = force_bn ? force_bn : env_bn;
    if (bn != 128 && bn != 192 && bn != 256) bn = ws_pick_bn(p.N, p.K, num_sms);
    if (!bn || p.N % bn != 0 || !ws_stages(bn, p.K / GEMM_BLOCK_K) || (num_sms / 2) / (p.N / bn) < 1) return cudaErrorInvalidValue;
#define BK_WS(BN_)                                                                                                                  \
    case BN_:                                                                                                                       \
        return p.wq_type == BK_Q4_0 ? launch_ws<1, BN_>(tmA, tmOut, p, num_sms, st) : launch_ws<2, BN_>(tmA, tmOut, p, num_sms, st)
    switch (bn) {
        BK_WS(256);
        BK_WS(192);
        default: return p.wq_type == BK_Q4_0 ? launch_ws<1, 128>(tmA, tmOut, p, num_sms, st) : launch_ws<2, 128>(tmA, tmOut, p, num_sms, st);
    }
#undef BK_WS
}

}  // namespace bk
