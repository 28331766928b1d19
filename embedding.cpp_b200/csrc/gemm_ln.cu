// gemm_ln.cu -- projection GEMM with the residual add AND the LayerNorm that follows it fused into the epilogue:
//
//   y[M][N] = LN( A[M][K] . W[N][K]^T + bias[N] + resid[M][N] ) * gamma + beta      (N = n_embd: 384 / 768 / 1024)
//   yq      = the same row Q8_0 / Q8_1-rounded per 32 elements (Q4 models: the next GEMM's input)
//
// Replaces ggml_mul_mat + add(bias) + add(residual) + ggml_norm + mul(gamma) + add(beta) of the attention-output
// and FFN-down blocks (reference bert.cpp:945-963, 974-992; ggml.c:9388-9597, 8959-9008), i.e. gemm_tcgen05's
// BK_EPI_BIAS_RESID launch plus add_ln_kernel: the pre-LN tensor never exists in HBM (2 of the 5 [T][d] passes
// of the unfused pair disappear) and is never rounded to f16 (the reference keeps it in fp32 too).
//
// A LayerNorm row needs all N columns, but one SM's TMEM holds 128 rows x 512 fp32 columns.  So a CLUSTER owns
// a 256-row block: NPAIRS CTA pairs side by side along N (cluster of 4 / 6 / 8 CTAs for N = 384 / 768 / 1024),
// each pair (cta_group::2) accumulating 256 rows x BNL columns (BNL = 192 or 256), two accumulator stages per
// CTA, so the epilogue of tile i runs under the main loop of tile i+1 exactly as in gemm_tcgen05.cu.
// The epilogue works on the accumulator in place:
//   pass 1  x = acc + bias + resid, written back to TMEM (tcgen05.st); shifted sums -> (mean_i, M2_i) per thread
//   stats   the 4 column groups of a CTA combine through shared memory; each CTA sends its (mean, M2) per row to
//           the CTAs of the other pairs that hold the same rows (st.async into their shared memory, completing
//           transaction bytes on their mbarrier); Chan's parallel-variance combination
//   pass 2  y = ((x - mean) * rstd) * gamma + beta -> f16 -> swizzled staging box -> TMA store (and yq)
// The residual tile arrives by TMA in the staging boxes the output later leaves from (in place).
// Warps: 0 TMA producer (A, W), 1 MMA issuer (leader CTA of each pair), 2 residual loader, 3 TMA-store issuer,
// 4-19 epilogue (two teams of 8 warps; team t owns the 64-column boxes t, t+2; warp = 32 rows x 32 columns).
#include <cstdio>
#include <cstdlib>

#include "common.cuh"
#include "kernels.h"
#include "../../include/bk_api.h"

namespace bk {

// every wait of this kernel goes through a bounded spin: a protocol error shows up as a message + trap (a launch
// failure the host reports) instead of a hung GPU
__device__ __noinline__ void ln_wait_timeout(int tag, uint32_t parity) {
    printf("gemm_ln: wait %d (parity %u) timed out: block %d warp %d lane %d\n", tag, parity, (int)blockIdx.x, (int)(threadIdx.x >> 5), (int)(threadIdx.x & 31));
    __trap();
}
__device__ __forceinline__ void ln_wait(uint64_t *bar, uint32_t parity, int tag) {
    if (mbar_try_wait(bar, parity)) return;
    const long long t0 = clock64();
    while (!mbar_try_wait(bar, parity))
        if (clock64() - t0 > 4000000000LL) ln_wait_timeout(tag, parity);  // ~2 s
}

constexpr int kLnThreads = 640;
constexpr int kLnEpiWarp0 = 4;
constexpr int kLnSmemLimit = 227 * 1024;
constexpr int kLnBoxBytes = GEMM_BLOCK_M * GEMM_OUT_BOX_N * 2;  // 16 KB

template <int NPAIRS, int BNL, int YQ>
struct GemmLnCfg {
    static_assert(BNL == 192 || BNL == 256, "columns per CTA pair = one UMMA N");
    static_assert(NPAIRS >= 2 && NPAIRS <= 4, "cluster of 4, 6 or 8 CTAs");
    static constexpr int B_ROWS = BNL / 2;                       // W rows this CTA loads per K block
    static constexpr int A_BYTES = GEMM_BLOCK_M * GEMM_BLOCK_K * 2;
    static constexpr int B_BYTES = B_ROWS * GEMM_BLOCK_K * 2;
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int NBOX = BNL / GEMM_OUT_BOX_N;            // 64-column boxes per CTA (3 or 4)
    // residual in, y / yq out, in place -- in a ring of NBOX + 1 physical boxes: use u = tile * NBOX + b of a box lives in
    // physical box u % NRING, so the residual of (tile + 1, b) only waits for the store of (tile, b - 1), not (tile, b), and
    // is already there when pass 1 of the next tile reaches it
    static constexpr int NRING = NBOX + 1;
    static constexpr int OUT_BYTES = NRING * kLnBoxBytes;
    // YQ = 1 (Q4 models, short K): the Q8-rounded copy gets a staging box per epilogue team, so y and yq leave in ONE sweep
    // over the accumulator (YQ = 0 stages yq in the y box after its store has been read: a second sweep).  The two boxes cost an
    // operand stage, which only the epilogue-bound short-K shape can spare.
    static constexpr int YQ_BYTES = YQ ? 2 * kLnBoxBytes : 0;
    static constexpr int PART_BYTES = 4 * GEMM_BLOCK_M * 8;      // (mean, M2) of the 4 column groups
    static constexpr int COMB_BYTES = GEMM_BLOCK_M * 8;          // this CTA's combined (mean, M2) per row
    static constexpr int RCV_BYTES = 2 * NPAIRS * GEMM_BLOCK_M * 8;  // [tile parity][sender pair][row]
    static constexpr int BAR_BYTES = 512;
    static constexpr int FIXED_BYTES = 1024 + BAR_BYTES + OUT_BYTES + YQ_BYTES + PART_BYTES + COMB_BYTES + RCV_BYTES;
    static constexpr int STAGES_FIT = (kLnSmemLimit - FIXED_BYTES) / STAGE_BYTES;
    static constexpr int STAGES = STAGES_FIT > 6 ? 6 : STAGES_FIT;
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + FIXED_BYTES;
    static_assert(STAGES >= 3, "operand ring too shallow");
    static_assert((2 * STAGES + 4 + 4 * NRING + 2 + 4) * 8 + 8 <= BAR_BYTES, "barrier area");
    static_assert(B_BYTES % 1024 == 0, "W tile ends on a swizzle atom");
};

struct GemmLnParams {
    int M, N, K;
    const float *bias, *gamma, *beta;
    float eps;
    int qmode;         // 0: y only; 1 / 2: also yq (Q8_0 / Q8_1 rounding)
    long long *trace;  // tuning aid: clock64 stamps of block 0's first epilogue thread, 8 per tile (null = off)
};

template <int NPAIRS, int BNL, int YQ>
__global__ void __launch_bounds__(kLnThreads, 1)
gemm_ln_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
               const __grid_constant__ CUtensorMap tmRes, const __grid_constant__ CUtensorMap tmY,
               const __grid_constant__ CUtensorMap tmYQ, const GemmLnParams p) {
    pdl_launch_dependents();
    using Cfg = GemmLnCfg<NPAIRS, BNL, YQ>;
    constexpr int STAGES = Cfg::STAGES;
    constexpr int NBOX = Cfg::NBOX;
    constexpr int NRING = Cfg::NRING;
    constexpr int CSIZE = 2 * NPAIRS;  // CTAs per cluster
    extern __shared__ uint8_t smem_raw[];
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t *ring = smem;
    uint8_t *obuf = smem + STAGES * Cfg::STAGE_BYTES;                       // NBOX boxes, 1024-aligned
    uint8_t *yqbuf = obuf + Cfg::OUT_BYTES;                                 // YQ: one box per epilogue team
    float2 *part = reinterpret_cast<float2 *>(yqbuf + Cfg::YQ_BYTES);       // [4][128]
    float2 *comb = part + 4 * GEMM_BLOCK_M;                                 // [128]
    float2 *rcv = comb + GEMM_BLOCK_M;                                      // [2][NPAIRS][128]
    uint64_t *full_bar = reinterpret_cast<uint64_t *>(rcv + 2 * NPAIRS * GEMM_BLOCK_M);
    uint64_t *empty_bar = full_bar + STAGES;
    uint64_t *tfull_bar = empty_bar + STAGES;    // [2] accumulator stage complete
    uint64_t *tempty_bar = tfull_bar + 2;        // [2] accumulator stage drained (the leader CTA's is waited on)
    uint64_t *rfull_bar = tempty_bar + 2;        // [NRING] (per physical box) residual landed
    uint64_t *rfree_bar = rfull_bar + NRING;     // [NRING] the box's last store has read it: next residual may land
    uint64_t *ofull_bar = rfree_bar + NRING;     // [NRING] output staged in the box by the 8 warps of its team
    uint64_t *oread_bar = ofull_bar + NRING;     // [NRING] the y store has read the box: the team may stage yq (Q4 models)
    uint64_t *stat_bar = oread_bar + NRING;      // [2] (by tile parity) row statistics of all pairs present
    uint64_t *yqfull_bar = stat_bar + 2;         // [2] (YQ) yq staged in the team's box by its 8 warps
    uint64_t *yqread_bar = yqfull_bar + 2;       // [2] (YQ) the yq store has read the team's box
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(yqread_bar + 2);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int crank = (int)cluster_ctarank();
    const int rank = crank & 1;        // half of the pair tile (rows)
    const int pair = crank >> 1;       // column block of the cluster tile
    const int cluster_id = (int)blockIdx.x / CSIZE;
    const int n_clusters = (int)gridDim.x / CSIZE;
    const int m_tiles = (p.M + 2 * GEMM_BLOCK_M - 1) / (2 * GEMM_BLOCK_M);
    const int num_kb = (p.K + GEMM_BLOCK_K - 1) / GEMM_BLOCK_K;
    const int ncol0 = pair * BNL;      // first output column of this CTA

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmA);
        tma_prefetch_desc(&tmB);
        tma_prefetch_desc(&tmRes);
        tma_prefetch_desc(&tmY);
        if (p.qmode) tma_prefetch_desc(&tmYQ);
        for (int i = 0; i < STAGES; i++) { mbar_init(&full_bar[i], 2); mbar_init(&empty_bar[i], 1); }
        for (int i = 0; i < 2; i++) { mbar_init(&tfull_bar[i], 1); mbar_init(&tempty_bar[i], 2 * 16); mbar_init(&stat_bar[i], GEMM_BLOCK_M); }
        for (int i = 0; i < NRING; i++) { mbar_init(&rfull_bar[i], 1); mbar_init(&rfree_bar[i], 1); mbar_init(&ofull_bar[i], 8); mbar_init(&oread_bar[i], 1); }
        for (int i = 0; i < 2; i++) { mbar_init(&yqfull_bar[i], 8); mbar_init(&yqread_bar[i], 1); }
        mbar_fence_init();
    }
    if (warp == 1) tmem_alloc<512, 2>(tmem_slot);
    tc_fence_before();
    cluster_sync_all();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();  // prologue done under the previous kernel's tail; global memory from here on

    if (warp == 0) {
        // ===================================================================== TMA producer (A, W)
        int s = 0; uint32_t ph = 0;
        for (int t = cluster_id; t < m_tiles; t += n_clusters) {
            const int m0 = t * 2 * GEMM_BLOCK_M + rank * GEMM_BLOCK_M;
            for (int kb = 0; kb < num_kb; kb++) {
                ln_wait(&empty_bar[s], ph ^ 1, 1);
                uint8_t *sa = ring + s * Cfg::STAGE_BYTES, *sb = sa + Cfg::A_BYTES;
                const uint32_t lead_full = smem_u32(&full_bar[s]) & kPeerBitMask;
                mbar_arrive_expect_tx_cluster_e(lead_full, Cfg::STAGE_BYTES);
                tma_load_2d_pair_e(sa, &tmA, lead_full, kb * GEMM_BLOCK_K, m0);
                tma_load_2d_pair_e(sb, &tmB, lead_full, kb * GEMM_BLOCK_K, ncol0 + rank * Cfg::B_ROWS);
                if (++s == STAGES) { s = 0; ph ^= 1; }
            }
        }
    } else if (warp == 1 && rank == 0) {
        // ======================================================================= MMA issuer (pair leader)
        constexpr uint32_t idesc = umma_idesc_f16(2 * GEMM_BLOCK_M, BNL);
        const uint16_t pair_mask = (uint16_t)(3u << (2 * pair));
        int s = 0; uint32_t ph = 0;
        int as = 0; uint32_t aph = 0;
        for (int t = cluster_id; t < m_tiles; t += n_clusters) {
            ln_wait(&tempty_bar[as], aph ^ 1, 2);  // both CTAs' epilogues have read this stage out of TMEM
            tc_fence_after();
            const uint32_t tmem_d = tmem_base + (uint32_t)(as * BNL);
            for (int kb = 0; kb < num_kb; kb++) {
                ln_wait(&full_bar[s], ph, 3);
                tc_fence_after();
                const uint32_t sa = smem_u32(ring + s * Cfg::STAGE_BYTES);
                const uint64_t adesc = umma_desc_sw128(sa);
                const uint64_t bdesc = umma_desc_sw128(sa + Cfg::A_BYTES);
#pragma unroll
                for (int k = 0; k < GEMM_BLOCK_K / 16; k++)
                    umma_f16_e<2>(tmem_d, adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), idesc, (kb | k) != 0);
                umma_commit_pair_mask_e(&empty_bar[s], pair_mask);
                if (kb == num_kb - 1) umma_commit_pair_mask_e(&tfull_bar[as], pair_mask);
                if (++s == STAGES) { s = 0; ph ^= 1; }
            }
            if (++as == 2) { as = 0; aph ^= 1; }
        }
    } else if (warp == 2) {
        // ================================================================= residual loader (TMA)
        uint32_t iter = 0;
        for (int t = cluster_id; t < m_tiles; t += n_clusters, iter++) {
            const int m0 = t * 2 * GEMM_BLOCK_M + rank * GEMM_BLOCK_M;
            for (int b = 0; b < NBOX; b++) {
                const uint32_t u = iter * NBOX + b, pb = u % NRING, k = u / NRING;  // k-th use of physical box pb
                if (k > 0) ln_wait(&rfree_bar[pb], (k - 1) & 1, 4);  // the stores of its previous use have read it
                mbar_expect_tx_e(&rfull_bar[pb], kLnBoxBytes);
                tma_load_2d_e(obuf + pb * kLnBoxBytes, &tmRes, &rfull_bar[pb], ncol0 + b * GEMM_OUT_BOX_N, m0);
            }
        }
    } else if (warp == 3) {
        // ================================================================= TMA-store issuer
        // a warp of its own, so that waiting for a store to finish reading its box (which releases the box to the
        // residual loader, or to the team for the yq copy) never stalls an epilogue warp
        const uint32_t n_pass = (p.qmode && !YQ) ? 2 : 1;
        uint32_t iter = 0;
        for (int t = cluster_id; t < m_tiles; t += n_clusters, iter++) {
            const int m0 = t * 2 * GEMM_BLOCK_M + rank * GEMM_BLOCK_M;
            if (YQ) {  // y from its (in-place) box and yq from the team's box, per column block
                for (int b = 0; b < NBOX; b++) {
                    const uint32_t u = iter * NBOX + b, pb = u % NRING, k = u / NRING;
                    const uint32_t team = (uint32_t)b & 1u;
                    const uint32_t uq = iter * (uint32_t)((NBOX + 1 - (int)team) / 2) + ((uint32_t)b >> 1);  // uses of the team's box so far
                    ln_wait(&ofull_bar[pb], k & 1, 10);
                    if (lane == 0) {
                        tma_store_2d(&tmY, obuf + pb * kLnBoxBytes, ncol0 + b * GEMM_OUT_BOX_N, m0);
                        bulk_commit_group();
                    }
                    ln_wait(&yqfull_bar[team], uq & 1, 11);
                    if (lane == 0) {
                        tma_store_2d(&tmYQ, yqbuf + team * kLnBoxBytes, ncol0 + b * GEMM_OUT_BOX_N, m0);
                        bulk_commit_group();
                        bulk_wait_group_read<0>();
                        mbar_arrive(&rfree_bar[pb]);
                        mbar_arrive(&yqread_bar[team]);
                    }
                    __syncwarp();
                }
                continue;
            }
            for (uint32_t pass = 0; pass < n_pass; pass++)
                for (int b = 0; b < NBOX; b++) {
                    const uint32_t u = iter * NBOX + b, pb = u % NRING, k = u / NRING;
                    ln_wait(&ofull_bar[pb], (k * n_pass + pass) & 1, 10);  // a box is staged n_pass times per use
                    if (lane == 0) {
                        tma_store_2d(pass == 0 ? &tmY : &tmYQ, obuf + pb * kLnBoxBytes, ncol0 + b * GEMM_OUT_BOX_N, m0);
                        bulk_commit_group();
                        bulk_wait_group_read<0>();
                        mbar_arrive(pass == n_pass - 1 ? &rfree_bar[pb] : &oread_bar[pb]);
                    }
                    __syncwarp();
                }
        }
        if (lane == 0) bulk_wait_group<0>();  // all stores complete before the CTA (and its smem) goes away
    } else if (warp >= kLnEpiWarp0) {
        // ========================================================================= epilogue
        const int ew = warp - kLnEpiWarp0;
        const int team = ew >> 3;         // boxes team, team+2
        const int quad = ew & 3;          // == warp % 4: TMEM lanes 32*quad .. 32*quad+31
        const int half = (ew >> 2) & 1;   // which 32 columns of the 64-column box
        const int grp = team * 2 + half;  // column group inside the CTA
        const int r = quad * 32 + lane;   // row inside this CTA's 128
        const uint32_t row_off = (uint32_t)r * 128u;
        const uint32_t sw = (uint32_t)(r & 7);  // 128-byte swizzle: 16-byte chunk index ^= row % 8
        const uint32_t tempty_lead0 = smem_u32(&tempty_bar[0]) & kPeerBitMask;
        constexpr float kN0 = 32.f * ((NBOX + 1) / 2), kN1 = 32.f * (NBOX / 2);  // columns a thread of team 0 / 1 owns
        const float n_mine = team == 0 ? kN0 : kN1;
        int as = 0; uint32_t aph = 0;
        uint32_t iter = 0;
        for (int t = cluster_id; t < m_tiles; t += n_clusters, iter++) {
            const int m0 = t * 2 * GEMM_BLOCK_M + rank * GEMM_BLOCK_M;
            const bool tracer = p.trace != nullptr && blockIdx.x == 0 && ew == 0 && lane == 0 && iter < 16;
            if (tracer) p.trace[iter * 8 + 0] = clock64();
            ln_wait(&tfull_bar[as], aph, 5);
            tc_fence_after();
            if (tracer) p.trace[iter * 8 + 1] = clock64();
            const uint32_t taddr = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(as * BNL);
            // ------------------------------------------ pass 1: x = acc + bias + resid -> TMEM; shifted sums
            float pivot = 0.f;
            float2 s1 = make_float2(0.f, 0.f), s2 = make_float2(0.f, 0.f);
#pragma unroll 1
            for (int b = team; b < NBOX; b += 2) {
                const int col = b * GEMM_OUT_BOX_N + half * 32;  // CTA-local column == TMEM column of the stage
                const uint32_t u = iter * NBOX + b, pb = u % NRING, k = u / NRING;
                const uint8_t *box = obuf + pb * kLnBoxBytes;
                uint32_t acc[32];
                tmem_ld_32x32(taddr + (uint32_t)col, acc);
                ln_wait(&rfull_bar[pb], k & 1, 6);
                tmem_ld_wait();
                const float4 *b4 = reinterpret_cast<const float4 *>(p.bias + ncol0 + col);
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    const uint4 u = *reinterpret_cast<const uint4 *>(box + row_off + ((((uint32_t)(half * 4 + j)) ^ sw) << 4));
                    const float4 ba = __ldg(b4 + 2 * j), bb = __ldg(b4 + 2 * j + 1);
                    const uint32_t uu[4] = {u.x, u.y, u.z, u.w};
                    const float2 bv[4] = {make_float2(ba.x, ba.y), make_float2(ba.z, ba.w), make_float2(bb.x, bb.y), make_float2(bb.z, bb.w)};
#pragma unroll
                    for (int i = 0; i < 4; i++) {
                        float2 x = make_float2(__uint_as_float(acc[8 * j + 2 * i]), __uint_as_float(acc[8 * j + 2 * i + 1]));
                        x = __fadd2_rn(__fadd2_rn(x, bv[i]), unpack_h2(uu[i]));  // (W.a + bias) + residual, bert.cpp:948-953
                        acc[8 * j + 2 * i] = __float_as_uint(x.x);
                        acc[8 * j + 2 * i + 1] = __float_as_uint(x.y);
                    }
                }
                tmem_st_32x32(taddr + (uint32_t)col, acc);
                // sums of (x - pivot) and (x - pivot)^2, pivot = the thread's first value: one pass, and no
                // cancellation beyond |mean - pivot| / sigma, which is O(1) for a value drawn from the row itself
                if (b == team) pivot = __uint_as_float(acc[0]);
                const float2 np2 = make_float2(-pivot, -pivot);
#pragma unroll
                for (int j = 0; j < 16; j++) {
                    const float2 dlt = __fadd2_rn(make_float2(__uint_as_float(acc[2 * j]), __uint_as_float(acc[2 * j + 1])), np2);
                    s1 = __fadd2_rn(s1, dlt);
                    s2 = __ffma2_rn(dlt, dlt, s2);
                }
            }
            tmem_st_wait();
            if (tracer) p.trace[iter * 8 + 2] = clock64();
            {
                const float S1 = s1.x + s1.y, S2 = s2.x + s2.y;
                part[grp * GEMM_BLOCK_M + r] = make_float2(pivot + S1 / n_mine, fmaxf(S2 - S1 * S1 / n_mine, 0.f));
            }
            // ------------------------------------------ row statistics: 4 groups -> CTA -> all pairs (Chan et al.)
            named_bar_sync(3, 512);
            uint64_t *stat = &stat_bar[iter & 1];
            float2 *rbuf = rcv + (size_t)(iter & 1) * NPAIRS * GEMM_BLOCK_M;
            if (grp == 0) {
                const float2 q0 = part[r], q1 = part[GEMM_BLOCK_M + r], q2 = part[2 * GEMM_BLOCK_M + r], q3 = part[3 * GEMM_BLOCK_M + r];
                const float mc = (kN0 * (q0.x + q1.x) + kN1 * (q2.x + q3.x)) * (1.0f / (float)BNL);
                const float d0 = q0.x - mc, d1 = q1.x - mc, d2 = q2.x - mc, d3 = q3.x - mc;
                const float m2c = (q0.y + q1.y) + (q2.y + q3.y) + kN0 * (d0 * d0 + d1 * d1) + kN1 * (d2 * d2 + d3 * d3);
                comb[r] = make_float2(mc, m2c);
                mbar_arrive_expect_tx_noelect(stat, 8u * (NPAIRS - 1));
                const uint32_t my_slot = smem_u32(&rbuf[pair * GEMM_BLOCK_M + r]), my_stat = smem_u32(stat);
#pragma unroll
                for (int pp = 0; pp < NPAIRS; pp++)
                    if (pp != pair) {  // the CTA of pair pp that holds the same rows
                        const uint32_t dst = (uint32_t)(2 * pp + rank);
                        st_async_f32x2(mapa_cluster(my_slot, dst), mc, m2c, mapa_cluster(my_stat, dst));
                    }
            }
            ln_wait(stat, (iter >> 1) & 1, 7);
            float mean, rstd;
            {
                float2 c[NPAIRS];
                float msum = 0.f, m2 = 0.f, dev = 0.f;
#pragma unroll
                for (int pp = 0; pp < NPAIRS; pp++) {
                    c[pp] = pp == pair ? comb[r] : rbuf[pp * GEMM_BLOCK_M + r];
                    msum += c[pp].x;
                    m2 += c[pp].y;
                }
                mean = msum * (1.0f / (float)NPAIRS);
#pragma unroll
                for (int pp = 0; pp < NPAIRS; pp++) { const float dm = c[pp].x - mean; dev = fmaf(dm, dm, dev); }
                m2 = fmaf((float)BNL, dev, m2);  // M2 = sum M2_i + n_i (mean_i - mean)^2, equal n_i
                rstd = 1.0f / sqrtf(m2 / (float)p.N + p.eps);  // ggml.c:8996-9000
            }
            if (tracer) p.trace[iter * 8 + 3] = clock64();
            // ------------------------------------------ pass 2: normalise, gamma, beta, store
            // y = ((x - mean) * rstd) * gamma + beta as two packed FMAs per element pair: x * rstd + (-mean * rstd), then
            // * gamma + beta.  The reference rounds after each of its four nodes (ggml.c:8996-9004, bert.cpp:958-963); the
            // fused form differs from that by at most an fp32 ulp, far below the f16 rounding of the stored result
            const float2 rs2 = make_float2(rstd, rstd), sh2 = make_float2(-mean * rstd, -mean * rstd);
            // Q4 models: sweep 0 stages y for every box of the team, sweep 1 re-reads x from TMEM, normalises again and
            // stages the Q8-rounded copy in the same boxes -- by then the y store of the box has normally been read, so
            // the team does not stall on it (keeping y in registers across the boxes would not fit in 96 registers)
            const int n_sweep = (p.qmode && !YQ) ? 2 : 1;
#pragma unroll 1
            for (int sweep = 0; sweep < n_sweep; sweep++) {
#pragma unroll 1
                for (int b = team; b < NBOX; b += 2) {
                    const int col = b * GEMM_OUT_BOX_N + half * 32;
                    const uint32_t u = iter * NBOX + b, pb = u % NRING, k = u / NRING;
                    uint8_t *box = obuf + pb * kLnBoxBytes;
                    uint32_t acc[32];
                    tmem_ld_32x32(taddr + (uint32_t)col, acc);
                    tmem_ld_wait();
                    if (sweep == n_sweep - 1 && b + 2 >= NBOX) {  // last TMEM read of this stage by this warp: hand the accumulator back
                        tc_fence_before();
                        __syncwarp();
                        if (lane == 0) mbar_arrive_cluster(tempty_lead0 + 8u * as);
                    }
                    const float4 *g4 = reinterpret_cast<const float4 *>(p.gamma + ncol0 + col);
                    const float4 *e4 = reinterpret_cast<const float4 *>(p.beta + ncol0 + col);
#pragma unroll
                    for (int j = 0; j < 8; j++) {
                        const float4 g = __ldg(g4 + j), be = __ldg(e4 + j);
                        float2 a = make_float2(__uint_as_float(acc[4 * j + 0]), __uint_as_float(acc[4 * j + 1]));
                        float2 c = make_float2(__uint_as_float(acc[4 * j + 2]), __uint_as_float(acc[4 * j + 3]));
                        a = __ffma2_rn(__ffma2_rn(a, rs2, sh2), make_float2(g.x, g.y), make_float2(be.x, be.y));
                        c = __ffma2_rn(__ffma2_rn(c, rs2, sh2), make_float2(g.z, g.w), make_float2(be.z, be.w));
                        acc[4 * j + 0] = __float_as_uint(a.x); acc[4 * j + 1] = __float_as_uint(a.y);
                        acc[4 * j + 2] = __float_as_uint(c.x); acc[4 * j + 3] = __float_as_uint(c.y);
                    }
                    if (sweep == 1) {
                        float amax = 0.f;  // of the fp32 LayerNorm output, as quantize_row_q8_0/1 see it (ggml.c:9493-9509)
#pragma unroll
                        for (int j = 0; j < 16; j++) amax = fmaxf(amax, fmaxf(fabsf(__uint_as_float(acc[2 * j])), fabsf(__uint_as_float(acc[2 * j + 1]))));
                        float id, dq;
                        q8_scales(amax, p.qmode, id, dq);
#pragma unroll
                        for (int j = 0; j < 16; j++) {
                            const float2 v = q8_round2(make_float2(__uint_as_float(acc[2 * j]), __uint_as_float(acc[2 * j + 1])), id, dq);
                            acc[2 * j] = __float_as_uint(v.x); acc[2 * j + 1] = __float_as_uint(v.y);
                        }
                        ln_wait(&oread_bar[pb], k & 1, 9);  // the y store has finished reading the box
                    }
                    // in place: this thread overwrites exactly the residual chunks it read in pass 1
#pragma unroll
                    for (int j = 0; j < 4; j++) {
                        uint4 u;
                        u.x = pack_h2(__uint_as_float(acc[8 * j + 0]), __uint_as_float(acc[8 * j + 1]));
                        u.y = pack_h2(__uint_as_float(acc[8 * j + 2]), __uint_as_float(acc[8 * j + 3]));
                        u.z = pack_h2(__uint_as_float(acc[8 * j + 4]), __uint_as_float(acc[8 * j + 5]));
                        u.w = pack_h2(__uint_as_float(acc[8 * j + 6]), __uint_as_float(acc[8 * j + 7]));
                        *reinterpret_cast<uint4 *>(box + row_off + ((((uint32_t)(half * 4 + j)) ^ sw) << 4)) = u;
                    }
                    fence_proxy_async_smem();  // my generic-proxy writes -> visible to the TMA store
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&ofull_bar[pb]);
                    if (YQ) {
                        // the Q8-rounded copy from the same registers (acc still holds the fp32 LayerNorm output) into the team's box
                        float amax = 0.f;  // as quantize_row_q8_0/1 see it (ggml.c:9493-9509)
#pragma unroll
                        for (int j = 0; j < 16; j++) amax = fmaxf(amax, fmaxf(fabsf(__uint_as_float(acc[2 * j])), fabsf(__uint_as_float(acc[2 * j + 1]))));
                        float id, dq;
                        q8_scales(amax, p.qmode, id, dq);
                        const uint32_t uq = iter * (uint32_t)((NBOX + 1 - team) / 2) + ((uint32_t)b >> 1);  // uses of the team's box so far
                        if (uq >= 1) ln_wait(&yqread_bar[team], (uq - 1) & 1, 12);  // the previous yq store has read it
                        uint8_t *qbox = yqbuf + team * kLnBoxBytes;
#pragma unroll
                        for (int j = 0; j < 4; j++) {
                            const float2 v0 = q8_round2(make_float2(__uint_as_float(acc[8 * j + 0]), __uint_as_float(acc[8 * j + 1])), id, dq);
                            const float2 v1 = q8_round2(make_float2(__uint_as_float(acc[8 * j + 2]), __uint_as_float(acc[8 * j + 3])), id, dq);
                            const float2 v2 = q8_round2(make_float2(__uint_as_float(acc[8 * j + 4]), __uint_as_float(acc[8 * j + 5])), id, dq);
                            const float2 v3 = q8_round2(make_float2(__uint_as_float(acc[8 * j + 6]), __uint_as_float(acc[8 * j + 7])), id, dq);
                            uint4 u4;
                            u4.x = pack_h2(v0.x, v0.y); u4.y = pack_h2(v1.x, v1.y); u4.z = pack_h2(v2.x, v2.y); u4.w = pack_h2(v3.x, v3.y);
                            *reinterpret_cast<uint4 *>(qbox + row_off + ((((uint32_t)(half * 4 + j)) ^ sw) << 4)) = u4;
                        }
                        fence_proxy_async_smem();
                        __syncwarp();
                        if (lane == 0) mbar_arrive(&yqfull_bar[team]);
                    }
                }
            }
            if (tracer) { p.trace[iter * 8 + 4] = clock64(); p.trace[iter * 8 + 5] = p.trace[iter * 8 + 4]; }
            if (++as == 2) { as = 0; aph ^= 1; }
        }
    }

    tc_fence_before();
    cluster_sync_all();  // peers may still be read by the leader's MMAs / written by st.async until here
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc<512, 2>(tmem_base);
    }
}

// ------------------------------------------------------------------------------ host side
bool gemm_ln_supported(int N, int K) { return (N == 384 || N == 768 || N == 1024) && K % 8 == 0 && K >= GEMM_BLOCK_K; }

template <int NPAIRS, int BNL, int YQ>
static cudaError_t launch_ln_cfg(const CUtensorMap &tmA, const CUtensorMap &tmB, const CUtensorMap &tmRes, const CUtensorMap &tmY,
                                 const CUtensorMap &tmYQ, const GemmLnParams &p, int num_sms, cudaStream_t st, int *clusters_out) {
    using Cfg = GemmLnCfg<NPAIRS, BNL, YQ>;
    constexpr int CSIZE = 2 * NPAIRS;
    static int max_clusters[64] = {0};  // per device: co-resident clusters (GPC boundaries can leave SMs unused)
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
    LaunchConfig lc(dim3(1), dim3(kLnThreads), Cfg::SMEM_BYTES, st, CSIZE);
    cudaLaunchConfig_t &cfg = lc.cfg;
    if (!max_clusters[dev]) {
        cudaError_t err = cudaFuncSetAttribute(gemm_ln_kernel<NPAIRS, BNL, YQ>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES);
        if (err != cudaSuccess) return err;
        cfg.gridDim = dim3((unsigned)(num_sms / CSIZE * CSIZE));
        int n = 0;
        err = cudaOccupancyMaxActiveClusters(&n, gemm_ln_kernel<NPAIRS, BNL, YQ>, &cfg);
        if (err != cudaSuccess) return err;
        if (n < 1) return cudaErrorLaunchOutOfResources;
        if (n > num_sms / CSIZE) n = num_sms / CSIZE;
        max_clusters[dev] = n;
        if (getenv("BERT_B200_DEBUG")) fprintf(stderr, "gemm_ln<%d,%d,%d>: %d co-resident clusters of %d CTAs, %d stages, %d B smem\n", NPAIRS, BNL, YQ, n, CSIZE, Cfg::STAGES, Cfg::SMEM_BYTES);
    }
    if (clusters_out) *clusters_out = max_clusters[dev];
    const int m_tiles = (p.M + 2 * GEMM_BLOCK_M - 1) / (2 * GEMM_BLOCK_M);
    if (m_tiles < 1) return cudaSuccess;
    const int clusters = m_tiles < max_clusters[dev] ? m_tiles : max_clusters[dev];
    cfg.gridDim = dim3((unsigned)(clusters * CSIZE));
    return cudaLaunchKernelEx(&cfg, gemm_ln_kernel<NPAIRS, BNL, YQ>, tmA, tmB, tmRes, tmY, tmYQ, p);
}

static long long *g_ln_trace = nullptr;
void gemm_ln_set_trace(long long *dev_buf) { g_ln_trace = dev_buf; }

cudaError_t launch_gemm_ln(const CUtensorMap &tmA, const WeightMaps &W, const CUtensorMap &tmRes, const CUtensorMap &tmY,
                           const CUtensorMap *tmYQ, const float *bias, const float *gamma, const float *beta, float eps, int qmode,
                           int M, int N, int K, int num_sms, cudaStream_t st) {
    if (!gemm_ln_supported(N, K) || W.N != N || W.K != K || (qmode && !tmYQ)) return cudaErrorInvalidValue;
    GemmLnParams p;
    p.M = M; p.N = N; p.K = K; p.bias = bias; p.gamma = gamma; p.beta = beta; p.eps = eps; p.qmode = qmode; p.trace = g_ln_trace;
    const CUtensorMap &tq = tmYQ ? *tmYQ : tmY;
    // WeightMaps::tm index: box heights 32/64/96/128/192/256 rows; a CTA loads BNL/2 rows of W per K block
    // Q4 models, short K (the attention-output projection): the epilogue is the bottleneck -> single-sweep variant
    static const int yq_max_k = [] { const char *v = getenv("BERT_B200_LN_YQ_K"); return v ? atoi(v) : 1024; }();
    if (qmode && K <= yq_max_k) {
        if (N == 384) return launch_ln_cfg<2, 192, 1>(tmA, W.tm[2], tmRes, tmY, tq, p, num_sms, st, nullptr);
        if (N == 768) return launch_ln_cfg<3, 256, 1>(tmA, W.tm[3], tmRes, tmY, tq, p, num_sms, st, nullptr);
        return launch_ln_cfg<4, 256, 1>(tmA, W.tm[3], tmRes, tmY, tq, p, num_sms, st, nullptr);
    }
    if (N == 384) return launch_ln_cfg<2, 192, 0>(tmA, W.tm[2], tmRes, tmY, tq, p, num_sms, st, nullptr);
    if (N == 768) return launch_ln_cfg<3, 256, 0>(tmA, W.tm[3], tmRes, tmY, tq, p, num_sms, st, nullptr);
    return launch_ln_cfg<4, 256, 0>(tmA, W.tm[3], tmRes, tmY, tq, p, num_sms, st, nullptr);
}

}  // namespace bk
