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
// a 256-row block: NPAIRS CTA pairs side by side along N, each pair (cta_group::2) accumulating 256 rows x BNL
// columns (BNL = 384 or 512; per CTA 128 rows x BNL fp32 = the CTA's whole accumulator, single stage).  The
// epilogue then runs on the accumulator in place:
//   pass 1  x = acc + bias + resid, written back to TMEM (tcgen05.st); row sum        -> local mean
//   pass 2  sum (x - local mean)^2                                                     -> local M2
//           exchange (mean_i, M2_i) between the 4 column groups of a CTA (shared memory) and the CTAs of the
//           other pair that hold the same rows (st.async into their shared memory, completing transaction
//           bytes on their mbarrier); combine with Chan's parallel-variance formula
//   pass 3  y = ((x - mean) * rstd) * gamma + beta -> f16 -> swizzled staging box -> TMA store (and yq)
// The residual arrives by TMA into the same staging slots (a ring of SL boxes per epilogue team).
// Warps: 0 TMA producer (A, W), 1 MMA issuer (leader CTA of each pair), 2-3 residual loaders (one per epilogue
// team), 4-19 epilogue (two teams of 8 warps; team t owns the 64-column boxes t, t+2, ...; warp = 32 rows x 32 cols).
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

template <int NPAIRS, int BNL>
struct GemmLnCfg {
    static_assert(BNL == 384 || BNL == 512, "columns per CTA pair");
    static constexpr int HALF_N = BNL / 2;                       // UMMA N: two MMAs per K step (N <= 256)
    static constexpr int B_HALF_ROWS = BNL / 4;                  // W rows this CTA loads per UMMA half and K block
    static constexpr int A_BYTES = GEMM_BLOCK_M * GEMM_BLOCK_K * 2;
    static constexpr int B_HALF_BYTES = B_HALF_ROWS * GEMM_BLOCK_K * 2;
    static constexpr int STAGE_BYTES = A_BYTES + 2 * B_HALF_BYTES;
    static constexpr int NBOX = BNL / GEMM_OUT_BOX_N;            // 64-column boxes per CTA
    static constexpr int BPT = NBOX / 2;                         // boxes per epilogue team
    static constexpr int SL = 2;                                 // staging slots per team
    static constexpr int OUT_BYTES = 2 * SL * kLnBoxBytes;
    static constexpr int NPART = 4 * NPAIRS;                     // (mean, M2) partials per row
    static constexpr int PART_BYTES = 2 * NPART * GEMM_BLOCK_M * 8;  // double buffered by tile parity
    static constexpr int BAR_BYTES = 512;
    static constexpr int STAGES_FIT = (kLnSmemLimit - 1024 - BAR_BYTES - OUT_BYTES - PART_BYTES) / STAGE_BYTES;
    static constexpr int STAGES = STAGES_FIT > 6 ? 6 : STAGES_FIT;
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + OUT_BYTES + PART_BYTES + 1024 + BAR_BYTES;
    static constexpr int CW = BPT * 32;                          // columns one epilogue thread owns
    static_assert(STAGES >= 3, "operand ring too shallow");
    static_assert((2 * STAGES + 2 + 6 * SL + 1) * 8 + 8 <= BAR_BYTES, "barrier area");
    static_assert(B_HALF_BYTES % 1024 == 0, "W half tiles start on a swizzle atom");
};

struct GemmLnParams {
    int M, N, K;
    const float *bias, *gamma, *beta;
    float eps;
    int qmode;  // 0: y only; 1 / 2: also yq (Q8_0 / Q8_1 rounding)
    long long *trace;  // tuning aid: clock64 stamps of block 0's first epilogue thread, 8 per tile (null = off)
};

template <int NPAIRS, int BNL>
__global__ void __launch_bounds__(kLnThreads, 1)
gemm_ln_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
               const __grid_constant__ CUtensorMap tmRes, const __grid_constant__ CUtensorMap tmY,
               const __grid_constant__ CUtensorMap tmYQ, const GemmLnParams p) {
    using Cfg = GemmLnCfg<NPAIRS, BNL>;
    constexpr int STAGES = Cfg::STAGES;
    constexpr int SL = Cfg::SL;
    constexpr int BPT = Cfg::BPT;
    constexpr int NPART = Cfg::NPART;
    constexpr int CSIZE = 2 * NPAIRS;  // CTAs per cluster
    extern __shared__ uint8_t smem_raw[];
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t *ring = smem;
    uint8_t *obuf = smem + STAGES * Cfg::STAGE_BYTES;                       // [team][slot] boxes, 1024-aligned
    float2 *part = reinterpret_cast<float2 *>(obuf + Cfg::OUT_BYTES);       // [2][NPART][128]
    uint64_t *full_bar = reinterpret_cast<uint64_t *>(obuf + Cfg::OUT_BYTES + Cfg::PART_BYTES);
    uint64_t *empty_bar = full_bar + STAGES;
    uint64_t *tfull_bar = empty_bar + STAGES;    // accumulator complete
    uint64_t *tempty_bar = tfull_bar + 1;        // accumulator drained (leader CTA's is the one waited on)
    uint64_t *rfull_bar = tempty_bar + 1;        // [2][SL] residual box landed
    // A slot is used alternately for residual boxes (filled by TMA, read by the team) and output boxes (written by the
    // team, read by a TMA store); the end of one use signals the barrier of the NEXT use's kind, so that each barrier
    // has one waiter that observes every one of its phases (a parity wait cannot tell phase n from phase n+2)
    uint64_t *rfree_bar = rfull_bar + 2 * SL;    // [2][SL] slot free for the next residual load (8 arrivals)
    uint64_t *ofree_bar = rfree_bar + 2 * SL;    // [2][SL] slot free for the next output box (8 arrivals)
    uint64_t *stat_bar = ofree_bar + 2 * SL;     // row statistics of all column groups present
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(stat_bar + 1);

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
    const int items_per_tile = BPT * (p.qmode ? 3 : 2);  // staging-slot uses per team and tile: BPT residual + BPT (or 2 BPT) output boxes

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmA);
        tma_prefetch_desc(&tmB);
        tma_prefetch_desc(&tmRes);
        tma_prefetch_desc(&tmY);
        if (p.qmode) tma_prefetch_desc(&tmYQ);
        for (int i = 0; i < STAGES; i++) { mbar_init(&full_bar[i], 2); mbar_init(&empty_bar[i], 1); }
        mbar_init(tfull_bar, 1);
        mbar_init(tempty_bar, 2 * 16);
        for (int i = 0; i < 2 * SL; i++) { mbar_init(&rfull_bar[i], 1); mbar_init(&rfree_bar[i], 8); mbar_init(&ofree_bar[i], 8); }
        mbar_init(stat_bar, 512);
        mbar_fence_init();
    }
    if (warp == 1) tmem_alloc<512, 2>(tmem_slot);
    tc_fence_before();
    cluster_sync_all();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===================================================================== TMA producer (A, W halves)
        int s = 0; uint32_t ph = 0;
        for (int t = cluster_id; t < m_tiles; t += n_clusters) {
            const int m0 = t * 2 * GEMM_BLOCK_M + rank * GEMM_BLOCK_M;
            for (int kb = 0; kb < num_kb; kb++) {
                ln_wait(&empty_bar[s], ph ^ 1, 1);
                uint8_t *sa = ring + s * Cfg::STAGE_BYTES, *sb = sa + Cfg::A_BYTES;
                const uint32_t lead_full = smem_u32(&full_bar[s]) & kPeerBitMask;
                mbar_arrive_expect_tx_cluster_e(lead_full, Cfg::STAGE_BYTES);
                tma_load_2d_pair_e(sa, &tmA, lead_full, kb * GEMM_BLOCK_K, m0);
#pragma unroll
                for (int h = 0; h < 2; h++)  // UMMA half h covers columns ncol0 + h*HALF_N ..; this CTA supplies its B_HALF_ROWS of them
                    tma_load_2d_pair_e(sb + h * Cfg::B_HALF_BYTES, &tmB, lead_full, kb * GEMM_BLOCK_K,
                                       ncol0 + h * Cfg::HALF_N + rank * Cfg::B_HALF_ROWS);
                if (++s == STAGES) { s = 0; ph ^= 1; }
            }
        }
    } else if (warp == 1 && rank == 0) {
        // ======================================================================= MMA issuer (pair leader)
        constexpr uint32_t idesc = umma_idesc_f16(2 * GEMM_BLOCK_M, Cfg::HALF_N);
        const uint16_t pair_mask = (uint16_t)(3u << (2 * pair));
        int s = 0; uint32_t ph = 0;
        uint32_t aph = 0;
        for (int t = cluster_id; t < m_tiles; t += n_clusters) {
            ln_wait(tempty_bar, aph ^ 1, 2);  // both CTAs' epilogues have read the previous tile out of TMEM
            aph ^= 1;
            tc_fence_after();
            for (int kb = 0; kb < num_kb; kb++) {
                ln_wait(&full_bar[s], ph, 3);
                tc_fence_after();
                const uint32_t sa = smem_u32(ring + s * Cfg::STAGE_BYTES);
                const uint64_t adesc = umma_desc_sw128(sa);
#pragma unroll
                for (int h = 0; h < 2; h++) {
                    const uint64_t bdesc = umma_desc_sw128(sa + Cfg::A_BYTES + h * Cfg::B_HALF_BYTES);
#pragma unroll
                    for (int k = 0; k < GEMM_BLOCK_K / 16; k++)
                        umma_f16_e<2>(tmem_base + (uint32_t)(h * Cfg::HALF_N), adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), idesc, (kb | k) != 0);
                }
                umma_commit_pair_mask_e(&empty_bar[s], pair_mask);
                if (kb == num_kb - 1) umma_commit_pair_mask_e(tfull_bar, pair_mask);
                if (++s == STAGES) { s = 0; ph ^= 1; }
            }
        }
    } else if (warp == 2 || warp == 3) {
        // ================================================== residual loader of epilogue team (warp - 2)
        const int team = warp - 2;
        uint32_t q = 0;  // running staging-item index of this team (same numbering as the team's own)
        uint32_t fph = 0;  // bit s = parity of the next wait on rfree_bar[s]
        for (int t = cluster_id; t < m_tiles; t += n_clusters) {
            const int m0 = t * 2 * GEMM_BLOCK_M + rank * GEMM_BLOCK_M;
            for (int k = 0; k < BPT; k++, q++) {
                const uint32_t slot = q % SL;
                uint64_t *rf = &rfull_bar[team * SL + slot];
                if (q >= (uint32_t)SL) {  // not the slot's first use: its previous use signals rfree when it ends
                    ln_wait(&rfree_bar[team * SL + slot], (fph >> slot) & 1, 4);
                    fph ^= 1u << slot;
                }
                mbar_expect_tx_e(rf, kLnBoxBytes);
                tma_load_2d_e(obuf + (team * SL + slot) * kLnBoxBytes, &tmRes, rf, ncol0 + (team + 2 * k) * GEMM_OUT_BOX_N, m0);
            }
            q += (uint32_t)(items_per_tile - BPT);
        }
    } else if (warp >= kLnEpiWarp0) {
        // ========================================================================= epilogue
        const int ew = warp - kLnEpiWarp0;
        const int team = ew >> 3;         // boxes team, team+2, ...
        const int quad = ew & 3;          // == warp % 4: TMEM lanes 32*quad .. 32*quad+31
        const int half = (ew >> 2) & 1;   // which 32 columns of the 64-column box
        const int grp = team * 2 + half;  // column group inside the CTA
        const int r = quad * 32 + lane;   // row inside this CTA's 128
        const bool issuer = (ew & 7) == 0 && lane == 0;
        const uint32_t row_off = (uint32_t)r * 128u;
        const uint32_t sw = (uint32_t)(r & 7);  // 128-byte swizzle: 16-byte chunk index ^= row % 8
        const uint32_t taddr_row = tmem_base + ((uint32_t)(quad * 32) << 16);
        const uint32_t tempty_lead = smem_u32(tempty_bar) & kPeerBitMask;
        uint64_t *my_rfull = rfull_bar + team * SL, *my_rfree = rfree_bar + team * SL, *my_ofree = ofree_bar + team * SL;
        // barrier to signal when use number q of a slot ends: the one its next use (q + SL) waits on
        auto next_use_bar = [&](uint32_t qq) -> uint64_t * {
            const uint32_t slot = qq % SL;
            return ((qq + SL) % (uint32_t)items_per_tile) < (uint32_t)BPT ? &my_rfree[slot] : &my_ofree[slot];
        };
        uint8_t *my_boxes = obuf + team * SL * kLnBoxBytes;
        uint32_t q = 0;      // running staging-item index of this team
        uint32_t rph = 0;    // bit s = parity of the next residual landing in slot s
        uint32_t tph = 0, sph = 0;
        int pending = -1;    // issuer: item index of the newest committed TMA store
        uint32_t oph = 0;    // bit s = parity of the next wait on ofree[s]
        uint32_t iter = 0;
        for (int t = cluster_id; t < m_tiles; t += n_clusters, iter++) {
            const int m0 = t * 2 * GEMM_BLOCK_M + rank * GEMM_BLOCK_M;
            const bool tracer = p.trace != nullptr && blockIdx.x == 0 && ew == 0 && lane == 0 && iter < 16;
            if (tracer) p.trace[iter * 8 + 0] = clock64();
            ln_wait(tfull_bar, tph, 5);
            tph ^= 1;
            tc_fence_after();
            if (tracer) p.trace[iter * 8 + 1] = clock64();
            // ---------------------------------------------------------------- pass 1: x = acc + bias + resid, row sum
            float2 sum2 = make_float2(0.f, 0.f);
#pragma unroll 1
            for (int k = 0; k < BPT; k++, q++) {
                const int col = (team + 2 * k) * GEMM_OUT_BOX_N + half * 32;  // CTA-local column == TMEM column
                const uint32_t slot = q % SL;
                const uint8_t *box = my_boxes + slot * kLnBoxBytes;
                uint32_t acc[32];
                tmem_ld_32x32(taddr_row + (uint32_t)col, acc);
                ln_wait(&my_rfull[slot], (rph >> slot) & 1, 6);
                rph ^= 1u << slot;
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
                        sum2 = __fadd2_rn(sum2, x);
                        acc[8 * j + 2 * i] = __float_as_uint(x.x);
                        acc[8 * j + 2 * i + 1] = __float_as_uint(x.y);
                    }
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(next_use_bar(q));  // this warp is done with the residual box
                tmem_st_32x32(taddr_row + (uint32_t)col, acc);
            }
            tmem_st_wait();
            if (tracer) p.trace[iter * 8 + 2] = clock64();
            const float mean_loc = (sum2.x + sum2.y) * (1.0f / (float)Cfg::CW);
            // ---------------------------------------------------------------- pass 2: local sum of squared deviations
            float2 m22 = make_float2(0.f, 0.f);
            {
                const float2 nm = make_float2(-mean_loc, -mean_loc);
#pragma unroll 1
                for (int k = 0; k < BPT; k++) {
                    const int col = (team + 2 * k) * GEMM_OUT_BOX_N + half * 32;
                    uint32_t acc[32];
                    tmem_ld_32x32(taddr_row + (uint32_t)col, acc);
                    tmem_ld_wait();
#pragma unroll
                    for (int j = 0; j < 16; j++) {
                        const float2 dlt = __fadd2_rn(make_float2(__uint_as_float(acc[2 * j]), __uint_as_float(acc[2 * j + 1])), nm);
                        m22 = __ffma2_rn(dlt, dlt, m22);
                    }
                }
            }
            const float m2_loc = m22.x + m22.y;
            if (tracer) p.trace[iter * 8 + 3] = clock64();
            // ---------------------------------------------------------------- exchange (mean_i, M2_i) and combine
            float2 *pbuf = part + (size_t)(iter & 1) * NPART * GEMM_BLOCK_M;
            float2 *mine = pbuf + (pair * 4 + grp) * GEMM_BLOCK_M + r;
            *mine = make_float2(mean_loc, m2_loc);
            mbar_arrive_expect_tx_noelect(stat_bar, 8u * (NPAIRS - 1));
            if (NPAIRS == 2) {
                const uint32_t other = (uint32_t)(crank ^ 2);  // the CTA of the other pair that holds the same rows
                st_async_f32x2(mapa_cluster(smem_u32(mine), other), mean_loc, m2_loc, mapa_cluster(smem_u32(stat_bar), other));
            }
            ln_wait(stat_bar, sph, 7);
            sph ^= 1;
            if (tracer) p.trace[iter * 8 + 4] = clock64();
            float mean = 0.f;
            float2 pr[NPART];
#pragma unroll
            for (int i = 0; i < NPART; i++) { pr[i] = pbuf[i * GEMM_BLOCK_M + r]; mean += pr[i].x; }
            mean *= 1.0f / (float)NPART;
            float M2 = 0.f, dev = 0.f;
#pragma unroll
            for (int i = 0; i < NPART; i++) { const float dm = pr[i].x - mean; M2 += pr[i].y; dev = fmaf(dm, dm, dev); }
            M2 = fmaf((float)Cfg::CW, dev, M2);  // Chan et al.: M2 = sum M2_i + n_i (mean_i - mean)^2, equal n_i
            const float var = M2 / (float)p.N;
            const float rstd = 1.0f / sqrtf(var + p.eps);  // ggml.c:8996-9000
            // ---------------------------------------------------------------- pass 3: normalise, gamma, beta, store
            const float2 nmean = make_float2(-mean, -mean), rs2 = make_float2(rstd, rstd);
#pragma unroll 1
            for (int k = 0; k < BPT; k++) {
                const int col = (team + 2 * k) * GEMM_OUT_BOX_N + half * 32;
                const int nb = ncol0 + (team + 2 * k) * GEMM_OUT_BOX_N;  // first global column of the box
                uint32_t acc[32];
                tmem_ld_32x32(taddr_row + (uint32_t)col, acc);
                tmem_ld_wait();
                if (k == BPT - 1) {  // last TMEM read of this tile by this warp: hand the accumulator back
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive_cluster(tempty_lead);
                }
                const float4 *g4 = reinterpret_cast<const float4 *>(p.gamma + ncol0 + col);
                const float4 *e4 = reinterpret_cast<const float4 *>(p.beta + ncol0 + col);
                float amax = 0.f;
#pragma unroll
                for (int j = 0; j < 8; j++) {
                    const float4 g = __ldg(g4 + j), be = __ldg(e4 + j);
                    float2 a = make_float2(__uint_as_float(acc[4 * j + 0]), __uint_as_float(acc[4 * j + 1]));
                    float2 b = make_float2(__uint_as_float(acc[4 * j + 2]), __uint_as_float(acc[4 * j + 3]));
                    a = __fmul2_rn(__fadd2_rn(a, nmean), rs2);
                    b = __fmul2_rn(__fadd2_rn(b, nmean), rs2);
                    // three roundings as the reference's norm, ggml_mul, ggml_add nodes (scalar products: ptxas would
                    // contract a packed mul + add into FFMA2)
                    a = __fadd2_rn(make_float2(__fmul_rn(g.x, a.x), __fmul_rn(g.y, a.y)), make_float2(be.x, be.y));
                    b = __fadd2_rn(make_float2(__fmul_rn(g.z, b.x), __fmul_rn(g.w, b.y)), make_float2(be.z, be.w));
                    acc[4 * j + 0] = __float_as_uint(a.x); acc[4 * j + 1] = __float_as_uint(a.y);
                    acc[4 * j + 2] = __float_as_uint(b.x); acc[4 * j + 3] = __float_as_uint(b.y);
                    amax = fmaxf(amax, fmaxf(fmaxf(fabsf(a.x), fabsf(a.y)), fmaxf(fabsf(b.x), fabsf(b.y))));
                }
#pragma unroll 1
                for (int pass = 0; pass < 2; pass++) {  // pass 0: y, pass 1: the Q8-rounded copy (Q4 models)
                    if (pass == 1) {
                        if (!p.qmode) break;
                        float id, dq;
                        q8_scales(amax, p.qmode, id, dq);
#pragma unroll
                        for (int j = 0; j < 16; j++) {
                            const float2 v = q8_round2(make_float2(__uint_as_float(acc[2 * j]), __uint_as_float(acc[2 * j + 1])), id, dq);
                            acc[2 * j] = __float_as_uint(v.x); acc[2 * j + 1] = __float_as_uint(v.y);
                        }
                    }
                    const uint32_t slot = q % SL, qcur = q;
                    q++;
                    uint8_t *box = my_boxes + slot * kLnBoxBytes;
                    ln_wait(&my_ofree[slot], (oph >> slot) & 1, 8);  // the slot's previous use (a residual box or an older store) is over
                    oph ^= 1u << slot;
#pragma unroll
                    for (int j = 0; j < 4; j++) {
                        uint4 u;
                        u.x = pack_h2(__uint_as_float(acc[8 * j + 0]), __uint_as_float(acc[8 * j + 1]));
                        u.y = pack_h2(__uint_as_float(acc[8 * j + 2]), __uint_as_float(acc[8 * j + 3]));
                        u.z = pack_h2(__uint_as_float(acc[8 * j + 4]), __uint_as_float(acc[8 * j + 5]));
                        u.w = pack_h2(__uint_as_float(acc[8 * j + 6]), __uint_as_float(acc[8 * j + 7]));
                        *reinterpret_cast<uint4 *>(box + row_off + ((((uint32_t)(half * 4 + j)) ^ sw) << 4)) = u;
                    }
                    fence_proxy_async_smem();
                    named_bar_sync(1 + team, 256);
                    if (issuer) {
                        tma_store_2d(pass == 0 ? &tmY : &tmYQ, box, nb, m0);
                        bulk_commit_group();
                        if (pending >= 0) { bulk_wait_group_read<1>(); mbar_arrive_n(next_use_bar((uint32_t)pending), 8); }
                        pending = (int)qcur;
                    }
                }
            }
            if (tracer) p.trace[iter * 8 + 5] = clock64();
            if (issuer && pending >= 0) {  // the next tile's residual loads need every slot back
                bulk_wait_group_read<0>();
                mbar_arrive_n(next_use_bar((uint32_t)pending), 8);
                pending = -1;
            }
            if (tracer) p.trace[iter * 8 + 6] = clock64();
        }
        if (issuer) bulk_wait_group<0>();  // all stores complete before the CTA (and its smem) goes away
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

template <int NPAIRS, int BNL>
static cudaError_t launch_ln_cfg(const CUtensorMap &tmA, const CUtensorMap &tmB, const CUtensorMap &tmRes, const CUtensorMap &tmY,
                                 const CUtensorMap &tmYQ, const GemmLnParams &p, int num_sms, cudaStream_t st) {
    using Cfg = GemmLnCfg<NPAIRS, BNL>;
    constexpr int CSIZE = 2 * NPAIRS;
    static int max_clusters[64] = {0};  // per device: co-resident clusters (GPC boundaries can leave SMs unused)
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
    cudaLaunchConfig_t cfg = {};
    cfg.blockDim = dim3(kLnThreads);
    cfg.dynamicSmemBytes = Cfg::SMEM_BYTES;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = CSIZE; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    if (!max_clusters[dev]) {
        cudaError_t err = cudaFuncSetAttribute(gemm_ln_kernel<NPAIRS, BNL>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES);
        if (err != cudaSuccess) return err;
        cfg.gridDim = dim3((unsigned)(num_sms / CSIZE * CSIZE));
        int n = 0;
        err = cudaOccupancyMaxActiveClusters(&n, gemm_ln_kernel<NPAIRS, BNL>, &cfg);
        if (err != cudaSuccess) return err;
        if (n < 1) return cudaErrorLaunchOutOfResources;
        if (n > num_sms / CSIZE) n = num_sms / CSIZE;
        max_clusters[dev] = n;
        if (getenv("BERT_B200_DEBUG")) fprintf(stderr, "gemm_ln<%d,%d>: %d co-resident clusters of %d CTAs, %d stages, %d B smem\n", NPAIRS, BNL, n, CSIZE, Cfg::STAGES, Cfg::SMEM_BYTES);
    }
    const int m_tiles = (p.M + 2 * GEMM_BLOCK_M - 1) / (2 * GEMM_BLOCK_M);
    if (m_tiles < 1) return cudaSuccess;
    const int clusters = m_tiles < max_clusters[dev] ? m_tiles : max_clusters[dev];
    cfg.gridDim = dim3((unsigned)(clusters * CSIZE));
    return cudaLaunchKernelEx(&cfg, gemm_ln_kernel<NPAIRS, BNL>, tmA, tmB, tmRes, tmY, tmYQ, p);
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
    // WeightMaps::tm index: box heights 32/64/96/128/192/256 rows
    if (N == 384) return launch_ln_cfg<1, 384>(tmA, W.tm[2], tmRes, tmY, tq, p, num_sms, st);
    if (N == 768) return launch_ln_cfg<2, 384>(tmA, W.tm[2], tmRes, tmY, tq, p, num_sms, st);
    return launch_ln_cfg<2, 512>(tmA, W.tm[3], tmRes, tmY, tq, p, num_sms, st);
}

}  // namespace bk
