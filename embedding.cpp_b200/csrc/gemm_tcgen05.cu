// gemm_tcgen05.cu -- the projection GEMMs of the encoder (QKV, attention-output, FFN up/down) as one
// persistent, warp-specialised sm_100a kernel: TMA -> 128B-swizzled shared memory -> tcgen05.mma
// (kind::f16, fp32 accumulators in TMEM) -> fused epilogue (bias | bias+residual | bias+GELU
// [+ Q8 activation rounding]) -> swizzled shared-memory staging -> TMA store.
//
// Replaces ggml_compute_forward_mul_mat + the bias/residual/GELU graph nodes around it
// (reference ggml.c:9388-9597; bert.cpp:907-926, 945-953, 967-977).
//
//   out[M][N] = epi( A[M][K] . W[N][K]^T + bias[N] (+ resid[M][N]) )          A, W f16 with K contiguous
//
// CTA = 640 threads, one CTA per SM:
//   warp 0      TMA producer of the A / W operand ring (one lane)
//   warp 1      TMEM allocator + tcgen05.mma issuer (one lane)
//   warp 2      TMA loader of the residual tile into the output staging buffer (one lane; RESID epilogue only)
//   warp 3      idle
//   warps 4-19  epilogue, two teams of 8 warps; team g handles output boxes g, g+2, ...; inside a team warp w
//               owns TMEM lanes 32*(w%4).. (a hardware rule) and one 32-column half of the 64-column box
// Tiles are BLOCK_M=128 x BN (BN = 64/128/192/256 picked per problem to minimise wave quantisation);
// K is consumed 64 halves (one 128-byte swizzle atom) per ring stage, 4 UMMA_K=16 steps each.
// Two TMEM accumulator stages let the epilogue of tile i overlap the main loop of tile i+1.
// The output tile is staged in shared memory as BN/64 boxes of 128 rows x 128 bytes in the layout TMA's
// SWIZZLE_128B expects, so global stores are full 128-byte lines issued by the TMA engine and the
// residual arrives the same way (loaded into the box it will be added to, in place).  Box b of tile i+1
// is refilled as soon as the store of box b of tile i has finished reading it (sfree barriers), i.e. the
// staging buffer is pipelined at box granularity across tiles.
// Grid = min(#tiles, #SMs); tile t -> (m_blk = t / n_tiles, n_blk = t % n_tiles) so CTAs running
// concurrently share A row-blocks through L2 and the (small) weight stays L2 resident.
#include <atomic>
#include <cstdlib>

#include "common.cuh"
#include "gemm_parts.cuh"
#include "kernels.h"
#include "../../include/bk_api.h"

namespace bk {

constexpr int kGemmThreads = 640;
constexpr int kEpiWarp0 = 4;
constexpr int kTeamThreads = 256;
constexpr int kSmemLimit = 227 * 1024;
constexpr int kBoxBytes = GEMM_BLOCK_M * GEMM_OUT_BOX_N * 2;  // 16 KB

// CG = 1: one CTA per tile (tcgen05.mma.cta_group::1, M = 128).  CG = 2: a CTA pair (cluster of 2) per 256 x BN
// tile: each CTA holds its own 128 rows of A and BN/2 rows of W, the leader issues cta_group::2 MMAs that read
// both CTAs' shared memory, so every CTA streams (128 + BN/2) instead of (128 + BN) operand rows per K block.
// WQ = 0: W is an f16 [N][K] operand copy fetched by TMA.  WQ = 1 / 2: W stays in its GGUF Q4_0 / Q4_1 blocks in
// HBM (re-tiled at load so that the two 32-weight blocks of 64 consecutive K of BN consecutive rows are
// contiguous); the producer streams the packed bytes with 1-D bulk copies and eight decode warps expand them to
// f16 directly into the swizzled UMMA operand tile in shared memory (block layout ggml-quants.h:10-55, decode
// ggml-quants.c:1074-1113) -- only packed weight bytes cross HBM / L2.  One epilogue team instead of two.
template <int BN, int CG, int WQ>
struct GemmCfg {
    static_assert(WQ == 0 || CG == 1, "in-smem Q4 decode is built for one-CTA tiles");
    static constexpr int A_BYTES = GEMM_BLOCK_M * GEMM_BLOCK_K * 2;  // 16 KB
    static constexpr int B_ROWS = BN / CG;                           // W rows this CTA loads per K block
    static constexpr int B_BYTES = B_ROWS * GEMM_BLOCK_K * 2;
    static constexpr int PK_ROWB = WQ == 1 ? 36 : WQ == 2 ? 40 : 0;  // packed bytes of 64 weights of one row
    static constexpr int PK_BYTES = (B_ROWS * PK_ROWB + 1023) / 1024 * 1024;
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES + PK_BYTES;
    static constexpr int NTEAMS = WQ ? 1 : 2;                        // epilogue teams of 8 warps
    static constexpr int NBOX = BN / GEMM_OUT_BOX_N;
    static constexpr int OUT_BYTES = NBOX * kBoxBytes;
    static constexpr int BAR_BYTES = 512;
    static constexpr int STAGES_FIT = (kSmemLimit - 1024 /*align*/ - BAR_BYTES - OUT_BYTES) / STAGE_BYTES;
    static constexpr int STAGES = STAGES_FIT > 8 ? 8 : STAGES_FIT;
    static constexpr int TMEM_COLS = 2 * BN <= 128 ? 128 : 2 * BN <= 256 ? 256 : 512;  // power of two >= 32
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + OUT_BYTES + 1024 + BAR_BYTES;
    static_assert(BN % 64 == 0 && BN <= 256, "tile width");
    static_assert(STAGES >= 3, "operand ring too shallow");
    static_assert((3 * STAGES + 4 + 2 * NBOX) * 8 + 8 <= BAR_BYTES, "barrier area");
};

template <int BN, int CG, int WQ>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_tcgen05_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                    const __grid_constant__ CUtensorMap tmOut, const __grid_constant__ CUtensorMap tmRes, const GemmParams p) {
    pdl_launch_dependents();
    using Cfg = GemmCfg<BN, CG, WQ>;
    constexpr int STAGES = Cfg::STAGES;
    constexpr int NBOX = Cfg::NBOX;
    constexpr int NTEAMS = Cfg::NTEAMS;
    constexpr int TILE_M = GEMM_BLOCK_M * CG;  // rows of one (pair) tile
    extern __shared__ uint8_t smem_raw[];
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t *ring = smem;
    uint8_t *obuf = smem + STAGES * Cfg::STAGE_BYTES;  // NBOX boxes, each 1024-aligned
    uint64_t *full_bar = reinterpret_cast<uint64_t *>(obuf + Cfg::OUT_BYTES);
    uint64_t *empty_bar = full_bar + STAGES;
    uint64_t *tfull_bar = empty_bar + STAGES;   // [2] accumulator ready
    uint64_t *tempty_bar = tfull_bar + 2;       // [2] accumulator drained
    uint64_t *rfull_bar = tempty_bar + 2;       // [NBOX] residual box landed
    uint64_t *sfree_bar = rfull_bar + NBOX;     // [NBOX] store of this box has finished reading it
    uint64_t *pkfull_bar = sfree_bar + NBOX;    // [STAGES] packed weight bytes landed (WQ only)
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(pkfull_bar + STAGES);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const bool has_resid = p.epi == BK_EPI_BIAS_RESID;

    const int m_tiles = (p.M + TILE_M - 1) / TILE_M;
    const int n_tiles = (p.N + BN - 1) / BN;
    const int num_tiles = m_tiles * n_tiles;
    const int num_kb = (p.K + GEMM_BLOCK_K - 1) / GEMM_BLOCK_K;
    // tile walk: unit u of U (a CTA, or a CTA pair) takes tiles u, u+U, ...; rank = this CTA's half of the pair tile
    const int rank = CG == 2 ? (int)cluster_ctarank() : 0;
    const int unit = CG == 2 ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;
    const int units = CG == 2 ? (int)(gridDim.x >> 1) : (int)gridDim.x;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmA);
        tma_prefetch_desc(&tmB);
        tma_prefetch_desc(&tmOut);
        if (has_resid) tma_prefetch_desc(&tmRes);
        // full: one producer arrival per CTA (+ one per decode warp that writes rows of the W tile)
        for (int i = 0; i < STAGES; i++) { mbar_init(&full_bar[i], CG + (WQ ? BN / 32 : 0)); mbar_init(&empty_bar[i], 1); mbar_init(&pkfull_bar[i], 1); }
        for (int i = 0; i < 2; i++) { mbar_init(&tfull_bar[i], 1); mbar_init(&tempty_bar[i], CG * NTEAMS * 8); }
        for (int i = 0; i < NBOX; i++) { mbar_init(&rfull_bar[i], 1); mbar_init(&sfree_bar[i], 1); }
        mbar_fence_init();
    }
    if (warp == 1) tmem_alloc<Cfg::TMEM_COLS, CG>(tmem_slot);
    tc_fence_before();
    if (CG == 2) cluster_sync_all();  // the peer's barriers are initialised before anything signals them
    else __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    if (warp != 0) pdl_wait();  // everything above ran under the previous kernel's tail; from here on global memory is touched

    // Warps 0-2 run their loops with all 32 lanes in uniform control flow and issue through the *_e
    // ("elected lane") wrappers -- see common.cuh for why.
    if (warp == 0) {
        // ===================================================================== TMA producer (A, W)
        // The weights do not depend on the previous kernel: the W tiles of the first K blocks of this CTA's first tile
        // are requested BEFORE the grid dependency resolves (programmatic dependent launch, common.cuh), the activations
        // after it.  In the one-sentence regime a projection is mostly the time to stream its weights.
        const int n_pre = num_kb < STAGES ? num_kb : STAGES;  // stages filled ahead (the ring is empty: no wait)
        {
            const int n0 = (unit % n_tiles) * BN + rank * Cfg::B_ROWS;
            for (int kb = 0; kb < n_pre; kb++) {
                uint8_t *sb = ring + kb * Cfg::STAGE_BYTES + Cfg::A_BYTES;
                if (WQ) {
                    const int rows = p.N - n0 < BN ? p.N - n0 : BN;  // multiple of 32 -> bytes multiple of 16
                    const uint32_t bytes = (uint32_t)(rows * Cfg::PK_ROWB);
                    mbar_expect_tx_e(&pkfull_bar[kb], bytes);
                    bulk_load_1d_e(sb + Cfg::B_BYTES, p.wq + ((size_t)kb * p.N + n0) * Cfg::PK_ROWB, bytes, &pkfull_bar[kb]);
                } else if (CG == 1) {
                    mbar_expect_tx_e(&full_bar[kb], Cfg::A_BYTES + Cfg::B_BYTES);
                    tma_load_2d_e(sb, &tmB, &full_bar[kb], kb * GEMM_BLOCK_K, n0);
                } else {
                    const uint32_t lead_full = smem_u32(&full_bar[kb]) & kPeerBitMask;
                    mbar_arrive_expect_tx_cluster_e(lead_full, Cfg::A_BYTES + Cfg::B_BYTES);
                    tma_load_2d_pair_e(sb, &tmB, lead_full, kb * GEMM_BLOCK_K, n0);
                }
            }
        }
        pdl_wait();
        int s = 0; uint32_t ph = 0;
        for (int t = unit; t < num_tiles; t += units) {
            const int m0 = (t / n_tiles) * TILE_M + rank * GEMM_BLOCK_M, n0 = (t % n_tiles) * BN + rank * Cfg::B_ROWS;
            for (int kb = 0; kb < num_kb; kb++) {
                const bool w_ahead = t == unit && kb < n_pre;  // W (and the barrier's byte count) already under way
                mbar_wait(&empty_bar[s], ph ^ 1);
                uint8_t *sa = ring + s * Cfg::STAGE_BYTES, *sb = sa + Cfg::A_BYTES;
                if (WQ) {
                    mbar_expect_tx_e(&full_bar[s], Cfg::A_BYTES);
                    tma_load_2d_e(sa, &tmA, &full_bar[s], kb * GEMM_BLOCK_K, m0);
                    if (!w_ahead) {
                        const int rows = p.N - n0 < BN ? p.N - n0 : BN;  // multiple of 32 -> bytes multiple of 16
                        const uint32_t bytes = (uint32_t)(rows * Cfg::PK_ROWB);
                        mbar_expect_tx_e(&pkfull_bar[s], bytes);
                        bulk_load_1d_e(sb + Cfg::B_BYTES, p.wq + ((size_t)kb * p.N + n0) * Cfg::PK_ROWB, bytes, &pkfull_bar[s]);
                    }
                } else if (CG == 1) {
                    if (!w_ahead) mbar_expect_tx_e(&full_bar[s], Cfg::A_BYTES + Cfg::B_BYTES);
                    tma_load_2d_e(sa, &tmA, &full_bar[s], kb * GEMM_BLOCK_K, m0);
                    if (!w_ahead) tma_load_2d_e(sb, &tmB, &full_bar[s], kb * GEMM_BLOCK_K, n0);
                } else {
                    // both CTAs report to the leader's barrier: one arrival + this CTA's bytes each
                    const uint32_t lead_full = smem_u32(&full_bar[s]) & kPeerBitMask;
                    if (!w_ahead) mbar_arrive_expect_tx_cluster_e(lead_full, Cfg::A_BYTES + Cfg::B_BYTES);
                    tma_load_2d_pair_e(sa, &tmA, lead_full, kb * GEMM_BLOCK_K, m0);
                    if (!w_ahead) tma_load_2d_pair_e(sb, &tmB, lead_full, kb * GEMM_BLOCK_K, n0);
                }
                if (++s == STAGES) { s = 0; ph ^= 1; }
            }
        }
    } else if (warp == 1 && rank == 0) {
        // ======================================================================= MMA issuer (leader CTA)
        constexpr uint32_t idesc = umma_idesc_f16(TILE_M, BN);
        int s = 0; uint32_t ph = 0;
        int as = 0; uint32_t aph = 0;
        for (int t = unit; t < num_tiles; t += units) {
            mbar_wait(&tempty_bar[as], aph ^ 1);  // epilogue has drained this accumulator stage
            tc_fence_after();
            const uint32_t tmem_d = tmem_base + (uint32_t)(as * BN);
            for (int kb = 0; kb < num_kb; kb++) {
                mbar_wait(&full_bar[s], ph);
                tc_fence_after();
                const uint32_t sa = smem_u32(ring + s * Cfg::STAGE_BYTES);
                const uint64_t adesc = umma_desc_sw128(sa);
                const uint64_t bdesc = umma_desc_sw128(sa + Cfg::A_BYTES);
#pragma unroll
                for (int k = 0; k < GEMM_BLOCK_K / 16; k++) {
                    // +32 bytes per UMMA_K step inside the swizzle atom: +2 in the (addr >> 4) field
                    umma_f16_e<CG>(tmem_d, adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), idesc, (kb | k) != 0);
                }
                umma_commit_e<CG>(&empty_bar[s]);                        // smem stage reusable when these MMAs retire (both CTAs of a pair)
                if (kb == num_kb - 1) umma_commit_e<CG>(&tfull_bar[as]);  // accumulator complete
                if (++s == STAGES) { s = 0; ph ^= 1; }
            }
            if (++as == 2) { as = 0; aph ^= 1; }
        }
    } else if (warp == 2) {
        // ================================================================= residual loader (TMA)
        if (has_resid) {
            uint32_t bph = 0;  // bit b = parity of the next use of box b
            for (int t = unit; t < num_tiles; t += units) {
                const int m0 = (t / n_tiles) * TILE_M + rank * GEMM_BLOCK_M, n0 = (t % n_tiles) * BN;
                for (int b = 0; b < NBOX; b++) {
                    if (n0 + b * GEMM_OUT_BOX_N >= p.N) break;
                    mbar_wait(&sfree_bar[b], ((bph >> b) & 1) ^ 1);  // the previous store of this box has read it
                    mbar_expect_tx_e(&rfull_bar[b], kBoxBytes);
                    tma_load_2d_e(obuf + b * kBoxBytes, &tmRes, &rfull_bar[b], n0 + b * GEMM_OUT_BOX_N, m0);
                    bph ^= 1u << b;
                }
            }
        }
    } else if (WQ != 0 && warp >= kEpiWarp0 + 8) {
        // ============================================================= Q4 decode warps (8 warps in NGRP groups)
        // group g takes every NGRP-th K block, so NGRP stages are being decoded at any time; thread = one W row
        constexpr int WPT = BN / 32 < 8 ? BN / 32 : 8;  // warps needed per W tile (one row per thread)
        constexpr int NGRP = 8 / WPT;                    // tiles (K blocks) being decoded concurrently
        const int dw = warp - (kEpiWarp0 + 8);
        const int grp = dw / WPT;
        const int r = (dw % WPT) * 32 + lane;  // row of the W tile
        const uint32_t sw = (uint32_t)(r & 7);
        int s = 0; uint32_t ph = 0;
        uint32_t kcount = 0;
        for (int t = unit; t < num_tiles; t += units) {
            const int n0 = (t % n_tiles) * BN;
            for (int kb = 0; kb < num_kb; kb++, kcount++) {
                if ((int)(kcount % NGRP) == grp && r < BN) {  // warp-uniform (BN is a multiple of 32)
                    uint8_t *sb = ring + s * Cfg::STAGE_BYTES + Cfg::A_BYTES;
                    mbar_wait(&pkfull_bar[s], ph);
                    uint4 o[8];
                    if (n0 + r < p.N) q4_decode_row<WQ>(reinterpret_cast<const uint32_t *>(sb + Cfg::B_BYTES + r * Cfg::PK_ROWB), o);
                    else {
#pragma unroll
                        for (int j = 0; j < 8; j++) o[j] = make_uint4(0, 0, 0, 0);
                    }
#pragma unroll
                    for (int j = 0; j < 8; j++) *reinterpret_cast<uint4 *>(sb + r * 128 + ((((uint32_t)j) ^ sw) << 4)) = o[j];
                    fence_proxy_async_smem();  // generic-proxy writes -> visible to the MMA's operand fetch
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&full_bar[s]);
                }
                if (++s == STAGES) { s = 0; ph ^= 1; }
            }
        }
    } else if (warp >= kEpiWarp0 && warp < kEpiWarp0 + 8 * NTEAMS) {
        // ========================================================================= epilogue
        const int ew = warp - kEpiWarp0;
        const int team = ew >> 3;         // boxes team, team+2, ...
        const int quad = ew & 3;          // == warp % 4: TMEM lanes 32*quad .. 32*quad+31
        const int half = (ew >> 2) & 1;   // which 32 columns of the 64-column box
        const int r = quad * 32 + lane;   // row inside the tile
        const bool issuer = (ew & 7) == 0 && lane == 0;
        const bool defer = (NBOX - team + NTEAMS - 1) / NTEAMS >= 2;  // this team owns >= 2 boxes: free box k when box k+1 is stored
        const uint32_t row_off = (uint32_t)r * 128u;
        const uint32_t sw = (uint32_t)(r & 7);      // 128-byte swizzle: 16-byte chunk index ^= row % 8
        int as = 0; uint32_t aph = 0;
        uint32_t bph = 0;
        int prev_box = -1;  // issuer: box of the newest committed store group
        // accumulator hand-back goes to the MMA warp of the leader CTA (own CTA when CG == 1)
        const uint32_t tempty_addr0 = smem_u32(&tempty_bar[0]) & (CG == 2 ? kPeerBitMask : 0xFFFFFFFFu);
        for (int t = unit; t < num_tiles; t += units) {
            const int m0 = (t / n_tiles) * TILE_M + rank * GEMM_BLOCK_M, n0 = (t % n_tiles) * BN;
            mbar_wait(&tfull_bar[as], aph);
            tc_fence_after();
            const uint32_t taddr = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(as * BN);
            if (team >= NBOX) {  // nothing to read for this team (BN = 64)
                tc_fence_before();
                if (lane == 0) mbar_arrive_cluster(tempty_addr0 + 8u * as);
            }
#pragma unroll 1
            for (int b = team; b < NBOX; b += NTEAMS) {
                const int nb = n0 + b * GEMM_OUT_BOX_N;
                const bool box_live = nb < p.N;
                uint32_t acc[32];
                tmem_ld_32x32(taddr + (uint32_t)(b * GEMM_OUT_BOX_N + half * 32), acc);
                tmem_ld_wait();
                if (b + NTEAMS >= NBOX) {  // last read of this stage by this warp: hand the accumulator back to the MMA warp
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive_cluster(tempty_addr0 + 8u * as);
                }
                if (!box_live) continue;  // uniform over the CTA
                uint8_t *box = obuf + b * kBoxBytes;
                if (has_resid) mbar_wait(&rfull_bar[b], (bph >> b) & 1);           // residual landed in the box
                else mbar_wait(&sfree_bar[b], ((bph >> b) & 1) ^ 1);               // previous store has read the box
                bph ^= 1u << b;
                const int n = nb + half * 32;
                if (n < p.N) {
                    // all epilogue arithmetic on element PAIRS with the packed fp32 instructions (f32x2): the epilogue
                    // warps are issue-slot bound, and each lane of a packed op is the scalar IEEE op
                    float2 v[16];
                    epi_add_bias(acc, p.bias + n, v);
                    if (has_resid) {
#pragma unroll
                        for (int j = 0; j < 4; j++) {
                            const uint4 u = *reinterpret_cast<const uint4 *>(box + row_off + ((((uint32_t)(half * 4 + j)) ^ sw) << 4));
                            v[4 * j + 0] = __fadd2_rn(v[4 * j + 0], unpack_h2(u.x));
                            v[4 * j + 1] = __fadd2_rn(v[4 * j + 1], unpack_h2(u.y));
                            v[4 * j + 2] = __fadd2_rn(v[4 * j + 2], unpack_h2(u.z));
                            v[4 * j + 3] = __fadd2_rn(v[4 * j + 3], unpack_h2(u.w));
                        }
                    } else if (p.epi >= BK_EPI_BIAS_GELU) {
                        epi_gelu_q8(v, p.epi);
                    }
                    epi_stage_f16(v, box + row_off, half, sw);
                }
                fence_proxy_async_smem();              // my generic-proxy writes -> visible to the TMA store
                named_bar_sync(1 + team, kTeamThreads);
                if (issuer) {
                    tma_store_2d(&tmOut, box, nb, m0);
                    bulk_commit_group();
                    if (!defer) {
                        bulk_wait_group_read<0>();
                        mbar_arrive(&sfree_bar[b]);
                    } else {
                        // all groups but the newest have finished reading -> this team's previous box is free
                        if (prev_box >= 0) { bulk_wait_group_read<1>(); mbar_arrive(&sfree_bar[prev_box]); }
                        prev_box = b;
                    }
                }
            }
            if (++as == 2) { as = 0; aph ^= 1; }
        }
        if (issuer) bulk_wait_group<0>();  // all stores complete before the CTA (and its smem) goes away
    }

    tc_fence_before();
    if (CG == 2) cluster_sync_all();  // the peer may still be read by the leader's MMAs / signalled until here
    else __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc<Cfg::TMEM_COLS, CG>(tmem_base);
    }
}

// ------------------------------------------------------------------------------ tile shape
static const int kBnChoices[4] = {64, 128, 192, 256};
static const int kBRows[6] = {32, 64, 96, 128, 192, 256};  // WeightMaps::tm index <-> W rows per TMA box
static int brow_index(int rows) {
    for (int i = 0; i < 6; i++) if (kBRows[i] == rows) return i;
    return -1;
}

void gemm_pick_tile(int M, int N, int num_sms, int *bn_out, int *cg_out) {
    // cost ~ waves x BN (a tile's main loop and epilogue are both proportional to BN), divided by an
    // efficiency: wider tiles and CTA pairs stream fewer operand bytes per FLOP through L2 -> smem
    static const double eff[4] = {0.70, 0.90, 0.97, 1.00};
    int best_bn = 64, best_cg = 1;
    double best_cost = 1e300;
    for (int cg = 1; cg <= 2; cg++) {
        const long units = num_sms / cg;
        const long m_tiles = (M + GEMM_BLOCK_M * cg - 1) / (GEMM_BLOCK_M * cg);
        for (int i = 0; i < 4; i++) {
            const int bn = kBnChoices[i];
            if (bn != 64 && N % bn != 0) continue;
            const long tiles = m_tiles * ((N + bn - 1) / bn);
            if (cg == 2 && tiles < 2 * units) continue;  // small problems: spread over all SMs instead of pairing them
            const long waves = (tiles + units - 1) / units;
            const double cost = (double)waves * bn / (eff[i] * (cg == 2 ? 1.12 : 1.0));
            if (cost < best_cost - 1e-9) { best_cost = cost; best_bn = bn; best_cg = cg; }
        }
    }
    *bn_out = best_bn; *cg_out = best_cg;
}

template <int BN, int CG, int WQ = 0>
static cudaError_t launch_cfg(const CUtensorMap &tmA, const WeightMaps &W, const CUtensorMap &tmOut, const CUtensorMap &tmRes,
                              const GemmParams &p, int num_sms, cudaStream_t st) {
    using Cfg = GemmCfg<BN, CG, WQ>;
    static std::atomic<bool> attr_set[64];  // per kernel instance and device (the attribute is per device)
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64 || !attr_set[dev]) {
        cudaError_t err = cudaFuncSetAttribute(gemm_tcgen05_kernel<BN, CG, WQ>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES);
        if (err != cudaSuccess) return err;
        if (dev >= 0 && dev < 64) attr_set[dev] = true;
    }
    const int bi = brow_index(Cfg::B_ROWS);
    if (bi < 0) return cudaErrorInvalidValue;
    const int m_tiles = (p.M + GEMM_BLOCK_M * CG - 1) / (GEMM_BLOCK_M * CG), n_tiles = (p.N + BN - 1) / BN;
    int grid = m_tiles * n_tiles * CG;
    const int max_grid = num_sms / CG * CG;
    if (grid > max_grid) grid = max_grid;
    if (grid < 1) return cudaSuccess;
    LaunchConfig lc(dim3((unsigned)grid), dim3(kGemmThreads), Cfg::SMEM_BYTES, st, CG);
    return cudaLaunchKernelEx(&lc.cfg, gemm_tcgen05_kernel<BN, CG, WQ>, tmA, W.tm[bi], tmOut, tmRes, p);
}

static inline const CUtensorMap &tmR0(const CUtensorMap *tmRes, const CUtensorMap &tmOut) { return tmRes ? *tmRes : tmOut; }

cudaError_t launch_gemm(const CUtensorMap &tmA, const WeightMaps &W, const CUtensorMap &tmOut, const CUtensorMap *tmRes,
                        const GemmParams &p, int num_sms, cudaStream_t st, int force_bn, int force_cg) {
    if (p.N % 32 != 0 || p.K % 8 != 0 || W.N != p.N || W.K != p.K) return cudaErrorInvalidValue;
    if (p.epi == BK_EPI_BIAS_RESID && !tmRes) return cudaErrorInvalidValue;
    static const int env_bn = [] { const char *v = getenv("BERT_B200_GEMM_BN"); return v ? atoi(v) : 0; }();
    static const int env_cg = [] { const char *v = getenv("BERT_B200_GEMM_CG"); return v ? atoi(v) : 0; }();
    int bn, cg;
    gemm_pick_tile(p.M, p.N, num_sms, &bn, &cg);
    const int want_bn = force_bn ? force_bn : env_bn, want_cg = force_cg ? force_cg : env_cg;
    if (want_cg == 1 || want_cg == 2) cg = want_cg;
    if (p.wq) {
        // W stays packed (Q4_0 / Q4_1 blocks) and is decoded in shared memory: one-CTA tiles, 128 or 64 columns
        if (p.wq_type != BK_Q4_0 && p.wq_type != BK_Q4_1) return cudaErrorInvalidValue;
        // 128-column tiles unless that leaves SMs without a tile (small batches: more, narrower tiles)
        const long tiles128 = (long)((p.M + GEMM_BLOCK_M - 1) / GEMM_BLOCK_M) * (p.N / 128);
        const bool wide = p.N % 128 == 0 && want_bn != 64 && (want_bn == 128 || tiles128 >= num_sms);
        if (p.wq_type == BK_Q4_0) return wide ? launch_cfg<128, 1, 1>(tmA, W, tmOut, tmR0(tmRes, tmOut), p, num_sms, st) : launch_cfg<64, 1, 1>(tmA, W, tmOut, tmR0(tmRes, tmOut), p, num_sms, st);
        return wide ? launch_cfg<128, 1, 2>(tmA, W, tmOut, tmR0(tmRes, tmOut), p, num_sms, st) : launch_cfg<64, 1, 2>(tmA, W, tmOut, tmR0(tmRes, tmOut), p, num_sms, st);
    }
    if (want_bn == 64 || ((want_bn == 128 || want_bn == 192 || want_bn == 256) && p.N % want_bn == 0)) bn = want_bn;
    const CUtensorMap &tmR = tmRes ? *tmRes : tmOut;
#define BK_GEMM_CASE(BN_)                                                                                   \
    case BN_:                                                                                               \
        return cg == 2 ? launch_cfg<BN_, 2>(tmA, W, tmOut, tmR, p, num_sms, st) : launch_cfg<BN_, 1>(tmA, W, tmOut, tmR, p, num_sms, st)
    switch (bn) {
        BK_GEMM_CASE(256);
        BK_GEMM_CASE(192);
        BK_GEMM_CASE(128);
        default: return cg == 2 ? launch_cfg<64, 2>(tmA, W, tmOut, tmR, p, num_sms, st) : launch_cfg<64, 1>(tmA, W, tmOut, tmR, p, num_sms, st);
    }
#undef BK_GEMM_CASE
}

// ------------------------------------------------------------------------------ tensor maps
typedef CUresult (*PFN_encodeTiled)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                    const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static PFN_encodeTiled get_encode_fn() {
    static PFN_encodeTiled fn = nullptr;
    if (!fn) {
        void *ptr = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<PFN_encodeTiled>(ptr);
    }
    return fn;
}

bool make_tmap_f16_2d(CUtensorMap *tm, const void *base, uint64_t inner, uint64_t rows, uint64_t row_stride_bytes,
                      uint32_t box_inner, uint32_t box_rows, int swizzle_bytes) {
    PFN_encodeTiled enc = get_encode_fn();
    if (!enc) return false;
    cuuint64_t dims[2] = {inner, rows};
    cuuint64_t strides[1] = {row_stride_bytes};
    cuuint32_t box[2] = {box_inner, box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<void *>(base), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle_bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_128B,
                     CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS;
}

bool make_weight_maps(WeightMaps *w, const void *base, int N, int K) {
    w->N = N; w->K = K;
    for (int i = 0; i < 6; i++)
        if (!make_tmap_f16_2d(&w->tm[i], base, (uint64_t)K, (uint64_t)N, (uint64_t)K * 2, GEMM_BLOCK_K, (uint32_t)kBRows[i])) return false;
    return true;
}

// ------------------------------------------------------------------------------ Q4 weight re-tiling
// GGUF order [n][K/32 blocks] -> [K/64][n][2 blocks]: the 2 blocks (64 weights) of one row are bb2 = 36 / 40
// contiguous bytes, and consecutive rows of the same 64-wide K slice follow each other, so a W tile of any
// height is ONE contiguous 16-byte-aligned run for cp.async.bulk.  Same bytes, same count; done once at load.
// The source may be one of several matrices stacked along N (Q, K, V): its rows go to rows n_off.. of N_total.
__global__ void q4_retile_kernel(const uint8_t *__restrict__ src, uint8_t *__restrict__ dst, int n_rows, int K, int bb, int N_total, int n_off) {
    const int kb = blockIdx.y, n = blockIdx.x * blockDim.y + threadIdx.y;
    if (n >= n_rows) return;
    const int bb2 = 2 * bb;
    const uint8_t *s = src + ((size_t)n * (K / 32) + 2 * kb) * bb;
    uint8_t *d = dst + ((size_t)kb * N_total + n_off + n) * bb2;
    for (int i = threadIdx.x; i < bb2; i += blockDim.x) d[i] = s[i];
}
cudaError_t launch_q4_retile(const void *src, int wtype, int n_rows, int K, void *dst, int N_total, int n_off, cudaStream_t st) {
    if ((wtype != BK_Q4_0 && wtype != BK_Q4_1) || K % 64 != 0 || N_total % 32 != 0 || n_off % 32 != 0) return cudaErrorInvalidValue;
    dim3 block(8, 32), grid((n_rows + 31) / 32, K / 64);
    q4_retile_kernel<<<grid, block, 0, st>>>(reinterpret_cast<const uint8_t *>(src), reinterpret_cast<uint8_t *>(dst), n_rows, K,
                                             wtype == BK_Q4_0 ? 18 : 20, N_total, n_off);
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------ GELU alone (tests)
__global__ void gelu_test_kernel(const float *__restrict__ x, float *__restrict__ y, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) {  // the packed form the GEMM epilogue uses (both lanes), checked against the scalar form
        const float2 g = gelu_fast2_unrounded(make_float2(x[i], x[i]));
        const float a = round_f16(g.x), b = round_f16(g.y);
        y[i] = (a == b && a == gelu_fast(x[i])) ? a : __int_as_float(0x7fc00000);
    }
}
cudaError_t launch_gelu_test(const float *x, float *y, int n, cudaStream_t st) {
    gelu_test_kernel<<<(n + 255) / 256, 256, 0, st>>>(x, y, n);
    return cudaGetLastError();
}

}  // namespace bk
