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
#include <cstdlib>

#include "common.cuh"
#include "kernels.h"
#include "../../include/bk_api.h"

namespace bk {

constexpr int kGemmThreads = 640;
constexpr int kEpiWarp0 = 4;
constexpr int kEpiThreads = 512;   // two teams
constexpr int kTeamThreads = 256;
constexpr int kSmemLimit = 227 * 1024;
constexpr int kBoxBytes = GEMM_BLOCK_M * GEMM_OUT_BOX_N * 2;  // 16 KB

template <int BN>
struct GemmCfg {
    static constexpr int A_BYTES = GEMM_BLOCK_M * GEMM_BLOCK_K * 2;  // 16 KB
    static constexpr int B_BYTES = BN * GEMM_BLOCK_K * 2;
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int NBOX = BN / GEMM_OUT_BOX_N;
    static constexpr int OUT_BYTES = NBOX * kBoxBytes;
    static constexpr int BAR_BYTES = 512;
    static constexpr int STAGES_FIT = (kSmemLimit - 1024 /*align*/ - BAR_BYTES - OUT_BYTES) / STAGE_BYTES;
    static constexpr int STAGES = STAGES_FIT > 8 ? 8 : STAGES_FIT;
    static constexpr int TMEM_COLS = 2 * BN <= 128 ? 128 : 2 * BN <= 256 ? 256 : 512;  // power of two >= 32
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + OUT_BYTES + 1024 + BAR_BYTES;
    static_assert(BN % 64 == 0 && BN <= 256, "tile width");
    static_assert(STAGES >= 3, "operand ring too shallow");
    static_assert((2 * STAGES + 4 + 2 * NBOX) * 8 + 8 <= BAR_BYTES, "barrier area");
};

template <int BN>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_tcgen05_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                    const __grid_constant__ CUtensorMap tmOut, const __grid_constant__ CUtensorMap tmRes, const GemmParams p) {
    using Cfg = GemmCfg<BN>;
    constexpr int STAGES = Cfg::STAGES;
    constexpr int NBOX = Cfg::NBOX;
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
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(sfree_bar + NBOX);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const bool has_resid = p.epi == BK_EPI_BIAS_RESID;

    const int m_tiles = (p.M + GEMM_BLOCK_M - 1) / GEMM_BLOCK_M;
    const int n_tiles = (p.N + BN - 1) / BN;
    const int num_tiles = m_tiles * n_tiles;
    const int num_kb = (p.K + GEMM_BLOCK_K - 1) / GEMM_BLOCK_K;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmA);
        tma_prefetch_desc(&tmB);
        tma_prefetch_desc(&tmOut);
        if (has_resid) tma_prefetch_desc(&tmRes);
        for (int i = 0; i < STAGES; i++) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], 1); }
        for (int i = 0; i < 2; i++) { mbar_init(&tfull_bar[i], 1); mbar_init(&tempty_bar[i], kEpiThreads / 32); }
        for (int i = 0; i < NBOX; i++) { mbar_init(&rfull_bar[i], 1); mbar_init(&sfree_bar[i], 1); }
        mbar_fence_init();
    }
    if (warp == 1) tmem_alloc<Cfg::TMEM_COLS>(tmem_slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===================================================================== TMA producer (A, W)
        if (lane == 0) {
            int s = 0; uint32_t ph = 0;
            for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
                const int m0 = (t / n_tiles) * GEMM_BLOCK_M, n0 = (t % n_tiles) * BN;
                for (int kb = 0; kb < num_kb; kb++) {
                    mbar_wait(&empty_bar[s], ph ^ 1);
                    uint8_t *sa = ring + s * Cfg::STAGE_BYTES, *sb = sa + Cfg::A_BYTES;
                    mbar_expect_tx(&full_bar[s], Cfg::STAGE_BYTES);
                    tma_load_2d(sa, &tmA, &full_bar[s], kb * GEMM_BLOCK_K, m0);
                    tma_load_2d(sb, &tmB, &full_bar[s], kb * GEMM_BLOCK_K, n0);
                    if (++s == STAGES) { s = 0; ph ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ======================================================================= MMA issuer
        constexpr uint32_t idesc = umma_idesc_f16(GEMM_BLOCK_M, BN);
        int s = 0; uint32_t ph = 0;
        int as = 0; uint32_t aph = 0;
        for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
            mbar_wait(&tempty_bar[as], aph ^ 1);  // epilogue has drained this accumulator stage
            tc_fence_after();
            const uint32_t tmem_d = tmem_base + (uint32_t)(as * BN);
            for (int kb = 0; kb < num_kb; kb++) {
                mbar_wait(&full_bar[s], ph);
                tc_fence_after();
                if (lane == 0) {
                    const uint32_t sa = smem_u32(ring + s * Cfg::STAGE_BYTES);
                    const uint64_t adesc = umma_desc_sw128(sa);
                    const uint64_t bdesc = umma_desc_sw128(sa + Cfg::A_BYTES);
#pragma unroll
                    for (int k = 0; k < GEMM_BLOCK_K / 16; k++) {
                        // +32 bytes per UMMA_K step inside the swizzle atom: +2 in the (addr >> 4) field
                        umma_f16(tmem_d, adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), idesc, (kb | k) != 0);
                    }
                    umma_commit(&empty_bar[s]);                        // smem stage reusable when these MMAs retire
                    if (kb == num_kb - 1) umma_commit(&tfull_bar[as]);  // accumulator complete
                }
                __syncwarp();
                if (++s == STAGES) { s = 0; ph ^= 1; }
            }
            if (++as == 2) { as = 0; aph ^= 1; }
        }
    } else if (warp == 2) {
        // ================================================================= residual loader (TMA)
        if (has_resid && lane == 0) {
            uint32_t bph = 0;  // bit b = parity of the next use of box b
            for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
                const int m0 = (t / n_tiles) * GEMM_BLOCK_M, n0 = (t % n_tiles) * BN;
                for (int b = 0; b < NBOX; b++) {
                    if (n0 + b * GEMM_OUT_BOX_N >= p.N) break;
                    mbar_wait(&sfree_bar[b], ((bph >> b) & 1) ^ 1);  // the previous store of this box has read it
                    mbar_expect_tx(&rfull_bar[b], kBoxBytes);
                    tma_load_2d(obuf + b * kBoxBytes, &tmRes, &rfull_bar[b], n0 + b * GEMM_OUT_BOX_N, m0);
                    bph ^= 1u << b;
                }
            }
        }
    } else if (warp >= kEpiWarp0) {
        // ========================================================================= epilogue
        const int ew = warp - kEpiWarp0;
        const int team = ew >> 3;         // boxes team, team+2, ...
        const int quad = ew & 3;          // == warp % 4: TMEM lanes 32*quad .. 32*quad+31
        const int half = (ew >> 2) & 1;   // which 32 columns of the 64-column box
        const int r = quad * 32 + lane;   // row inside the tile
        const bool issuer = (ew & 7) == 0 && lane == 0;
        const bool defer = (NBOX - team + 1) / 2 >= 2;  // this team owns >= 2 boxes: free box k when box k+1 is stored
        const uint32_t row_off = (uint32_t)r * 128u;
        const uint32_t sw = (uint32_t)(r & 7);      // 128-byte swizzle: 16-byte chunk index ^= row % 8
        int as = 0; uint32_t aph = 0;
        uint32_t bph = 0;
        int prev_box = -1;  // issuer: box of the newest committed store group
        for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
            const int m0 = (t / n_tiles) * GEMM_BLOCK_M, n0 = (t % n_tiles) * BN;
            mbar_wait(&tfull_bar[as], aph);
            tc_fence_after();
            const uint32_t taddr = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(as * BN);
            if (team >= NBOX) {  // nothing to read for this team (BN = 64)
                tc_fence_before();
                if (lane == 0) mbar_arrive(&tempty_bar[as]);
            }
#pragma unroll 1
            for (int b = team; b < NBOX; b += 2) {
                const int nb = n0 + b * GEMM_OUT_BOX_N;
                const bool box_live = nb < p.N;
                uint32_t acc[32];
                tmem_ld_32x32(taddr + (uint32_t)(b * GEMM_OUT_BOX_N + half * 32), acc);
                tmem_ld_wait();
                if (b + 2 >= NBOX) {  // last read of this stage by this warp: hand the accumulator back to the MMA warp
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&tempty_bar[as]);
                }
                if (!box_live) continue;  // uniform over the CTA
                uint8_t *box = obuf + b * kBoxBytes;
                if (has_resid) mbar_wait(&rfull_bar[b], (bph >> b) & 1);           // residual landed in the box
                else mbar_wait(&sfree_bar[b], ((bph >> b) & 1) ^ 1);               // previous store has read the box
                bph ^= 1u << b;
                const int n = nb + half * 32;
                if (n < p.N) {
                    const float4 *b4 = reinterpret_cast<const float4 *>(p.bias + n);
                    float v[32];
#pragma unroll
                    for (int j = 0; j < 8; j++) {
                        const float4 bb = __ldg(b4 + j);
                        v[4 * j + 0] = __uint_as_float(acc[4 * j + 0]) + bb.x;
                        v[4 * j + 1] = __uint_as_float(acc[4 * j + 1]) + bb.y;
                        v[4 * j + 2] = __uint_as_float(acc[4 * j + 2]) + bb.z;
                        v[4 * j + 3] = __uint_as_float(acc[4 * j + 3]) + bb.w;
                    }
                    if (has_resid) {
#pragma unroll
                        for (int j = 0; j < 4; j++) {
                            const uint4 u = *reinterpret_cast<const uint4 *>(box + row_off + ((((uint32_t)(half * 4 + j)) ^ sw) << 4));
                            float2 f;
                            f = unpack_h2(u.x); v[8 * j + 0] += f.x; v[8 * j + 1] += f.y;
                            f = unpack_h2(u.y); v[8 * j + 2] += f.x; v[8 * j + 3] += f.y;
                            f = unpack_h2(u.z); v[8 * j + 4] += f.x; v[8 * j + 5] += f.y;
                            f = unpack_h2(u.w); v[8 * j + 6] += f.x; v[8 * j + 7] += f.y;
                        }
                    } else if (p.epi >= BK_EPI_BIAS_GELU) {
#pragma unroll
                        for (int j = 0; j < 32; j++) v[j] = gelu_fast(v[j]);
                        if (p.epi != BK_EPI_BIAS_GELU) {
                            // the next GEMM has Q4 weights: its input row is quantised per 32 (ggml.c:9493-9509)
                            float amax = 0.f;
#pragma unroll
                            for (int j = 0; j < 32; j++) amax = fmaxf(amax, fabsf(v[j]));
                            float id, dq;
                            q8_scales(amax, p.epi == BK_EPI_BIAS_GELU_Q8_0 ? 1 : 2, id, dq);
#pragma unroll
                            for (int j = 0; j < 32; j++) v[j] = q8_round(v[j], id, dq);
                        }
                    }
#pragma unroll
                    for (int j = 0; j < 4; j++) {
                        uint4 u;
                        u.x = pack_h2(v[8 * j + 0], v[8 * j + 1]);
                        u.y = pack_h2(v[8 * j + 2], v[8 * j + 3]);
                        u.z = pack_h2(v[8 * j + 4], v[8 * j + 5]);
                        u.w = pack_h2(v[8 * j + 6], v[8 * j + 7]);
                        *reinterpret_cast<uint4 *>(box + row_off + ((((uint32_t)(half * 4 + j)) ^ sw) << 4)) = u;
                    }
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
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc<Cfg::TMEM_COLS>(tmem_base);
    }
}

// ------------------------------------------------------------------------------ tile width
static const int kBnChoices[4] = {64, 128, 192, 256};

int gemm_pick_bn(int M, int N, int num_sms) {
    // cost ~ waves x BN (a tile's main loop and epilogue are both proportional to BN), divided by a
    // per-width efficiency (wider tiles re-read less of A per FLOP and amortise per-tile overheads)
    static const double eff[4] = {0.70, 0.90, 0.97, 1.00};
    const long m_tiles = (M + GEMM_BLOCK_M - 1) / GEMM_BLOCK_M;
    int best = 64;
    double best_cost = 1e300;
    for (int i = 0; i < 4; i++) {
        const int bn = kBnChoices[i];
        if (bn != 64 && N % bn != 0) continue;
        const long tiles = m_tiles * ((N + bn - 1) / bn);
        const long waves = (tiles + num_sms - 1) / num_sms;
        const double cost = (double)waves * bn / eff[i];
        if (cost < best_cost - 1e-9) { best_cost = cost; best = bn; }
    }
    return best;
}

template <int BN>
static cudaError_t launch_bn(const CUtensorMap &tmA, const CUtensorMap &tmB, const CUtensorMap &tmOut, const CUtensorMap &tmRes,
                             const GemmParams &p, int num_sms, cudaStream_t st) {
    using Cfg = GemmCfg<BN>;
    cudaError_t err = cudaFuncSetAttribute(gemm_tcgen05_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES);
    if (err != cudaSuccess) return err;
    const int m_tiles = (p.M + GEMM_BLOCK_M - 1) / GEMM_BLOCK_M, n_tiles = (p.N + BN - 1) / BN;
    int grid = m_tiles * n_tiles;
    if (grid > num_sms) grid = num_sms;
    if (grid < 1) return cudaSuccess;
    gemm_tcgen05_kernel<BN><<<grid, kGemmThreads, Cfg::SMEM_BYTES, st>>>(tmA, tmB, tmOut, tmRes, p);
    return cudaGetLastError();
}

cudaError_t launch_gemm(const CUtensorMap &tmA, const WeightMaps &W, const CUtensorMap &tmOut, const CUtensorMap *tmRes,
                        const GemmParams &p, int num_sms, cudaStream_t st, int force_bn) {
    if (p.N % 32 != 0 || p.K % 8 != 0 || W.N != p.N || W.K != p.K) return cudaErrorInvalidValue;
    if (p.epi == BK_EPI_BIAS_RESID && !tmRes) return cudaErrorInvalidValue;
    static const int env_bn = [] { const char *v = getenv("BERT_B200_GEMM_BN"); return v ? atoi(v) : 0; }();
    int bn = force_bn ? force_bn : env_bn;
    if (!(bn == 64 || ((bn == 128 || bn == 192 || bn == 256) && p.N % bn == 0))) bn = gemm_pick_bn(p.M, p.N, num_sms);
    const CUtensorMap &tmR = tmRes ? *tmRes : tmOut;
    switch (bn) {
        case 256: return launch_bn<256>(tmA, W.tm[3], tmOut, tmR, p, num_sms, st);
        case 192: return launch_bn<192>(tmA, W.tm[2], tmOut, tmR, p, num_sms, st);
        case 128: return launch_bn<128>(tmA, W.tm[1], tmOut, tmR, p, num_sms, st);
        default: return launch_bn<64>(tmA, W.tm[0], tmOut, tmR, p, num_sms, st);
    }
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
                      uint32_t box_inner, uint32_t box_rows) {
    PFN_encodeTiled enc = get_encode_fn();
    if (!enc) return false;
    cuuint64_t dims[2] = {inner, rows};
    cuuint64_t strides[1] = {row_stride_bytes};
    cuuint32_t box[2] = {box_inner, box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<void *>(base), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS;
}

bool make_weight_maps(WeightMaps *w, const void *base, int N, int K) {
    w->N = N; w->K = K;
    for (int i = 0; i < 4; i++)
        if (!make_tmap_f16_2d(&w->tm[i], base, (uint64_t)K, (uint64_t)N, (uint64_t)K * 2, GEMM_BLOCK_K, (uint32_t)kBnChoices[i])) return false;
    return true;
}

// ------------------------------------------------------------------------------ GELU alone (tests)
__global__ void gelu_test_kernel(const float *__restrict__ x, float *__restrict__ y, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) y[i] = gelu_fast(x[i]);
}
cudaError_t launch_gelu_test(const float *x, float *y, int n, cudaStream_t st) {
    gelu_test_kernel<<<(n + 255) / 256, 256, 0, st>>>(x, y, n);
    return cudaGetLastError();
}

}  // namespace bk
