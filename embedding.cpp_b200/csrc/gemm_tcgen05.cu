// gemm_tcgen05.cu -- the projection GEMMs of the encoder (QKV, attention-output, FFN up/down) as one
// persistent, warp-specialised sm_100a kernel: TMA -> 128B-swizzled shared memory -> tcgen05.mma
// (kind::f16, fp32 accumulators in TMEM) -> fused epilogue (bias | bias+residual | bias+GELU
// [+ Q8 activation rounding]) -> f16 global stores.
//
// Replaces ggml_compute_forward_mul_mat + the bias/residual/GELU graph nodes around it
// (reference ggml.c:9388-9597; bert.cpp:907-926, 945-953, 967-977).
//
//   out[M][N] = epi( A[M][K] . W[N][K]^T + bias[N] )          A, W f16 with K contiguous
//
// CTA = 192 threads: warp 0 = TMA producer, warp 1 = TMEM allocator + MMA issuer (one lane),
// warps 2..5 = epilogue (each owns one 32-lane TMEM quadrant).  Tiles are BLOCK_M=128 x BN,
// K is consumed 64 halves (one 128-byte swizzle atom) per pipeline stage, 4 UMMA_K=16 steps each.
// Two TMEM accumulator stages (2 x BN columns) let the epilogue of tile i overlap the main loop
// of tile i+1.  Grid = min(#tiles, #SMs); tile t -> (m_blk = t / n_tiles, n_blk = t % n_tiles) so
// CTAs running concurrently share A row-blocks through L2 and the (small) weight stays L2 resident.
#include "common.cuh"
#include "kernels.h"
#include "../../include/bk_api.h"

namespace bk {

constexpr int kGemmThreads = 192;
constexpr int kSmemBudget = 192 * 1024;  // operand ring

template <int BN>
struct GemmCfg {
    static constexpr int A_BYTES = GEMM_BLOCK_M * GEMM_BLOCK_K * 2;  // 16 KB
    static constexpr int B_BYTES = BN * GEMM_BLOCK_K * 2;
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int STAGES = kSmemBudget / STAGE_BYTES;
    static constexpr int TMEM_COLS = 2 * BN;  // 128 / 256 / 512: power of two >= 32
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 /*align*/ + 256 /*barriers*/;
};

template <int BN>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_tcgen05_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const GemmParams p) {
    using Cfg = GemmCfg<BN>;
    constexpr int STAGES = Cfg::STAGES;
    extern __shared__ uint8_t smem_raw[];
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t *ring = smem;
    uint64_t *full_bar = reinterpret_cast<uint64_t *>(smem + STAGES * Cfg::STAGE_BYTES);
    uint64_t *empty_bar = full_bar + STAGES;
    uint64_t *tfull_bar = empty_bar + STAGES;   // [2] accumulator ready
    uint64_t *tempty_bar = tfull_bar + 2;       // [2] accumulator drained
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(tempty_bar + 2);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    const int m_tiles = (p.M + GEMM_BLOCK_M - 1) / GEMM_BLOCK_M;
    const int n_tiles = (p.N + BN - 1) / BN;
    const int num_tiles = m_tiles * n_tiles;
    const int num_kb = (p.K + GEMM_BLOCK_K - 1) / GEMM_BLOCK_K;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmA);
        tma_prefetch_desc(&tmB);
        for (int i = 0; i < STAGES; i++) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], 1); }
        for (int i = 0; i < 2; i++) { mbar_init(&tfull_bar[i], 1); mbar_init(&tempty_bar[i], 4); }
        mbar_fence_init();
    }
    if (warp == 1) tmem_alloc<Cfg::TMEM_COLS>(tmem_slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===================================================================== TMA producer
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
    } else {
        // ========================================================================= epilogue
        const int quad = warp & 3;  // TMEM lanes 32*quad .. 32*quad+31
        int as = 0; uint32_t aph = 0;
        for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
            const int m0 = (t / n_tiles) * GEMM_BLOCK_M, n0 = (t % n_tiles) * BN;
            const int row = m0 + quad * 32 + lane;
            mbar_wait(&tfull_bar[as], aph);
            tc_fence_after();
            const uint32_t taddr = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(as * BN);
#pragma unroll 1
            for (int c = 0; c < BN; c += 32) {
                uint32_t r[32];
                tmem_ld_32x32(taddr + (uint32_t)c, r);
                tmem_ld_wait();
                if (c + 32 >= BN) {  // last read of this stage: hand the accumulator back to the MMA warp
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&tempty_bar[as]);
                }
                const int n = n0 + c;
                if (row < p.M && n < p.N) {
                    float v[32];
                    const float4 *b4 = reinterpret_cast<const float4 *>(p.bias + n);
#pragma unroll
                    for (int j = 0; j < 8; j++) {
                        const float4 b = __ldg(b4 + j);
                        v[4 * j + 0] = __uint_as_float(r[4 * j + 0]) + b.x;
                        v[4 * j + 1] = __uint_as_float(r[4 * j + 1]) + b.y;
                        v[4 * j + 2] = __uint_as_float(r[4 * j + 2]) + b.z;
                        v[4 * j + 3] = __uint_as_float(r[4 * j + 3]) + b.w;
                    }
                    if (p.epi == BK_EPI_BIAS_RESID) {
                        const uint4 *r4 = reinterpret_cast<const uint4 *>(p.resid + (size_t)row * p.ldr + n);
#pragma unroll
                        for (int j = 0; j < 4; j++) {
                            const uint4 u = __ldg(r4 + j);
                            float2 f;
                            f = unpack_h2(u.x); v[8 * j + 0] += f.x; v[8 * j + 1] += f.y;
                            f = unpack_h2(u.y); v[8 * j + 2] += f.x; v[8 * j + 3] += f.y;
                            f = unpack_h2(u.z); v[8 * j + 4] += f.x; v[8 * j + 5] += f.y;
                            f = unpack_h2(u.w); v[8 * j + 6] += f.x; v[8 * j + 7] += f.y;
                        }
                    } else if (p.epi >= BK_EPI_BIAS_GELU) {
#pragma unroll
                        for (int j = 0; j < 32; j++) v[j] = gelu_ref(v[j]);
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
                    uint4 *o4 = reinterpret_cast<uint4 *>(p.out + (size_t)row * p.ldo + n);
#pragma unroll
                    for (int j = 0; j < 4; j++) {
                        uint4 u;
                        u.x = pack_h2(v[8 * j + 0], v[8 * j + 1]);
                        u.y = pack_h2(v[8 * j + 2], v[8 * j + 3]);
                        u.z = pack_h2(v[8 * j + 4], v[8 * j + 5]);
                        u.w = pack_h2(v[8 * j + 6], v[8 * j + 7]);
                        o4[j] = u;
                    }
                }
            }
            if (++as == 2) { as = 0; aph ^= 1; }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc<Cfg::TMEM_COLS>(tmem_base);
    }
}

int gemm_block_n(int N) {
    if (N % 256 == 0) return 256;
    if (N % 128 == 0) return 128;
    return 64;
}

template <int BN>
static cudaError_t launch_bn(const CUtensorMap &tmA, const CUtensorMap &tmB, const GemmParams &p, int num_sms, cudaStream_t st) {
    using Cfg = GemmCfg<BN>;
    static bool attr_set = false;  // per template instance; the attribute is per function and per device (set each time a device is first used below)
    cudaError_t err = cudaFuncSetAttribute(gemm_tcgen05_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES);
    if (err != cudaSuccess) return err;
    (void)attr_set;
    const int m_tiles = (p.M + GEMM_BLOCK_M - 1) / GEMM_BLOCK_M, n_tiles = (p.N + BN - 1) / BN;
    int grid = m_tiles * n_tiles;
    if (grid > num_sms) grid = num_sms;
    if (grid < 1) return cudaSuccess;
    gemm_tcgen05_kernel<BN><<<grid, kGemmThreads, Cfg::SMEM_BYTES, st>>>(tmA, tmB, p);
    return cudaGetLastError();
}

cudaError_t launch_gemm(const CUtensorMap &tmA, const CUtensorMap &tmB, const GemmParams &p, int num_sms, cudaStream_t st) {
    if (p.N % 32 != 0 || p.K % 8 != 0 || p.ldo % 8 != 0 || (p.resid && p.ldr % 8 != 0)) return cudaErrorInvalidValue;
    switch (gemm_block_n(p.N)) {
        case 256: return launch_bn<256>(tmA, tmB, p, num_sms, st);
        case 128: return launch_bn<128>(tmA, tmB, p, num_sms, st);
        default: return launch_bn<64>(tmA, tmB, p, num_sms, st);
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

}  // namespace bk
