// attention_tc.cu -- fused masked-softmax attention for variable-length packed batches on tcgen05.
//
// Replaces the per-sentence ggml nodes of bert.cpp:928-943 (mul_mat K.Q, scale, soft_max, transpose +
// cont V, mul_mat, permute, cpy).  One CTA = one (sentence, head, 128-query block):
//
//   TMA      Q[128 x dh], K[Lk x dh], V[Lk x dh] head slices of the packed qkv[T][3d] -> swizzled smem
//   tcgen05  S = Q.K^T   (M = 128 queries, N = Lk <= 512 keys, fp32) -> the whole score block lives in TMEM,
//            one query row per TMEM lane = per softmax thread
//   softmax  pass 1: row max over the sentence's keys (a padded / foreign key never enters: "masking" is the
//            key bound L_b of this sentence, identical to the reference's unpadded per-sentence softmax);
//            pass 2: p = f16(exp(f16(s*scale - max))) -- the reference's f16 exp table, ggml.c:10536-10558 --
//            written as the f16 A operand of the second MMA, 64 keys per chunk, double buffered
//   tcgen05  O += P_chunk . V_chunk  (V consumed MN-major straight from its [key][dh] layout), O in TMEM
//            columns [0, dh) which alias the already-consumed first score columns
//   tcgen05  rowsum += P_chunk . 1  (an all-ones 16-column B tile): sum p lands in TMEM next to O
//   epilogue O * (1 / sum p)  [+ Q8 rounding per 32 features for Q4 models, ggml.c:9493-9509] -> f16 rows
//
// Because the full score row is resident in TMEM, the true row maximum is known before any exponential
// is taken (the reference's rounding points: separate scale multiply, subtract, f16 round, exp, f16
// round) with a single QK^T pass -- no online-softmax rescaling, no second score computation.
// Warps 0 .. 4G-1 are softmax warps (warp w owns TMEM lanes 32*(w%4)..; group g = w/4 owns every G-th
// 32-key slice), warp 4G is the control warp (TMA, MMA issue, TMEM allocation).  G = 1/2/4 for
// Lk <= 128/256/512 so a CTA always has one softmax warp per 32 rows x 128 keys.
#include <cmath>
#include <atomic>
#include <cstdlib>

#include "common.cuh"
#include "kernels.h"

namespace bk {

constexpr int kAttQ = 128;  // queries per CTA
constexpr bool kAttXsDefault = false;    // attention_xs.cuh for launches with sentences > 128 tokens (A/B: BERT_B200_ATT_XS)
constexpr bool kAttPtDefault = true;     // attention_pt.cuh for launches with sentences > 128 tokens (A/B: BERT_B200_ATT_PT)
constexpr bool kAttDuoDefault = false;   // attention_duo.cuh for launches with sentences > 128 tokens (A/B: BERT_B200_ATT_DUO)
constexpr bool kAttLoopDefault = true;   // attention_loop.cuh for launches of sentences <= 128 tokens (A/B: BERT_B200_ATT_LOOP)

// K-major operand tile whose rows are ROWB = 64 or 128 bytes (dh = 32 / 64 halves): TMA SWIZZLE_64B / 128B
template <int ROWB>
__device__ __forceinline__ uint64_t umma_desc_kmajor(uint32_t saddr) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr >> 4) & 0x3FFF);
    d |= (uint64_t)1 << 16;                       // LBO: unused for swizzled K-major
    d |= (uint64_t)((8 * ROWB) >> 4) << 32;       // SBO: 8-row groups
    d |= (uint64_t)1 << 46;                       // descriptor version (Blackwell)
    d |= (uint64_t)(ROWB == 128 ? 2 : 4) << 61;   // SWIZZLE_128B / SWIZZLE_64B
    return d;
}
// MN-major operand (V: rows = keys = the MMA's K dimension, dh contiguous): canonical layout
// ((8,n),(8,k)) : ((1,LBO),(8,SBO)) in 16-byte units for 128-byte rows (cute/atom/mma_traits_sm100.hpp);
// one swizzle atom = 8 keys x ROWB bytes, consecutive atoms along K are 8*ROWB bytes apart (SBO); dh <= one atom (n = 1).
template <int ROWB>
__device__ __forceinline__ uint64_t umma_desc_mnmajor(uint32_t saddr) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr >> 4) & 0x3FFF);
    d |= (uint64_t)((8 * ROWB) >> 4) << 16;       // LBO: next atom along MN (not reached: N = dh fits one atom)
    d |= (uint64_t)((8 * ROWB) >> 4) << 32;       // SBO: next 8-key atom
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)(ROWB == 128 ? 2 : 4) << 61;
    return d;
}
// kind::f16 instruction descriptor, F16 x F16 -> F32, M = 128; b_mn = 1: B is MN-major
__device__ __forceinline__ uint32_t att_idesc(uint32_t N, uint32_t b_mn) {
    return (1u << 4) | (b_mn << 16) | ((N >> 3) << 17) | ((128u >> 4) << 24);
}

// one TMEM column: lane i of the warp reads TMEM lane (taddr.lane + i)
__device__ __forceinline__ uint32_t tmem_ld_32x1(uint32_t taddr) {
    uint32_t v;
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x1.b32 {%0}, [%1];" : "=r"(v) : "r"(taddr) : "memory");
    return v;
}
// K-major operand without swizzle, used only for the all-ones tile of the row-sum MMA: every byte the MMA can
// address from it (LBO / SBO inside the 512-byte tile) is the same f16 1.0, so the core-matrix order is immaterial
__device__ __forceinline__ uint64_t umma_desc_ones(uint32_t saddr) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr >> 4) & 0x3FFF);
    d |= (uint64_t)(128 >> 4) << 16;
    d |= (uint64_t)(256 >> 4) << 32;
    d |= (uint64_t)1 << 46;
    return d;
}
constexpr int kSumCols = 16;  // TMEM columns [dh, dh + 16) take the row sums (every one of the 16 holds the same value)

template <int DH, int G>
__global__ void __launch_bounds__(32 + 128 * G) attention_tc_kernel(const __grid_constant__ CUtensorMap tmQKV, const int32_t *__restrict__ cu,
                                                                    int d, __half *__restrict__ out, int qmode, float scale, int multi) {
    constexpr int ROWB = DH * 2;           // bytes per smem row of Q / K / V
    constexpr int CW = 4 * G;              // control warp
    constexpr int NSOFT = 128 * G;         // softmax threads
    constexpr int PBUF = kAttQ * 128;      // one P chunk: 128 rows x 64 keys f16
    extern __shared__ __align__(1024) uint8_t att_raw[];  // swizzle atoms need 1024-byte alignment; no static shared memory in this kernel
    uint8_t *smem = att_raw;
    pdl_launch_dependents();
    pdl_wait();  // the sequence bounds are read first thing

    const int seq = blockIdx.z, head = blockIdx.y;
    const int t0 = __ldg(cu + seq);
    const int L = __ldg(cu + seq + 1) - t0;
    if (L <= 0) return;
    const int nqb = (L + kAttQ - 1) / kAttQ;  // 128-query blocks of this sentence, processed one after the other by this CTA
    const int Lk16 = (L + 15) & ~15;       // keys entering the MMAs
    const int Lk64 = (L + 63) & ~63;       // keys staged in smem (whole 64-row TMA boxes)
    const int n_chunks = Lk64 >> 6;
    const int n_slices = (L + 31) >> 5;    // 32-key slices holding at least one real key

    // multi == 0 (launches with max_len <= 128: one query block per sentence): region 0 = Q | K while S is computed, then
    // the two P chunk buffers; region 1 = V.  48 KB for dh = 64, four CTAs per SM.
    // multi == 1: K and V of the (sentence, head) stay resident for all its query blocks, Q is double buffered (block
    // i+1 is fetched while block i is in the softmax) and P has buffers of its own: Q0 | Q1 | P0 | P1 | K | V.
    const int qk_bytes = (kAttQ + Lk64) * ROWB;
    const int r0_bytes = qk_bytes > 2 * PBUF ? qk_bytes : 2 * PBUF;
    uint8_t *sQ = smem;
    uint8_t *sP = multi ? smem + 2 * kAttQ * ROWB : smem;
    const uint32_t npb = multi ? 4u : 2u;  // P chunk buffers: with four, a softmax group never waits for the MMA of its own previous chunk
    uint8_t *sK = multi ? sP + 4 * PBUF : smem + kAttQ * ROWB;
    uint8_t *sV = multi ? sK + Lk64 * ROWB : smem + r0_bytes;
    uint8_t *tail = sV + Lk64 * ROWB;
    uint64_t *bar_q = reinterpret_cast<uint64_t *>(tail);  // [2] Q buffer landed
    uint64_t *bar_k = bar_q + 2, *bar_v = bar_q + 3, *bar_s = bar_q + 4, *bar_o = bar_q + 5, *pfull = bar_q + 6, *pfree = bar_q + 10;
    uint64_t *bar_tfree = bar_q + 14;                      // every softmax warp has read O: TMEM may take the next score block
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bar_q + 15);
    float *red = reinterpret_cast<float *>(bar_q + 16);    // [G][128] partial row max
    uint32_t *ones = reinterpret_cast<uint32_t *>(red + G * 128);  // 16 x 16 f16 1.0: B operand of the row-sum MMA (512 B)

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp == CW) {
        if (lane == 0) {
            tma_prefetch_desc(&tmQKV);
            mbar_init(&bar_q[0], 1); mbar_init(&bar_q[1], 1); mbar_init(bar_k, 1); mbar_init(bar_v, 1); mbar_init(bar_s, 1); mbar_init(bar_o, 1);
            mbar_init(bar_tfree, 4 * G);
            for (int i = 0; i < 4; i++) { mbar_init(&pfull[i], 4 * (G >= 2 ? 2 : 1)); mbar_init(&pfree[i], 1); }
            mbar_fence_init();
        }
        __syncwarp();
        {
            // the loads first: they need the barriers only, and their latency then runs under the TMEM allocation and the CTA-wide
            // synchronisation below
            const int cq = head * DH, ck = d + head * DH, cv = 2 * d + head * DH;
            mbar_expect_tx_e(&bar_q[0], (uint32_t)(kAttQ * ROWB));
            tma_load_2d_e(sQ, &tmQKV, &bar_q[0], cq, t0);
            tma_load_2d_e(sQ + 64 * ROWB, &tmQKV, &bar_q[0], cq, t0 + 64);
            mbar_expect_tx_e(bar_k, (uint32_t)(Lk64 * ROWB));
            for (int c = 0; c < n_chunks; c++) tma_load_2d_e(sK + c * 64 * ROWB, &tmQKV, bar_k, ck, t0 + c * 64);
            mbar_expect_tx_e(bar_v, (uint32_t)(Lk64 * ROWB));
            for (int c = 0; c < n_chunks; c++) tma_load_2d_e(sV + c * 64 * ROWB, &tmQKV, bar_v, cv, t0 + c * 64);
        }
        for (int i = lane; i < 128; i += 32) ones[i] = 0x3C003C00u;
        fence_proxy_async_smem();
        __syncwarp();
        // S needs Lk16 columns; O (dh columns) and the row sums alias columns [0, dh + 16).  Power of two >= 32.
        const int need = Lk16 > DH + kSumCols ? Lk16 : DH + kSumCols;
        if (need <= 32) tmem_alloc<32>(tmem_slot);
        else if (need <= 64) tmem_alloc<64>(tmem_slot);
        else if (need <= 128) tmem_alloc<128>(tmem_slot);
        else if (need <= 256) tmem_alloc<256>(tmem_slot);
        else tmem_alloc<512>(tmem_slot);
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == CW) {
        // ================================================================ control warp (uniform, elected issue)
        const int cq = head * DH;  // (Q, K, V of the first block were requested in the prologue)

        const uint32_t idesc_pv = att_idesc(DH, 1);
        // S = Q . K^T for key columns [n_lo, n_hi): instructions of <= 256 columns
        auto issue_s = [&](int qb, int n_lo, int n_hi) {
            const uint64_t dq = umma_desc_kmajor<ROWB>(smem_u32(sQ + (qb & 1) * kAttQ * ROWB));
            for (int n0 = n_lo; n0 < n_hi; n0 += 256) {
                const int nn = n_hi - n0 < 256 ? n_hi - n0 : 256;
                const uint64_t dk = umma_desc_kmajor<ROWB>(smem_u32(sK + n0 * ROWB));
                const uint32_t idesc = att_idesc((uint32_t)nn, 0);
#pragma unroll
                for (int k = 0; k < DH / 16; k++)
                    umma_f16_e<1>(tmem_base + (uint32_t)n0, dq + (uint64_t)(2 * k), dk + (uint64_t)(2 * k), idesc, k != 0);
            }
        };
        const int n_split = Lk16 < DH + kSumCols ? Lk16 : DH + kSumCols;  // score columns [0, n_split) are the ones O and the row sums alias
        const uint32_t idesc_sum = att_idesc(kSumCols, 0);
        const uint64_t done = umma_desc_ones(smem_u32(ones));
        // the row sums land in the score columns of chunk kLag, so the row-sum MMAs of chunk c are issued after chunk
        // c + kLag has been consumed by the softmax warps (its P delivered)
        constexpr int kLag = DH / 64;
        mbar_wait(&bar_q[0], 0);
        mbar_wait(bar_k, 0);
        tc_fence_after();
        issue_s(0, 0, Lk16);
        umma_commit_e<1>(bar_s);
        if (nqb > 1) {  // fetch the second query block while the first is computed
            uint8_t *qn = sQ + kAttQ * ROWB;
            mbar_expect_tx_e(&bar_q[1], (uint32_t)(kAttQ * ROWB));
            tma_load_2d_e(qn, &tmQKV, &bar_q[1], cq, t0 + kAttQ);
            tma_load_2d_e(qn + 64 * ROWB, &tmQKV, &bar_q[1], cq, t0 + kAttQ + 64);
        }
        mbar_wait(bar_v, 0);
        uint32_t cg = 0;  // P chunks issued so far over all query blocks (buffer = cg % npb)
        for (int qb = 0; qb < nqb; qb++) {
            // ---- O += P_c . V_c and rowsum += P_c . 1 as the softmax warps deliver P chunks
            auto issue_sum = [&](int c, uint32_t cgc) {  // row sums of chunk c (global number cgc), then its P buffer is free
                const uint32_t pbi = cgc % npb;
                const uint64_t dp = umma_desc_sw128(smem_u32(sP + pbi * PBUF));
                const int rem = Lk16 - c * 64;
                const int ksteps = rem >= 64 ? 4 : rem >> 4;
                for (int ks = 0; ks < ksteps; ks++)
                    umma_f16_e<1>(tmem_base + (uint32_t)DH, dp + (uint64_t)(2 * ks), done, idesc_sum, (c | ks) != 0);
                umma_commit_e<1>(&pfree[pbi]);
            };
            for (int c = 0; c < n_chunks; c++, cg++) {
                const uint32_t pbi = cg % npb;
                mbar_wait(&pfull[pbi], (cg / npb) & 1);
                tc_fence_after();
                const uint64_t dp = umma_desc_sw128(smem_u32(sP + pbi * PBUF));
                const int rem = Lk16 - c * 64;
                const int ksteps = rem >= 64 ? 4 : rem >> 4;
                for (int ks = 0; ks < ksteps; ks++) {
                    const uint64_t dv = umma_desc_mnmajor<ROWB>(smem_u32(sV + (c * 64 + ks * 16) * ROWB));
                    umma_f16_e<1>(tmem_base, dp + (uint64_t)(2 * ks), dv, idesc_pv, (c | ks) != 0);
                }
                if (c >= kLag) issue_sum(c - kLag, cg - kLag);
            }
            for (int c = n_chunks > kLag ? n_chunks - kLag : 0; c < n_chunks; c++) issue_sum(c, cg - (uint32_t)(n_chunks - c));
            umma_commit_e<1>(bar_o);
            if (qb + 1 < nqb) {
                // every P chunk of this block has been delivered, so every softmax read of its scores is over: the
                // next block's scores can already go to the columns O does not alias, behind the P.V MMAs in the pipe
                mbar_wait(&bar_q[(qb + 1) & 1], (uint32_t)(((qb + 1) >> 1) & 1));
                tc_fence_after();
                issue_s(qb + 1, n_split, Lk16);
                if (qb + 2 < nqb) {  // Q buffer of block qb is free (its S MMAs retired long ago): fetch block qb+2
                    uint8_t *qn = sQ + (qb & 1) * kAttQ * ROWB;
                    mbar_expect_tx_e(&bar_q[qb & 1], (uint32_t)(kAttQ * ROWB));
                    tma_load_2d_e(qn, &tmQKV, &bar_q[qb & 1], cq, t0 + (qb + 2) * kAttQ);
                    tma_load_2d_e(qn + 64 * ROWB, &tmQKV, &bar_q[qb & 1], cq, t0 + (qb + 2) * kAttQ + 64);
                }
                mbar_wait(bar_tfree, (uint32_t)(qb & 1));  // the softmax warps have read O: columns [0, dh) are free too
                tc_fence_after();
                issue_s(qb + 1, 0, n_split);
                umma_commit_e<1>(bar_s);
            }
        }
    } else {
        // ======================================================================= softmax warps
        const int quad = warp & 3, g = warp >> 2;
        const int r = quad * 32 + lane;                       // query row in the block == TMEM lane
        const uint32_t taddr = tmem_base + ((uint32_t)(quad * 32) << 16);
        const uint32_t row_off = (uint32_t)r * 128u, sw = (uint32_t)(r & 7);
        // slice j (32 keys) -> chunk j/2, half j%2; owner group: G=1: 0, G=2: half, G=4: 2*(chunk%2) + half
        auto owner = [](int j) { return G == 1 ? 0 : G == 2 ? (j & 1) : (((j >> 1) & 1) << 1) | (j & 1); };

        for (int qb = 0; qb < nqb; qb++) {
        const int q0 = qb * kAttQ;
        mbar_wait(bar_s, (uint32_t)(qb & 1));
        tc_fence_after();
        // ---- pass 1: row maximum of the raw scores over the real keys
        float mx = -INFINITY;
        for (int j = 0; j < n_slices; j++) {
            if (owner(j) != g) continue;
            uint32_t a[32];
            tmem_ld_32x32(taddr + (uint32_t)(j * 32), a);
            tmem_ld_wait();
            const int nv = L - j * 32;  // real keys in this slice (>= 1)
            if (nv < 32) {              // last slice: keys past the sentence never enter (CTA-uniform branch)
#pragma unroll
                for (int i = 0; i < 32; i++) if (i >= nv) a[i] = 0xff800000u;  // -inf
            }
            float m0 = mx, m1 = -INFINITY;
#pragma unroll
            for (int i = 0; i < 32; i += 4) {
                m0 = fmaxf(m0, fmaxf(__uint_as_float(a[i]), __uint_as_float(a[i + 1])));
                m1 = fmaxf(m1, fmaxf(__uint_as_float(a[i + 2]), __uint_as_float(a[i + 3])));
            }
            mx = fmaxf(m0, m1);
        }
        if (G > 1) {
            red[g * 128 + r] = mx;
            named_bar_sync(1, NSOFT);
#pragma unroll
            for (int gg = 0; gg < G; gg++) mx = fmaxf(mx, red[gg * 128 + r]);
            // (red is next written for the following query block, after every warp has passed bar_tfree / bar_s)
        }
        mx *= scale;  // max(s*scale) == max(s)*scale for scale > 0 (rounding is monotonic)

        // ---- pass 2: p = f16(exp(f16(s*scale - max))) -> P chunk buffers
        for (int c = 0; c < n_chunks; c++) {
            const uint32_t cg = (uint32_t)(qb * n_chunks + c);  // chunk number over all query blocks: buffer cg & 1
            // groups working on chunk cg: G=1: {0}; G=2: {0,1}; G=4: {2*(cg%2), 2*(cg%2)+1}
            if (G == 4 && (uint32_t)(g >> 1) != (cg & 1)) continue;
            const uint32_t pbi = cg % npb;
            if (cg >= npb) mbar_wait(&pfree[pbi], ((cg / npb) - 1) & 1);  // the MMAs that read this buffer retired
            uint8_t *pb = sP + pbi * PBUF;
#pragma unroll 1
            for (int h = 0; h < 2; h++) {
                if (G >= 2 && h != (g & 1)) continue;
                const int j = 2 * c + h;
                if (j >= n_slices) continue;  // no real key: these columns are not read by the MMA
                uint32_t a[32];
                tmem_ld_32x32(taddr + (uint32_t)(j * 32), a);
                tmem_ld_wait();
                const int nv = L - j * 32;
                if (nv < 32) {  // s = -inf -> x = -inf -> p = 2^-inf = 0 exactly: a masked key contributes nothing
#pragma unroll
                    for (int i = 0; i < 32; i++) if (i >= nv) a[i] = 0xff800000u;
                }
                uint32_t pk[16];
                const float2 sc2 = make_float2(scale, scale), nmx2 = make_float2(-mx, -mx), l2e = make_float2(1.4426950408889634f, 1.4426950408889634f);
#pragma unroll
                for (int i = 0; i < 32; i += 2) {
                    // ggml_scale (separate multiply), x - max, f16 round | exp | f16 round   (ggml.c:10536-10558), on element
                    // pairs with the packed fp32 instructions (each lane is the scalar IEEE op; the product is formed with
                    // scalar multiplies because ptxas contracts a packed mul + add into one FFMA2, which would skip the
                    // reference's rounding of s*scale)
                    float2 x;
                    if (DH == 64)  // scale = 1/8: s*scale is exact, so one FMA gives fl(fl(s*scale) - max) bit for bit
                        x = __ffma2_rn(make_float2(__uint_as_float(a[i]), __uint_as_float(a[i + 1])), sc2, nmx2);
                    else
                        x = __fadd2_rn(make_float2(__fmul_rn(__uint_as_float(a[i]), scale), __fmul_rn(__uint_as_float(a[i + 1]), scale)), nmx2);
                    const float2 z = __fmul2_rn(__half22float2(__floats2half2_rn(x.x, x.y)), l2e);
                    float2 e;
                    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e.x) : "f"(z.x));
                    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e.y) : "f"(z.y));
                    const __half2 hp = __floats2half2_rn(e.x, e.y);
                    pk[i >> 1] = *reinterpret_cast<const uint32_t *>(&hp);
                }
#pragma unroll
                for (int q = 0; q < 4; q++)
                    *reinterpret_cast<uint4 *>(pb + row_off + ((((uint32_t)(h * 4 + q)) ^ sw) << 4)) =
                        make_uint4(pk[4 * q], pk[4 * q + 1], pk[4 * q + 2], pk[4 * q + 3]);
            }
            fence_proxy_async_smem();  // P writes -> visible to the MMA (async proxy)
            tc_fence_before();         // my TMEM reads of these score columns precede the MMA that may overwrite them (O aliases chunk 0)
            __syncwarp();
            if (lane == 0) mbar_arrive(&pfull[pbi]);
        }
        // ---- epilogue: O * (1 / sum p), 32 features per (thread, pass).  The row sum comes out of the tensor core too:
        // rowsum = P . 1 accumulated next to O (TMEM column dh), i.e. the sum of the f16-rounded p over ALL keys in fp32,
        // in one fixed key order whatever the number of softmax groups -- a sentence's result does not depend on the batch
        // it is in -- and without the f16 -> f32 widening + add per element that the issue-bound softmax loop had
        mbar_wait(bar_o, (uint32_t)(qb & 1));
        tc_fence_after();
        const float sum = __uint_as_float(tmem_ld_32x1(taddr + (uint32_t)DH));
        tmem_ld_wait();
        const float inv = 1.0f / sum;
        if (g >= DH / 32) {  // this group reads no O columns: its part of TMEM is free for the next query block right away
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(bar_tfree);
        }
        for (int hc = g; hc < DH / 32; hc += G) {
            uint32_t a[32];
            tmem_ld_32x32(taddr + (uint32_t)(hc * 32), a);
            tmem_ld_wait();
            if (hc + G >= DH / 32) {  // last TMEM read of this block by this warp: the next block's last score columns may be
                tc_fence_before();    // issued while the scaling / rounding / stores below run
                __syncwarp();
                if (lane == 0) mbar_arrive(bar_tfree);
            }
            float o[32];
#pragma unroll
            for (int i = 0; i < 32; i++) o[i] = __uint_as_float(a[i]) * inv;
            if (qmode) {  // the attention-output GEMM of a Q4 model quantises its input per 32 features
                float amax = 0.f;
#pragma unroll
                for (int i = 0; i < 32; i++) amax = fmaxf(amax, fabsf(o[i]));
                float id, dq;
                q8_scales(amax, qmode, id, dq);
#pragma unroll
                for (int i = 0; i < 32; i++) o[i] = q8_round(o[i], id, dq);
            }
            if (q0 + r < L) {
                uint4 *dst = reinterpret_cast<uint4 *>(out + (size_t)(t0 + q0 + r) * d + (size_t)head * DH + hc * 32);
#pragma unroll
                for (int q = 0; q < 4; q++)
                    dst[q] = make_uint4(pack_h2(o[8 * q], o[8 * q + 1]), pack_h2(o[8 * q + 2], o[8 * q + 3]),
                                        pack_h2(o[8 * q + 4], o[8 * q + 5]), pack_h2(o[8 * q + 6], o[8 * q + 7]));
            }
        }
        }  // query blocks
    }

    tc_fence_before();
    __syncthreads();
    if (warp == CW) {
        tc_fence_after();
        const int need = Lk16 > DH + kSumCols ? Lk16 : DH + kSumCols;
        if (need <= 32) tmem_dealloc<32>(tmem_base);
        else if (need <= 64) tmem_dealloc<64>(tmem_base);
        else if (need <= 128) tmem_dealloc<128>(tmem_base);
        else if (need <= 256) tmem_dealloc<256>(tmem_base);
        else tmem_dealloc<512>(tmem_base);
    }
}

#include "attention_persist.cuh"
#include "attention_loop.cuh"
#include "attention_pack.cuh"
#include "attention_stream.cuh"
#include "attention_duo.cuh"
#include "attention_pt.cuh"
#include "attention_xs.cuh"

template <int DH, int G>
static cudaError_t launch_att(const CUtensorMap &tm, const int32_t *cu, int n_seq, int max_len, int d, int n_head, __half *out,
                              int qmode, cudaStream_t st) {
    const int Lk64 = (max_len + 63) & ~63;
    const int qk = (kAttQ + Lk64) * DH * 2;
    const int multi = max_len > kAttQ ? 1 : 0;  // sentences span several 128-query blocks: one CTA per (sentence, head) keeps K, V resident
    const size_t body = multi ? (size_t)2 * kAttQ * DH * 2 + 4 * kAttQ * 128 + (size_t)2 * Lk64 * DH * 2
                              : (size_t)(qk > 2 * kAttQ * 128 ? qk : 2 * kAttQ * 128) + (size_t)Lk64 * DH * 2;
    const size_t smem = body + 256 /*barriers*/ + (size_t)G * 128 * 4 + 512 /*ones tile*/;
    static std::atomic<bool> attr_set[64];  // per kernel instance and device (the attribute is per device)
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64 || !attr_set[dev]) {
        cudaError_t err = cudaFuncSetAttribute(attention_tc_kernel<DH, G>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
        if (err != cudaSuccess) return err;
        if (dev >= 0 && dev < 64) attr_set[dev] = true;
    }
    dim3 grid(1, n_head, n_seq);
    const float scale = (float)(1.0 / sqrt((double)DH));  // bert.cpp:934: 1.0f/sqrt((float)dh) evaluated in double
    LaunchConfig lc(grid, dim3(32 + 128 * G), smem, st);
    return cudaLaunchKernelEx(&lc.cfg, attention_tc_kernel<DH, G>, tm, cu, d, out, qmode, scale, multi);
}

cudaError_t launch_attention(const CUtensorMap &tmQKV, const int32_t *cu_seqlens, int n_seq, int max_len, int d, int n_head,
                             __half *out, int qmode, cudaStream_t st) {
    if (n_seq <= 0) return cudaSuccess;
    if (d % n_head != 0 || max_len < 1 || max_len > 512) return cudaErrorInvalidValue;
    const int dh = d / n_head;
    static const int env_g = [] { const char *v = getenv("BERT_B200_ATT_G"); return v ? atoi(v) : 0; }();
    int G = max_len <= 128 ? 1 : max_len <= 256 ? 2 : 4;
    if (env_g == 1 || env_g == 2 || env_g == 4) G = env_g;
    // launches whose longest sentence has <= 128 tokens: the per-item kernel with the CTA kept on its SM, four per SM, each walking over
    // items (attention_loop.cuh).  Per 65 536 tokens of BERT-base F16 (tools/att_len_probe.py), loop / persistent two-stage / one CTA
    // per item: L = 16: 286 / 378 / 529 us, 32: 169 / 204 / 295, 48: 148 / 203 / 229, 64: 122 / 178 / 184, 96: 129 / - / 171,
    // 128: 128 / - / 138 (Q4_0: 379 / 504 / 527 at 16, 137 / - / 188 at 64, 136 / - / 137 at 128).
    // BERT_B200_ATT_LOOP=0: one CTA per (sentence, head); BERT_B200_ATT_PERSIST=n: the persistent two-stage kernel
    // (attention_persist.cuh) for launches up to n tokens -- both kept for the A/B tests, read per launch.
    const char *pv = getenv("BERT_B200_ATT_PERSIST");
    const int persist_max = pv ? atoi(pv) : 0;
    if (max_len <= kAttQ && max_len <= persist_max && env_g == 0) {
        if (dh == 64) return launch_att_persist<64>(tmQKV, cu_seqlens, n_seq, d, n_head, out, qmode, st);
        if (dh == 32) return launch_att_persist<32>(tmQKV, cu_seqlens, n_seq, d, n_head, out, qmode, st);
    }
    const char *lv = getenv("BERT_B200_ATT_LOOP");
    if (max_len <= kAttQ && env_g == 0 && (lv ? atoi(lv) != 0 : kAttLoopDefault)) {
        if (dh == 64) return launch_att_loop<64>(tmQKV, cu_seqlens, n_seq, d, n_head, out, qmode, st);
        if (dh == 32) return launch_att_loop<32>(tmQKV, cu_seqlens, n_seq, d, n_head, out, qmode, st);
    }
    // launches with sentences of more than 128 tokens: scores streamed through a TMEM ring (BERT_B200_ATT_STREAM=0: whole
    // score block resident, attention_tc_kernel)
    const char *sv = getenv("BERT_B200_ATT_STREAM");
    // exp / service warp split (attention_xs.cuh); BERT_B200_ATT_XS=0: attention_pt.cuh
    const char *xv = getenv("BERT_B200_ATT_XS");
    if (max_len > kAttQ && (!sv || atoi(sv) != 0) && env_g == 0 && (xv ? atoi(xv) != 0 : kAttXsDefault)) {
        if (dh == 64) return launch_att_xs<64>(tmQKV, cu_seqlens, n_seq, max_len, d, n_head, out, qmode, st);
        if (dh == 32) return launch_att_xs<32>(tmQKV, cu_seqlens, n_seq, max_len, d, n_head, out, qmode, st);
    }
    // P kept in tensor memory (attention_pt.cuh); BERT_B200_ATT_PT=0: through shared memory (attention_stream.cuh)
    const char *tv = getenv("BERT_B200_ATT_PT");
    if (max_len > kAttQ && (!sv || atoi(sv) != 0) && env_g == 0 && (tv ? atoi(tv) != 0 : kAttPtDefault)) {
        if (dh == 64) return launch_att_pt<64>(tmQKV, cu_seqlens, n_seq, max_len, d, n_head, out, qmode, st);
        if (dh == 32) return launch_att_pt<32>(tmQKV, cu_seqlens, n_seq, max_len, d, n_head, out, qmode, st);
    }
    // two query blocks in flight per CTA (attention_duo.cuh, BERT_B200_ATT_DUO=1; measured no faster than one block at a time)
    const char *dv = getenv("BERT_B200_ATT_DUO");
    if (max_len > kAttQ && (!sv || atoi(sv) != 0) && env_g == 0 && (dv ? atoi(dv) != 0 : kAttDuoDefault)) {
        if (dh == 64) return launch_att_duo<64>(tmQKV, cu_seqlens, n_seq, max_len, d, n_head, out, qmode, st);
        if (dh == 32) return launch_att_duo<32>(tmQKV, cu_seqlens, n_seq, max_len, d, n_head, out, qmode, st);
    }
    if (max_len > kAttQ && (!sv || atoi(sv) != 0) && env_g == 0) {
        if (dh == 64) return launch_att_stream<64>(tmQKV, cu_seqlens, n_seq, max_len, d, n_head, out, qmode, st);
        if (dh == 32) return launch_att_stream<32>(tmQKV, cu_seqlens, n_seq, max_len, d, n_head, out, qmode, st);
    }
#define BK_ATT(DH_)                                                                                      \
    switch (G) {                                                                                         \
        case 1: return launch_att<DH_, 1>(tmQKV, cu_seqlens, n_seq, max_len, d, n_head, out, qmode, st); \
        case 2: return launch_att<DH_, 2>(tmQKV, cu_seqlens, n_seq, max_len, d, n_head, out, qmode, st); \
        default: return launch_att<DH_, 4>(tmQKV, cu_seqlens, n_seq, max_len, d, n_head, out, qmode, st); \
    }
    if (dh == 64) { BK_ATT(64) }
    if (dh == 32) { BK_ATT(32) }
#undef BK_ATT
    return cudaErrorInvalidValue;
}

cudaError_t launch_attention_packed(const CUtensorMap &tm64, const CUtensorMap &tm32, const CUtensorMap &tm16, const int32_t *cu_seqlens,
                                    int n64, int n32, int n16, int d, int n_head, __half *out, int qmode, cudaStream_t st) {
    if (n64 < 0 || n32 < 0 || n16 < 0 || d % n_head != 0) return cudaErrorInvalidValue;
    if (n64 + n32 + n16 == 0) return cudaSuccess;
    const PackCounts pc{n64, n32, n16};
    const int dh = d / n_head;
    if (dh == 64) return launch_att_pack<64>(tm64, tm32, tm16, cu_seqlens, pc, d, n_head, out, qmode, st);
    if (dh == 32) return launch_att_pack<32>(tm64, tm32, tm16, cu_seqlens, pc, d, n_head, out, qmode, st);
    return cudaErrorInvalidValue;
}

// box {dh, box_rows rows} over qkv[rows][3d]; 128-byte swizzle for dh = 64, 64-byte for dh = 32
bool make_tmap_attention(CUtensorMap *tm, const void *qkv, int d, int n_head, uint64_t rows, int box_rows) {
    const int dh = d / n_head;
    return make_tmap_f16_2d(tm, qkv, (uint64_t)3 * d, rows, (uint64_t)3 * d * 2, (uint32_t)dh, (uint32_t)box_rows, dh == 64 ? 128 : 64);
}

}  // namespace bk

#ifdef BK_EXP_TRACE
extern "C" __attribute__((visibility("default"))) int bk_debug_att_trace(unsigned long long *out) {
    return (int)cudaMemcpyFromSymbol(out, bk::g_att_trace, sizeof(unsigned long long) * 8 * 1024);
}
#endif
