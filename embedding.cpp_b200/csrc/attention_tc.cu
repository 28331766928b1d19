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
//   epilogue O * (1 / sum p)  [+ Q8 rounding per 32 features for Q4 models, ggml.c:9493-9509] -> f16 rows
//
// Because the full score row is resident in TMEM, the true row maximum is known before any exponential
// is taken (the reference's rounding points: separate scale multiply, subtract, f16 round, exp, f16
// round) with a single QK^T pass -- no online-softmax rescaling, no second score computation.
// Warps 0 .. 4G-1 are softmax warps (warp w owns TMEM lanes 32*(w%4)..; group g = w/4 owns every G-th
// 32-key slice), warp 4G is the control warp (TMA, MMA issue, TMEM allocation).  G = 1/2/4 for
// Lk <= 128/256/512 so a CTA always has one softmax warp per 32 rows x 128 keys.
#include <cmath>
#include <cstdlib>

#include "common.cuh"
#include "kernels.h"

namespace bk {

constexpr int kAttQ = 128;  // queries per CTA

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

template <int DH, int G>
__global__ void __launch_bounds__(32 + 128 * G) attention_tc_kernel(const __grid_constant__ CUtensorMap tmQKV, const int32_t *__restrict__ cu,
                                                                    int d, __half *__restrict__ out, int qmode, float scale) {
    constexpr int ROWB = DH * 2;           // bytes per smem row of Q / K / V
    constexpr int CW = 4 * G;              // control warp
    constexpr int NSOFT = 128 * G;         // softmax threads
    constexpr int PBUF = kAttQ * 128;      // one P chunk: 128 rows x 64 keys f16
    extern __shared__ uint8_t att_raw[];
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(att_raw) + 1023) & ~uintptr_t(1023));

    const int seq = blockIdx.z, head = blockIdx.y, q0 = blockIdx.x * kAttQ;
    const int t0 = __ldg(cu + seq);
    const int L = __ldg(cu + seq + 1) - t0;
    if (q0 >= L) return;
    const int Lk16 = (L + 15) & ~15;       // keys entering the MMAs
    const int Lk64 = (L + 63) & ~63;       // keys staged in smem (whole 64-row TMA boxes)
    const int n_chunks = Lk64 >> 6;
    const int n_slices = (L + 31) >> 5;    // 32-key slices holding at least one real key

    // region 0: Q | K while S is computed, then the two P chunk buffers; region 1: V
    const int qk_bytes = (kAttQ + Lk64) * ROWB;
    const int r0_bytes = qk_bytes > 2 * PBUF ? qk_bytes : 2 * PBUF;
    uint8_t *sQ = smem, *sK = smem + kAttQ * ROWB, *sP = smem, *sV = smem + r0_bytes;
    uint8_t *tail = sV + Lk64 * ROWB;
    uint64_t *bar_qk = reinterpret_cast<uint64_t *>(tail);
    uint64_t *bar_v = bar_qk + 1, *bar_s = bar_qk + 2, *bar_o = bar_qk + 3, *pfull = bar_qk + 4, *pfree = bar_qk + 6;
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bar_qk + 8);
    float *red = reinterpret_cast<float *>(bar_qk + 10);  // [G][128] partial row max, then partial row sums

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp == CW) {
        if (lane == 0) {
            tma_prefetch_desc(&tmQKV);
            mbar_init(bar_qk, 1); mbar_init(bar_v, 1); mbar_init(bar_s, 1); mbar_init(bar_o, 1);
            mbar_init(&pfull[0], 4 * (G >= 2 ? 2 : 1)); mbar_init(&pfull[1], 4 * (G >= 2 ? 2 : 1));
            mbar_init(&pfree[0], 1); mbar_init(&pfree[1], 1);
            mbar_fence_init();
        }
        __syncwarp();
        // S needs Lk16 columns; O (dh columns) aliases columns [0, dh).  Power of two >= 32.
        const int need = Lk16 > DH ? Lk16 : DH;
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
        const int cq = head * DH, ck = d + head * DH, cv = 2 * d + head * DH;
        mbar_expect_tx_e(bar_qk, (uint32_t)qk_bytes);
        tma_load_2d_e(sQ, &tmQKV, bar_qk, cq, t0 + q0);
        tma_load_2d_e(sQ + 64 * ROWB, &tmQKV, bar_qk, cq, t0 + q0 + 64);
        for (int c = 0; c < n_chunks; c++) tma_load_2d_e(sK + c * 64 * ROWB, &tmQKV, bar_qk, ck, t0 + c * 64);
        mbar_expect_tx_e(bar_v, (uint32_t)(Lk64 * ROWB));
        for (int c = 0; c < n_chunks; c++) tma_load_2d_e(sV + c * 64 * ROWB, &tmQKV, bar_v, cv, t0 + c * 64);

        // ---- S = Q . K^T : N = Lk16 split into instructions of <= 256 columns
        mbar_wait(bar_qk, 0);
        tc_fence_after();
        const uint64_t dq = umma_desc_kmajor<ROWB>(smem_u32(sQ));
        for (int n0 = 0; n0 < Lk16; n0 += 256) {
            const int nn = Lk16 - n0 < 256 ? Lk16 - n0 : 256;
            const uint64_t dk = umma_desc_kmajor<ROWB>(smem_u32(sK + n0 * ROWB));
            const uint32_t idesc = att_idesc((uint32_t)nn, 0);
#pragma unroll
            for (int k = 0; k < DH / 16; k++)
                umma_f16_e<1>(tmem_base + (uint32_t)n0, dq + (uint64_t)(2 * k), dk + (uint64_t)(2 * k), idesc, k != 0);
        }
        umma_commit_e<1>(bar_s);

        // ---- O += P_c . V_c as the softmax warps deliver P chunks
        mbar_wait(bar_v, 0);
        const uint32_t idesc_pv = att_idesc(DH, 1);
        for (int c = 0; c < n_chunks; c++) {
            mbar_wait(&pfull[c & 1], (uint32_t)((c >> 1) & 1));
            tc_fence_after();
            const uint64_t dp = umma_desc_sw128(smem_u32(sP + (c & 1) * PBUF));
            const int rem = Lk16 - c * 64;
            const int ksteps = rem >= 64 ? 4 : rem >> 4;
            for (int ks = 0; ks < ksteps; ks++) {
                const uint64_t dv = umma_desc_mnmajor<ROWB>(smem_u32(sV + (c * 64 + ks * 16) * ROWB));
                umma_f16_e<1>(tmem_base, dp + (uint64_t)(2 * ks), dv, idesc_pv, (c | ks) != 0);
            }
            umma_commit_e<1>(&pfree[c & 1]);
        }
        umma_commit_e<1>(bar_o);
    } else {
        // ======================================================================= softmax warps
        const int quad = warp & 3, g = warp >> 2;
        const int r = quad * 32 + lane;                       // query row in the block == TMEM lane
        const uint32_t taddr = tmem_base + ((uint32_t)(quad * 32) << 16);
        const uint32_t row_off = (uint32_t)r * 128u, sw = (uint32_t)(r & 7);
        // slice j (32 keys) -> chunk j/2, half j%2; owner group: G=1: 0, G=2: half, G=4: 2*(chunk%2) + half
        auto owner = [](int j) { return G == 1 ? 0 : G == 2 ? (j & 1) : (((j >> 1) & 1) << 1) | (j & 1); };

        mbar_wait(bar_s, 0);
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
            named_bar_sync(1, NSOFT);  // red is reused for the sums
        }
        mx *= scale;  // max(s*scale) == max(s)*scale for scale > 0 (rounding is monotonic)

        // ---- pass 2: p = f16(exp(f16(s*scale - max))) -> P chunk buffers
        float sum = 0.f;
        for (int c = 0; c < n_chunks; c++) {
            // groups working on chunk c: G=1: {0}; G=2: {0,1}; G=4: {2*(c%2), 2*(c%2)+1}
            if (G == 4 && (g >> 1) != (c & 1)) continue;
            if (c >= 2) mbar_wait(&pfree[c & 1], (uint32_t)(((c >> 1) - 1) & 1));  // the MMAs that read this buffer retired
            uint8_t *pb = sP + (c & 1) * PBUF;
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
                float2 s2 = make_float2(0.f, 0.f);
                const float2 sc2 = make_float2(scale, scale), nmx2 = make_float2(-mx, -mx), l2e = make_float2(1.4426950408889634f, 1.4426950408889634f);
#pragma unroll
                for (int i = 0; i < 32; i += 2) {
                    // ggml_scale (separate multiply), x - max, f16 round | exp | f16 round   (ggml.c:10536-10558), on element
                    // pairs with the packed fp32 instructions (each lane is the scalar IEEE op; the product is formed with
                    // scalar multiplies because ptxas contracts a packed mul + add into one FFMA2, which would skip the
                    // reference's rounding of s*scale)
                    const float2 x = __fadd2_rn(make_float2(__fmul_rn(__uint_as_float(a[i]), scale), __fmul_rn(__uint_as_float(a[i + 1]), scale)), nmx2);
                    const float2 z = __fmul2_rn(__half22float2(__floats2half2_rn(x.x, x.y)), l2e);
                    float e0, e1;
                    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e0) : "f"(z.x));
                    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e1) : "f"(z.y));
                    const __half2 hp = __floats2half2_rn(e0, e1);
                    s2 = __fadd2_rn(s2, __half22float2(hp));
                    pk[i >> 1] = *reinterpret_cast<const uint32_t *>(&hp);
                }
                const float s0 = s2.x, s1 = s2.y;
                (void)sc2;
                sum += s0 + s1;
#pragma unroll
                for (int q = 0; q < 4; q++)
                    *reinterpret_cast<uint4 *>(pb + row_off + ((((uint32_t)(h * 4 + q)) ^ sw) << 4)) =
                        make_uint4(pk[4 * q], pk[4 * q + 1], pk[4 * q + 2], pk[4 * q + 3]);
            }
            fence_proxy_async_smem();  // P writes -> visible to the MMA (async proxy)
            tc_fence_before();         // my TMEM reads of these score columns precede the MMA that may overwrite them (O aliases chunk 0)
            __syncwarp();
            if (lane == 0) mbar_arrive(&pfull[c & 1]);
        }
        if (G > 1) {
            red[g * 128 + r] = sum;
            named_bar_sync(1, NSOFT);
            sum = 0.f;
#pragma unroll
            for (int gg = 0; gg < G; gg++) sum += red[gg * 128 + r];
        }
        const float inv = 1.0f / sum;

        // ---- epilogue: O * inv, 32 features per (thread, pass)
        mbar_wait(bar_o, 0);
        tc_fence_after();
        for (int hc = g; hc < DH / 32; hc += G) {
            uint32_t a[32];
            tmem_ld_32x32(taddr + (uint32_t)(hc * 32), a);
            tmem_ld_wait();
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
    }

    tc_fence_before();
    __syncthreads();
    if (warp == CW) {
        tc_fence_after();
        const int need = Lk16 > DH ? Lk16 : DH;
        if (need <= 32) tmem_dealloc<32>(tmem_base);
        else if (need <= 64) tmem_dealloc<64>(tmem_base);
        else if (need <= 128) tmem_dealloc<128>(tmem_base);
        else if (need <= 256) tmem_dealloc<256>(tmem_base);
        else tmem_dealloc<512>(tmem_base);
    }
}

template <int DH, int G>
static cudaError_t launch_att(const CUtensorMap &tm, const int32_t *cu, int n_seq, int max_len, int d, int n_head, __half *out,
                              int qmode, cudaStream_t st) {
    const int Lk64 = (max_len + 63) & ~63;
    const int qk = (kAttQ + Lk64) * DH * 2;
    const size_t smem = (size_t)(qk > 2 * kAttQ * 128 ? qk : 2 * kAttQ * 128) + (size_t)Lk64 * DH * 2 + 1024 /*align*/ + 128 + (size_t)G * 128 * 4;
    static bool attr_set[64] = {false};  // per kernel instance and device (the attribute is per device)
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64 || !attr_set[dev]) {
        cudaError_t err = cudaFuncSetAttribute(attention_tc_kernel<DH, G>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
        if (err != cudaSuccess) return err;
        if (dev >= 0 && dev < 64) attr_set[dev] = true;
    }
    dim3 grid((max_len + kAttQ - 1) / kAttQ, n_head, n_seq);
    const float scale = (float)(1.0 / sqrt((double)DH));  // bert.cpp:934: 1.0f/sqrt((float)dh) evaluated in double
    attention_tc_kernel<DH, G><<<grid, 32 + 128 * G, smem, st>>>(tm, cu, d, out, qmode, scale);
    return cudaGetLastError();
}

cudaError_t launch_attention(const CUtensorMap &tmQKV, const int32_t *cu_seqlens, int n_seq, int max_len, int d, int n_head,
                             __half *out, int qmode, cudaStream_t st) {
    if (n_seq <= 0) return cudaSuccess;
    if (d % n_head != 0 || max_len < 1 || max_len > 512) return cudaErrorInvalidValue;
    const int dh = d / n_head;
    static const int env_g = [] { const char *v = getenv("BERT_B200_ATT_G"); return v ? atoi(v) : 0; }();
    int G = max_len <= 128 ? 1 : max_len <= 256 ? 2 : 4;
    if (env_g == 1 || env_g == 2 || env_g == 4) G = env_g;
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

// box {dh, 64 rows} over qkv[rows][3d]; 128-byte swizzle for dh = 64, 64-byte for dh = 32
bool make_tmap_attention(CUtensorMap *tm, const void *qkv, int d, int n_head, uint64_t rows) {
    const int dh = d / n_head;
    return make_tmap_f16_2d(tm, qkv, (uint64_t)3 * d, rows, (uint64_t)3 * d * 2, (uint32_t)dh, 64, dh == 64 ? 128 : 64);
}

}  // namespace bk
