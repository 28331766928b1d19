// attention.cu -- fused masked-softmax attention for variable-length packed batches.
//
// Replaces the per-sentence ggml nodes of bert.cpp:928-943 (mul_mat K.Q, scale, soft_max, transpose+
// cont V, mul_mat, permute, cpy).  One CTA = one (sentence, head, 128-query block); the sentence's K
// and V head slices live in shared memory, so "masking" is simply the key-range bound L_b of this
// sentence (a padded key would get p = 0, identical to the reference's unpadded softmax).
//
// Numerics follow ggml_compute_forward_soft_max_f32 (ggml.c:10498-10567): S = (Q.K^T) * (1/sqrt(dh))
// as a separate fp32 multiply, row max over the whole row, p = f16(exp(f16(s - max))) (the reference's
// f16 exp table), row sum of the rounded p, P = p / sum.  To use the true row max (not a running one)
// the score tile is computed twice: pass 1 finds the max, pass 2 recomputes S, forms p and
// accumulates P.V.  p is exactly representable in f16, so it feeds the second MMA unrounded.
// Tensor path: mma.sync m16n8k16 f16 -> fp32 (the kernel is HBM/L2-bound at L<=128).
#include "common.cuh"
#include "kernels.h"

namespace bk {

constexpr int kAttnThreads = 256;  // 8 warps x 16 query rows
constexpr int kAttnQBlock = 128;

__device__ __forceinline__ void cp_async16(void *smem_dst, const void *gsrc) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(smem_dst)), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }
__device__ __forceinline__ void ldsm_x4(uint32_t addr, uint32_t &r0, uint32_t &r1, uint32_t &r2, uint32_t &r3) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];" : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t addr, uint32_t &r0, uint32_t &r1, uint32_t &r2, uint32_t &r3) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];" : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void mma_16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

template <int DH>
__global__ void __launch_bounds__(kAttnThreads) attention_kernel(const __half *__restrict__ qkv, const int32_t *__restrict__ cu,
                                                                 int d, __half *__restrict__ out, int qmode, float scale) {
    constexpr int LDS = DH + 8;       // padded row (halves): 16 B shift per row -> conflict-free ldmatrix
    constexpr int CPR = DH / 8;       // 16-byte chunks per row
    constexpr int KSTEPS = DH / 16;   // k-steps of Q.K^T
    constexpr int NT_O = DH / 8;      // output n-tiles
    extern __shared__ __align__(16) uint8_t att_smem[];

    const int seq = blockIdx.z, head = blockIdx.y, qb = blockIdx.x;
    const int t0 = __ldg(cu + seq);
    const int L = __ldg(cu + seq + 1) - t0;
    const int q0 = qb * kAttnQBlock;
    if (q0 >= L) return;
    const int Lk = (L + 15) & ~15;
    __half *sK = reinterpret_cast<__half *>(att_smem);
    __half *sV = sK + (size_t)Lk * LDS;
    __half *sQ = sV + (size_t)Lk * LDS;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const size_t ld = (size_t)3 * d;
    const __half *gq = qkv + (size_t)t0 * ld + (size_t)head * DH;
    const __half *gk = gq + d, *gv = gq + 2 * d;

    // ---- stage K, V (whole sentence) and this block's Q rows; rows past L are zero ----
    for (int i = tid; i < Lk * CPR; i += kAttnThreads) {
        const int r = i / CPR, c = i % CPR;
        __half *dk = sK + (size_t)r * LDS + c * 8, *dv = sV + (size_t)r * LDS + c * 8;
        if (r < L) {
            cp_async16(dk, gk + (size_t)r * ld + c * 8);
            cp_async16(dv, gv + (size_t)r * ld + c * 8);
        } else {
            *reinterpret_cast<uint4 *>(dk) = make_uint4(0, 0, 0, 0);
            *reinterpret_cast<uint4 *>(dv) = make_uint4(0, 0, 0, 0);
        }
    }
    for (int i = tid; i < kAttnQBlock * CPR; i += kAttnThreads) {
        const int r = i / CPR, c = i % CPR;
        __half *dq = sQ + (size_t)r * LDS + c * 8;
        if (q0 + r < L) cp_async16(dq, gq + (size_t)(q0 + r) * ld + c * 8);
        else *reinterpret_cast<uint4 *>(dq) = make_uint4(0, 0, 0, 0);
    }
    cp_async_wait_all();
    __syncthreads();

    const int qrow0 = q0 + warp * 16;  // first query row of this warp
    if (qrow0 >= L) return;

    // ---- Q fragments (A operand), kept in registers ----
    uint32_t qf[KSTEPS][4];
    {
        const int r = warp * 16 + (lane & 7) + ((lane >> 3) & 1) * 8;
        const int cb = (lane >> 4) * 8;
#pragma unroll
        for (int kk = 0; kk < KSTEPS; kk++)
            ldsm_x4(smem_u32(sQ + (size_t)r * LDS + kk * 16 + cb), qf[kk][0], qf[kk][1], qf[kk][2], qf[kk][3]);
    }
    const int kcol = (lane >> 3) * 8;   // K ldmatrix: lane -> (key row lane%8, 8-col group lane/8)
    const int krow = lane & 7;

    // ---- pass 1: row maxima of the raw scores ----
    float mx0 = -INFINITY, mx1 = -INFINITY;  // rows lane/4 and lane/4+8
    for (int j = 0; j < Lk / 8; j++) {
        float c[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int kk = 0; kk < KSTEPS; kk += 2) {
            uint32_t b0, b1, b2, b3;
            ldsm_x4(smem_u32(sK + (size_t)(j * 8 + krow) * LDS + kk * 16 + kcol), b0, b1, b2, b3);
            mma_16816(c, qf[kk], b0, b1);
            mma_16816(c, qf[kk + 1], b2, b3);
        }
        const int key = j * 8 + (lane & 3) * 2;
        if (key < L) { mx0 = fmaxf(mx0, c[0]); mx1 = fmaxf(mx1, c[2]); }
        if (key + 1 < L) { mx0 = fmaxf(mx0, c[1]); mx1 = fmaxf(mx1, c[3]); }
    }
    mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1));
    mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
    mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1));
    mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
    mx0 *= scale;  // max(s*scale) == max(s)*scale for scale > 0 (rounding is monotonic)
    mx1 *= scale;

    // ---- pass 2: p = f16(exp(f16(s*scale - max))), O += p.V ----
    float o[NT_O][4];
#pragma unroll
    for (int n = 0; n < NT_O; n++) { o[n][0] = o[n][1] = o[n][2] = o[n][3] = 0.f; }
    float sum0 = 0.f, sum1 = 0.f;
    const int vrow = (lane & 7) + ((lane >> 3) & 1) * 8;  // V ldmatrix.trans: lane -> key row, 8-col group lane/16
    const int vcol = (lane >> 4) * 8;
    for (int jj = 0; jj < Lk / 16; jj++) {
        uint32_t pa[4];
#pragma unroll
        for (int half_ = 0; half_ < 2; half_++) {
            const int j = jj * 2 + half_;
            float c[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int kk = 0; kk < KSTEPS; kk += 2) {
                uint32_t b0, b1, b2, b3;
                ldsm_x4(smem_u32(sK + (size_t)(j * 8 + krow) * LDS + kk * 16 + kcol), b0, b1, b2, b3);
                mma_16816(c, qf[kk], b0, b1);
                mma_16816(c, qf[kk + 1], b2, b3);
            }
            const int key = j * 8 + (lane & 3) * 2;
            float p[4];
#pragma unroll
            for (int e = 0; e < 4; e++) {
                const float m = (e < 2) ? mx0 : mx1;
                const float x = round_f16(__fsub_rn(__fmul_rn(c[e], scale), m));  // separate multiply and subtract roundings, as ggml_scale + soft_max
                const float pe = round_f16(__expf(x));
                p[e] = (key + (e & 1) < L) ? pe : 0.f;
            }
            sum0 += p[0] + p[1];
            sum1 += p[2] + p[3];
            pa[half_ * 2 + 0] = pack_h2(p[0], p[1]);  // a0 / a2: row lane/4
            pa[half_ * 2 + 1] = pack_h2(p[2], p[3]);  // a1 / a3: row lane/4 + 8
        }
#pragma unroll
        for (int n = 0; n < NT_O; n += 2) {
            uint32_t b0, b1, b2, b3;
            ldsm_x4_t(smem_u32(sV + (size_t)(jj * 16 + vrow) * LDS + n * 8 + vcol), b0, b1, b2, b3);
            mma_16816(o[n], pa, b0, b1);
            mma_16816(o[n + 1], pa, b2, b3);
        }
    }
    sum0 += __shfl_xor_sync(0xffffffffu, sum0, 1);
    sum0 += __shfl_xor_sync(0xffffffffu, sum0, 2);
    sum1 += __shfl_xor_sync(0xffffffffu, sum1, 1);
    sum1 += __shfl_xor_sync(0xffffffffu, sum1, 2);
    const float inv0 = 1.0f / sum0, inv1 = 1.0f / sum1;
#pragma unroll
    for (int n = 0; n < NT_O; n++) { o[n][0] *= inv0; o[n][1] *= inv0; o[n][2] *= inv1; o[n][3] *= inv1; }

    // ---- Q4 models: the attention-output GEMM quantises its input per 32 features (ggml.c:9493-9509) ----
    if (qmode) {
#pragma unroll
        for (int b = 0; b < DH / 32; b++) {
            float a0 = 0.f, a1 = 0.f;
#pragma unroll
            for (int n = 4 * b; n < 4 * b + 4; n++) {
                a0 = fmaxf(a0, fmaxf(fabsf(o[n][0]), fabsf(o[n][1])));
                a1 = fmaxf(a1, fmaxf(fabsf(o[n][2]), fabsf(o[n][3])));
            }
            a0 = fmaxf(a0, __shfl_xor_sync(0xffffffffu, a0, 1));
            a0 = fmaxf(a0, __shfl_xor_sync(0xffffffffu, a0, 2));
            a1 = fmaxf(a1, __shfl_xor_sync(0xffffffffu, a1, 1));
            a1 = fmaxf(a1, __shfl_xor_sync(0xffffffffu, a1, 2));
            float id0, dq0, id1, dq1;
            q8_scales(a0, qmode, id0, dq0);
            q8_scales(a1, qmode, id1, dq1);
#pragma unroll
            for (int n = 4 * b; n < 4 * b + 4; n++) {
                o[n][0] = q8_round(o[n][0], id0, dq0); o[n][1] = q8_round(o[n][1], id0, dq0);
                o[n][2] = q8_round(o[n][2], id1, dq1); o[n][3] = q8_round(o[n][3], id1, dq1);
            }
        }
    }

    // ---- stage the 16 x DH tile through this warp's own Q rows, then coalesced 16-byte stores ----
    __syncwarp();
    {
        __half *sw = sQ + (size_t)(warp * 16) * LDS;
        const int r = lane >> 2, cc = (lane & 3) * 2;
#pragma unroll
        for (int n = 0; n < NT_O; n++) {
            *reinterpret_cast<uint32_t *>(sw + (size_t)r * LDS + n * 8 + cc) = pack_h2(o[n][0], o[n][1]);
            *reinterpret_cast<uint32_t *>(sw + (size_t)(r + 8) * LDS + n * 8 + cc) = pack_h2(o[n][2], o[n][3]);
        }
        __syncwarp();
        __half *go = out + (size_t)(t0 + qrow0) * d + (size_t)head * DH;
        for (int i = lane; i < 16 * CPR; i += 32) {
            const int rr = i / CPR, c = i % CPR;
            if (qrow0 + rr < L)
                *reinterpret_cast<uint4 *>(go + (size_t)rr * d + c * 8) = *reinterpret_cast<const uint4 *>(sw + (size_t)rr * LDS + c * 8);
        }
    }
}

template <int DH>
static cudaError_t launch_dh(const __half *qkv, const int32_t *cu, int n_seq, int max_len, int d, int n_head, __half *out,
                             int qmode, cudaStream_t st) {
    const int Lk = (max_len + 15) & ~15;
    const size_t smem = (size_t)(2 * Lk + kAttnQBlock) * (DH + 8) * sizeof(__half);
    cudaError_t err = cudaFuncSetAttribute(attention_kernel<DH>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (err != cudaSuccess) return err;
    dim3 grid((max_len + kAttnQBlock - 1) / kAttnQBlock, n_head, n_seq);
    const float scale = (float)(1.0 / sqrt((double)DH));  // bert.cpp:934: 1.0f/sqrt((float)dh) evaluated in double
    attention_kernel<DH><<<grid, kAttnThreads, smem, st>>>(qkv, cu, d, out, qmode, scale);
    return cudaGetLastError();
}

cudaError_t launch_attention(const __half *qkv, const int32_t *cu_seqlens, int n_seq, int max_len, int d, int n_head,
                             __half *out, int qmode, cudaStream_t st) {
    if (n_seq <= 0) return cudaSuccess;
    if (d % n_head != 0 || d % 8 != 0) return cudaErrorInvalidValue;
    const int dh = d / n_head;
    if (dh == 64) return launch_dh<64>(qkv, cu_seqlens, n_seq, max_len, d, n_head, out, qmode, st);
    if (dh == 32) return launch_dh<32>(qkv, cu_seqlens, n_seq, max_len, d, n_head, out, qmode, st);
    return cudaErrorInvalidValue;
}

}  // namespace bk
