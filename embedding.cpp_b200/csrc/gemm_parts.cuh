// gemm_parts.cuh -- pieces shared by the projection GEMM kernels (gemm_tcgen05.cu, gemm_ws.cu): the Q4 block decoder and the
// per-thread epilogue arithmetic on one 32-column slice of an accumulator row.
#pragma once
#include "common.cuh"
#include "../../include/bk_api.h"

namespace bk {

// 64 consecutive weights of one row = two GGUF blocks (w = the row's 36 / 40 packed bytes, 4-byte aligned) -> 64
// halves o[0..7].  Q4_0: (q - 8) * d, Q4_1: q * d + m; element j of a block is the low nibble of byte j, element
// j + 16 the high nibble (ggml-quants.c:1074-1113).  Bit-identical to the load-time expansion (rowwise.cu
// load_row8 + one f16 rounding): Q4_0 products are exact before the single f16 rounding; Q4_1 goes through the
// same fp32 fma.
template <int WQ>
__device__ __forceinline__ void q4_decode_row(const uint32_t *w, uint4 (&o)[8]) {
    constexpr int NW = WQ == 1 ? 9 : 10;
    uint32_t x[NW];
#pragma unroll
    for (int i = 0; i < NW; i++) x[i] = w[i];
#pragma unroll
    for (int blk = 0; blk < 2; blk++) {
        uint32_t qs[4];
        uint32_t dm;  // d in the low half (and m in the high half for Q4_1)
        if (WQ == 1) {
            if (blk == 0) {  // bytes 0..17: d | 16 nibble bytes starting at byte 2
                dm = x[0] & 0xFFFFu;
#pragma unroll
                for (int i = 0; i < 4; i++) qs[i] = __funnelshift_r(x[i], x[i + 1], 16);
            } else {         // bytes 18..35
                dm = x[4] >> 16;
#pragma unroll
                for (int i = 0; i < 4; i++) qs[i] = x[5 + i];
            }
        } else {
            dm = x[5 * blk];
#pragma unroll
            for (int i = 0; i < 4; i++) qs[i] = x[5 * blk + 1 + i];
        }
        const __half dh = __ushort_as_half((unsigned short)(dm & 0xFFFFu));
        uint32_t lo[8], hi[8];  // half2 words: K 0..15 from the low nibbles, K 16..31 from the high nibbles
#pragma unroll
        for (int i = 0; i < 4; i++) {
            const uint32_t l = qs[i] & 0x0F0F0F0Fu, h = (qs[i] >> 4) & 0x0F0F0F0Fu;
            // bytes [v, 0x64] = the half 1024 + v
            const uint32_t l01 = __byte_perm(l, 0x64646464u, 0x4140), l23 = __byte_perm(l, 0x64646464u, 0x4342);
            const uint32_t h01 = __byte_perm(h, 0x64646464u, 0x4140), h23 = __byte_perm(h, 0x64646464u, 0x4342);
            if (WQ == 1) {
                const __half2 d2 = __half2half2(dh), off = __float2half2_rn(1032.f);
                auto f = [&](uint32_t v) { __half2 r = __hmul2(__hsub2(*reinterpret_cast<const __half2 *>(&v), off), d2); return *reinterpret_cast<uint32_t *>(&r); };
                lo[2 * i] = f(l01); lo[2 * i + 1] = f(l23); hi[2 * i] = f(h01); hi[2 * i + 1] = f(h23);
            } else {
                const float dd = __half2float(dh), mm = __half2float(__ushort_as_half((unsigned short)(dm >> 16)));
                const __half2 off = __float2half2_rn(1024.f);
                auto f = [&](uint32_t v) {
                    const float2 q = __half22float2(__hsub2(*reinterpret_cast<const __half2 *>(&v), off));
                    return pack_h2(fmaf(q.x, dd, mm), fmaf(q.y, dd, mm));
                };
                lo[2 * i] = f(l01); lo[2 * i + 1] = f(l23); hi[2 * i] = f(h01); hi[2 * i + 1] = f(h23);
            }
        }
        o[4 * blk + 0] = make_uint4(lo[0], lo[1], lo[2], lo[3]);
        o[4 * blk + 1] = make_uint4(lo[4], lo[5], lo[6], lo[7]);
        o[4 * blk + 2] = make_uint4(hi[0], hi[1], hi[2], hi[3]);
        o[4 * blk + 3] = make_uint4(hi[4], hi[5], hi[6], hi[7]);
    }
}

// ---------------------------------------------------------------------------------- epilogue arithmetic
// All of it on element PAIRS with the packed fp32 instructions (f32x2): the epilogue warps are issue-slot bound, and each lane of
// a packed op is the scalar IEEE op (bit-identical results).
// v = acc + bias  (32 accumulator columns of one row; the bias add is a separate fp32 add in the reference, bert.cpp:909-910)
__device__ __forceinline__ void epi_add_bias(const uint32_t (&acc)[32], const float *bias_n, float2 (&v)[16]) {
    const float4 *b4 = reinterpret_cast<const float4 *>(bias_n);
#pragma unroll
    for (int j = 0; j < 8; j++) {
        const float4 bb = __ldg(b4 + j);
        v[2 * j + 0] = __fadd2_rn(make_float2(__uint_as_float(acc[4 * j + 0]), __uint_as_float(acc[4 * j + 1])), make_float2(bb.x, bb.y));
        v[2 * j + 1] = __fadd2_rn(make_float2(__uint_as_float(acc[4 * j + 2]), __uint_as_float(acc[4 * j + 3])), make_float2(bb.z, bb.w));
    }
}
// GELU through the reference's f16 table semantics (ggml.c:1418-1441) and, for BK_EPI_BIAS_GELU_Q8_0 / _Q8_1, the Q8 rounding of the
// 32 values (= one quantisation block of the next GEMM's input row, ggml.c:9493-9509): the quantiser sees the f16-rounded GELU value
__device__ __forceinline__ void epi_gelu_q8(float2 (&v)[16], int epi) {
#pragma unroll
    for (int j = 0; j < 16; j++) v[j] = gelu_fast2_unrounded(v[j]);
    if (epi != BK_EPI_BIAS_GELU) {
        float amax = 0.f;
#pragma unroll
        for (int j = 0; j < 16; j++) {
            v[j] = unpack_h2(pack_h2(v[j].x, v[j].y));
            amax = fmaxf(amax, fmaxf(fabsf(v[j].x), fabsf(v[j].y)));
        }
        float id, dq;
        q8_scales(amax, epi == BK_EPI_BIAS_GELU_Q8_0 ? 1 : 2, id, dq);
#pragma unroll
        for (int j = 0; j < 16; j++) v[j] = q8_round2(v[j], id, dq);
    }
}
// 32 values -> f16 -> this row's 64 bytes of a SWIZZLE_128B staging box (row = the box row's first byte; half = which 32 columns)
__device__ __forceinline__ void epi_stage_f16(const float2 (&v)[16], uint8_t *row, int half, uint32_t sw) {
#pragma unroll
    for (int j = 0; j < 4; j++) {
        uint4 u;
        u.x = pack_h2(v[4 * j + 0].x, v[4 * j + 0].y);
        u.y = pack_h2(v[4 * j + 1].x, v[4 * j + 1].y);
        u.z = pack_h2(v[4 * j + 2].x, v[4 * j + 2].y);
        u.w = pack_h2(v[4 * j + 3].x, v[4 * j + 3].y);
        *reinterpret_cast<uint4 *>(row + ((((uint32_t)(half * 4 + j)) ^ sw) << 4)) = u;
    }
}

}  // namespace bk
