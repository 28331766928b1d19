// rowwise.cu -- the HBM-bound kernels of the encoder, one warp per token row, warp-shuffle
// reductions, 16-byte vector accesses:
//   batch_sort/pack   the on-device batcher: sentences sorted by length, packed; cu_seqlens, positions, output order
//   embed_ln_kernel   gather(word,type,pos) + add + LayerNorm            (bert.cpp:868-898)
//   add_ln_kernel     LayerNorm of the pre-LN sum the GEMM epilogue wrote (bert.cpp:958-963, 981-992)
//   pool_norm_kernel  mean over the sentence's tokens + L2 normalise      (bert.cpp:995-1006)
//   weight_to_f16     GGUF storage (F32/F16/Q4_0/Q4_1) -> f16 UMMA operand copy at load time
// LayerNorm follows ggml_compute_forward_norm_f32 (ggml.c:8976-9004): two-pass mean / biased
// variance, y = (x-mean) * (1/sqrtf(var+eps)), then *gamma, +beta as separate roundings.
// For Q4 models the LN output is also written Q8_0/Q8_1-rounded per 32 elements (what
// quantize_row_q8_0/1 feed the reference's integer dot products, ggml-quants.c:702-747, 944-985).
#include "common.cuh"
#include "kernels.h"
#include "../../include/bk_api.h"

namespace bk {

constexpr int kMaxChunks = 8;  // 8 halves per chunk, 32 lanes -> d <= 2048

// ------------------------------------------------------------------------------------- batcher
// The on-device batcher.  Input: the sentences of one micro-batch in the CALLER'S order -- ids concatenated (tok_in) and
// lens[n_seq].  batch_sort_kernel (one CTA): stable counting sort of the sentences by length, longest first (keys 1..512:
// histogram -> descending bin starts -> ranks), then for the sorted order s: order[s] = caller index, cu[s+1] = packed token
// offset (prefix sum of the sorted lengths), src_off[s] = where that sentence starts in tok_in.  batch_pack_kernel (all
// CTAs): packed token t -> its sentence s (binary search over cu), pos[t] = t - cu[s], tok[t] = tok_in[src_off[s] + pos].
// Everything downstream works on the packed, length-sorted layout (the attention kernels take the sentences of more than 128
// tokens as a prefix); pool_norm writes sentence s to output row order[s], so results leave in the caller's order.
// The reference has no batcher: bert_eval_batch loops over the sentences one graph at a time (bert.cpp:1065).
constexpr int kMaxLen = 512;

// inclusive block scan of one value per thread (1024 threads); returns the inclusive prefix, *total = sum over the block
__device__ __forceinline__ int block_scan_incl(int v, int32_t *warp_tot, int *total) {
    int incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int n = __shfl_up_sync(0xffffffffu, incl, o);
        if ((int)lane_id() >= o) incl += n;
    }
    __syncthreads();  // warp_tot reuse
    if (lane_id() == 31) warp_tot[threadIdx.x >> 5] = incl;
    __syncthreads();
    if (threadIdx.x < 32) {
        const int w = (threadIdx.x < (blockDim.x >> 5)) ? warp_tot[threadIdx.x] : 0;
        int wi = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int n = __shfl_up_sync(0xffffffffu, wi, o);
            if ((int)lane_id() >= o) wi += n;
        }
        warp_tot[threadIdx.x] = wi - w;  // exclusive warp offsets
        if (threadIdx.x == 31) warp_tot[32] = wi;
    }
    __syncthreads();
    *total = warp_tot[32];
    return incl + warp_tot[threadIdx.x >> 5];
}

__global__ void __launch_bounds__(1024) batch_sort_kernel(const int32_t *__restrict__ lens, int n_seq, int32_t *__restrict__ order,
                                                          int32_t *__restrict__ cu, int32_t *__restrict__ src_off,
                                                          int32_t *__restrict__ caller_off) {
    pdl_launch_dependents();
    pdl_wait();
    __shared__ int32_t cursor[kMaxLen + 2];  // histogram, then running output position per length
    __shared__ int32_t warp_tot[33];
    const int tid = threadIdx.x;
    if (n_seq == 1) {  // latency path: nothing to sort
        if (tid == 0) { order[0] = 0; src_off[0] = 0; cu[0] = 0; cu[1] = min(max(lens[0], 1), kMaxLen); }
        return;
    }
    for (int k = tid; k < kMaxLen + 2; k += blockDim.x) cursor[k] = 0;
    __syncthreads();
    // histogram of lengths + caller-order token offsets (exclusive scan of lens in the caller's order)
    int carry = 0;
    for (int base = 0; base < n_seq; base += blockDim.x) {
        const int i = base + tid;
        const int v = (i < n_seq) ? min(max(lens[i], 1), kMaxLen) : 0;
        if (i < n_seq) atomicAdd(&cursor[v], 1);
        int tot;
        const int incl = block_scan_incl(v, warp_tot, &tot);
        if (i < n_seq) caller_off[i] = carry + incl - v;
        carry += tot;
    }
    __syncthreads();
    // descending bin starts: start[len] = number of sentences longer than len  (bins kMaxLen .. 1 scanned in that order)
    {
        const int len = kMaxLen - tid;  // thread 0 <-> longest
        const int v = (len >= 1) ? cursor[len] : 0;
        int tot;
        const int incl = block_scan_incl(v, warp_tot, &tot);
        __syncthreads();
        if (len >= 1) cursor[len] = incl - v;
    }
    __syncthreads();
    // stable ranks: chunks of 1024 sentences, the warps of a chunk take their turn in order; inside a warp the lanes with the same
    // length are ranked by lane (match_any), the lowest of them advances the bin's cursor
    for (int base = 0; base < n_seq; base += blockDim.x) {
        const int i = base + tid;
        const bool ok = i < n_seq;
        const int key = ok ? min(max(lens[i], 1), kMaxLen) : 0;  // key 0: idle lanes group among themselves
        const unsigned peers = __match_any_sync(0xffffffffu, key);
        const int rank = __popc(peers & ((1u << lane_id()) - 1u));
        const int cnt = __popc(peers);
        const int n_warps = (min(n_seq - base, (int)blockDim.x) + 31) >> 5;
        for (int w = 0; w < n_warps; w++) {
            if ((tid >> 5) == w && ok) {
                const int dst = cursor[key] + rank;
                order[dst] = i;
                __syncwarp(peers);
                if (rank == 0) cursor[key] += cnt;
            }
            __syncthreads();
        }
    }
    __threadfence_block();
    __syncthreads();
    // packed offsets in sorted order
    if (tid == 0) cu[0] = 0;
    carry = 0;
    for (int base = 0; base < n_seq; base += blockDim.x) {
        const int s = base + tid;
        int v = 0, src = 0;
        if (s < n_seq) {
            src = order[s];
            v = min(max(lens[src], 1), kMaxLen);
        }
        int tot;
        const int incl = block_scan_incl(v, warp_tot, &tot);
        if (s < n_seq) {
            cu[s + 1] = carry + incl;
            src_off[s] = caller_off[src];
        }
        carry += tot;
    }
}
__global__ void batch_pack_kernel(const int32_t *__restrict__ tok_in, const int32_t *__restrict__ cu, const int32_t *__restrict__ src_off,
                                  int n_seq, int32_t *__restrict__ tok, int32_t *__restrict__ pos, int T) {
    pdl_launch_dependents();
    pdl_wait();
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= T) return;
    int lo = 0, hi = n_seq;  // find s with cu[s] <= t < cu[s+1]
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (__ldg(cu + mid) <= t) lo = mid; else hi = mid;
    }
    const int p = t - __ldg(cu + lo);
    pos[t] = p;
    tok[t] = __ldg(tok_in + __ldg(src_off + lo) + p);
}
cudaError_t launch_batcher(const BatcherParams &b, cudaStream_t st) {
    {
        LaunchConfig lc(dim3(1), dim3(1024), 0, st);
        cudaError_t err = cudaLaunchKernelEx(&lc.cfg, batch_sort_kernel, b.lens, b.n_seq, b.order, b.cu, b.src_off, b.caller_off);
        if (err != cudaSuccess) return err;
    }
    {
        LaunchConfig lc(dim3((unsigned)((b.T + 255) / 256)), dim3(256), 0, st);
        cudaError_t err = cudaLaunchKernelEx(&lc.cfg, batch_pack_kernel, b.tok_in, (const int32_t *)b.cu, (const int32_t *)b.src_off, b.n_seq, b.tok,
                                             b.pos, b.T);
        if (err != cudaSuccess) return err;
    }
    return cudaGetLastError();
}

// ----------------------------------------------------------------------- table row decoding
// 8 consecutive elements [8c, 8c+8) of row `r` of a [rows][d] table in GGUF storage
// (ggml_compute_forward_get_rows_*, ggml.c:10108-10224; Q4 decode ggml-quants.c:1074-1113).
__device__ __forceinline__ void load_row8(const void *tab, int wtype, int64_t r, int d, int c, float (&v)[8]) {
    if (wtype == BK_F16) {
        const uint4 u = __ldg(reinterpret_cast<const uint4 *>(reinterpret_cast<const __half *>(tab) + r * d) + c);
        float2 f;
        f = unpack_h2(u.x); v[0] = f.x; v[1] = f.y;
        f = unpack_h2(u.y); v[2] = f.x; v[3] = f.y;
        f = unpack_h2(u.z); v[4] = f.x; v[5] = f.y;
        f = unpack_h2(u.w); v[6] = f.x; v[7] = f.y;
    } else if (wtype == BK_F32) {
        const float4 *p = reinterpret_cast<const float4 *>(reinterpret_cast<const float *>(tab) + r * d) + 2 * c;
        const float4 a = __ldg(p), b = __ldg(p + 1);
        v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
    } else {
        // block of 32 = f16 d (+ f16 m for Q4_1) + 16 bytes; element j<16 low nibble of byte j, j+16 high nibble
        const int bb = (wtype == BK_Q4_0) ? 18 : 20;
        const int blk = c >> 2, o = (c & 3) * 8;
        const uint8_t *bp = reinterpret_cast<const uint8_t *>(tab) + (r * (d >> 5) + blk) * bb;
        const uint16_t *hp = reinterpret_cast<const uint16_t *>(bp);  // blocks are 2-byte aligned
        const float dd = __half2float(__ushort_as_half(__ldg(hp)));
        const float mm = (wtype == BK_Q4_1) ? __half2float(__ushort_as_half(__ldg(hp + 1))) : 0.f;
        const uint16_t *qp = reinterpret_cast<const uint16_t *>(bp + (bb - 16) + (o & 15));
        const int sh = (o >= 16) ? 4 : 0;
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const uint32_t w = __ldg(qp + j);
            const int q0 = (w >> sh) & 0xF, q1 = (w >> (8 + sh)) & 0xF;
            if (wtype == BK_Q4_0) { v[2 * j] = __fmul_rn(small_int_to_float(q0, 8.f), dd); v[2 * j + 1] = __fmul_rn(small_int_to_float(q1, 8.f), dd); }
            else { v[2 * j] = fmaf(small_int_to_float(q0, 0.f), dd, mm); v[2 * j + 1] = fmaf(small_int_to_float(q1, 0.f), dd, mm); }
        }
    }
}

// --------------------------------------------------------------- LN core on registers (one warp)
// v[it][8] holds chunk c = lane + 32*it of the row (valid iff c < d/8).
template <bool kWriteQ, int NIT>
__device__ __forceinline__ void ln_store(float (&v)[NIT][8], int d, float eps, const float *__restrict__ gamma,
                                         const float *__restrict__ beta, __half *__restrict__ y, __half *__restrict__ yq, int qmode) {
    const int lane = lane_id();
    const int nch = d >> 3;
    float s = 0.f;
#pragma unroll
    for (int it = 0; it < NIT; it++)
        if (lane + 32 * it < nch)
#pragma unroll
            for (int j = 0; j < 8; j++) s += v[it][j];
    const float mean = warp_sum(s) / (float)d;
    float s2 = 0.f;
#pragma unroll
    for (int it = 0; it < NIT; it++)
        if (lane + 32 * it < nch)
#pragma unroll
            for (int j = 0; j < 8; j++) { const float t = v[it][j] - mean; v[it][j] = t; s2 += t * t; }
    const float var = warp_sum(s2) / (float)d;
    const float scale = 1.0f / sqrtf(var + eps);
#pragma unroll
    for (int it = 0; it < NIT; it++) {
        const int c = lane + 32 * it;
        const bool ok = c < nch;
        float o[8];
        if (ok) {
            const float4 g0 = __ldg(reinterpret_cast<const float4 *>(gamma) + 2 * c), g1 = __ldg(reinterpret_cast<const float4 *>(gamma) + 2 * c + 1);
            const float4 b0 = __ldg(reinterpret_cast<const float4 *>(beta) + 2 * c), b1 = __ldg(reinterpret_cast<const float4 *>(beta) + 2 * c + 1);
            const float g[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
            const float b[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
            for (int j = 0; j < 8; j++) o[j] = __fadd_rn(__fmul_rn(g[j], __fmul_rn(v[it][j], scale)), b[j]);  // three roundings: norm, ggml_mul, ggml_add
            uint4 u;
            u.x = pack_h2(o[0], o[1]); u.y = pack_h2(o[2], o[3]); u.z = pack_h2(o[4], o[5]); u.w = pack_h2(o[6], o[7]);
            reinterpret_cast<uint4 *>(y)[c] = u;
        } else {
#pragma unroll
            for (int j = 0; j < 8; j++) o[j] = 0.f;
        }
        if (kWriteQ) {
            // one Q8 block = 32 elements = chunks 4b..4b+3 = four adjacent lanes of this iteration
            float amax = 0.f;
#pragma unroll
            for (int j = 0; j < 8; j++) amax = fmaxf(amax, fabsf(o[j]));
            amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 1));
            amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 2));
            if (ok) {
                float id, dq;
                q8_scales(amax, qmode, id, dq);
                uint4 u;
                u.x = pack_h2(q8_round(o[0], id, dq), q8_round(o[1], id, dq));
                u.y = pack_h2(q8_round(o[2], id, dq), q8_round(o[3], id, dq));
                u.z = pack_h2(q8_round(o[4], id, dq), q8_round(o[5], id, dq));
                u.w = pack_h2(q8_round(o[6], id, dq), q8_round(o[7], id, dq));
                reinterpret_cast<uint4 *>(yq)[c] = u;
            }
        }
    }
}

// ------------------------------------------------------------------------------- embed + LN
// The gather is latency bound, not byte bound (ncu r01e: DRAM 20 % busy, 25 warps resident, two dependent round trips per
// token: id -> row): persistent warps walk over the tokens with the NEXT token's table rows already requested (raw bytes
// held in registers) and the ids of the one after that on their way, so a warp always has a row in flight while it
// normalises another.  The token-type row (always row 0, bert.cpp:871-872) is decoded once per CTA into shared memory.
template <int WT, int NIT>
struct RawRows {  // undecoded bytes of chunks lane, lane+32, .. of the word row and the position row
    // F16: 16 B per chunk; Q4: scale word (d | m << 16) + the 8 bytes holding the chunk's nibbles; F32: not prefetched
    uint4 w[WT == BK_F16 ? NIT : 1], p[WT == BK_F16 ? NIT : 1];
    uint32_t wd[(WT == BK_Q4_0 || WT == BK_Q4_1) ? NIT : 1], pd[(WT == BK_Q4_0 || WT == BK_Q4_1) ? NIT : 1];
    uint2 wq[(WT == BK_Q4_0 || WT == BK_Q4_1) ? NIT : 1], pq[(WT == BK_Q4_0 || WT == BK_Q4_1) ? NIT : 1];
};
template <int WT>
__device__ __forceinline__ void q4_raw(const void *tab, int64_t r, int d, int c, uint32_t &dm, uint2 &qs) {
    constexpr int bb = (WT == BK_Q4_0) ? 18 : 20;
    const int blk = c >> 2, o = (c & 3) * 8;
    const uint8_t *bp = reinterpret_cast<const uint8_t *>(tab) + (r * (d >> 5) + blk) * bb;
    const uint16_t *hp = reinterpret_cast<const uint16_t *>(bp);  // blocks are 2-byte aligned
    dm = __ldg(hp);
    if (WT == BK_Q4_1) dm |= (uint32_t)__ldg(hp + 1) << 16;
    const uint16_t *qp = reinterpret_cast<const uint16_t *>(bp + (bb - 16) + (o & 15));
    qs.x = (uint32_t)__ldg(qp) | ((uint32_t)__ldg(qp + 1) << 16);
    qs.y = (uint32_t)__ldg(qp + 2) | ((uint32_t)__ldg(qp + 3) << 16);
}
template <int WT>
__device__ __forceinline__ void q4_decode8(uint32_t dm, uint2 qs, int c, float (&v)[8]) {
    // element j < 16 of a block is the low nibble of byte j, element j + 16 the high nibble (ggml-quants.c:1074-1113)
    const float dd = __half2float(__ushort_as_half((unsigned short)(dm & 0xFFFF)));
    const float mm = (WT == BK_Q4_1) ? __half2float(__ushort_as_half((unsigned short)(dm >> 16))) : 0.f;
    const int sh = ((c & 3) * 8 >= 16) ? 4 : 0;
    const uint32_t w2[2] = {qs.x >> sh, qs.y >> sh};
#pragma unroll
    for (int j = 0; j < 8; j++) {
        const int q = (int)((w2[j >> 2] >> (8 * (j & 3))) & 0xF);
        v[j] = (WT == BK_Q4_0) ? __fmul_rn(small_int_to_float(q, 8.f), dd) : fmaf(small_int_to_float(q, 0.f), dd, mm);
    }
}
template <int WT, int NIT>
__device__ __forceinline__ void load_raw(const EmbedParams &p, int tok, int ps, int lane, int nch, RawRows<WT, NIT> &r) {
#pragma unroll
    for (int it = 0; it < NIT; it++) {
        const int c = lane + 32 * it;
        if (c < nch) {
            if (WT == BK_F16) {
                r.w[it] = __ldg(reinterpret_cast<const uint4 *>(reinterpret_cast<const __half *>(p.word) + (int64_t)tok * p.d) + c);
                r.p[it] = __ldg(reinterpret_cast<const uint4 *>(reinterpret_cast<const __half *>(p.posw) + (int64_t)ps * p.d) + c);
            } else if (WT == BK_Q4_0 || WT == BK_Q4_1) {
                q4_raw<WT>(p.word, tok, p.d, c, r.wd[it], r.wq[it]);
                q4_raw<WT>(p.posw, ps, p.d, c, r.pd[it], r.pq[it]);
            }
        }
    }
}
template <int WT, bool kWriteQ, int NIT>
__global__ void __launch_bounds__(256) embed_ln_kernel(const EmbedParams p) {
    pdl_launch_dependents();
    extern __shared__ float s_type[];  // [d] token-type row 0, decoded
    const int lane = lane_id();
    const int nch = p.d >> 3;
    for (int c = threadIdx.x; c < nch; c += blockDim.x) {  // weights: independent of the previous kernel
        float ty[8];
        load_row8(p.type, WT, 0, p.d, c, ty);
#pragma unroll
        for (int j = 0; j < 8; j++) s_type[c * 8 + j] = ty[j];
    }
    pdl_wait();
    __syncthreads();
    const int wpg = gridDim.x * (blockDim.x >> 5);
    int t = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (t >= p.T) return;
    auto ids = [&](int tt, int &tok, int &ps) {
        tok = __ldg(p.tokens + tt);
        tok = min(max(tok, 0), p.n_vocab - 1);
        ps = __ldg(p.pos + tt);
    };
    int tok, ps, tok_n = 0, ps_n = 0;
    ids(t, tok, ps);
    RawRows<WT, NIT> cur, nxt;
    load_raw<WT, NIT>(p, tok, ps, lane, nch, cur);
    int t_n = t + wpg;
    if (t_n < p.T) ids(t_n, tok_n, ps_n);
    for (;;) {
        const bool more = t_n < p.T;
        int tok_nn = 0, ps_nn = 0;
        if (more) {
            load_raw<WT, NIT>(p, tok_n, ps_n, lane, nch, nxt);   // next token's rows: in flight while this one is normalised
            if (t_n + wpg < p.T) ids(t_n + wpg, tok_nn, ps_nn);
        }
        float v[NIT][8];
#pragma unroll
        for (int it = 0; it < NIT; it++) {
            const int c = lane + 32 * it;
            if (c < nch) {
                float w[8], po[8];
                if (WT == BK_F16) {
                    float2 f;
                    f = unpack_h2(cur.w[it].x); w[0] = f.x; w[1] = f.y; f = unpack_h2(cur.w[it].y); w[2] = f.x; w[3] = f.y;
                    f = unpack_h2(cur.w[it].z); w[4] = f.x; w[5] = f.y; f = unpack_h2(cur.w[it].w); w[6] = f.x; w[7] = f.y;
                    f = unpack_h2(cur.p[it].x); po[0] = f.x; po[1] = f.y; f = unpack_h2(cur.p[it].y); po[2] = f.x; po[3] = f.y;
                    f = unpack_h2(cur.p[it].z); po[4] = f.x; po[5] = f.y; f = unpack_h2(cur.p[it].w); po[6] = f.x; po[7] = f.y;
                } else if (WT == BK_Q4_0 || WT == BK_Q4_1) {
                    q4_decode8<WT>(cur.wd[it], cur.wq[it], c, w);
                    q4_decode8<WT>(cur.pd[it], cur.pq[it], c, po);
                } else {
                    load_row8(p.word, WT, tok, p.d, c, w);
                    load_row8(p.posw, WT, ps, p.d, c, po);
                }
                const float4 t0 = *reinterpret_cast<const float4 *>(s_type + c * 8), t1 = *reinterpret_cast<const float4 *>(s_type + c * 8 + 4);
                const float ty[8] = {t0.x, t0.y, t0.z, t0.w, t1.x, t1.y, t1.z, t1.w};
#pragma unroll
                for (int j = 0; j < 8; j++) { const float a = ty[j] + w[j]; v[it][j] = po[j] + a; }  // (type + word) + pos, bert.cpp:880-887
            }
        }
        ln_store<kWriteQ, NIT>(v, p.d, p.eps, p.gamma, p.beta, p.x + (size_t)t * p.d, kWriteQ ? p.xq + (size_t)t * p.d : nullptr, p.qmode);
        if (!more) break;
        t = t_n; tok = tok_n; ps = ps_n;
        t_n += wpg; tok_n = tok_nn; ps_n = ps_nn;
        cur = nxt;
    }
}
// resident CTAs of one instantiation on the current device (queried once): the persistent grid is exactly one wave
template <int WT, bool Q, int NIT>
static int embed_ln_resident(size_t smem) {
    static int n = 0;
    if (n == 0) {
        int dev = 0, sms = 148, per = 1;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per, embed_ln_kernel<WT, Q, NIT>, 256, smem) != cudaSuccess || per < 1) per = 1;
        n = sms * per;
    }
    return n;
}
cudaError_t launch_embed_ln(const EmbedParams &p, cudaStream_t st) {
    if (p.d % 32 != 0 || p.d > kMaxChunks * 256) return cudaErrorInvalidValue;
    const int wpb = 8;
    const int want = (p.T + wpb - 1) / wpb;
    const int nit = (p.d + 255) / 256;  // 16-byte chunks per lane
    const size_t smem = (size_t)p.d * sizeof(float);
#define BK_EMBED_WT(WT_, NIT)                                                                          \
    do { /* more tokens than one wave of warps: persistent warps, each walks over T / (grid * 8) tokens */ \
        const int cap = p.qmode ? embed_ln_resident<WT_, true, NIT>(smem) : embed_ln_resident<WT_, false, NIT>(smem); \
        LaunchConfig lc(dim3((unsigned)(want < cap ? want : cap)), dim3((unsigned)(wpb * 32)), smem, st); \
        if (p.qmode) cudaLaunchKernelEx(&lc.cfg, embed_ln_kernel<WT_, true, NIT>, p);                  \
        else cudaLaunchKernelEx(&lc.cfg, embed_ln_kernel<WT_, false, NIT>, p);                         \
    } while (0)
#define BK_EMBED(NIT)                                                                                  \
    do {                                                                                               \
        switch (p.wtype) {                                                                             \
            case BK_F32: BK_EMBED_WT(BK_F32, NIT); break;                                              \
            case BK_F16: BK_EMBED_WT(BK_F16, NIT); break;                                              \
            case BK_Q4_0: BK_EMBED_WT(BK_Q4_0, NIT); break;                                            \
            default: BK_EMBED_WT(BK_Q4_1, NIT); break;                                                 \
        }                                                                                              \
    } while (0)
    if (nit <= 1) BK_EMBED(1);
    else if (nit == 2) BK_EMBED(2);
    else if (nit == 3) BK_EMBED(3);
    else if (nit == 4) BK_EMBED(4);
    else BK_EMBED(kMaxChunks);
#undef BK_EMBED
#undef BK_EMBED_WT
    return cudaGetLastError();
}

// --------------------------------------------------------------------------------- add + LN
template <bool kWriteQ, int NIT>
__global__ void __launch_bounds__(256) add_ln_kernel(const __half *__restrict__ pre, const float *__restrict__ gamma,
                                                     const float *__restrict__ beta, int T, int d, float eps,
                                                     __half *__restrict__ y, __half *__restrict__ yq, int qmode) {
    pdl_launch_dependents();
    pdl_wait();
    const int t = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (t >= T) return;
    const int lane = lane_id();
    const int nch = d >> 3;
    const uint4 *src = reinterpret_cast<const uint4 *>(pre + (size_t)t * d);
    float v[NIT][8];
    uint4 u[NIT];
#pragma unroll
    for (int it = 0; it < NIT; it++)
        if (lane + 32 * it < nch) u[it] = __ldg(src + lane + 32 * it);
#pragma unroll
    for (int it = 0; it < NIT; it++)
        if (lane + 32 * it < nch) {
            float2 f;
            f = unpack_h2(u[it].x); v[it][0] = f.x; v[it][1] = f.y;
            f = unpack_h2(u[it].y); v[it][2] = f.x; v[it][3] = f.y;
            f = unpack_h2(u[it].z); v[it][4] = f.x; v[it][5] = f.y;
            f = unpack_h2(u[it].w); v[it][6] = f.x; v[it][7] = f.y;
        }
    ln_store<kWriteQ, NIT>(v, d, eps, gamma, beta, y + (size_t)t * d, kWriteQ ? yq + (size_t)t * d : nullptr, qmode);
}
cudaError_t launch_add_ln(const __half *pre, const float *gamma, const float *beta, int T, int d, float eps, __half *y,
                          __half *yq, int qmode, cudaStream_t st) {
    if (d % 32 != 0 || d > kMaxChunks * 256) return cudaErrorInvalidValue;
    const int wpb = 8;
    const int grid = (T + wpb - 1) / wpb;
    const int nit = (d + 255) / 256;  // 16-byte chunks per lane
#define BK_ADDLN(NIT)                                                                                          \
    do {                                                                                                       \
        LaunchConfig lc(dim3((unsigned)grid), dim3((unsigned)(wpb * 32)), 0, st);                                     \
        if (qmode) cudaLaunchKernelEx(&lc.cfg, add_ln_kernel<true, NIT>, pre, gamma, beta, T, d, eps, y, yq, qmode); \
        else cudaLaunchKernelEx(&lc.cfg, add_ln_kernel<false, NIT>, pre, gamma, beta, T, d, eps, y, yq, 0);         \
    } while (0)
    if (nit <= 1) BK_ADDLN(1);
    else if (nit == 2) BK_ADDLN(2);
    else if (nit == 3) BK_ADDLN(3);
    else if (nit == 4) BK_ADDLN(4);
    else BK_ADDLN(kMaxChunks);
#undef BK_ADDLN
    return cudaGetLastError();
}

// ------------------------------------------------------------------------ mean-pool + L2 norm
// One CTA per sentence.  e[k] = sum_t x[t][k] * (1/L) (the reference multiplies by a vector of 1/L,
// bert.cpp:999-1001), out = e * (1/sqrt(sum e^2)) (bert.cpp:1004-1006).
__global__ void __launch_bounds__(512) pool_norm_kernel(const __half *__restrict__ x, const int32_t *__restrict__ cu,
                                                         const int32_t *__restrict__ order, int d, float *__restrict__ out) {
    pdl_launch_dependents();
    pdl_wait();
    extern __shared__ float sacc[];  // [ngroups][d]
    __shared__ float red[32];
    const int s = blockIdx.x;
    const int t0 = __ldg(cu + s), t1 = __ldg(cu + s + 1);
    const int L = t1 - t0;
    const int nch = d >> 3;
    const int nthr = (int)blockDim.x;
    // thread layout: column-chunk c = threadIdx.x % cpr, row group g = threadIdx.x / cpr (the groups split the sentence's tokens)
    const int cpr = nch < nthr ? nch : nthr;  // chunks handled per pass
    const int groups = nthr / cpr > 0 ? nthr / cpr : 1;
    const int g = threadIdx.x / cpr, c0 = threadIdx.x % cpr;
    const float invL = 1.0f / (float)L;
    for (int c = c0; c < nch; c += cpr) {
        float a[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        if (g < groups) {
#pragma unroll 8  // the loads of 8 tokens in flight together (same summation order): a one-sentence launch is latency bound
            for (int t = t0 + g; t < t1; t += groups) {
                const uint4 u = __ldg(reinterpret_cast<const uint4 *>(x + (size_t)t * d) + c);
                float2 f;
                f = unpack_h2(u.x); a[0] = fmaf(f.x, invL, a[0]); a[1] = fmaf(f.y, invL, a[1]);
                f = unpack_h2(u.y); a[2] = fmaf(f.x, invL, a[2]); a[3] = fmaf(f.y, invL, a[3]);
                f = unpack_h2(u.z); a[4] = fmaf(f.x, invL, a[4]); a[5] = fmaf(f.y, invL, a[5]);
                f = unpack_h2(u.w); a[6] = fmaf(f.x, invL, a[6]); a[7] = fmaf(f.y, invL, a[7]);
            }
#pragma unroll
            for (int j = 0; j < 8; j++) sacc[(size_t)g * d + c * 8 + j] = a[j];
        }
    }
    __syncthreads();
    float ss = 0.f;
    for (int k = threadIdx.x; k < d; k += blockDim.x) {
        float e = 0.f;
        for (int gg = 0; gg < groups; gg++) e += sacc[(size_t)gg * d + k];
        sacc[k] = e;  // group 0 slot now holds the mean (each k is touched by one thread only)
        ss += e * e;
    }
    ss = warp_sum(ss);
    if (lane_id() == 0) red[threadIdx.x >> 5] = ss;
    __syncthreads();
    float tot = 0.f;
    for (int w = 0; w < (nthr >> 5); w++) tot += red[w];
    const float inv = 1.0f / sqrtf(tot);
    const int row = order ? __ldg(order + s) : s;  // the batcher's sorted position -> the caller's sentence index
    for (int k = threadIdx.x; k < d; k += blockDim.x) out[(size_t)row * d + k] = sacc[k] * inv;
}
cudaError_t launch_pool_norm(const __half *x, const int32_t *cu_seqlens, const int32_t *order, int n_seq, int d, float *out, cudaStream_t st) {
    if (d % 8 != 0) return cudaErrorInvalidValue;
    // threads per sentence (a multiple of 32): the sentence's tokens are split over threads / (d / 8) row groups.
    // BERT_B200_POOL_THREADS overrides (tuning); the summation order, hence the bits of the result, depend on it
    static const int env_thr = [] { const char *v = getenv("BERT_B200_POOL_THREADS"); return v ? atoi(v) : 0; }();
    int nthr = env_thr >= 32 && env_thr <= 512 ? (env_thr & ~31) : 512;  // measured on 512 sentences x 128 tokens: 256 threads 32 us, 512 threads 30 us
    const int nch = d >> 3;
    const int cpr = nch < nthr ? nch : nthr;
    const int groups = nthr / cpr > 0 ? nthr / cpr : 1;
    const size_t smem = (size_t)groups * d * sizeof(float);
    LaunchConfig lc(dim3((unsigned)n_seq), dim3((unsigned)nthr), smem, st);
    cudaLaunchKernelEx(&lc.cfg, pool_norm_kernel, x, cu_seqlens, order, d, out);
    return cudaGetLastError();
}

// ------------------------------------------------------------------- weight storage -> f16
__global__ void weight_to_f16_kernel(const void *__restrict__ src, int wtype, int64_t rows, int K, __half *__restrict__ dst) {
    const int64_t nch = rows * (int64_t)(K >> 3);
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < nch; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = i / (K >> 3);
        const int c = (int)(i % (K >> 3));
        float v[8];
        load_row8(src, wtype, r, K, c, v);
        uint4 u;
        u.x = pack_h2(v[0], v[1]); u.y = pack_h2(v[2], v[3]); u.z = pack_h2(v[4], v[5]); u.w = pack_h2(v[6], v[7]);
        reinterpret_cast<uint4 *>(dst)[i] = u;
    }
}
cudaError_t launch_weight_to_f16(const void *src, int wtype, int64_t rows, int64_t K, __half *dst, cudaStream_t st) {
    if (K % 32 != 0) return cudaErrorInvalidValue;
    const int64_t nch = rows * (K >> 3);
    int grid = (int)((nch + 255) / 256);
    if (grid > 148 * 16) grid = 148 * 16;
    weight_to_f16_kernel<<<grid, 256, 0, st>>>(src, wtype, rows, (int)K, dst);
    return cudaGetLastError();
}

}  // namespace bk
