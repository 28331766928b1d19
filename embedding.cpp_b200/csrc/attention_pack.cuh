// attention_pack.cuh -- attention for SHORT sentences (<= 64 tokens): several sentences share one 128-row tile.
// Included by attention_tc.cu, inside namespace bk, after attention_loop.cuh.
//
// The looping kernel (attention_loop.cuh) spends a whole 128 x 128 tile -- loads, S = Q.K^T, a P tile, P.V, the per-item
// dependency chain -- on every (sentence, head), whatever the sentence's length: 16-token sentences use 1/8 of the rows and cost
// 2.2x as much per token as 128-token ones (286 vs 128 us per 65 536 tokens).  Real embedding traffic is mostly short texts.
// Here a tile holds 2 / 4 / 8 sentences of the same length class in slots of w = 64 / 32 / 16 rows (lengths 33..64 / 17..32 /
// 1..16; the batcher has sorted the sentences by length, so a class is a contiguous run).  Each slot is loaded by its own TMA box
// (w rows of Q, K, V starting at the sentence's first token), S is the full 128 x 128 block, and a query row only looks at the
// keys of its own slot: the other scores are never read, their P entries are written as exact zeros.
//
// Bit-identical to the looping kernel: a row's scores, maximum and exponentials are the same numbers; slots start on multiples of
// 16 keys, so the sentence's keys fall into the same 16-key MMA steps as when it has a tile to itself, and the steps of the other
// slots add exact zeros (p = 0 times finite V) to the fp32 accumulators of O and of the row sum.
// Same structure as attention_loop.cuh otherwise: 4 CTAs per SM walking over items (tile, head), set-up once per CTA, next item's
// loads requested when the current P.V MMAs have retired.
#pragma once

struct PackCounts { int n64, n32, n16; };  // sentences of 33..64 / 17..32 / 1..16 tokens, in this order from cu[0] on
__host__ __device__ inline int pack_tiles(const PackCounts &c) { return (c.n64 + 1) / 2 + (c.n32 + 3) / 4 + (c.n16 + 7) / 8; }
// tile -> slot width, first sentence, number of sentences
__device__ __forceinline__ void pack_tile(const PackCounts &c, int tile, int &w, int &s0, int &ns) {
    const int t64 = (c.n64 + 1) / 2, t32 = (c.n32 + 3) / 4;
    if (tile < t64) { w = 64; s0 = 2 * tile; ns = min(2, c.n64 - s0); }
    else if (tile < t64 + t32) { w = 32; const int i = 4 * (tile - t64); s0 = c.n64 + i; ns = min(4, c.n32 - i); }
    else { w = 16; const int i = 8 * (tile - t64 - t32); s0 = c.n64 + c.n32 + i; ns = min(8, c.n16 - i); }
}

template <int DH>
__global__ void __launch_bounds__(kLoopThreads, 4)
attention_pack_kernel(const __grid_constant__ CUtensorMap tm64, const __grid_constant__ CUtensorMap tm32, const __grid_constant__ CUtensorMap tm16,
                      const int32_t *__restrict__ cu, const PackCounts pc, int d, int n_head, __half *__restrict__ out, int qmode, float scale) {
    constexpr int ROWB = DH * 2;
    constexpr int CW = 4;
    constexpr int PBUF = kAttQ * 128;
    constexpr int QK_BYTES = 2 * kAttQ * ROWB;
    constexpr int R0 = QK_BYTES > 2 * PBUF ? QK_BYTES : 2 * PBUF;  // Q | K, later the two P chunks
    extern __shared__ __align__(1024) uint8_t att_raw[];
    uint8_t *smem = att_raw;
    uint8_t *sQ = smem, *sK = smem + kAttQ * ROWB, *sP = smem, *sV = smem + R0;
    uint8_t *tail = sV + kAttQ * ROWB;
    uint64_t *bar_qk = reinterpret_cast<uint64_t *>(tail);
    uint64_t *bar_v = bar_qk + 1, *bar_s = bar_qk + 2, *bar_o = bar_qk + 3, *bar_tfree = bar_qk + 4, *pfull = bar_qk + 5;  // pfull[2]
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bar_qk + 7);
    uint32_t *ones = reinterpret_cast<uint32_t *>(bar_qk + 8);  // 512 B
    uint8_t *stage = tail + 64 + 512;                           // [4 warps][1 KB] output staging (store_rows64_1k, attention_loop.cuh)

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n_items = pack_tiles(pc) * n_head;
    pdl_launch_dependents();
    // V (and Q | K) start finite: a slot without a sentence keeps whatever the stage held, and 0 * NaN would poison O
    for (int i = threadIdx.x; i < (R0 + kAttQ * ROWB) / 16; i += kLoopThreads) reinterpret_cast<uint4 *>(smem)[i] = make_uint4(0, 0, 0, 0);
    if (warp == CW) {
        if (lane == 0) {
            tma_prefetch_desc(&tm64); tma_prefetch_desc(&tm32); tma_prefetch_desc(&tm16);
            mbar_init(bar_qk, 1); mbar_init(bar_v, 1); mbar_init(bar_s, 1); mbar_init(bar_o, 1); mbar_init(bar_tfree, 4);
            mbar_init(&pfull[0], 4); mbar_init(&pfull[1], 4);
            mbar_fence_init();
        }
        for (int i = lane; i < 128; i += 32) ones[i] = 0x3C003C00u;
        __syncwarp();
        tmem_alloc<128>(tmem_slot);
    }
    fence_proxy_async_smem();  // the zero fill and the ones tile -> visible to TMA / MMA
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();

    if (warp == CW) {
        // ================================================================ control warp (uniform, elected issue)
        const uint32_t idesc_pv = att_idesc(DH, 1), idesc_sum = att_idesc(kSumCols, 0), idesc_s = att_idesc(128, 0);
        const uint64_t done = umma_desc_ones(smem_u32(ones));
        const uint64_t dq = umma_desc_kmajor<ROWB>(smem_u32(sQ)), dk = umma_desc_kmajor<ROWB>(smem_u32(sK));
        auto load = [&](int it) {  // one box of w rows per slot and operand
            const int tile = it / n_head, head = it - tile * n_head;
            int w, s0, ns;
            pack_tile(pc, tile, w, s0, ns);
            const CUtensorMap *tm = w == 64 ? &tm64 : w == 32 ? &tm32 : &tm16;
            const int cq = head * DH, ck = d + head * DH, cv = 2 * d + head * DH;
            mbar_expect_tx_e(bar_qk, (uint32_t)(2 * ns * w * ROWB));
            mbar_expect_tx_e(bar_v, (uint32_t)(ns * w * ROWB));
            for (int k = 0; k < ns; k++) {
                const int t0 = __ldg(cu + s0 + k);
                tma_load_2d_e(sQ + k * w * ROWB, tm, bar_qk, cq, t0);
                tma_load_2d_e(sK + k * w * ROWB, tm, bar_qk, ck, t0);
                tma_load_2d_e(sV + k * w * ROWB, tm, bar_v, cv, t0);
            }
        };
        if ((int)blockIdx.x < n_items) load((int)blockIdx.x);
        uint32_t k = 0;
        for (int it = (int)blockIdx.x; it < n_items; it += (int)gridDim.x, k++) {
            mbar_wait(bar_qk, k & 1);
            if (k) mbar_wait(bar_tfree, (k - 1) & 1);  // the softmax warps have read the previous item's O and row sums
            tc_fence_after();
#pragma unroll
            for (int kk = 0; kk < DH / 16; kk++) umma_f16_e<1>(tmem_base, dq + (uint64_t)(2 * kk), dk + (uint64_t)(2 * kk), idesc_s, kk != 0);
            umma_commit_e<1>(bar_s);
            mbar_wait(bar_v, k & 1);
            // O += P_c . V_c, rowsum += P_c . 1 over both 64-key chunks (every key of the tile: the other slots' P entries are zeros);
            // the row sums land in score columns [dh, dh + 16): issued once the chunk holding those columns has been consumed
            constexpr int kLag = DH / 64;
            auto issue_sum = [&](int c) {
                const uint64_t dp = umma_desc_sw128(smem_u32(sP + c * PBUF));
                for (int ks = 0; ks < 4; ks++) umma_f16_e<1>(tmem_base + (uint32_t)DH, dp + (uint64_t)(2 * ks), done, idesc_sum, (c | ks) != 0);
            };
            for (int c = 0; c < 2; c++) {
                mbar_wait(&pfull[c], k & 1);
                tc_fence_after();
                const uint64_t dp = umma_desc_sw128(smem_u32(sP + c * PBUF));
                for (int ks = 0; ks < 4; ks++) {
                    const uint64_t dv = umma_desc_mnmajor<ROWB>(smem_u32(sV + (c * 64 + ks * 16) * ROWB));
                    umma_f16_e<1>(tmem_base, dp + (uint64_t)(2 * ks), dv, idesc_pv, (c | ks) != 0);
                }
                if (c >= kLag) issue_sum(c - kLag);
            }
            for (int c = 2 - kLag; c < 2; c++) issue_sum(c);
            umma_commit_e<1>(bar_o);
            if (it + (int)gridDim.x < n_items) {  // next item's loads under this item's epilogue
                mbar_wait(bar_o, k & 1);
                load(it + (int)gridDim.x);
            }
        }
    } else {
        // ======================================================================= softmax warps (one per 32 query rows)
        const int quad = warp & 3;
        const int r = quad * 32 + lane;
        const uint32_t taddr = tmem_base + ((uint32_t)(quad * 32) << 16);
        const uint32_t row_off = (uint32_t)r * 128u, sw = (uint32_t)(r & 7);
        uint32_t k = 0;
        for (int it = (int)blockIdx.x; it < n_items; it += (int)gridDim.x, k++) {
            const int tile = it / n_head, head = it - tile * n_head;
            int w, s0, ns;
            pack_tile(pc, tile, w, s0, ns);
            const int slot = r / w;                       // my sentence's slot; keys [lo, lo + L) of the tile are its tokens
            const bool have = slot < ns;
            const int t0 = have ? __ldg(cu + s0 + slot) : 0;
            const int L = have ? __ldg(cu + s0 + slot + 1) - t0 : 0;
            const int lo = slot * w, qi = r - lo;         // qi: my token inside the sentence
            // the 32-key slices my warp's rows look at: one (w = 16, 32: slice = quad) or two (w = 64: the slot's two halves) -- warp-uniform
            const int j0 = w == 64 ? 2 * (quad >> 1) : quad, nj = w == 64 ? 2 : 1;
            mbar_wait(bar_s, k & 1);
            tc_fence_after();
            // ---- pass 1: row maximum over my sentence's keys
            float mx = -INFINITY;
            for (int jj = 0; jj < nj; jj++) {
                const int j = j0 + jj;
                uint32_t a[32];
                tmem_ld_32x32(taddr + (uint32_t)(j * 32), a);
                tmem_ld_wait();
                const int klo = lo - j * 32, khi = klo + L;   // my keys inside this slice: [klo, khi)
#pragma unroll
                for (int i = 0; i < 32; i++) if (i < klo || i >= khi) a[i] = 0xff800000u;  // -inf: other slots' keys never enter
                float m0 = mx, m1 = -INFINITY;
#pragma unroll
                for (int i = 0; i < 32; i += 4) {
                    m0 = fmaxf(m0, fmaxf(__uint_as_float(a[i]), __uint_as_float(a[i + 1])));
                    m1 = fmaxf(m1, fmaxf(__uint_as_float(a[i + 2]), __uint_as_float(a[i + 3])));
                }
                mx = fmaxf(m0, m1);
            }
            mx = L > 0 ? mx * scale : 0.f;  // rows of an empty slot: every p below becomes 2^-inf = 0
            // ---- pass 2: my full P row (128 keys = both chunks): p for my slices, exact zeros elsewhere
            const float2 sc2 = make_float2(scale, scale), nmx2 = make_float2(-mx, -mx), l2e = make_float2(1.4426950408889634f, 1.4426950408889634f);
#pragma unroll 1
            for (int j = 0; j < 4; j++) {
                uint8_t *pb = sP + (j >> 1) * PBUF;
                const int h = j & 1;
                uint32_t pk[16];
                if (j >= j0 && j < j0 + nj) {
                    uint32_t a[32];
                    tmem_ld_32x32(taddr + (uint32_t)(j * 32), a);
                    tmem_ld_wait();
                    const int klo = lo - j * 32, khi = klo + L;
#pragma unroll
                    for (int i = 0; i < 32; i++) if (i < klo || i >= khi) a[i] = 0xff800000u;  // s = -inf -> p = 0 exactly
#pragma unroll
                    for (int i = 0; i < 32; i += 2) {
                        // ggml_scale, x - max, f16 round | exp | f16 round (ggml.c:10536-10558): the instruction sequence of attention_tc_kernel
                        float2 x;
                        if (DH == 64) x = __ffma2_rn(make_float2(__uint_as_float(a[i]), __uint_as_float(a[i + 1])), sc2, nmx2);
                        else x = __fadd2_rn(make_float2(__fmul_rn(__uint_as_float(a[i]), scale), __fmul_rn(__uint_as_float(a[i + 1]), scale)), nmx2);
                        const float2 z = __fmul2_rn(__half22float2(__floats2half2_rn(x.x, x.y)), l2e);
                        float2 e;
                        asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e.x) : "f"(z.x));
                        asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e.y) : "f"(z.y));
                        const __half2 hp = __floats2half2_rn(e.x, e.y);
                        pk[i >> 1] = *reinterpret_cast<const uint32_t *>(&hp);
                    }
                } else {
#pragma unroll
                    for (int i = 0; i < 16; i++) pk[i] = 0u;
                }
#pragma unroll
                for (int q = 0; q < 4; q++)
                    *reinterpret_cast<uint4 *>(pb + row_off + ((((uint32_t)(h * 4 + q)) ^ sw) << 4)) =
                        make_uint4(pk[4 * q], pk[4 * q + 1], pk[4 * q + 2], pk[4 * q + 3]);
            }
            fence_proxy_async_smem();  // P writes -> visible to the MMA (async proxy)
            tc_fence_before();         // all my TMEM reads of the scores precede the MMAs that overwrite them (O and the row sums alias them)
            __syncwarp();
            if (lane == 0) { mbar_arrive(&pfull[0]); mbar_arrive(&pfull[1]); }
            // ---- epilogue: O * (1 / sum p)
            mbar_wait(bar_o, k & 1);
            tc_fence_after();
            const float sum = __uint_as_float(tmem_ld_32x1(taddr + (uint32_t)DH));
            tmem_ld_wait();
            const float inv = 1.0f / sum;
#pragma unroll 1
            for (int hc = 0; hc < DH / 32; hc++) {
                uint32_t a[32];
                tmem_ld_32x32(taddr + (uint32_t)(hc * 32), a);
                tmem_ld_wait();
                if (hc + 1 == DH / 32) {  // last TMEM read of this item by this warp: the next item's scores may be issued
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(bar_tfree);
                }
                float o[32];
#pragma unroll
                for (int i = 0; i < 32; i++) o[i] = __uint_as_float(a[i]) * inv;
                if (qmode) {
                    float amax = 0.f;
#pragma unroll
                    for (int i = 0; i < 32; i++) amax = fmaxf(amax, fabsf(o[i]));
                    float id, dq8;
                    q8_scales(amax, qmode, id, dq8);
#pragma unroll
                    for (int i = 0; i < 32; i++) o[i] = q8_round(o[i], id, dq8);
                }
                uint32_t pk[16];
#pragma unroll
                for (int i = 0; i < 16; i++) pk[i] = pack_h2(o[2 * i], o[2 * i + 1]);
                store_rows64_1k(stage + quad * 1024, pk, qi < L ? out + (size_t)(t0 + qi) * d + (size_t)head * DH + hc * 32 : nullptr, lane);
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == CW) {
        tc_fence_after();
        tmem_dealloc<128>(tmem_base);
    }
}

// tm64 / tm32 / tm16: qkv maps with boxes of 64 / 32 / 16 rows (make_tmap_attention); cu points at the first short sentence
template <int DH>
static cudaError_t launch_att_pack(const CUtensorMap &tm64, const CUtensorMap &tm32, const CUtensorMap &tm16, const int32_t *cu,
                                   const PackCounts &pc, int d, int n_head, __half *out, int qmode, cudaStream_t st) {
    constexpr int ROWB = DH * 2;
    constexpr int R0 = 2 * kAttQ * ROWB > 2 * kAttQ * 128 ? 2 * kAttQ * ROWB : 2 * kAttQ * 128;
    constexpr size_t smem = R0 + kAttQ * ROWB + 64 + 512 + kLoopStage;
    static int sms[64] = {0};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
    if (!sms[dev]) {
        cudaError_t err = cudaFuncSetAttribute(attention_pack_kernel<DH>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (err != cudaSuccess) return err;
        int n = 0;
        err = cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
        if (err != cudaSuccess) return err;
        sms[dev] = n;
    }
    const int n_items = pack_tiles(pc) * n_head;
    if (n_items <= 0) return cudaSuccess;
    const int grid = n_items < 4 * sms[dev] ? n_items : 4 * sms[dev];
    const float scale = (float)(1.0 / sqrt((double)DH));  // bert.cpp:934
    LaunchConfig lc(dim3((unsigned)grid), dim3(kLoopThreads), smem, st);
    return cudaLaunchKernelEx(&lc.cfg, attention_pack_kernel<DH>, tm64, tm32, tm16, cu, pc, d, n_head, out, qmode, scale);
}
