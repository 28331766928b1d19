// attention_loop.cuh -- the per-item attention kernel (attention_tc_kernel<DH, 1>, sentences of at most 128 tokens) with the CTA
// kept on its SM: four CTAs per SM (the TMEM ceiling of a resident-score design: 128 of 512 columns per item) each walk over
// items (sentence, head) = blockIdx.x, + gridDim.x, ...  Included by attention_tc.cu, inside namespace bk.
//
// Why: ncu (profiles/r02a_stalls.md) puts 54 % of the per-item kernel's warp samples into waits of its dependency chain, a third of
// them in what a CTA does once per item only because it is launched once per item -- the CTA-wide barrier behind barrier
// initialisation + tensor-map prefetch + TMEM allocation (9 %), special-register / TMEM-slot reads after it (11 %), the exit barrier
// before the TMEM de-allocation (6 %).  Here that set-up happens once per CTA; per item only the chain itself remains, and the next
// item's Q / K / V loads are requested the moment the current item's P.V MMAs have retired (they are the last readers of the
// shared-memory stage), i.e. under the current item's epilogue.  No second shared-memory or TMEM stage (that is
// attention_persist.cuh: two stages, two CTAs per SM -- slower than this kernel at every length, tools/att_len_probe.py).
//
// Arithmetic, rounding points, MMA sequence and P layout are those of attention_tc_kernel: bit-identical results.
#pragma once

constexpr int kLoopThreads = 160;  // 4 softmax warps + control warp
constexpr int kLoopStage = 4 * 1024; // output staging, 1 KB per softmax warp

// A warp's 32 rows x 64 bytes (32 f16 features of 32 query rows, one row per lane, pk = 16 f16 pairs) -> global memory through a 1 KB
// staging tile, 16 rows at a time, so that one store instruction covers 8 rows x 64 contiguous bytes (full 32-byte sectors) instead
// of 32 rows x 16 bytes.  Stores of the direct form -- every lane its own row, 16 bytes per instruction -- are 32 half-sector
// transactions each and back up the SM's load/store path, which the mbarrier operations of every resident CTA share (measured on the
// kernel for longer sentences, attention_pt.cuh: 13 %).  `my_dst`: where this lane's row goes (its 64 bytes), nullptr for a row past
// the sentence.  Chunk c of staged row R sits at 16-byte position (c + R / 2) % 4 of the row: conflict-free both ways.
__device__ __forceinline__ void store_rows64_1k(uint8_t *stage, const uint32_t (&pk)[16], __half *my_dst, int lane) {
    const unsigned long long mine = reinterpret_cast<unsigned long long>(my_dst);
#pragma unroll
    for (int hh = 0; hh < 2; hh++) {
        if ((lane >> 4) == hh) {
            const int rr = lane & 15;
#pragma unroll
            for (int q = 0; q < 4; q++)
                *reinterpret_cast<uint4 *>(stage + rr * 64 + (((q + (rr >> 1)) & 3) << 4)) = make_uint4(pk[4 * q], pk[4 * q + 1], pk[4 * q + 2], pk[4 * q + 3]);
        }
        __syncwarp();
#pragma unroll
        for (int j = 0; j < 2; j++) {
            const int R = 8 * j + (lane >> 2), c = lane & 3;
            const uint4 v = *reinterpret_cast<const uint4 *>(stage + R * 64 + (((c + (R >> 1)) & 3) << 4));
            const unsigned long long dst = __shfl_sync(0xffffffffu, mine, hh * 16 + R);
            if (dst) *reinterpret_cast<uint4 *>(reinterpret_cast<__half *>(dst) + c * 8) = v;
        }
        __syncwarp();
    }
}

template <int DH>
__global__ void __launch_bounds__(kLoopThreads, 4)
attention_loop_kernel(const __grid_constant__ CUtensorMap tmQKV, const int32_t *__restrict__ cu, int d, int n_head, int n_items,
                      __half *__restrict__ out, int qmode, float scale) {
    constexpr int ROWB = DH * 2;
    constexpr int CW = 4;
    constexpr int PBUF = kAttQ * 128;
    constexpr int QK_BYTES = 2 * kAttQ * ROWB;
    constexpr int R0 = QK_BYTES > 2 * PBUF ? QK_BYTES : 2 * PBUF;  // Q | K, later the two P chunks
    extern __shared__ __align__(1024) uint8_t att_raw[];
    uint8_t *smem = att_raw;
    uint8_t *sQ = smem, *sK = smem + kAttQ * ROWB, *sP = smem, *sV = smem + R0;
    uint8_t *tail = sV + kAttQ * ROWB;
    uint64_t *bar_qk = reinterpret_cast<uint64_t *>(tail);  // Q and K landed
    uint64_t *bar_v = bar_qk + 1, *bar_s = bar_qk + 2, *bar_o = bar_qk + 3, *bar_tfree = bar_qk + 4, *pfull = bar_qk + 5;  // pfull[2]
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bar_qk + 7);
    uint32_t *ones = reinterpret_cast<uint32_t *>(bar_qk + 8);  // 512 B
    uint8_t *stage = tail + 64 + 512;                           // [4 warps][1 KB] output staging

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    pdl_launch_dependents();
    if (warp == CW) {
        if (lane == 0) {
            tma_prefetch_desc(&tmQKV);
            mbar_init(bar_qk, 1); mbar_init(bar_v, 1); mbar_init(bar_s, 1); mbar_init(bar_o, 1); mbar_init(bar_tfree, 4);
            mbar_init(&pfull[0], 4); mbar_init(&pfull[1], 4);
            mbar_fence_init();
        }
        for (int i = lane; i < 128; i += 32) ones[i] = 0x3C003C00u;
        fence_proxy_async_smem();
        __syncwarp();
        tmem_alloc<128>(tmem_slot);
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();  // set-up done under the previous kernel's tail; global memory (sequence bounds, qkv, out) from here on

    if (warp == CW) {
        // ================================================================ control warp (uniform, elected issue)
        const uint32_t idesc_pv = att_idesc(DH, 1), idesc_sum = att_idesc(kSumCols, 0);
        const uint64_t done = umma_desc_ones(smem_u32(ones));
        const uint64_t dq = umma_desc_kmajor<ROWB>(smem_u32(sQ));
        auto load = [&](int it) {  // Q | K then V of item `it`: the stage is free (the previous item's MMAs have retired)
            const int seq = it / n_head, head = it - seq * n_head;
            const int t0 = __ldg(cu + seq), L = __ldg(cu + seq + 1) - t0;
            const int n_chunks = (L + 63) >> 6;
            const int cq = head * DH, ck = d + head * DH, cv = 2 * d + head * DH;
            mbar_expect_tx_e(bar_qk, (uint32_t)((kAttQ + n_chunks * 64) * ROWB));
            tma_load_2d_e(sQ, &tmQKV, bar_qk, cq, t0);
            tma_load_2d_e(sQ + 64 * ROWB, &tmQKV, bar_qk, cq, t0 + 64);
            for (int c = 0; c < n_chunks; c++) tma_load_2d_e(sK + c * 64 * ROWB, &tmQKV, bar_qk, ck, t0 + c * 64);
            mbar_expect_tx_e(bar_v, (uint32_t)(n_chunks * 64 * ROWB));
            for (int c = 0; c < n_chunks; c++) tma_load_2d_e(sV + c * 64 * ROWB, &tmQKV, bar_v, cv, t0 + c * 64);
        };
        if ((int)blockIdx.x < n_items) load((int)blockIdx.x);
        uint32_t k = 0;   // items done by this CTA: parity of the once-per-item barriers
        uint32_t cg = 0;  // P chunks so far
        for (int it = (int)blockIdx.x; it < n_items; it += (int)gridDim.x, k++) {
            const int seq = it / n_head;
            const int L = __ldg(cu + seq + 1) - __ldg(cu + seq);
            const int Lk16 = (L + 15) & ~15, n_chunks = (L + 63) >> 6;
            mbar_wait(bar_qk, k & 1);
            if (k) mbar_wait(bar_tfree, (k - 1) & 1);  // the softmax warps have read the previous item's O and row sums
            tc_fence_after();
            {   // S = Q . K^T, columns [0, Lk16)
                const uint64_t dk = umma_desc_kmajor<ROWB>(smem_u32(sK));
                const uint32_t idesc = att_idesc((uint32_t)Lk16, 0);
#pragma unroll
                for (int kk = 0; kk < DH / 16; kk++) umma_f16_e<1>(tmem_base, dq + (uint64_t)(2 * kk), dk + (uint64_t)(2 * kk), idesc, kk != 0);
                umma_commit_e<1>(bar_s);
            }
            mbar_wait(bar_v, k & 1);
            // O += P_c . V_c, then rowsum += P_c . 1 (the row sums land in the score columns of chunk 0 for dh = 32 / chunk 1 for
            // dh = 64: issued after that chunk's P has been delivered, exactly as in attention_tc_kernel)
            constexpr int kLag = DH / 64;
            auto issue_sum = [&](int c, uint32_t cgc) {
                const uint64_t dp = umma_desc_sw128(smem_u32(sP + (cgc & 1) * PBUF));
                const int rem = Lk16 - c * 64;
                const int ksteps = rem >= 64 ? 4 : rem >> 4;
                for (int ks = 0; ks < ksteps; ks++) umma_f16_e<1>(tmem_base + (uint32_t)DH, dp + (uint64_t)(2 * ks), done, idesc_sum, (c | ks) != 0);
            };
            for (int c = 0; c < n_chunks; c++, cg++) {
                mbar_wait(&pfull[cg & 1], (cg >> 1) & 1);
                tc_fence_after();
                const uint64_t dp = umma_desc_sw128(smem_u32(sP + (cg & 1) * PBUF));
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
            // the next item's loads go out as soon as these MMAs (the last readers of P and V) have retired: under this item's epilogue
            if (it + (int)gridDim.x < n_items) {
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
        uint32_t k = 0, cg = 0;
        for (int it = (int)blockIdx.x; it < n_items; it += (int)gridDim.x, k++) {
            const int seq = it / n_head, head = it - seq * n_head;
            const int t0 = __ldg(cu + seq), L = __ldg(cu + seq + 1) - t0;
            const int n_chunks = (L + 63) >> 6, n_slices = (L + 31) >> 5;
            mbar_wait(bar_s, k & 1);
            tc_fence_after();
            // ---- pass 1: row maximum of the raw scores over the real keys
            float mx = -INFINITY;
            for (int j = 0; j < n_slices; j++) {
                uint32_t a[32];
                tmem_ld_32x32(taddr + (uint32_t)(j * 32), a);
                tmem_ld_wait();
                const int nv = L - j * 32;
                if (nv < 32) {
#pragma unroll
                    for (int i = 0; i < 32; i++) if (i >= nv) a[i] = 0xff800000u;  // -inf: keys past the sentence never enter
                }
                float m0 = mx, m1 = -INFINITY;
#pragma unroll
                for (int i = 0; i < 32; i += 4) {
                    m0 = fmaxf(m0, fmaxf(__uint_as_float(a[i]), __uint_as_float(a[i + 1])));
                    m1 = fmaxf(m1, fmaxf(__uint_as_float(a[i + 2]), __uint_as_float(a[i + 3])));
                }
                mx = fmaxf(m0, m1);
            }
            mx *= scale;
            // ---- pass 2: p = f16(exp(f16(s*scale - max))) -> P chunk buffers (they alias Q | K, which the S MMAs have consumed).
            // (Issuing the tcgen05.ld of slice j + 1 ahead of the arithmetic of slice j was measured: no gain, 96 instead of 60 registers.)
            for (int c = 0; c < n_chunks; c++, cg++) {
                uint8_t *pb = sP + (cg & 1) * PBUF;
#pragma unroll 1
                for (int h = 0; h < 2; h++) {
                    const int j = 2 * c + h;
                    if (j >= n_slices) continue;
                    uint32_t a[32];
                    tmem_ld_32x32(taddr + (uint32_t)(j * 32), a);
                    tmem_ld_wait();
                    const int nv = L - j * 32;
                    if (nv < 32) {  // s = -inf -> p = 2^-inf = 0 exactly: a masked key contributes nothing
#pragma unroll
                        for (int i = 0; i < 32; i++) if (i >= nv) a[i] = 0xff800000u;
                    }
                    uint32_t pk[16];
                    const float2 sc2 = make_float2(scale, scale), nmx2 = make_float2(-mx, -mx), l2e = make_float2(1.4426950408889634f, 1.4426950408889634f);
#pragma unroll
                    for (int i = 0; i < 32; i += 2) {
                        // ggml_scale (separate multiply), x - max, f16 round | exp | f16 round (ggml.c:10536-10558); see attention_tc_kernel
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
#pragma unroll
                    for (int q = 0; q < 4; q++)
                        *reinterpret_cast<uint4 *>(pb + row_off + ((((uint32_t)(h * 4 + q)) ^ sw) << 4)) =
                            make_uint4(pk[4 * q], pk[4 * q + 1], pk[4 * q + 2], pk[4 * q + 3]);
                }
                fence_proxy_async_smem();  // P writes -> visible to the MMA (async proxy)
                tc_fence_before();         // my TMEM reads of these score columns precede the MMAs that overwrite them (O aliases chunk 0)
                __syncwarp();
                if (lane == 0) mbar_arrive(&pfull[cg & 1]);
            }
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
                store_rows64_1k(stage + quad * 1024, pk, r < L ? out + (size_t)(t0 + r) * d + (size_t)head * DH + hc * 32 : nullptr, lane);
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

template <int DH>
static cudaError_t launch_att_loop(const CUtensorMap &tm, const int32_t *cu, int n_seq, int d, int n_head, __half *out, int qmode,
                                   cudaStream_t st) {
    constexpr int ROWB = DH * 2;
    constexpr int R0 = 2 * kAttQ * ROWB > 2 * kAttQ * 128 ? 2 * kAttQ * ROWB : 2 * kAttQ * 128;
    constexpr size_t smem = R0 + kAttQ * ROWB + 64 + 512 + kLoopStage;
    static int sms[64] = {0};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
    if (!sms[dev]) {
        cudaError_t err = cudaFuncSetAttribute(attention_loop_kernel<DH>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (err != cudaSuccess) return err;
        int n = 0;
        err = cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
        if (err != cudaSuccess) return err;
        sms[dev] = n;
    }
    const int n_items = n_seq * n_head;
    const int grid = n_items < 4 * sms[dev] ? n_items : 4 * sms[dev];
    const float scale = (float)(1.0 / sqrt((double)DH));  // bert.cpp:934
    LaunchConfig lc(dim3((unsigned)grid), dim3(kLoopThreads), smem, st);
    return cudaLaunchKernelEx(&lc.cfg, attention_loop_kernel<DH>, tm, cu, d, n_head, n_items, out, qmode, scale);
}
