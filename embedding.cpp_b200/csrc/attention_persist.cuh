// attention_persist.cuh -- persistent variant of the attention kernel for launches whose longest sentence has <= 128
// tokens (one 128-query block and at most two 64-key chunks per (sentence, head)).  Included by attention_tc.cu.
//
// The one-CTA-per-(sentence, head) kernel pays its whole dependency chain -- barrier init, TMEM allocation, sequence
// bounds, TMA of Q/K/V, S MMA, softmax, P.V MMA, epilogue -- once per item and hides it only behind the other three
// CTAs on the SM; for short sentences (a few rows of the 128-row tile) that chain is nearly all of an item's life.
// Here a CTA stays on its SM and walks over items (sentence, head) = blockIdx.x, blockIdx.x + gridDim.x, ...:
//
//   * shared memory holds TWO item stages (Q | K, later overwritten by the two P chunks; V), so the TMA of item
//     n + 2 is in flight while item n + 1 is in the softmax;
//   * TMEM holds TWO score/output stages of 128 columns, so S(n+1) = Q.K^T is computed while the softmax warps are
//     still on item n, and they find their next scores ready;
//   * the control warp never blocks on one event: it polls "scores can be issued", "a P chunk has been delivered" and
//     "a stage can be refilled" and serves whichever is ready.
//
// Arithmetic, rounding points, key order of the row sums and the layout of P are those of attention_tc_kernel, so
// the result of a sentence is bit-identical in both kernels (tests/test_gpu_kernels.py).  Two CTAs per SM
// (2 x 256 TMEM columns, 2 x ~98 KB).
#pragma once
// (included inside namespace bk, after the descriptor helpers of attention_tc.cu)

constexpr int kPersistThreads = 160;  // 4 softmax warps + control warp

template <int DH>
struct AttPersistCfg {
    static constexpr int ROWB = DH * 2;
    static constexpr int PBUF = kAttQ * 128;                                     // one P chunk: 128 rows x 64 keys f16
    static constexpr int R0 = 2 * kAttQ * ROWB > 2 * PBUF ? 2 * kAttQ * ROWB : 2 * PBUF;  // Q | K, then P0 | P1
    static constexpr int VB = kAttQ * ROWB;
    static constexpr int STAGE = R0 + VB;
    static constexpr int TAIL = 256 /*barriers, TMEM slot, item ring*/ + 512 /*ones tile*/;
    static constexpr int SMEM = 2 * STAGE + TAIL;
    static constexpr int TMEM_STAGE = 128;                                       // score columns of one item
};

template <int DH>
__global__ void __launch_bounds__(kPersistThreads, 2)
attention_persist_kernel(const __grid_constant__ CUtensorMap tmQKV, const int32_t *__restrict__ cu, int d, int n_head, int n_items,
                         __half *__restrict__ out, int qmode, float scale) {
    using Cfg = AttPersistCfg<DH>;
    constexpr int ROWB = Cfg::ROWB, PBUF = Cfg::PBUF;
    extern __shared__ __align__(1024) uint8_t att_raw[];
    uint8_t *smem = att_raw;
    uint8_t *tail = smem + 2 * Cfg::STAGE;
    uint64_t *bar_qk = reinterpret_cast<uint64_t *>(tail);  // [2] Q and K of the stage landed
    uint64_t *bar_v = bar_qk + 2;                           // [2] V landed
    uint64_t *bar_s = bar_qk + 4;                           // [2] scores complete in the TMEM stage
    uint64_t *bar_o = bar_qk + 6;                           // [2] O and row sums complete; shared-memory stage no longer read
    uint64_t *bar_tfree = bar_qk + 8;                       // [2] the softmax warps have read O: TMEM stage free
    uint64_t *pfull = bar_qk + 10;                          // [2][2] P chunk delivered
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bar_qk + 14);
    int2 *ring = reinterpret_cast<int2 *>(bar_qk + 16);     // [4] (first token, length) of items n % 4, written by the control warp
    uint32_t *ones = reinterpret_cast<uint32_t *>(tail + 256);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n_local = (n_items - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;  // grid <= n_items
    constexpr int CW = 4;
    pdl_launch_dependents();
    if (warp == CW) {
        if (lane == 0) {
            tma_prefetch_desc(&tmQKV);
            for (int i = 0; i < 2; i++) {
                mbar_init(&bar_qk[i], 1); mbar_init(&bar_v[i], 1); mbar_init(&bar_s[i], 1); mbar_init(&bar_o[i], 1);
                mbar_init(&bar_tfree[i], 4); mbar_init(&pfull[2 * i], 4); mbar_init(&pfull[2 * i + 1], 4);
            }
            mbar_fence_init();
        }
        for (int i = lane; i < 128; i += 32) ones[i] = 0x3C003C00u;
        fence_proxy_async_smem();
        __syncwarp();
        tmem_alloc<2 * Cfg::TMEM_STAGE>(tmem_slot);
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
        constexpr int kLag = DH / 64;  // the row sums alias the score columns of chunk kLag (see attention_tc_kernel)
        auto ready = [&](uint64_t *bar, uint32_t parity) { return __any_sync(0xffffffffu, mbar_test_wait(bar, parity)) != 0; };
        auto keys16 = [&](int n) {  // keys entering the MMAs of item n (multiple of 16)
            int L = ring[n & 3].y;
            L = L < 1 ? 1 : L > kAttQ ? kAttQ : L;
            return (L + 15) & ~15;
        };
        // fetch item n into shared-memory stage n % 2
        auto load = [&](int n) {
            const int it = (int)blockIdx.x + n * (int)gridDim.x;
            const int seq = it / n_head, head = it - seq * n_head;
            const int t0 = __ldg(cu + seq), L = __ldg(cu + seq + 1) - t0;
            const int Lc = L < 1 ? 1 : L > kAttQ ? kAttQ : L;
            const int Lk64 = (Lc + 63) & ~63, nch = Lk64 >> 6;
            const int s = n & 1;
            uint8_t *sQ = smem + s * Cfg::STAGE, *sK = sQ + kAttQ * ROWB, *sV = sQ + Cfg::R0;
            if (lane == 0) ring[n & 3] = make_int2(t0, L);
            __syncwarp();
            const int cq = head * DH, ck = d + head * DH, cv = 2 * d + head * DH;
            mbar_expect_tx_e(&bar_qk[s], (uint32_t)((kAttQ + Lk64) * ROWB));
            tma_load_2d_e(sQ, &tmQKV, &bar_qk[s], cq, t0);
            tma_load_2d_e(sQ + 64 * ROWB, &tmQKV, &bar_qk[s], cq, t0 + 64);
            for (int c = 0; c < nch; c++) tma_load_2d_e(sK + c * 64 * ROWB, &tmQKV, &bar_qk[s], ck, t0 + c * 64);
            mbar_expect_tx_e(&bar_v[s], (uint32_t)(Lk64 * ROWB));
            for (int c = 0; c < nch; c++) tma_load_2d_e(sV + c * 64 * ROWB, &tmQKV, &bar_v[s], cv, t0 + c * 64);
        };
        // S(n) = Q . K^T into TMEM stage n % 2
        auto issue_s = [&](int n) {
            const int s = n & 1;
            const uint8_t *sQ = smem + s * Cfg::STAGE;
            const uint64_t dq = umma_desc_kmajor<ROWB>(smem_u32(sQ)), dk = umma_desc_kmajor<ROWB>(smem_u32(sQ + kAttQ * ROWB));
            const uint32_t idesc = att_idesc((uint32_t)keys16(n), 0);
            tc_fence_after();
#pragma unroll
            for (int k = 0; k < DH / 16; k++)
                umma_f16_e<1>(tmem_base + (uint32_t)(s * Cfg::TMEM_STAGE), dq + (uint64_t)(2 * k), dk + (uint64_t)(2 * k), idesc, k != 0);
            umma_commit_e<1>(&bar_s[s]);
        };
        // rowsum += P_c . 1 for chunk c of item n
        auto issue_sum = [&](int n, int c) {
            const int s = n & 1;
            const uint64_t dp = umma_desc_sw128(smem_u32(smem + s * Cfg::STAGE + c * PBUF));
            const int rem = keys16(n) - c * 64;
            const int ksteps = rem >= 64 ? 4 : rem >> 4;
            for (int ks = 0; ks < ksteps; ks++)
                umma_f16_e<1>(tmem_base + (uint32_t)(s * Cfg::TMEM_STAGE + DH), dp + (uint64_t)(2 * ks), done, idesc_sum, (c | ks) != 0);
        };
        // O += P_c . V_c
        auto issue_pv = [&](int n, int c) {
            const int s = n & 1;
            const uint8_t *sb = smem + s * Cfg::STAGE;
            const uint64_t dp = umma_desc_sw128(smem_u32(sb + c * PBUF));
            const int rem = keys16(n) - c * 64;
            const int ksteps = rem >= 64 ? 4 : rem >> 4;
            tc_fence_after();
            for (int ks = 0; ks < ksteps; ks++) {
                const uint64_t dv = umma_desc_mnmajor<ROWB>(smem_u32(sb + Cfg::R0 + (c * 64 + ks * 16) * ROWB));
                umma_f16_e<1>(tmem_base + (uint32_t)(s * Cfg::TMEM_STAGE), dp + (uint64_t)(2 * ks), dv, idesc_pv, (c | ks) != 0);
            }
        };

        load(0);
        if (n_local > 1) load(1);
        mbar_wait(&bar_qk[0], 0);
        issue_s(0);
        int n_pv = 0, c_pv = 0;  // item / chunk whose P.V MMAs are issued next
        int n_s = 1;             // item whose scores are issued next
        int n_ld = 2;            // item fetched next
        while (n_pv < n_local) {
            // scores of the next item: its Q, K have landed and the softmax warps are done with the TMEM stage (item n_s - 2)
            if (n_s < n_local && n_s <= n_pv + 1 && ready(&bar_qk[n_s & 1], (uint32_t)((n_s >> 1) & 1)) &&
                (n_s < 2 || ready(&bar_tfree[n_s & 1], (uint32_t)(((n_s - 2) >> 1) & 1)))) {
                issue_s(n_s);
                n_s++;
            }
            // a P chunk of the current item has been delivered
            if (ready(&pfull[2 * (n_pv & 1) + c_pv], (uint32_t)((n_pv >> 1) & 1))) {
                if (c_pv == 0) mbar_wait(&bar_v[n_pv & 1], (uint32_t)((n_pv >> 1) & 1));
                const int nch = (keys16(n_pv) + 63) >> 6;
                issue_pv(n_pv, c_pv);
                if (c_pv >= kLag) issue_sum(n_pv, c_pv - kLag);
                if (++c_pv == nch) {
                    for (int c = nch > kLag ? nch - kLag : 0; c < nch; c++) issue_sum(n_pv, c);
                    umma_commit_e<1>(&bar_o[n_pv & 1]);
                    c_pv = 0;
                    n_pv++;
                }
            }
            // refill the shared-memory stage of a finished item (its MMAs have retired)
            if (n_ld < n_local && n_ld - 2 < n_pv && ready(&bar_o[n_ld & 1], (uint32_t)(((n_ld - 2) >> 1) & 1))) {
                load(n_ld);
                n_ld++;
            }
        }
    } else {
        // ======================================================================= softmax warps (one query row per thread)
        const int r = warp * 32 + lane;  // query row == TMEM lane
        const uint32_t row_off = (uint32_t)r * 128u, sw = (uint32_t)(r & 7);
        for (int n = 0; n < n_local; n++) {
            const int s = n & 1;
            const uint32_t ph = (uint32_t)((n >> 1) & 1);
            const int it = (int)blockIdx.x + n * (int)gridDim.x;
            const int head = it % n_head;
            const uint32_t taddr = tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)(s * Cfg::TMEM_STAGE);
            uint8_t *sP = smem + s * Cfg::STAGE;
            mbar_wait(&bar_s[s], ph);
            tc_fence_after();
            const int2 info = ring[n & 3];
            const int t0 = info.x, Lr = info.y;                     // Lr: rows actually stored
            const int L = Lr < 1 ? 1 : Lr > kAttQ ? kAttQ : Lr;     // keys (as the control warp clamps them)
            const int n_chunks = (L + 63) >> 6, n_slices = (L + 31) >> 5;
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
            mx *= scale;  // max(s*scale) == max(s)*scale for scale > 0 (rounding is monotonic)
            // ---- pass 2: p = f16(exp(f16(s*scale - max))) -> P chunks (over Q | K, which the S MMAs have finished reading)
            for (int c = 0; c < n_chunks; c++) {
                uint8_t *pb = sP + c * PBUF;
#pragma unroll 1
                for (int h = 0; h < 2; h++) {
                    const int j = 2 * c + h;
                    if (j >= n_slices) continue;  // no real key: these columns are not read by the MMA
                    uint32_t a[32];
                    tmem_ld_32x32(taddr + (uint32_t)(j * 32), a);
                    tmem_ld_wait();
                    const int nv = L - j * 32;
                    if (nv < 32) {  // s = -inf -> p = 0 exactly
#pragma unroll
                        for (int i = 0; i < 32; i++) if (i >= nv) a[i] = 0xff800000u;
                    }
                    uint32_t pk[16];
                    const float2 sc2 = make_float2(scale, scale), nmx2 = make_float2(-mx, -mx), l2e = make_float2(1.4426950408889634f, 1.4426950408889634f);
#pragma unroll
                    for (int i = 0; i < 32; i += 2) {  // same instruction sequence as attention_tc_kernel (rounding points of ggml.c:10536-10558)
                        float2 x;
                        if (DH == 64)
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
                tc_fence_before();         // my reads of these score columns precede the MMA that overwrites them (O aliases chunk 0)
                __syncwarp();
                if (lane == 0) mbar_arrive(&pfull[2 * s + c]);
            }
            // every barrier of a stage completes exactly once per item (its parity is the item's): an item of one chunk
            // passes the second chunk's barrier empty-handed
            if (n_chunks == 1 && lane == 0) mbar_arrive(&pfull[2 * s + 1]);
            // ---- epilogue: O * (1 / sum p); the row sum was accumulated by the tensor core next to O
            mbar_wait(&bar_o[s], ph);
            tc_fence_after();
            const float sum = __uint_as_float(tmem_ld_32x1(taddr + (uint32_t)DH));
            tmem_ld_wait();
            const float inv = 1.0f / sum;
            for (int hc = 0; hc < DH / 32; hc++) {
                uint32_t a[32];
                tmem_ld_32x32(taddr + (uint32_t)(hc * 32), a);
                tmem_ld_wait();
                if (hc + 1 == DH / 32) {  // last TMEM read of this item by this warp: the stage may take the scores of item n + 2
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&bar_tfree[s]);
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
                if (r < Lr) {
                    uint4 *dst = reinterpret_cast<uint4 *>(out + (size_t)(t0 + r) * d + (size_t)head * DH + hc * 32);
#pragma unroll
                    for (int q = 0; q < 4; q++)
                        dst[q] = make_uint4(pack_h2(o[8 * q], o[8 * q + 1]), pack_h2(o[8 * q + 2], o[8 * q + 3]),
                                            pack_h2(o[8 * q + 4], o[8 * q + 5]), pack_h2(o[8 * q + 6], o[8 * q + 7]));
                }
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == CW) {
        tc_fence_after();
        tmem_dealloc<2 * Cfg::TMEM_STAGE>(tmem_base);
    }
}

template <int DH>
static cudaError_t launch_att_persist(const CUtensorMap &tm, const int32_t *cu, int n_seq, int d, int n_head, __half *out, int qmode,
                                      cudaStream_t st) {
    using Cfg = AttPersistCfg<DH>;
    static int sms[64] = {0};  // per device: SM count; the shared-memory attribute is set when it is first read
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
    if (!sms[dev]) {
        cudaError_t err = cudaFuncSetAttribute(attention_persist_kernel<DH>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM);
        if (err != cudaSuccess) return err;
        int n = 0;
        err = cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
        if (err != cudaSuccess) return err;
        sms[dev] = n;
    }
    const int n_items = n_seq * n_head;
    const int grid = n_items < 2 * sms[dev] ? n_items : 2 * sms[dev];
    const float scale = (float)(1.0 / sqrt((double)DH));  // bert.cpp:934
    LaunchConfig lc(dim3((unsigned)grid), dim3(kPersistThreads), Cfg::SMEM, st);
    return cudaLaunchKernelEx(&lc.cfg, attention_persist_kernel<DH>, tm, cu, d, n_head, n_items, out, qmode, scale);
}

