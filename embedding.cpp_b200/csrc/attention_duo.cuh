// attention_duo.cuh -- attention for launches with sentences of more than 128 tokens, two query blocks in flight per CTA.
// Included by attention_tc.cu (inside namespace bk, after attention_stream.cuh whose constants it shares).
//
// attention_stream_kernel (one query block at a time, 16 softmax warps in lock-step) leaves every unit under 40 % busy at
// L = 512 (ncu, profiles/r02l_*: MUFU 36 %, tensor pipe 29 %, issue slots 45 %): all 16 warps wait for the same score chunk,
// then all contend for the MUFU, then all store P and fence, and the whole CTA drains at every block boundary.  The
// exponentials are the floor of this kernel (L^2 per head at 16 per clock per SM), so the unit to keep busy is the MUFU.
//
// Here the 16 softmax warps form TWO TEAMS of 8 that work on different query blocks of the same (sentence, head) -- team t
// takes blocks t, t + 2, ... -- each with its own issuing warps, score slots, P slots, O accumulator and barriers, so that
// one team's waits (score round trips, P hand-over, O drain, Q load, the cheap maximum pass) sit under the other team's
// exponentials.  Same arithmetic as attention_stream_kernel, per team and block:
//
//   pass A   S_c = Q . K_c^T for every 64-key chunk c -> row maximum only
//   (the two warps that share a row exchange their partial maxima; the previous block's O is scaled and stored here)
//   pass B   S_c again -> p = f16(exp(f16(s*scale - max))) (ggml.c:10536-10558) -> 64-key P slot in shared memory
//            -> O += P . V, rowsum += P . 1 on the tensor core
//
// Score chunks are 64 keys (one P slot = one chunk = the 8 warps of a team), two TMEM slots per team: a slot is handed back
// as soon as it is in registers, so the next chunk's MMAs run under this chunk's exponentials.  No A/B interleave inside a
// team (one Q buffer per team is all the shared memory left next to K, V and the P slots at L = 512): the other team is the
// overlap.  TMEM: team t scores [128 t, 128 t + 128), O_t and its row sums [256 + 80 t, 256 + 80 t + dh + 16).
// Warps: w < 16 softmax (team w / 8, key half (w / 4) % 2 of every chunk, TMEM lanes 32 (w % 4) ..); 16 / 17 score + TMA
// and P.V issuers of team 0, 18 / 19 of team 1.
//
// The 16-key MMA steps that accumulate O and the row sums come in the same order as in attention_stream_kernel /
// attention_tc_kernel and every rounding point is the same: bit-identical results (tests/test_gpu_kernels.py).
#pragma once

constexpr int kDuoThreads = 512 + 128;  // 16 softmax warps + 2 x (score warp, P.V warp)
constexpr int kDuoKeys = 64;            // keys per score chunk / P slot
constexpr int kDuoOCol = 256;           // TMEM: first O column
constexpr int kDuoOStride = 80;         // dh + 16 <= 80 columns per team
// shared memory behind Q | P | K | V: barriers + TMEM slot (256 B), all-ones tile (512 B), partial row maxima (2 KB)
constexpr int kDuoTail = 256 + 512 + 2048;

template <int DH>
__global__ void __launch_bounds__(kDuoThreads, 1)
attention_duo_kernel(const __grid_constant__ CUtensorMap tmQKV, const int32_t *__restrict__ cu, int d, __half *__restrict__ out,
                     int qmode, float scale) {
    constexpr int ROWB = DH * 2;
    constexpr int CW = 16;                 // first control warp
    constexpr int PSUB = kAttQ * 128;      // one P slot: 128 rows x 64 keys f16
    extern __shared__ __align__(1024) uint8_t att_raw[];
    uint8_t *smem = att_raw;

    pdl_launch_dependents();
    pdl_wait();  // the sequence bounds are read first thing
    const int seq = blockIdx.z, head = blockIdx.y;
    const int t0 = __ldg(cu + seq);
    const int L = __ldg(cu + seq + 1) - t0;
    if (L <= 0) return;
    const int nqb = (L + kAttQ - 1) / kAttQ;
    const int Lk16 = (L + 15) & ~15;
    const int Lk64 = (L + 63) & ~63;
    const int nc = Lk64 >> 6;              // 64-key score chunks per pass

    uint8_t *sQ = smem;                    // [2 teams][128 rows]
    uint8_t *sP = sQ + 2 * kAttQ * ROWB;   // [2 teams][2 slots]
    uint8_t *sK = sP + 4 * PSUB;
    uint8_t *sV = sK + Lk64 * ROWB;
    uint8_t *tail = sV + Lk64 * ROWB;
    uint64_t *bars = reinterpret_cast<uint64_t *>(tail);
    uint64_t *bar_q = bars;         // [2]    team's Q buffer landed
    uint64_t *bar_k = bars + 2;     // [4]    keys [128 c, 128 c + 128) landed
    uint64_t *bar_v = bars + 6;     //        V landed
    uint64_t *sfull = bars + 7;     // [2][2] score chunk complete in the team's slot
    uint64_t *sfree = bars + 11;    // [2][2] the team's 8 warps have read the slot
    uint64_t *pfull = bars + 15;    // [2][2] P slot written by the team's 8 warps
    uint64_t *pfree = bars + 19;    // [2][2] the MMAs that read the P slot have retired
    uint64_t *bar_o = bars + 23;    // [2]    O and row sums of the team's query block complete
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 25);
    uint32_t *ones = reinterpret_cast<uint32_t *>(tail + 256);  // 512 B, 16 x 16 f16 1.0
    float *red = reinterpret_cast<float *>(tail + 256 + 512);   // [16 warps][32] partial row maxima

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int cq = head * DH, ck = d + head * DH, cv = 2 * d + head * DH;
    auto load_q = [&](int team, int qb) {  // query block qb -> the team's Q buffer
        uint8_t *q = sQ + team * kAttQ * ROWB;
        mbar_expect_tx_e(&bar_q[team], (uint32_t)(kAttQ * ROWB));
        tma_load_2d_e(q, &tmQKV, &bar_q[team], cq, t0 + qb * kAttQ);
        tma_load_2d_e(q + 64 * ROWB, &tmQKV, &bar_q[team], cq, t0 + qb * kAttQ + 64);
    };
    if (warp == CW) {
        if (lane == 0) {
            tma_prefetch_desc(&tmQKV);
            for (int i = 0; i < 4; i++) mbar_init(&bar_k[i], 1);
            mbar_init(bar_v, 1);
            for (int t = 0; t < 2; t++) {
                mbar_init(&bar_q[t], 1); mbar_init(&bar_o[t], 1);
                for (int i = 0; i < 2; i++) {
                    mbar_init(&sfull[2 * t + i], 1); mbar_init(&sfree[2 * t + i], 8);
                    mbar_init(&pfull[2 * t + i], 8); mbar_init(&pfree[2 * t + i], 1);
                }
            }
            mbar_fence_init();
        }
        __syncwarp();
        {
            // the loads first (they need the barriers only): their latency runs under the TMEM allocation and the CTA-wide sync
            const int n_box = Lk64 >> 6;
            load_q(0, 0);
            for (int c = 0; 2 * c < n_box; c++) {  // K in chunks of 128 keys (two 64-row boxes, the last chunk possibly one)
                const int boxes = n_box - 2 * c < 2 ? n_box - 2 * c : 2;
                mbar_expect_tx_e(&bar_k[c], (uint32_t)(boxes * 64 * ROWB));
                for (int b = 0; b < boxes; b++) tma_load_2d_e(sK + (2 * c + b) * 64 * ROWB, &tmQKV, &bar_k[c], ck, t0 + (2 * c + b) * 64);
            }
            if (nqb > 1) load_q(1, 1);
            mbar_expect_tx_e(bar_v, (uint32_t)(Lk64 * ROWB));
            for (int c = 0; c < n_box; c++) tma_load_2d_e(sV + c * 64 * ROWB, &tmQKV, bar_v, cv, t0 + c * 64);
        }
        for (int i = lane; i < 128; i += 32) ones[i] = 0x3C003C00u;
        fence_proxy_async_smem();
        __syncwarp();
        tmem_alloc<512>(tmem_slot);
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp >= CW) {
        const int team = (warp - CW) >> 1;
        if (((warp - CW) & 1) == 0) {
            // ============================================================ score warp of the team: Q loads, S = Q . K^T jobs
            uint64_t *sf = sfull + 2 * team, *sr = sfree + 2 * team;
            const uint64_t dk0 = umma_desc_kmajor<ROWB>(smem_u32(sK));
            const uint64_t dq = umma_desc_kmajor<ROWB>(smem_u32(sQ + team * kAttQ * ROWB));
            const uint32_t tS = tmem_base + (uint32_t)(team * 2 * kDuoKeys);
            uint32_t js = 0;  // score jobs issued: slot js % 2, use js / 2 of that slot
            int use = 0;      // blocks of this team so far
            for (int qb = team; qb < nqb; qb += 2, use++) {
                if (use > 0) {
                    // the Q buffer was last read by the previous block's final score job: wait for its MMAs, then refill
                    const uint32_t pj = js - 1;
                    mbar_wait(&sf[pj & 1], (pj >> 1) & 1);
                    load_q(team, qb);
                }
                mbar_wait(&bar_q[team], (uint32_t)(use & 1));
                for (int pass = 0; pass < 2; pass++) {
                    for (int c = 0; c < nc; c++, js++) {
                        if (use == 0 && pass == 0) mbar_wait(&bar_k[c >> 1], 0);  // the team's first use of these keys
                        const uint32_t slot = js & 1;
                        if (js >= 2) mbar_wait(&sr[slot], ((js >> 1) - 1) & 1);  // read by the team's 8 warps (two jobs ago)
                        const int n0 = c * kDuoKeys;
                        const int nn = Lk16 - n0 < kDuoKeys ? Lk16 - n0 : kDuoKeys;
                        const uint64_t dk = dk0 + (uint64_t)((n0 * ROWB) >> 4);
                        const uint32_t idesc = att_idesc((uint32_t)nn, 0);
                        tc_fence_after();
#pragma unroll
                        for (int k = 0; k < DH / 16; k++)
                            umma_f16_e<1>(tS + slot * kDuoKeys, dq + (uint64_t)(2 * k), dk + (uint64_t)(2 * k), idesc, k != 0);
                        umma_commit_e<1>(&sf[slot]);
                    }
                }
            }
        } else {
            // ============================================================ P.V warp of the team: O += P . V, rowsum += P . 1
            const uint32_t idesc_pv = att_idesc(DH, 1), idesc_sum = att_idesc(kSumCols, 0);
            const uint64_t done = umma_desc_ones(smem_u32(ones));
            const uint64_t dp0 = umma_desc_sw128(smem_u32(sP + team * 2 * PSUB));
            const uint64_t dv0 = umma_desc_mnmajor<ROWB>(smem_u32(sV));
            const uint32_t tO = tmem_base + (uint32_t)(kDuoOCol + team * kDuoOStride);
            uint64_t *pf = pfull + 2 * team, *pr = pfree + 2 * team;
            if (team < nqb) mbar_wait(bar_v, 0);
            uint32_t pv = 0;  // P slots consumed
            for (int qb = team; qb < nqb; qb += 2) {
                for (int c = 0; c < nc; c++, pv++) {
                    const uint32_t ps = pv & 1;
                    mbar_wait(&pf[ps], (pv >> 1) & 1);
                    tc_fence_after();
                    const int key0 = c * kDuoKeys;
                    const int rem = Lk16 - key0;
                    const int ksteps = rem >= kDuoKeys ? 4 : rem >> 4;
                    const uint64_t dp = dp0 + (uint64_t)((ps * PSUB) >> 4);
                    const uint64_t dv = dv0 + (uint64_t)((key0 * ROWB) >> 4);
                    for (int ks = 0; ks < ksteps; ks++)
                        umma_f16_e<1>(tO, dp + (uint64_t)(2 * ks), dv + (uint64_t)((ks * 16 * ROWB) >> 4), idesc_pv, (c | ks) != 0);
                    for (int ks = 0; ks < ksteps; ks++)
                        umma_f16_e<1>(tO + (uint32_t)DH, dp + (uint64_t)(2 * ks), done, idesc_sum, (c | ks) != 0);
                    umma_commit_e<1>(&pr[ps]);
                }
                umma_commit_e<1>(&bar_o[team]);
            }
        }
    } else {
        // ======================================================================= softmax warps
        const int team = warp >> 3, half = (warp >> 2) & 1, quad = warp & 3;
        const int r = quad * 32 + lane;                       // query row in the block == TMEM lane
        const uint32_t tlane = tmem_base + ((uint32_t)(quad * 32) << 16);
        const uint32_t tS = tlane + (uint32_t)(team * 2 * kDuoKeys + half * 32);
        const uint32_t tO = tlane + (uint32_t)(kDuoOCol + team * kDuoOStride);
        const uint32_t row_off = (uint32_t)r * 128u, sw = (uint32_t)(r & 7);
        uint64_t *sf = sfull + 2 * team, *sr = sfree + 2 * team, *pf = pfull + 2 * team, *pr = pfree + 2 * team;
        uint8_t *sPt = sP + team * 2 * PSUB;
        const uint32_t pair_bar = 1u + (uint32_t)(team * 4 + quad);  // named barrier of the two warps that share these 32 rows
        uint32_t js = 0;  // score jobs consumed
        uint32_t nb = 0;  // P slots filled
        uint32_t ne = 0;  // epilogues done

        // O * (1 / sum p) of query block qb -> out rows; warps of key half h scale and store features [32 h, 32 h + 32)
        auto epilogue = [&](int qb) {
            if (half < DH / 32) {
                mbar_wait(&bar_o[team], ne & 1);
                tc_fence_after();
                const float sum = __uint_as_float(tmem_ld_32x1(tO + (uint32_t)DH));
                uint32_t a[32];
                tmem_ld_32x32(tO + (uint32_t)(half * 32), a);
                tmem_ld_wait();
                const float inv = 1.0f / sum;
                float o[32];
#pragma unroll
                for (int i = 0; i < 32; i++) o[i] = __uint_as_float(a[i]) * inv;
                if (qmode) {  // the attention-output GEMM of a Q4 model quantises its input per 32 features (ggml.c:9493-9509)
                    float amax = 0.f;
#pragma unroll
                    for (int i = 0; i < 32; i++) amax = fmaxf(amax, fabsf(o[i]));
                    float id, dq;
                    q8_scales(amax, qmode, id, dq);
#pragma unroll
                    for (int i = 0; i < 32; i++) o[i] = q8_round(o[i], id, dq);
                }
                const int q0 = qb * kAttQ;
                if (q0 + r < L) {
                    uint4 *dst = reinterpret_cast<uint4 *>(out + (size_t)(t0 + q0 + r) * d + (size_t)head * DH + half * 32);
#pragma unroll
                    for (int q = 0; q < 4; q++)
                        dst[q] = make_uint4(pack_h2(o[8 * q], o[8 * q + 1]), pack_h2(o[8 * q + 2], o[8 * q + 3]),
                                            pack_h2(o[8 * q + 4], o[8 * q + 5]), pack_h2(o[8 * q + 6], o[8 * q + 7]));
                }
                tc_fence_before();  // the O reads are ordered before this warp's next P hand-over (the MMA behind it overwrites O)
            }
            ne++;
        };

        int prev = -1;
        for (int qb = team; qb < nqb; qb += 2) {
            // ---- pass A: row maximum over my 32 keys of every chunk
            float mx = -INFINITY;
            for (int c = 0; c < nc; c++, js++) {
                const uint32_t slot = js & 1;
                mbar_wait(&sf[slot], (js >> 1) & 1);
                tc_fence_after();
                const int key0 = c * kDuoKeys + half * 32;
                if (key0 < L) {
                    uint32_t a[32];
                    tmem_ld_32x32(tS + slot * kDuoKeys, a);
                    tmem_ld_wait();
                    const int nv = L - key0;
                    if (nv < 32) {  // keys past the sentence never enter
#pragma unroll
                        for (int i = 0; i < 32; i++) if (i >= nv) a[i] = 0xff800000u;
                    }
                    float m0 = mx, m1 = -INFINITY;
#pragma unroll
                    for (int i = 0; i < 32; i += 4) {
                        m0 = fmaxf(m0, fmaxf(__uint_as_float(a[i]), __uint_as_float(a[i + 1])));
                        m1 = fmaxf(m1, fmaxf(__uint_as_float(a[i + 2]), __uint_as_float(a[i + 3])));
                    }
                    mx = fmaxf(m0, m1);
                }
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&sr[slot]);
            }
            // ---- the other key half of these rows
            red[warp * 32 + lane] = mx;
            named_bar_sync(pair_bar, 64);
            mx = fmaxf(mx, red[(warp ^ 4) * 32 + lane]);
            named_bar_sync(pair_bar, 64);  // both have read before either publishes its next block's maximum
            mx *= scale;  // max(s*scale) == max(s)*scale for scale > 0 (rounding is monotonic)
            // ---- the team's previous block: its P.V MMAs were issued before this block's maximum pass
            if (prev >= 0) epilogue(prev);
            // ---- pass B: p = f16(exp(f16(s*scale - max))) for my 32 keys of every chunk
            for (int c = 0; c < nc; c++, js++, nb++) {
                const uint32_t slot = js & 1;
                mbar_wait(&sf[slot], (js >> 1) & 1);
                tc_fence_after();
                const int key0 = c * kDuoKeys + half * 32;
                const bool mine = key0 < L;  // otherwise no real key: the MMA does not read these P columns
                uint32_t pk[16];
                if (mine) {
                    uint32_t a[32];
                    tmem_ld_32x32(tS + slot * kDuoKeys, a);
                    tmem_ld_wait();
                    const int nv = L - key0;
                    if (nv < 32) {  // s = -inf -> p = 0 exactly
#pragma unroll
                        for (int i = 0; i < 32; i++) if (i >= nv) a[i] = 0xff800000u;
                    }
                    const float2 sc2 = make_float2(scale, scale), nmx2 = make_float2(-mx, -mx), l2e = make_float2(1.4426950408889634f, 1.4426950408889634f);
#pragma unroll
                    for (int i = 0; i < 32; i += 2) {  // same instruction sequence as attention_tc_kernel
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
                }
                // the score slot is free as soon as it is in registers
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&sr[slot]);
                const uint32_t ps = nb & 1;
                if (nb >= 2) mbar_wait(&pr[ps], ((nb - 2) >> 1) & 1);  // the MMAs on the previous contents of this P slot have retired
                if (mine) {
                    uint8_t *pb = sPt + ps * PSUB;
#pragma unroll
                    for (int q = 0; q < 4; q++)
                        *reinterpret_cast<uint4 *>(pb + row_off + ((((uint32_t)(half * 4 + q)) ^ sw) << 4)) =
                            make_uint4(pk[4 * q], pk[4 * q + 1], pk[4 * q + 2], pk[4 * q + 3]);
                    fence_proxy_async_smem();  // P writes -> visible to the MMA (async proxy)
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(&pf[ps]);
            }
            prev = qb;
        }
        if (prev >= 0) epilogue(prev);
    }

    tc_fence_before();
    __syncthreads();
    if (warp == CW) {
        tc_fence_after();
        tmem_dealloc<512>(tmem_base);
    }
}

template <int DH>
static cudaError_t launch_att_duo(const CUtensorMap &tm, const int32_t *cu, int n_seq, int max_len, int d, int n_head, __half *out,
                                  int qmode, cudaStream_t st) {
    const int Lk64 = (max_len + 63) & ~63;
    const size_t smem = (size_t)2 * kAttQ * DH * 2 + 4 * kAttQ * 128 + (size_t)2 * Lk64 * DH * 2 + kDuoTail;
    static std::atomic<bool> attr_set[64];
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64 || !attr_set[dev]) {
        cudaError_t err = cudaFuncSetAttribute(attention_duo_kernel<DH>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
        if (err != cudaSuccess) return err;
        if (dev >= 0 && dev < 64) attr_set[dev] = true;
    }
    dim3 grid(1, n_head, n_seq);
    const float scale = (float)(1.0 / sqrt((double)DH));  // bert.cpp:934
    LaunchConfig lc(grid, dim3(kDuoThreads), smem, st);
    return cudaLaunchKernelEx(&lc.cfg, attention_duo_kernel<DH>, tm, cu, d, out, qmode, scale);
}
