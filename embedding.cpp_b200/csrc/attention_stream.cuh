// attention_stream.cuh -- attention for launches with sentences of more than 128 tokens.  Included by attention_tc.cu
// (inside namespace bk, after its descriptor helpers).
//
// attention_tc_kernel keeps the whole score block S[128 x Lk] of a query block in TMEM (up to all 512 columns), so O
// has to alias consumed score columns, only one query block can be in flight, and its 16 softmax warps move in
// lock-step through "wait for S / row max / exponentials / wait for O".  Here the scores are STREAMED in chunks of 128
// keys through a three-slot TMEM ring and computed twice:
//
//   pass A   S_c = Q . K_c^T for every chunk c -> running row maximum only (no exponentials; ~0.6 instructions/score)
//   pass B   S_c again (the tensor pipe has the slack: the kernel is issue-bound on the softmax side) ->
//            p = f16(exp(f16(s*scale - max))) with the TRUE row maximum, as the reference computes it
//            (ggml.c:10536-10558) -> 64-key P slots in shared memory (two per half chunk) -> O += P . V, rowsum += P . 1
//
// Pass A of query block q + 1 is interleaved chunk by chunk with pass B of block q (job order A(0,*), then
// B(q,0) A(q+1,0) B(q,1) A(q+1,1) ...), so the round trip "slot read -> next score MMA -> slot full" of the cheap
// maximum pass hides behind the exponentials of the other pass.  O and the row sums own TMEM columns
// [384, 384 + dh + 16): nothing aliases, the control warp issues score chunks as soon as a ring slot has been read,
// also across query blocks, and the epilogue of block q runs at the start of block q + 1, when its last P.V MMAs
// have long retired.  Two issuing warps -- one for the score jobs and the loads, one for the P.V / row-sum MMAs -- because a
// single control warp (about 740 instructions per pair of chunks, sharing its scheduler with four softmax warps) was
// the bottleneck of the whole kernel.  One CTA = one (sentence, head), K and V resident in shared memory, two Q buffers, 16 softmax
// warps: warp w owns TMEM lanes 32 * (w % 4) .., group g = w / 4 owns keys [32 g, 32 g + 32) of every 128-key chunk;
// groups {0,1} / {2,3} fill the P slots of the first / second 64-key half of a chunk.
//
// The sequence of 16-key MMA steps that accumulate O and the row sums is the same as in attention_tc_kernel, and so are
// all rounding points: results are bit-identical to it (tests/test_gpu_kernels.py).
#pragma once

constexpr int kStreamThreads = 64 + 512;  // 16 softmax warps + score warp + P.V warp
constexpr int kStreamKeys = 128;          // keys per score chunk
constexpr int kStreamRing = 3;            // score slots in TMEM
constexpr int kStreamQ = 2;               // Q buffers

// shared memory behind Q | P | K | V: barriers + TMEM slot (192 B), all-ones tile (512 B), then one or two row-maximum
// exchange buffers of 2 KB
constexpr int kStreamTail = 192 + 512;

template <int DH>
__global__ void __launch_bounds__(kStreamThreads, 1)
attention_stream_kernel(const __grid_constant__ CUtensorMap tmQKV, const int32_t *__restrict__ cu, int d, __half *__restrict__ out,
                        int qmode, float scale) {
    constexpr int ROWB = DH * 2;
    constexpr int CW = 16;                 // score warp (TMA, S MMAs, TMEM allocation); CW + 1: P.V warp
    constexpr int PSUB = kAttQ * 128;      // one P slot: 128 rows x 64 keys f16
    constexpr uint32_t kOCol = kStreamRing * kStreamKeys;  // TMEM: 3 score slots | O [384, 384+dh) | row sums [384+dh, +16)
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
    const int nc = (L + kStreamKeys - 1) / kStreamKeys;  // score chunks per pass

    uint8_t *sQ = smem;                          // [2][128 rows]
    uint8_t *sP = sQ + kStreamQ * kAttQ * ROWB;  // [2 group pairs][2 slots]
    uint8_t *sK = sP + 4 * PSUB;
    uint8_t *sV = sK + Lk64 * ROWB;
    uint8_t *tail = sV + Lk64 * ROWB;
    uint64_t *bar_q = reinterpret_cast<uint64_t *>(tail);  // [2] Q buffer landed
    uint64_t *bar_k = bar_q + 20;   // [4] keys [128 c, 128 c + 128) landed: the first score chunk does not wait for all of K
    uint64_t *bar_v = bar_q + 3;
    uint64_t *sfull = bar_q + 4;    // [3] score chunk complete in the ring slot
    uint64_t *sfree = bar_q + 7;    // [3] all 16 warps have read the slot
    uint64_t *pfull = bar_q + 10;   // [2 pairs][2 slots] P slot written by the 8 warps of its group pair
    uint64_t *pfree = bar_q + 14;   // [2 pairs][2 slots] the MMAs that read the P slot have retired
    uint64_t *bar_o = bar_q + 18;   // O and row sums of a query block complete
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bar_q + 19);
    uint32_t *ones = reinterpret_cast<uint32_t *>(bar_q + 24);  // 512 B, 16 x 16 f16 1.0
    // [1 or 2][4][128] partial row maxima.  Two buffers (by block parity) when this sentence has <= 2 chunks per pass: a
    // warp ahead of the others could otherwise publish the next block's maximum while a slow one still reads this
    // block's.  With >= 3 chunks a group pair cannot get more than two P slots ahead of the other pair (the control warp
    // consumes the slots in order), so it cannot finish a block alone -- and only then is there no room for a second buffer.
    float *red = reinterpret_cast<float *>(ones + 128);
    const bool red2 = Lk64 <= 256;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp == CW) {
        if (lane == 0) {
            tma_prefetch_desc(&tmQKV);
            for (int i = 0; i < 4; i++) mbar_init(&bar_k[i], 1);
            mbar_init(bar_v, 1); mbar_init(bar_o, 1);
            mbar_init(&bar_q[0], 1); mbar_init(&bar_q[1], 1);
            for (int i = 0; i < 3; i++) { mbar_init(&sfull[i], 1); mbar_init(&sfree[i], 16); }
            for (int i = 0; i < 4; i++) { mbar_init(&pfull[i], 8); mbar_init(&pfree[i], 1); }
            mbar_fence_init();
        }
        __syncwarp();
        {
            // the loads first (they need the barriers only): their latency runs under the TMEM allocation and the CTA-wide sync
            const int cq = head * DH, ck = d + head * DH, cv = 2 * d + head * DH;
            const int n_box = Lk64 >> 6;
            mbar_expect_tx_e(&bar_q[0], (uint32_t)(kAttQ * ROWB));
            tma_load_2d_e(sQ, &tmQKV, &bar_q[0], cq, t0);
            tma_load_2d_e(sQ + 64 * ROWB, &tmQKV, &bar_q[0], cq, t0 + 64);
            for (int c = 0; c < nc; c++) {  // K in chunks of 128 keys (two 64-row boxes, the last chunk possibly one)
                const int boxes = n_box - 2 * c < 2 ? n_box - 2 * c : 2;
                mbar_expect_tx_e(&bar_k[c], (uint32_t)(boxes * 64 * ROWB));
                for (int b = 0; b < boxes; b++) tma_load_2d_e(sK + (2 * c + b) * 64 * ROWB, &tmQKV, &bar_k[c], ck, t0 + (2 * c + b) * 64);
            }
            if (nqb > 1) {
                uint8_t *q1 = sQ + kAttQ * ROWB;
                mbar_expect_tx_e(&bar_q[1], (uint32_t)(kAttQ * ROWB));
                tma_load_2d_e(q1, &tmQKV, &bar_q[1], cq, t0 + kAttQ);
                tma_load_2d_e(q1 + 64 * ROWB, &tmQKV, &bar_q[1], cq, t0 + kAttQ + 64);
            }
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

    if (warp == CW) {
        // ================================================================ score warp: TMA loads, S = Q . K^T jobs
        const int cq = head * DH;  // (Q(0), Q(1), K and V were requested in the prologue)
        auto load_q = [&](int qb) {  // query block qb -> Q buffer qb % 2
            uint8_t *q = sQ + (qb % kStreamQ) * kAttQ * ROWB;
            uint64_t *bar = &bar_q[qb % kStreamQ];
            mbar_expect_tx_e(bar, (uint32_t)(kAttQ * ROWB));
            tma_load_2d_e(q, &tmQKV, bar, cq, t0 + qb * kAttQ);
            tma_load_2d_e(q + 64 * ROWB, &tmQKV, bar, cq, t0 + qb * kAttQ + 64);
        };

        // score jobs in the order the softmax warps consume them: A(0,0..nc-1), then B(q,c) A(q+1,c) for c = 0..nc-1, q = 0..
        const int n_jobs = 2 * nqb * nc;
        int slot = 0;
        uint32_t sph = 0;                 // parity of the current use of ring slot `slot`
        bool prologue = true, job_a = true;
        int jq = 0, jc = 0;               // B jobs: block jq, chunk jc; A jobs: block jq + 1 (block 0 in the prologue), chunk jc
        const uint64_t dk0 = umma_desc_kmajor<ROWB>(smem_u32(sK));
        const uint64_t dq0 = umma_desc_kmajor<ROWB>(smem_u32(sQ));
        for (int js = 0; js < n_jobs; js++) {
            if (js < nc) mbar_wait(&bar_k[js], 0);  // prologue job A(0, js): the first user of key chunk js
            if (js >= kStreamRing) mbar_wait(&sfree[slot], sph ^ 1);  // the slot has been read by all softmax warps (three jobs ago)
            const int blk = job_a ? (prologue ? 0 : jq + 1) : jq;
            if (job_a && jc == 0) {
                // first use of Q(blk).  Its buffer held Q(blk - 2), last read by B(blk - 2, nc - 1) -- three jobs back, retired
                // (slot hand-back above)
                if (blk >= 2) load_q(blk);
                mbar_wait(&bar_q[blk % kStreamQ], (uint32_t)((blk / kStreamQ) & 1));
            }
            const int n0 = jc * kStreamKeys;
            const int nn = Lk16 - n0 < kStreamKeys ? Lk16 - n0 : kStreamKeys;
            const uint64_t dq = dq0 + (uint64_t)(((blk % kStreamQ) * kAttQ * ROWB) >> 4);
            const uint64_t dk = dk0 + (uint64_t)((n0 * ROWB) >> 4);
            const uint32_t idesc = att_idesc((uint32_t)nn, 0);
            tc_fence_after();
#pragma unroll
            for (int k = 0; k < DH / 16; k++)
                umma_f16_e<1>(tmem_base + (uint32_t)(slot * kStreamKeys), dq + (uint64_t)(2 * k), dk + (uint64_t)(2 * k), idesc, k != 0);
            umma_commit_e<1>(&sfull[slot]);
            if (++slot == kStreamRing) { slot = 0; sph ^= 1; }
            if (prologue) {
                if (++jc == nc) { prologue = false; job_a = false; jc = 0; }
            } else if (!job_a) {
                if (jq + 1 < nqb) job_a = true;
                else if (++jc == nc) { jc = 0; jq++; }
            } else {
                job_a = false;
                if (++jc == nc) { jc = 0; jq++; }
            }
        }
    } else if (warp == CW + 1) {
        // ================================================================ P.V warp: O += P . V, rowsum += P . 1 per delivered P slot
        const uint32_t idesc_pv = att_idesc(DH, 1), idesc_sum = att_idesc(kSumCols, 0);
        const uint64_t done = umma_desc_ones(smem_u32(ones));
        const uint64_t dp0 = umma_desc_sw128(smem_u32(sP));
        const uint64_t dv0 = umma_desc_mnmajor<ROWB>(smem_u32(sV));
        mbar_wait(bar_v, 0);
        uint32_t pv_b = 0;  // pass-B jobs consumed
        for (int qb = 0; qb < nqb; qb++) {
            for (int c = 0; c < nc; c++, pv_b++) {
                for (int sub = 0; sub < 2; sub++) {
                    const uint32_t ps = (uint32_t)sub * 2 + (pv_b & 1);
                    mbar_wait(&pfull[ps], (pv_b >> 1) & 1);
                    tc_fence_after();
                    const int key0 = c * kStreamKeys + sub * 64;
                    const int rem = Lk16 - key0;
                    const int ksteps = rem >= 64 ? 4 : rem > 0 ? rem >> 4 : 0;
                    const uint64_t dp = dp0 + (uint64_t)((ps * PSUB) >> 4);
                    const uint64_t dv = dv0 + (uint64_t)((key0 * ROWB) >> 4);
                    for (int ks = 0; ks < ksteps; ks++)
                        umma_f16_e<1>(tmem_base + kOCol, dp + (uint64_t)(2 * ks), dv + (uint64_t)((ks * 16 * ROWB) >> 4), idesc_pv, (c | sub | ks) != 0);
                    for (int ks = 0; ks < ksteps; ks++)
                        umma_f16_e<1>(tmem_base + kOCol + (uint32_t)DH, dp + (uint64_t)(2 * ks), done, idesc_sum, (c | sub | ks) != 0);
                    umma_commit_e<1>(&pfree[ps]);
                }
            }
            umma_commit_e<1>(bar_o);
        }
    } else {
        // ======================================================================= softmax warps
        const int quad = warp & 3, g = warp >> 2;
        const int r = quad * 32 + lane;                       // query row in the block == TMEM lane
        const uint32_t tlane = tmem_base + ((uint32_t)(quad * 32) << 16);
        const uint32_t row_off = (uint32_t)r * 128u, sw = (uint32_t)(r & 7);
        const int psub = g >> 1, phalf = g & 1;               // my 32 keys inside a chunk: P slot (64-key half), 32-key half of it
        uint32_t slot = 0, sph = 0;                           // score ring position of the next job
        uint32_t nb = 0;                                      // pass-B jobs done (P slot uses)

        // one pass-A job: fold this group's 32 keys of chunk c into the running row maximum
        auto step_a = [&](int c, float mx) {
            mbar_wait(&sfull[slot], sph);
            tc_fence_after();
            const int key0 = c * kStreamKeys + g * 32;
            if (key0 < L) {
                uint32_t a[32];
                tmem_ld_32x32(tlane + slot * kStreamKeys + (uint32_t)(g * 32), a);
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
            if (lane == 0) mbar_arrive(&sfree[slot]);
            if (++slot == kStreamRing) { slot = 0; sph ^= 1; }
            return mx;
        };
        // O * (1 / sum p) of query block qb -> out rows
        auto epilogue = [&](int qb) {
            if (g >= DH / 32) return;  // groups 0 .. dh/32 - 1 each scale and store 32 features
            mbar_wait(bar_o, (uint32_t)(qb & 1));
            tc_fence_after();
            const float sum = __uint_as_float(tmem_ld_32x1(tlane + kOCol + (uint32_t)DH));
            uint32_t a[32];
            tmem_ld_32x32(tlane + kOCol + (uint32_t)(g * 32), a);
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
                uint4 *dst = reinterpret_cast<uint4 *>(out + (size_t)(t0 + q0 + r) * d + (size_t)head * DH + g * 32);
#pragma unroll
                for (int q = 0; q < 4; q++)
                    dst[q] = make_uint4(pack_h2(o[8 * q], o[8 * q + 1]), pack_h2(o[8 * q + 2], o[8 * q + 3]),
                                        pack_h2(o[8 * q + 4], o[8 * q + 5]), pack_h2(o[8 * q + 6], o[8 * q + 7]));
            }
        };

        float mx_next = -INFINITY;
        for (int c = 0; c < nc; c++) mx_next = step_a(c, mx_next);  // pass A of block 0 (not overlapped)
        for (int qb = 0; qb < nqb; qb++) {
            // ---- row maximum of this block over the four groups
            float mx = mx_next;
            {
                float *rb = red + (red2 ? (qb & 1) * 512 : 0);
                rb[g * 128 + r] = mx;
                named_bar_sync(1, 512);
#pragma unroll
                for (int gg = 0; gg < 4; gg++) mx = fmaxf(mx, rb[gg * 128 + r]);
            }
            mx *= scale;  // max(s*scale) == max(s)*scale for scale > 0 (rounding is monotonic)
            mx_next = -INFINITY;
            // the previous block's O: its last P.V MMAs were issued a whole pass-A chunk ago.  Read before this block's first
            // P slot is handed over (the MMA on it overwrites O): the groups that read O are the ones that fill P slot 0
            if (qb > 0) epilogue(qb - 1);
            for (int c = 0; c < nc; c++) {
                // ---- pass B job: p = f16(exp(f16(s*scale - max))) for my 32 keys of chunk c
                mbar_wait(&sfull[slot], sph);
                tc_fence_after();
                const int key0 = c * kStreamKeys + g * 32;
                const bool mine = key0 < L;  // otherwise no real key: the MMA does not read these P columns
                uint32_t pk[16];
                if (mine) {
                    uint32_t a[32];
                    tmem_ld_32x32(tlane + slot * kStreamKeys + (uint32_t)(g * 32), a);
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
                if (lane == 0) mbar_arrive(&sfree[slot]);
                if (++slot == kStreamRing) { slot = 0; sph ^= 1; }
                const uint32_t pslot = (uint32_t)psub * 2 + (nb & 1);
                if (nb >= 2) mbar_wait(&pfree[pslot], ((nb - 2) >> 1) & 1);  // the MMAs on the previous contents of this P slot have retired
                if (mine) {
                    uint8_t *pb = sP + pslot * PSUB;
#pragma unroll
                    for (int q = 0; q < 4; q++)
                        *reinterpret_cast<uint4 *>(pb + row_off + ((((uint32_t)(phalf * 4 + q)) ^ sw) << 4)) =
                            make_uint4(pk[4 * q], pk[4 * q + 1], pk[4 * q + 2], pk[4 * q + 3]);
                    fence_proxy_async_smem();  // P writes -> visible to the MMA (async proxy)
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(&pfull[pslot]);
                nb++;
                // ---- pass A job of the next block, same chunk
                if (qb + 1 < nqb) mx_next = step_a(c, mx_next);
            }
        }
        epilogue(nqb - 1);
    }

    tc_fence_before();
    __syncthreads();
    if (warp == CW) {
        tc_fence_after();
        tmem_dealloc<512>(tmem_base);
    }
}

template <int DH>
static cudaError_t launch_att_stream(const CUtensorMap &tm, const int32_t *cu, int n_seq, int max_len, int d, int n_head, __half *out,
                                     int qmode, cudaStream_t st) {
    const int Lk64 = (max_len + 63) & ~63;
    // a shorter sentence inside a longer launch finds its second exchange buffer in the K / V space it does not use
    const size_t smem = (size_t)kStreamQ * kAttQ * DH * 2 + 4 * kAttQ * 128 + (size_t)2 * Lk64 * DH * 2 + kStreamTail + (Lk64 <= 256 ? 2 : 1) * 2048;
    static std::atomic<bool> attr_set[64];
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64 || !attr_set[dev]) {
        cudaError_t err = cudaFuncSetAttribute(attention_stream_kernel<DH>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
        if (err != cudaSuccess) return err;
        if (dev >= 0 && dev < 64) attr_set[dev] = true;
    }
    dim3 grid(1, n_head, n_seq);
    const float scale = (float)(1.0 / sqrt((double)DH));  // bert.cpp:934
    LaunchConfig lc(grid, dim3(kStreamThreads), smem, st);
    return cudaLaunchKernelEx(&lc.cfg, attention_stream_kernel<DH>, tm, cu, d, out, qmode, scale);
}
