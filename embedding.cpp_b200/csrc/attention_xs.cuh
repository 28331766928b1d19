// attention_xs.cuh -- attention for launches with sentences of more than 128 tokens: attention_pt_kernel (P in tensor memory,
// persistent CTAs) with the warps split by RESOURCE instead of by key group.  Included by attention_tc.cu after attention_pt.cuh,
// whose helpers (PtItem, pt_item, pt_ahead, store_rows64, umma_f16_ts_e, umma_desc_mnmajor_ones, tmem_st_32x16) it uses.
//
// What bounds these kernels is the MUFU: L^2 exponentials per head at 16 per clock per SM = 1024 cycles per 128 x 128 chunk.  In
// attention_pt_kernel every one of the 16 softmax warps does everything for its 32 rows x 32 keys of a chunk -- wait, TMEM load,
// exponentials, TMEM store, hand-over, then the same for the maximum pass -- about as many cycles of latencies as of MUFU time, and
// the four warps of a scheduler do it nearly in step (tools/att_trace.py): the MUFU idles half of the time.  Here
//
//   * 8 EXP warps (TMEM lane quarter q, 64-key half h) do nothing but pass B: 32 rows x 64 keys per chunk, two per scheduler;
//   * 8 SERVICE warps (q, h) do everything else: the maximum pass over their 64 keys of every A-slot chunk, the exchange of the two
//     halves' maxima and their publication to the exp warps, and the output of finished query blocks (features [32 h, 32 h + 32));
//   * the score jobs of the two passes are issued by TWO warps (pass B + all TMA loads of Q and K; pass A), so that a late hand-back of
//     the A slot -- the service warps are busy with an output -- never holds up a pass-B job behind it; a third issues P.V (+ V loads).
//
// Job order per item, slots, the next item's first maximum pass in the idle A slot of the last block, Q-buffer alternation, arithmetic,
// rounding points and the order of the 16-key MMA steps are those of attention_pt_kernel: bit-identical results.
//
// MEASURED: no faster than attention_pt_kernel (L = 512: 274 vs 268 us per 65 536 tokens, 256: 161 vs 168; profiles/
// r02l_attention_long.md) -- two exp warps per scheduler run the exponential loop at 25 cycles per pair of scores and scheduler,
// four symmetric warps at 13: the loop is bound by issue and dependency latency, not by the MUFU, and wants more warps, not fewer.
// Opt-in (BERT_B200_ATT_XS=1), kept for the bit-equality test and as the record of the experiment.
#pragma once

constexpr int kXsThreads = 19 * 32;    // 8 exp + 8 service warps, pass-B score warp, P.V warp, pass-A score warp
constexpr int kXsTail = 256 + 3 * 512 + 2 * 1024 + 8 * 2048;  // barriers; row maxima of 3 blocks; 2 exchange buffers; output staging

template <int DH>
__global__ void __launch_bounds__(kXsThreads, 1)
attention_xs_kernel(const __grid_constant__ CUtensorMap tmQKV, const int32_t *__restrict__ cu, int n_items, int n_head, int lk_max, int d,
                    __half *__restrict__ out, int qmode, float scale) {
    constexpr int ROWB = DH * 2;
    constexpr int WB = 16, WPV = 17, WA = 18;              // pass-B score warp (+ Q / K loads), P.V warp (+ V loads, TMEM allocation), pass-A score warp
    constexpr uint32_t kOCol = 3 * kStreamKeys;            // TMEM: pass-B slots [0,128) [128,256), A slot [256,384), O [384,384+dh) | row sums [.., +16)
    extern __shared__ __align__(1024) uint8_t att_raw[];
    uint8_t *smem = att_raw;

    pdl_launch_dependents();
    pdl_wait();  // the sequence bounds are read first thing

    // the layout is that of the longest sentence of the launch (lk_max = its length rounded up to 64)
    uint8_t *sQ = smem;                          // [2][128 rows]
    uint8_t *sK = sQ + kStreamQ * kAttQ * ROWB;
    uint8_t *sV = sK + lk_max * ROWB;
    uint8_t *sOnes = sV + lk_max * ROWB;         // 16 keys x ROWB of f16 1.0: the second MN atom of every P.V step
    uint8_t *tail = sOnes + 16 * ROWB;
    uint64_t *bar_q = reinterpret_cast<uint64_t *>(tail);  // [2] Q buffer landed
    uint64_t *bar_v = bar_q + 2;
    uint64_t *bar_k = bar_q + 3;    // [4] keys [128 c, 128 c + 128) landed
    uint64_t *sfull = bar_q + 7;    // [3] score chunk complete in the slot (0, 1: pass B by chunk parity; 2: pass A)
    uint64_t *sfree = bar_q + 10;   // [3] slot may be overwritten (0, 1: the MMAs that read its P have retired; 2: read by the 8 service warps)
    uint64_t *pfull = bar_q + 13;   // [2 slots][2 halves] P of a 64-key half written by its 4 exp warps
    uint64_t *bar_o = bar_q + 17;   // O and row sums of a query block complete
    uint64_t *ofree = bar_q + 18;   // O and the row sums are in the registers of the service warps
    uint64_t *mxfull = bar_q + 19;  // [3] row maxima of a query block published (block number % 3)
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bar_q + 22);
    float *mxbuf = reinterpret_cast<float *>(tail + 256);              // [3][128] row maxima of a query block
    float *red = reinterpret_cast<float *>(tail + 256 + 3 * 512);      // [2][2 halves][128] partial maxima (by publication parity)
    uint8_t *stage = tail + 256 + 3 * 512 + 2 * 1024;                  // [8 warps][2 KB] output staging

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#ifdef BK_EXP_TRACE
    const bool trace_on = blockIdx.x == 5;
    int tr_n = 0;
    const int tw = warp == 0 ? 0 : warp == 4 ? 1 : warp == 8 ? 2 : warp == 12 ? 3 : warp == WB ? 4 : warp == WPV ? 5 : warp == WA ? 6 : 7;
    BK_TR(tw, 0);
#endif
    // every warp walks the same item list; the bounds of the next item are fetched ahead
    const int item0 = (int)blockIdx.x, istep = (int)gridDim.x;
    auto fetch = [&](int i, int &t0, int &L) {
        if (i < n_items) { const int s = i / n_head; t0 = __ldg(cu + s); L = __ldg(cu + s + 1) - t0; } else { t0 = 0; L = 0; }
    };
    int nx_t0, nx_L;
    fetch(item0, nx_t0, nx_L);

    auto load_q = [&](const PtItem &it, int blk, int buf) {  // query block blk -> Q buffer buf
        if (blk >= it.nqb) return;
        const int cq = it.head * DH;
        uint8_t *q = sQ + buf * kAttQ * ROWB;
        mbar_expect_tx_e(&bar_q[buf], (uint32_t)(kAttQ * ROWB));
        tma_load_2d_e(q, &tmQKV, &bar_q[buf], cq, it.t0 + blk * kAttQ);
        tma_load_2d_e(q + 64 * ROWB, &tmQKV, &bar_q[buf], cq, it.t0 + blk * kAttQ + 64);
    };
    auto load_k = [&](const PtItem &it, int c) {  // keys [128 c, 128 c + 128): two 64-row boxes, the last chunk possibly one
        if (c >= it.nc) return;
        const int ck = d + it.head * DH;
        const int n_box = it.Lk64 >> 6;
        const int boxes = n_box - 2 * c < 2 ? n_box - 2 * c : 2;
        mbar_expect_tx_e(&bar_k[c], (uint32_t)(boxes * 64 * ROWB));
        for (int b = 0; b < boxes; b++) tma_load_2d_e(sK + (2 * c + b) * 64 * ROWB, &tmQKV, &bar_k[c], ck, it.t0 + (2 * c + b) * 64);
    };
    auto load_v = [&](const PtItem &it) {
        const int cv = 2 * d + it.head * DH;
        const int n_box = it.Lk64 >> 6;
        mbar_expect_tx_e(bar_v, (uint32_t)(it.Lk64 * ROWB));
        for (int c = 0; c < n_box; c++) tma_load_2d_e(sV + c * 64 * ROWB, &tmQKV, bar_v, cv, it.t0 + c * 64);
    };

    if (warp == WB) {
        if (lane == 0) {
            tma_prefetch_desc(&tmQKV);
            for (int i = 0; i < 4; i++) mbar_init(&bar_k[i], 1);
            mbar_init(bar_v, 1); mbar_init(bar_o, 1); mbar_init(ofree, 4 * (DH / 32));
            mbar_init(&bar_q[0], 1); mbar_init(&bar_q[1], 1);
            for (int i = 0; i < 3; i++) { mbar_init(&sfull[i], 1); mbar_init(&sfree[i], i < 2 ? 1 : 8); mbar_init(&mxfull[i], 4); }
            for (int i = 0; i < 4; i++) mbar_init(&pfull[i], 4);
            mbar_fence_init();
        }
        __syncwarp();
        // the first item's loads (they need the barriers only): their latency runs under the TMEM allocation and the CTA-wide sync
        const PtItem it = pt_item(nx_t0, nx_L, item0 % n_head);
        load_q(it, 0, 0);
        for (int c = 0; c < 4; c++) load_k(it, c);
        load_q(it, 1, 1);
        load_v(it);
    } else if (warp == WPV) {  // meanwhile: the all-ones tile and the TMEM allocation
        for (int i = lane; i < 16 * ROWB / 4; i += 32) reinterpret_cast<uint32_t *>(sOnes)[i] = 0x3C003C00u;
        fence_proxy_async_smem();
        __syncwarp();
        tmem_alloc<512>(tmem_slot);
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == WB || warp == WA) {
        // ================================================================ score warps: S = Q . K^T jobs (WB also: the TMA loads of Q and K)
        // Q buffers: block q of an item sits in buffer (q + qoff) % 2, qoff chosen per item so that its Q(0) can be fetched into the
        // buffer the previous item's last block does NOT read.  Both warps wait for every Q / K load they read (each exactly once).
        uint32_t used = 0, par = 0;       // per score slot (bit): has been used / number of uses so far is odd
        uint32_t kpar = 0, qpar = 0;      // per K-chunk / Q-buffer barrier (bit): parity of the next phase to wait for
        uint32_t nbj = 0;                 // pass-B jobs issued
        int pre_in = 0;                   // chunks of this item's A(0,*) done during the previous item
        int qoff = 0;
        const uint64_t dk0 = umma_desc_kmajor<ROWB>(smem_u32(sK));
        const uint64_t dq0 = umma_desc_kmajor<ROWB>(smem_u32(sQ));
        auto wait_k = [&](int c) { mbar_wait(&bar_k[c], (kpar >> c) & 1); kpar ^= 1u << c; };
        auto wait_q = [&](int b) { mbar_wait(&bar_q[b], (qpar >> b) & 1); qpar ^= 1u << b; };
        auto wait_scores = [&](uint32_t slot) { mbar_wait(&sfull[slot], ((par >> slot) & 1) ^ 1); };  // the latest job issued on `slot` has retired
        auto issue = [&](uint32_t slot, int qbuf, int chunk, int Lk16) {  // S[slot] = Q[qbuf] . K[chunk]^T
            BK_TR(tw, 20);
            if ((used >> slot) & 1) mbar_wait(&sfree[slot], ((par >> slot) & 1) ^ 1);  // the previous contents have been consumed
            used |= 1u << slot;
            par ^= 1u << slot;
            BK_TR(tw, 21);
            const int n0 = chunk * kStreamKeys;
            const int nn = Lk16 - n0 < kStreamKeys ? Lk16 - n0 : kStreamKeys;
            const uint64_t dq = dq0 + (uint64_t)((qbuf * kAttQ * ROWB) >> 4);
            const uint64_t dk = dk0 + (uint64_t)((n0 * ROWB) >> 4);
            const uint32_t idesc = att_idesc((uint32_t)nn, 0);
            tc_fence_after();
#pragma unroll
            for (int k = 0; k < DH / 16; k++)
                umma_f16_e<1>(tmem_base + slot * kStreamKeys, dq + (uint64_t)(2 * k), dk + (uint64_t)(2 * k), idesc, k != 0);
            umma_commit_e<1>(&sfull[slot]);
            BK_TR(tw, 22);
        };
        int n2_t0, n2_L;  // the next item's bounds are needed during the current one: fetched two items ahead
        fetch(item0 + istep, n2_t0, n2_L);
        for (int item = item0; item < n_items; item += istep) {
            const PtItem it = pt_item(nx_t0, nx_L, item % n_head);
            nx_t0 = n2_t0; nx_L = n2_L;
            fetch(item + 2 * istep, n2_t0, n2_L);
            const bool more = item + istep < n_items;
            const PtItem nxt = pt_item(nx_t0, nx_L, (item + istep) % n_head);  // (valid if `more`)
            const int pre_out = pt_ahead(it, nxt, more);
            const int nqb = it.nqb, nc = it.nc;
            const int lastbuf = (nqb - 1 + qoff) & 1;  // the buffer the last query block reads
            if (warp == WB) {
                // ---- pass B: B(q, c) -> slot (chunk number) % 2, handed back by the P.V warp
                for (int q = 0; q < nqb; q++) {
                    const int qbuf = (q + qoff) & 1;
                    for (int c = 0; c < nc; c++) {
                        if (q == 0) wait_k(c);
                        if (c == 0) wait_q(qbuf);
                        const uint32_t sb = nbj & 1;
                        nbj++;
                        issue(sb, qbuf, c, it.Lk16);
                        if (q + 1 < nqb) {
                            if (c == nc - 1 && q + 2 < nqb) {
                                // B(q, nc - 1) was the last reader of Q(q): refill its buffer with Q(q + 2), a round of pass-B jobs before
                                // B(q + 2, 0) and as early as pass A of block q + 2 can use it
                                wait_scores(sb);
                                load_q(it, q + 2, qbuf);
                            }
                        } else if (more) {
                            // last query block: B(last, c) was the last reader of key chunk c -- refill it with the next item's keys as
                            // soon as its MMAs have retired; with the first refill also the next Q(0), into the buffer this block does not read
                            wait_scores(sb);
                            load_k(nxt, c);
                            if (c == 0) load_q(nxt, 0, lastbuf ^ 1);
                        }
                    }
                }
                if (more) {  // (every score MMA of this item has retired: see the last refill)
                    for (int c = nc; c < 4; c++) load_k(nxt, c);  // chunks the next item has beyond this one's
                    load_q(nxt, 1, lastbuf);
                }
            } else {
                // ---- pass A: slot 2, handed back by the service warps.  What is left of A(0, *), then A(q + 1, *) during block q; during
                // the last block the next item's A'(0, c - 1), as soon as key chunk c - 1 holds that item's keys
                for (int c = pre_in; c < nc; c++) {
                    wait_k(c);
                    if (c == 0) wait_q(qoff);
                    issue(2u, qoff, c, it.Lk16);
                }
                for (int q = 0; q < nqb; q++) {
                    const int qbuf = (q + qoff) & 1;
                    if (q + 1 < nqb) {
                        for (int c = 0; c < nc; c++) {
                            if (c == 0) wait_q(qbuf ^ 1);
                            issue(2u, qbuf ^ 1, c, it.Lk16);
                        }
                    } else {
                        for (int c = 0; c < pre_out; c++) {
                            wait_k(c);
                            if (c == 0) wait_q(lastbuf ^ 1);
                            issue(2u, lastbuf ^ 1, c, nxt.Lk16);
                        }
                    }
                }
            }
            pre_in = pre_out;
            qoff = lastbuf ^ 1;
        }
    } else if (warp == WPV) {
        // ================================================================ P.V warp: [O | row sums] += P . [V | 1], P read from the score slot
        // one MMA per 16 keys, N = dh + 16: the row sums come out of the same instruction
        const uint32_t idesc_pv = att_idesc(DH + kSumCols, 1);
        const uint32_t v0 = smem_u32(sV), o1 = smem_u32(sOnes);
        uint32_t pv_b = 0;  // pass-B jobs consumed
        uint32_t gb = 0;    // query blocks done by this CTA
        uint32_t vpar = 0;
        for (int item = item0; item < n_items; item += istep) {
            const PtItem it = pt_item(nx_t0, nx_L, item % n_head);
            fetch(item + istep, nx_t0, nx_L);
            mbar_wait(bar_v, vpar);
            vpar ^= 1;
            for (int qb = 0; qb < it.nqb; qb++, gb++) {
                for (int c = 0; c < it.nc; c++, pv_b++) {
                    const uint32_t slot = pv_b & 1;
                    for (int sub = 0; sub < 2; sub++) {
                        BK_TR(tw, 30 + sub);
                        mbar_wait(&pfull[slot * 2 + sub], (pv_b >> 1) & 1);
                        if (gb > 0 && (c | sub) == 0) mbar_wait(ofree, (gb - 1) & 1);  // the first MMA of a block overwrites O
                        BK_TR(tw, 32 + sub);
                        tc_fence_after();
                        const int key0 = c * kStreamKeys + sub * 64;
                        const int rem = it.Lk16 - key0;
                        const int ksteps = rem >= 64 ? 4 : rem > 0 ? rem >> 4 : 0;
                        const uint32_t pa = tmem_base + slot * kStreamKeys + (uint32_t)(sub * 64);  // keys 16 ks ..: 32-key group 2 sub + ks / 2, its columns 8 (ks % 2) ..
                        for (int ks = 0; ks < ksteps; ks++) {
                            const uint32_t va = v0 + (uint32_t)((key0 + 16 * ks) * ROWB);
                            umma_f16_ts_e(tmem_base + kOCol, pa + (uint32_t)((ks >> 1) * 32 + (ks & 1) * 8), umma_desc_mnmajor_ones<ROWB>(va, o1 - va), idesc_pv,
                                          (c | sub | ks) != 0);
                        }
                    }
                    umma_commit_e<1>(&sfree[slot]);
                    BK_TR(tw, 34);
                }
                umma_commit_e<1>(bar_o);
            }
            // the next item's V: when this item's last P.V MMAs have retired
            if (item + istep < n_items) {
                mbar_wait(bar_o, (gb - 1) & 1);
                load_v(pt_item(nx_t0, nx_L, (item + istep) % n_head));
            }
        }
    } else if (warp >= 8) {
        // ======================================================================= service warps (q, h): row maxima and output
        const int quad = warp & 3, half = (warp >> 2) & 1;
        const int r = quad * 32 + lane;   // query row in the block == TMEM lane
        const uint32_t tlane = tmem_base + ((uint32_t)(quad * 32) << 16);
        uint8_t *st = stage + (warp - 8) * 2048;
        uint32_t gb = 0;      // query blocks started by this CTA
        uint32_t np = 0;      // maxima published
        uint32_t apar = 0;    // parity of the next A-slot chunk
        int pre_in = 0;
        float mx = -INFINITY; // running maximum, over my 64 keys per chunk, of the block whose pass A is under way
        int pv_t0 = 0, pv_L = 0, pv_head = 0, pv_qb = -1;  // the block whose output is pending
        auto consume = [&](int c, int L) {  // one pass-A chunk: my 64 keys of it
            BK_TR(tw, 7);
            mbar_wait(&sfull[2], apar);
            apar ^= 1;
            BK_TR(tw, 8);
            tc_fence_after();
#pragma unroll
            for (int j = 0; j < 2; j++) {
                const int key0 = c * kStreamKeys + half * 64 + j * 32;
                if (key0 < L) {
                    uint32_t a[32];
                    tmem_ld_32x32(tlane + 2 * kStreamKeys + (uint32_t)(half * 64 + j * 32), a);
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
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&sfree[2]);
            BK_TR(tw, 9);
        };
        // the maxima of query block g (CTA-wide count) are complete: combine the two key halves, hand them to the exp warps.
        // (mxbuf / mxfull are rings of 3: block g + 3 can be published only when every exp warp is past block g -- its Q is fetched
        // after a pass-B job of block g + 1 whose slot the P.V warp hands back only with all of their P of block g.)
        auto publish = [&](uint32_t g) {
            float *rb = red + (np & 1) * 256;
            rb[half * 128 + r] = mx;
            named_bar_sync(1, 256);
            const float m = fmaxf(rb[r], rb[128 + r]);
            if (half == 0) {
                mxbuf[(g % 3) * 128 + r] = m;
                __syncwarp();
                if (lane == 0) mbar_arrive(&mxfull[g % 3]);
            }
            mx = -INFINITY;
            np++;
        };
        auto output = [&]() {  // the pending block (CTA-wide number gb - 1): O * (1 / sum p) -> out rows, features [32 half, 32 half + 32)
            if (pv_qb < 0) return;
            if (half < DH / 32) {
                mbar_wait(bar_o, (gb - 1) & 1);
                tc_fence_after();
                const float sum = __uint_as_float(tmem_ld_32x1(tlane + kOCol + (uint32_t)DH));
                uint32_t a[32];
                tmem_ld_32x32(tlane + kOCol + (uint32_t)(half * 32), a);
                tmem_ld_wait();
                tc_fence_before();  // O is in registers: the P.V warp's first MMA of the next block may overwrite it
                __syncwarp();
                if (lane == 0) mbar_arrive(ofree);
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
                uint32_t pk[16];
#pragma unroll
                for (int i = 0; i < 16; i++) pk[i] = pack_h2(o[2 * i], o[2 * i + 1]);
                const int row0 = pv_qb * kAttQ + quad * 32;
                store_rows64(st, pk, out + (size_t)(pv_t0 + row0) * d + (size_t)pv_head * DH + half * 32, (size_t)d, pv_L - row0, lane);
            }
            pv_qb = -1;
        };
        int n2_t0, n2_L;  // the next item's length is needed during the current one: fetched two items ahead
        fetch(item0 + istep, n2_t0, n2_L);
        for (int item = item0; item < n_items; item += istep) {
            const PtItem it = pt_item(nx_t0, nx_L, item % n_head);
            nx_t0 = n2_t0; nx_L = n2_L;
            fetch(item + 2 * istep, n2_t0, n2_L);
            const bool more = item + istep < n_items;
            const PtItem nxt = pt_item(nx_t0, nx_L, (item + istep) % n_head);  // (valid if `more`)
            const int pre_out = pt_ahead(it, nxt, more);
            for (int c = pre_in; c < it.nc; c++) consume(c, it.L);  // what is left of pass A of block 0
            publish(gb);
            for (int q = 0; q < it.nqb; q++) {
                // block q runs in the exp warps now.  First the previous block's output (the P.V warp needs `ofree` before this block's
                // first MMA, a chunk of exponentials from now), then the A slot's jobs of this round
                output();
                if (q + 1 < it.nqb) {
                    for (int c = 0; c < it.nc; c++) consume(c, it.L);
                    publish(gb + 1);
                } else {
                    for (int c = 0; c < pre_out; c++) consume(c, nxt.L);
                }
                pv_t0 = it.t0; pv_L = it.L; pv_head = it.head; pv_qb = q;
                gb++;
            }
            pre_in = pre_out;
        }
        output();
    } else {
        // ======================================================================= exp warps (q, h): pass B, 32 rows x 64 keys per chunk
        const int quad = warp & 3, half = warp >> 2;
        const int r = quad * 32 + lane;                       // query row in the block == TMEM lane
        const uint32_t tlane = tmem_base + ((uint32_t)(quad * 32) << 16);
        const uint32_t tcol = tlane + (uint32_t)(half * 64);  // my 64 columns of a slot
        uint32_t nb = 0;                                      // pass-B jobs done
        uint32_t gb = 0;                                      // query blocks done by this CTA
        auto softmax32 = [&](uint32_t (&a)[32], int nv, float mx, uint32_t (&pk)[16]) {  // p = f16(exp(f16(s*scale - max))), 32 keys
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
        };
        for (int item = item0; item < n_items; item += istep) {
            const PtItem it = pt_item(nx_t0, nx_L, item % n_head);
            fetch(item + istep, nx_t0, nx_L);
            const int L = it.L, nc = it.nc, nqb = it.nqb;
            for (int qb = 0; qb < nqb; qb++, gb++) {
                // ---- row maximum of this block, from the service warps
                mbar_wait(&mxfull[gb % 3], (gb / 3) & 1);
                const float mx = mxbuf[(gb % 3) * 128 + r] * scale;  // max(s*scale) == max(s)*scale for scale > 0 (rounding is monotonic)
                for (int c = 0; c < nc; c++, nb++) {
                    const uint32_t slot = nb & 1;
                    BK_TR(tw, 1);
                    mbar_wait(&sfull[slot], (nb >> 1) & 1);
                    BK_TR(tw, 2);
                    tc_fence_after();
                    const int key0 = c * kStreamKeys + half * 64;
                    const uint32_t ts = tcol + slot * kStreamKeys;
                    if (key0 < L) {  // otherwise no real key in my half: the MMA does not read these P columns
                        const bool two = key0 + 32 < L;
                        uint32_t a0[32], a1[32], pk[16];
                        tmem_ld_32x32(ts, a0);
                        if (two) tmem_ld_32x32(ts + 32u, a1);  // (in flight under the first 32 keys' exponentials)
                        tmem_ld_wait();
                        BK_TR(tw, 3);
                        softmax32(a0, L - key0, mx, pk);
                        tmem_st_32x16(ts, pk);  // written back over the scores: keys 16 k .. of a 32-key group in its columns 8 (k % 2) ..
                        if (two) {
                            softmax32(a1, L - key0 - 32, mx, pk);
                            tmem_st_32x16(ts + 32u, pk);
                        }
                        BK_TR(tw, 4);
                        tmem_st_wait();
                        BK_TR(tw, 5);
                    }
                    tc_fence_before();  // P in tensor memory -> visible to the MMAs issued behind the barrier
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&pfull[slot * 2 + half]);
                    BK_TR(tw, 6);
                }
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == WPV) {
        tc_fence_after();
        tmem_dealloc<512>(tmem_base);
    }
}

template <int DH>
static cudaError_t launch_att_xs(const CUtensorMap &tm, const int32_t *cu, int n_seq, int max_len, int d, int n_head, __half *out,
                                 int qmode, cudaStream_t st) {
    const int Lk64 = (max_len + 63) & ~63;
    const size_t smem = (size_t)kStreamQ * kAttQ * DH * 2 + (size_t)2 * Lk64 * DH * 2 + 16 * DH * 2 + kXsTail;
    static std::atomic<bool> attr_set[64];
    static std::atomic<int> n_sm[64];
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
    if (!attr_set[dev]) {
        cudaError_t err = cudaFuncSetAttribute(attention_xs_kernel<DH>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
        if (err != cudaSuccess) return err;
        int n = 0;
        err = cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
        if (err != cudaSuccess) return err;
        n_sm[dev] = n;
        attr_set[dev] = true;
    }
    // one CTA per SM (tensor memory: all 512 columns), each walking over items (sentence, head)
    const long long n_items = (long long)n_seq * n_head;
    const int sms = n_sm[dev];
    dim3 grid((unsigned)(n_items < sms ? n_items : sms));
    const float scale = (float)(1.0 / sqrt((double)DH));  // bert.cpp:934
    LaunchConfig lc(grid, dim3(kXsThreads), smem, st);
    return cudaLaunchKernelEx(&lc.cfg, attention_xs_kernel<DH>, tm, cu, (int)n_items, n_head, Lk64, d, out, qmode, scale);
}
