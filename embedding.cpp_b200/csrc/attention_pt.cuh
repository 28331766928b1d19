// attention_pt.cuh -- attention for launches with sentences of more than 128 tokens, P kept in tensor memory.
// Included by attention_tc.cu (inside namespace bk, after attention_stream.cuh whose constants and structure it shares).
//
// attention_stream_kernel hands P to the tensor core through shared memory.  At L = 512 the shared-memory data pipe is then the
// busiest unit of the SM (ncu, profiles/r02l_*: 49 % -- tensor-core operand reads 38 %, of which 22 points are the P.V and
// row-sum MMAs re-reading P, plus 9 % for the P stores; MUFU 36 %, tensor pipe 29 %), and every softmax warp pays a
// st.shared + fence.proxy.async round trip per chunk (12 % of its time).  Here a warp writes the f16 pairs of its 32 keys
// back over the first 16 of the 32 score columns it has just read (tcgen05.st), and the P.V and row-sum MMAs take their A
// operand from tensor memory (tcgen05.mma [d], [a], b-desc): P never touches shared memory.
//
// Consequences for the ring: a pass-B score slot now stays occupied until the MMAs that read its P have retired, so the slots
// are dedicated instead of cycled -- two pass-B slots (chunk parity), handed back by the P.V warp's commit, and one pass-A slot,
// handed back by the 16 softmax warps as soon as it is in registers.  Job order, arithmetic, rounding points and the order of
// the 16-key MMA steps that accumulate O and the row sums are those of attention_stream_kernel: bit-identical results
// (tests/test_gpu_kernels.py).  TMEM: pass-B slots [0, 128) [128, 256), pass-A slot [256, 384), O [384, 384 + dh), row sums
// [384 + dh, +16).  A warp of key group g owns columns [32 g, 32 g + 32) of a slot on its 32 lanes: scores in, P (keys 32 g ..
// 32 g + 15 in columns 32 g .. 32 g + 7, the next 16 keys in the next 8 columns) out; nobody else touches them in between.
#pragma once

// registers -> TMEM: lane i of the warp writes TMEM lane (taddr.lane + i), 16 consecutive columns
__device__ __forceinline__ void tmem_st_32x16(uint32_t taddr, const uint32_t (&r)[16]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
        "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"
        ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]),
          "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
        : "memory");
}
// D[tmem] (+)= A[tmem] . B[smem]: A is 128 lanes x 8 columns of f16 pairs (K = 16), K-major by construction
__device__ __forceinline__ void umma_f16_ts_e(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(BK_ELECT_PROLOGUE ".reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "@e tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
                 ::"r"(tmem_d), "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}

// MN-major B operand of N = dh + 16 columns: the first atom (dh columns) is the V tile at `saddr`, the second atom lies `lbo_bytes`
// further -- an all-ones tile, so that columns [dh, dh + 16) of the product are the row sums of A (canonical layout
// ((8,n),(8,k)) : ((1,LBO),(8,SBO)), cute/atom/mma_traits_sm100.hpp; atoms along MN are LBO apart)
template <int ROWB>
__device__ __forceinline__ uint64_t umma_desc_mnmajor_ones(uint32_t saddr, uint32_t lbo_bytes) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr >> 4) & 0x3FFF);
    d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
    d |= (uint64_t)((8 * ROWB) >> 4) << 32;       // SBO: next 8-key atom
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)(ROWB == 128 ? 2 : 4) << 61;
    return d;
}

// A warp's 32 rows x 64 bytes (its 32 features of 32 query rows, one row per lane, pk = 16 f16 pairs) -> global memory through a
// 2 KB staging tile, so that one store instruction covers 8 rows x 64 contiguous bytes (full 32-byte sectors) instead of 32 rows x
// 16 bytes.  The direct form -- every lane storing its own row 16 bytes at a time -- made 1024 half-sector transactions per
// query block and backed up the SM's load/store path, which the mbarrier operations of every other warp share: 13 % of the
// kernel at L = 512 (tools/att_trace.py: all warps of the SM slow down for ~2000 cycles around each epilogue).
// Chunk c of row R sits at 16-byte position (c + R / 2) % 4 of the row: conflict-free both ways.
__device__ __forceinline__ void store_rows64(uint8_t *stage, const uint32_t (&pk)[16], __half *row0, size_t row_stride, int rows_valid, int lane) {
#pragma unroll
    for (int q = 0; q < 4; q++)
        *reinterpret_cast<uint4 *>(stage + lane * 64 + (((q + (lane >> 1)) & 3) << 4)) = make_uint4(pk[4 * q], pk[4 * q + 1], pk[4 * q + 2], pk[4 * q + 3]);
    __syncwarp();
#pragma unroll
    for (int j = 0; j < 4; j++) {
        const int R = 8 * j + (lane >> 2), c = lane & 3;
        const uint4 v = *reinterpret_cast<const uint4 *>(stage + R * 64 + (((c + (R >> 1)) & 3) << 4));
        if (R < rows_valid) *reinterpret_cast<uint4 *>(row0 + (size_t)R * row_stride + c * 8) = v;
    }
    __syncwarp();
}

#ifdef BK_EXP_TRACE
__device__ unsigned long long g_att_trace[8][1024];
#define BK_TR(w, tag) do { if (trace_on && lane == 0) { g_att_trace[w][tr_n < 1023 ? tr_n++ : 1023] = ((unsigned long long)(tag) << 48) | ((unsigned long long)clock64() & 0xFFFFFFFFFFFFull); } } while (0)
#else
#define BK_TR(w, tag)
#endif

#ifndef BK_PT_STAGGER
#define BK_PT_STAGGER 1
#endif
constexpr bool kPtStagger = BK_PT_STAGGER != 0;
constexpr int kPtThreads = kStreamThreads + 128;  // 16 softmax warps, score warp, P.V warp, 4 output warps
constexpr int kPtTail = 192 + 2 * 2048 + 4 * 2048;  // barriers + TMEM slot, two row-maximum exchange buffers, output staging

// One (sentence, head) of the launch: item i = sentence * n_head + head.  A CTA stays on its SM and takes items
// blockIdx.x, blockIdx.x + gridDim.x, ... (the batcher has sorted the sentences longest first, so this deals the work out evenly).
struct PtItem {
    int t0, L, head;   // first token row, tokens, head
    int nqb, nc;       // 128-query blocks, 128-key score chunks per pass
    int Lk16, Lk64;
};
__device__ __forceinline__ PtItem pt_item(int t0, int L, int head) {
    PtItem it;
    it.t0 = t0; it.L = L; it.head = head;
    it.nqb = (L + kAttQ - 1) / kAttQ;
    it.nc = (L + kStreamKeys - 1) / kStreamKeys;
    it.Lk16 = (L + 15) & ~15;
    it.Lk64 = (L + 63) & ~63;
    return it;
}

// Chunks of the NEXT item's first maximum pass A(0,*) that run during this item's last query block (whose A slot is idle): chunk c - 1
// behind B(last, c), once key chunk c - 1 holds the next item's keys.
__device__ __forceinline__ int pt_ahead(const PtItem &it, const PtItem &nxt, bool more) {
    if (!more || it.nc < 2) return 0;
    return it.nc - 1 < nxt.nc ? it.nc - 1 : nxt.nc;
}

template <int DH>
__global__ void __launch_bounds__(kPtThreads, 1)
attention_pt_kernel(const __grid_constant__ CUtensorMap tmQKV, const int32_t *__restrict__ cu, int n_items, int n_head, int lk_max, int d,
                    __half *__restrict__ out, int qmode, float scale) {
    constexpr int ROWB = DH * 2;
    constexpr int CW = 16;                 // score warp (TMA, S MMAs); CW + 1: P.V warp (TMEM allocation); CW + 2 .. CW + 5: output warps
    constexpr uint32_t kOCol = 3 * kStreamKeys;            // O [384, 384+dh) | row sums [384+dh, +16)
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
    uint64_t *bar_k = bar_q + 3;    // [4] keys [128 c, 128 c + 128) landed: the first score chunk does not wait for all of K
    uint64_t *sfull = bar_q + 7;    // [3] score chunk complete in the slot (0, 1: pass B by chunk parity; 2: pass A)
    uint64_t *sfree = bar_q + 10;   // [3] slot may be overwritten (0, 1: the MMAs that read its P have retired; 2: read by all 16 warps)
    uint64_t *pfull = bar_q + 13;   // [2 slots][2 halves] P of a 64-key half written by the 8 warps of its two key groups
    uint64_t *bar_o = bar_q + 17;   // O and row sums of a query block complete
    uint64_t *ofree = bar_q + 18;   // O and the row sums are in the registers of the 4 output warps
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bar_q + 19);
    // [2][4][128] partial row maxima, by block parity: a warp ahead of the others may publish the next block's maximum while a slow
    // one still reads this block's (it cannot get further: pass-B job b + 2 is issued only when the P of job b, all 16 warps', was consumed)
    float *red = reinterpret_cast<float *>(tail + 192);
    uint8_t *stage = tail + 192 + 2 * 2048;                     // [4 warps][2 KB] output staging

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#ifdef BK_EXP_TRACE
    const bool trace_on = blockIdx.x == 5;
    int tr_n = 0;
    const int tw = warp == 20 ? 6 : warp >= CW + 2 ? 7 : warp >= CW ? 4 + (warp - CW) : warp == 0 ? 0 : warp == 5 ? 1 : warp == 10 ? 2 : warp == 15 ? 3 : 7;
    BK_TR(tw, 0);
#endif
    // every warp walks the same item list; the bounds of the next item are fetched one item ahead
    const int item0 = (int)blockIdx.x, istep = (int)gridDim.x;
    auto fetch = [&](int i, int &t0, int &L) {
        if (i < n_items) { const int s = i / n_head; t0 = __ldg(cu + s); L = __ldg(cu + s + 1) - t0; } else { t0 = 0; L = 0; }
    };
    int nx_t0, nx_L;
    fetch(item0, nx_t0, nx_L);

    // Q(0), Q(1) and K of an item (score warp), V of an item (score warp for the first item, afterwards the P.V warp)
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

    if (warp == CW) {
        if (lane == 0) {
            tma_prefetch_desc(&tmQKV);
            for (int i = 0; i < 4; i++) mbar_init(&bar_k[i], 1);
            mbar_init(bar_v, 1); mbar_init(bar_o, 1); mbar_init(ofree, 4);
            mbar_init(&bar_q[0], 1); mbar_init(&bar_q[1], 1);
            for (int i = 0; i < 3; i++) { mbar_init(&sfull[i], 1); mbar_init(&sfree[i], i < 2 ? 1 : 16); }
            for (int i = 0; i < 4; i++) mbar_init(&pfull[i], 8);
            mbar_fence_init();
        }
        __syncwarp();
        // the first item's loads (they need the barriers only): their latency runs under the TMEM allocation and the CTA-wide sync
        const PtItem it = pt_item(nx_t0, nx_L, item0 % n_head);
        load_q(it, 0, 0);
        for (int c = 0; c < 4; c++) load_k(it, c);
        load_q(it, 1, 1);
        load_v(it);
    } else if (warp == CW + 1) {  // meanwhile: the all-ones tile and the TMEM allocation
        for (int i = lane; i < 16 * ROWB / 4; i += 32) reinterpret_cast<uint32_t *>(sOnes)[i] = 0x3C003C00u;
        fence_proxy_async_smem();
        __syncwarp();
        tmem_alloc<512>(tmem_slot);
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    BK_TR(tw, 13);

    if (warp == CW) {
        // ================================================================ score warp: TMA loads, S = Q . K^T jobs
        // Score jobs of an item in the order the softmax warps consume them: A(0,c) for the chunks not done ahead, then B(q,c) A(q+1,c)
        // for c = 0..nc-1, q = 0..; in the LAST query block, which has no pass A of its own, the A slot takes the next item's
        // first block: A'(0,c-1) behind B(last,c) -- key chunk c-1 has just been refilled with the next item's keys (pt_ahead).
        // Slots: pass B of chunk number b (counted over the whole CTA) -> slot b % 2, pass A -> slot 2; the prologue A(0,c), which has
        // the ring to itself, cycles through all three (c -> (c + 2) % 3) so that its MMAs do not wait for the round trip of one slot.
        uint32_t used = 0, par = 0;       // per score slot (bit): has been used / number of uses so far is odd
        uint32_t kpar = 0, qpar = 0;      // per K-chunk / Q-buffer barrier (bit): parity of the next phase to wait for
        uint32_t nbj = 0;                 // pass-B jobs issued
        int pre_in = 0;                   // chunks of this item's A(0,*) done during the previous item
        int qoff = 0;                     // block q of the item reads Q buffer (q + qoff) % 2: Q(0) goes where the previous item's last block is not
        const uint64_t dk0 = umma_desc_kmajor<ROWB>(smem_u32(sK));
        const uint64_t dq0 = umma_desc_kmajor<ROWB>(smem_u32(sQ));
        auto wait_k = [&](int c) { mbar_wait(&bar_k[c], (kpar >> c) & 1); kpar ^= 1u << c; };
        auto wait_q = [&](int b) { mbar_wait(&bar_q[b], (qpar >> b) & 1); qpar ^= 1u << b; };
        auto wait_scores = [&](uint32_t slot) { mbar_wait(&sfull[slot], ((par >> slot) & 1) ^ 1); };  // the latest job issued on `slot` has retired
        // S[slot] = Q[qbuf] . K[chunk]^T
        auto issue = [&](uint32_t slot, int qbuf, int chunk, int Lk16) {
            // the previous contents have been consumed: slot 2 and the prologue's slots -- in the registers of all 16 warps; a pass-B
            // slot -- the MMAs that read its P have retired
            if ((used >> slot) & 1) mbar_wait(&sfree[slot], ((par >> slot) & 1) ^ 1);
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
        int n2_t0, n2_L;  // this warp needs the next item's bounds during the current one: it fetches two items ahead
        fetch(item0 + istep, n2_t0, n2_L);
        for (int item = item0; item < n_items; item += istep) {
            const PtItem it = pt_item(nx_t0, nx_L, item % n_head);
            nx_t0 = n2_t0; nx_L = n2_L;
            fetch(item + 2 * istep, n2_t0, n2_L);
            BK_TR(tw, 19);
            const bool more = item + istep < n_items;
            const PtItem nxt = pt_item(nx_t0, nx_L, (item + istep) % n_head);  // (valid if `more`)
            const int pre_out = pt_ahead(it, nxt, more);
            const int nqb = it.nqb, nc = it.nc;
            const int lastbuf = (nqb - 1 + qoff) & 1;  // the buffer the last query block reads
            for (int c = pre_in; c < nc; c++) {  // prologue: pass A of block 0
                BK_TR(tw, 20);
                wait_k(c);
                if (c == 0) wait_q(qoff);
                issue((uint32_t)((c + 2) % 3), qoff, c, it.Lk16);
            }
            for (int q = 0; q < nqb; q++) {
                const int qbuf = (q + qoff) & 1;
                for (int c = 0; c < nc; c++) {
                    BK_TR(tw, 20);
                    const uint32_t sb = nbj & 1;
                    nbj++;
                    issue(sb, qbuf, c, it.Lk16);
                    if (q + 1 < nqb) {
                        if (c == 0) wait_q(qbuf ^ 1);  // first use of Q(q + 1)
                        issue(2u, qbuf ^ 1, c, it.Lk16);
                        if (c == nc - 1 && q + 2 < nqb) {
                            // B(q, nc - 1) was the last reader of Q(q): refill its buffer with Q(q + 2), a whole round of pass-B jobs
                            // before the first use
                            wait_scores(sb);
                            load_q(it, q + 2, qbuf);
                        }
                    } else if (more) {
                        // last query block: B(last, c) was the last reader of key chunk c -- refill it with the next item's keys as soon
                        // as its MMAs have retired; with the first refill also the next Q(0), into the buffer this block does not read
                        wait_scores(sb);
                        load_k(nxt, c);
                        if (c == 0) load_q(nxt, 0, lastbuf ^ 1);
                        if (c >= 1 && c - 1 < pre_out) {  // pass A of the next item's first block, one chunk behind
                            wait_k(c - 1);
                            if (c == 1) wait_q(lastbuf ^ 1);
                            issue(2u, lastbuf ^ 1, c - 1, nxt.Lk16);
                        }
                    }
                }
            }
            if (more) {  // (every score MMA of this item has retired: see the last refill)
                for (int c = nc; c < 4; c++) load_k(nxt, c);  // chunks the next item has beyond this one's
                load_q(nxt, 1, lastbuf);
            }
            pre_in = pre_out;
            qoff = lastbuf ^ 1;
        }
    } else if (warp == CW + 1) {
        // ================================================================ P.V warp: [O | row sums] += P . [V | 1], P read from the score slot
        // one MMA per 16 keys, N = dh + 16: the row sums come out of the same instruction
        const uint32_t idesc_pv = att_idesc(DH + kSumCols, 1);
        const uint32_t v0 = smem_u32(sV), o1 = smem_u32(sOnes);
        uint32_t ppar = 0;  // per pass-B slot (bit): number of hand-overs so far is odd
        uint32_t pv_b = 0;  // pass-B jobs consumed
        uint32_t gb = 0;    // query blocks done by this CTA
        uint32_t vpar = 0;
        int pre_in = 0;     // chunks of this item's A(0,*) done during the previous item (they went through slot 2)
        for (int item = item0; item < n_items; item += istep) {
            const PtItem it = pt_item(nx_t0, nx_L, item % n_head);
            fetch(item + istep, nx_t0, nx_L);
            BK_TR(tw, 19);
            // the prologue's chunks 1 and 2 went through the pass-B slots: the 16 warps report "read" on the P barriers, and the slot is
            // handed back the way it always is -- by a commit from this warp (nothing outstanding on it: it arrives at once)
            for (int c = pre_in > 1 ? pre_in : 1; c < it.nc && c < 3; c++) {
                const uint32_t slot = (uint32_t)((c + 2) % 3);
                mbar_wait(&pfull[slot * 2], (ppar >> slot) & 1);
                mbar_wait(&pfull[slot * 2 + 1], (ppar >> slot) & 1);
                umma_commit_e<1>(&sfree[slot]);
                ppar ^= 1u << slot;
            }
            mbar_wait(bar_v, vpar);
            vpar ^= 1;
            for (int qb = 0; qb < it.nqb; qb++, gb++) {
                for (int c = 0; c < it.nc; c++, pv_b++) {
                    const uint32_t slot = pv_b & 1;
                    for (int sub = 0; sub < 2; sub++) {
                        BK_TR(tw, 30 + sub);
                        mbar_wait(&pfull[slot * 2 + sub], (ppar >> slot) & 1);
                        if (gb > 0 && (c | sub) == 0) mbar_wait(ofree, (gb - 1) & 1);  // the first MMA of a block overwrites O
                        BK_TR(tw, 32 + sub);
                        tc_fence_after();
                        const int key0 = c * kStreamKeys + sub * 64;
                        const int rem = it.Lk16 - key0;
                        const int ksteps = rem >= 64 ? 4 : rem > 0 ? rem >> 4 : 0;
                        const uint32_t pa = tmem_base + slot * kStreamKeys + (uint32_t)(sub * 64);  // keys 16 ks ..: group 2 sub + ks / 2, its columns 8 (ks % 2) ..
                        for (int ks = 0; ks < ksteps; ks++) {
                            const uint32_t va = v0 + (uint32_t)((key0 + 16 * ks) * ROWB);
                            umma_f16_ts_e(tmem_base + kOCol, pa + (uint32_t)((ks >> 1) * 32 + (ks & 1) * 8), umma_desc_mnmajor_ones<ROWB>(va, o1 - va), idesc_pv,
                                          (c | sub | ks) != 0);
                        }
                    }
                    umma_commit_e<1>(&sfree[slot]);
                    ppar ^= 1u << slot;
                    BK_TR(tw, 34);
                }
                umma_commit_e<1>(bar_o);
            }
            // the next item's V: when this item's last P.V MMAs have retired (the first P of the next item is a prologue and a chunk
            // of exponentials away)
            const bool more = item + istep < n_items;
            const PtItem nxt = pt_item(nx_t0, nx_L, (item + istep) % n_head);
            if (more) {
                mbar_wait(bar_o, (gb - 1) & 1);
                load_v(nxt);
            }
            pre_in = pt_ahead(it, nxt, more);
        }
    } else if (warp >= CW + 2) {
        // ======================================================================= output warps: O * (1 / sum p) -> out rows
        // One warp per TMEM lane quarter.  The softmax warps never touch O: while they are in the next block's first chunk (or the
        // next item's prologue), these warps take the finished block out of tensor memory (the P.V warp's first MMA of the next
        // block waits for `ofree`), scale, quantise for Q4 models and store.
        const int quad = warp & 3;
        const uint32_t tlane = tmem_base + ((uint32_t)(quad * 32) << 16);
        uint8_t *st = stage + (warp - (CW + 2)) * 2048;
        uint32_t gb = 0;  // query blocks done by this CTA
        for (int item = item0; item < n_items; item += istep) {
            const PtItem it = pt_item(nx_t0, nx_L, item % n_head);
            fetch(item + istep, nx_t0, nx_L);
            for (int qb = 0; qb < it.nqb; qb++, gb++) {
                mbar_wait(bar_o, gb & 1);
                BK_TR(tw, 16);
                tc_fence_after();
                const float sum = __uint_as_float(tmem_ld_32x1(tlane + kOCol + (uint32_t)DH));
                const float inv = 1.0f / sum;
                const int row0 = qb * kAttQ + quad * 32;
#pragma unroll
                for (int hf = 0; hf < DH / 32; hf++) {
                    uint32_t a[32];
                    tmem_ld_32x32(tlane + kOCol + (uint32_t)(hf * 32), a);
                    tmem_ld_wait();
                    if (hf == DH / 32 - 1) {  // all of O is in registers (the P.V warp is at least a chunk of exponentials away from needing it)
                        tc_fence_before();
                        __syncwarp();
                        if (lane == 0) mbar_arrive(ofree);
                        BK_TR(tw, 17);
                    }
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
                    store_rows64(st, pk, out + (size_t)(it.t0 + row0) * d + (size_t)it.head * DH + hf * 32, (size_t)d, it.L - row0, lane);
                }
                BK_TR(tw, 18);
            }
        }
    } else {
        // ======================================================================= softmax warps
        const int quad = warp & 3, g = warp >> 2;
        const int r = quad * 32 + lane;                       // query row in the block == TMEM lane
        const uint32_t tlane = tmem_base + ((uint32_t)(quad * 32) << 16);
        const uint32_t tcol = tlane + (uint32_t)(g * 32);     // my 32 columns of a slot
        const int psub = g >> 1;                              // my 64-key half of a chunk
        uint32_t spar = 0;                                    // per score slot (bit): number of chunks consumed from it so far is odd
        uint32_t nb = 0;                                      // pass-B jobs done
        uint32_t gb = 0;                                      // query blocks done by this CTA
        int pre_in = 0;                                       // chunks of this item's A(0,*) done during the previous item
        float mx_next = -INFINITY;                            // running maximum of the block whose pass A is under way
        int n2_t0, n2_L;  // the next item's length is needed during the current one: fetched two items ahead
        fetch(item0 + istep, n2_t0, n2_L);
        for (int item = item0; item < n_items; item += istep) {
            const PtItem it = pt_item(nx_t0, nx_L, item % n_head);
            nx_t0 = n2_t0; nx_L = n2_L;
            fetch(item + 2 * istep, n2_t0, n2_L);
            BK_TR(tw, 19);
            const int nc = it.nc, nqb = it.nqb;
            const bool more = item + istep < n_items;
            const PtItem nxt = pt_item(nx_t0, nx_L, (item + istep) % n_head);  // (valid if `more`)
            const int pre_out = pt_ahead(it, nxt, more);

            // one pass-A job: fold this group's 32 keys of chunk c (of a sentence of L tokens) into the running row maximum
            auto step_a = [&](int c, float mx, uint32_t slot, int L) {
                BK_TR(tw, 7);
                mbar_wait(&sfull[slot], (spar >> slot) & 1);
                spar ^= 1u << slot;
                BK_TR(tw, 8);
                tc_fence_after();
                const int key0 = c * kStreamKeys + g * 32;
                if (key0 < L) {
                    uint32_t a[32];
                    tmem_ld_32x32(tcol + slot * kStreamKeys, a);
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
                if (lane == 0) mbar_arrive(slot == 2 ? &sfree[2] : &pfull[slot * 2 + psub]);  // (a prologue chunk in a pass-B slot: see the P.V warp)
                BK_TR(tw, 9);
                return mx;
            };

            const int L = it.L;
            for (int c = pre_in; c < nc; c++) mx_next = step_a(c, mx_next, (uint32_t)((c + 2) % 3), L);  // rest of pass A of block 0
            for (int qb = 0; qb < nqb; qb++, gb++) {
                // ---- row maximum of this block over the four groups
                float mx = mx_next;
                {
                    float *rb = red + (gb & 1) * 512;
                    rb[g * 128 + r] = mx;
                    BK_TR(tw, 10);
                    named_bar_sync(1, 512);
                    BK_TR(tw, 11);
#pragma unroll
                    for (int gg = 0; gg < 4; gg++) mx = fmaxf(mx, rb[gg * 128 + r]);
                }
                mx *= scale;  // max(s*scale) == max(s)*scale for scale > 0 (rounding is monotonic)
                mx_next = -INFINITY;
                // ---- the A slot's job of this step: pass A of the next block, same chunk; in the last block: of the next item's first
                // block, one chunk behind.  The warps of the first 64-key half take it after their pass-B job, those of the second half
                // BEFORE it: the two halves -- two warps each on every scheduler -- then run about half a step apart, one half's
                // exponentials (the MUFU is what bounds this kernel) under the other half's loads, stores, barriers and maximum pass,
                // instead of four warps queueing for the MUFU and then leaving it idle together.
                auto pass_a = [&](int c) {
                    if (qb + 1 < nqb) mx_next = step_a(c, mx_next, 2u, L);
                    else if (c >= 1 && c - 1 < pre_out) mx_next = step_a(c - 1, mx_next, 2u, nxt.L);
                };
                for (int c = 0; c < nc; c++) {
                    if (kPtStagger && psub == 1) pass_a(c);
                    // ---- pass B job: p = f16(exp(f16(s*scale - max))) for my 32 keys of chunk c, written back over the scores
                    const uint32_t slot = nb & 1;
                    BK_TR(tw, 1);
                    mbar_wait(&sfull[slot], (spar >> slot) & 1);
                    spar ^= 1u << slot;
                    BK_TR(tw, 2);
                    tc_fence_after();
                    const int key0 = c * kStreamKeys + g * 32;
                    if (key0 < L) {  // otherwise no real key: the MMA does not read these P columns
                        uint32_t a[32];
                        tmem_ld_32x32(tcol + slot * kStreamKeys, a);
                        tmem_ld_wait();
                        BK_TR(tw, 3);
                        const int nv = L - key0;
                        if (nv < 32) {  // s = -inf -> p = 0 exactly
#pragma unroll
                            for (int i = 0; i < 32; i++) if (i >= nv) a[i] = 0xff800000u;
                        }
                        uint32_t pk[16];
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
                        BK_TR(tw, 4);
                        tmem_st_32x16(tcol + slot * kStreamKeys, pk);
                        tmem_st_wait();
                        BK_TR(tw, 5);
                    }
                    tc_fence_before();  // P in tensor memory -> visible to the MMAs issued behind the barrier
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&pfull[slot * 2 + psub]);
                    BK_TR(tw, 6);
                    nb++;
                    if (!kPtStagger || psub == 0) pass_a(c);
                }
            }
            pre_in = pre_out;
        }
    }
    BK_TR(tw, 14);

    tc_fence_before();
    __syncthreads();
    BK_TR(tw, 15);
    if (warp == CW + 1) {
        tc_fence_after();
        tmem_dealloc<512>(tmem_base);
    }
}

template <int DH>
static cudaError_t launch_att_pt(const CUtensorMap &tm, const int32_t *cu, int n_seq, int max_len, int d, int n_head, __half *out,
                                 int qmode, cudaStream_t st) {
    const int Lk64 = (max_len + 63) & ~63;
    const size_t smem = (size_t)kStreamQ * kAttQ * DH * 2 + (size_t)2 * Lk64 * DH * 2 + 16 * DH * 2 + kPtTail;
    static std::atomic<bool> attr_set[64];
    static std::atomic<int> n_sm[64];
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
    if (!attr_set[dev]) {
        cudaError_t err = cudaFuncSetAttribute(attention_pt_kernel<DH>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
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
    LaunchConfig lc(grid, dim3(kPtThreads), smem, st);
    return cudaLaunchKernelEx(&lc.cfg, attention_pt_kernel<DH>, tm, cu, (int)n_items, n_head, Lk64, d, out, qmode, scale);
}
