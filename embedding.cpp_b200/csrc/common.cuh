// common.cuh -- device-side helpers shared by the sm_100a kernels: PTX wrappers for mbarrier, TMA,
// tcgen05 (MMA / TMEM), and the f16 / Q8 rounding primitives that mirror the reference numerics.
#pragma once
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace bk {

// ------------------------------------------------------------------------------------------ misc
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint32_t lane_id() { return threadIdx.x & 31; }
__device__ __forceinline__ bool elect_one() {
    uint32_t pred = 0;
    asm volatile(
        "{\n\t.reg .pred P;\n\t"
        "elect.sync _|P, 0xffffffff;\n\t"
        "selp.u32 %0, 1, 0, P;\n\t}\n"
        : "=r"(pred));
    return pred != 0;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// ---------------------------------------------------------------- programmatic dependent launch
// With cudaLaunchAttributeProgrammaticStreamSerialization (the engine sets it for the small-batch launch plan, see
// kernels.h) the next kernel of the stream may become resident once every CTA of this one has executed
// launch_dependents, and runs its prologue (barrier init, TMEM allocation, descriptor prefetch) under this kernel's
// tail; griddepcontrol.wait then blocks until this grid has completed and its memory is visible.  Rule in this code
// base: launch_dependents first thing, wait before the first access to global memory other than kernel parameters --
// reads (a chain of prologues can be in flight: data of ANY earlier kernel may still be on its way) and writes (an earlier
// kernel may still be reading the buffer).  Both are no-ops in a launch without the attribute.
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

// ------------------------------------------------------------------------------------- mbarrier
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// BK_WAIT_HINT_NS > 0: try_wait carries a suspend-time hint, so a waiting thread sleeps up to that long (it still wakes when the phase
// completes) instead of returning after the short default and polling again.  Measured with 1 us and 20 us hints on the 20-step
// power-capped bench (make BUILD=... DEFS=-DBK_WAIT_HINT_NS=..., BERT_B200_LIB): 42.8-43.2 k emb/s for all three builds -- the
// polling is not what the power budget is spent on; left off
#ifndef BK_WAIT_HINT_NS
#define BK_WAIT_HINT_NS 0
#endif
__device__ __forceinline__ bool mbar_try_wait(uint64_t *bar, uint32_t parity) {
    uint32_t ok;
#if BK_WAIT_HINT_NS > 0
    asm volatile(
        "{\n\t.reg .pred P;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 P, [%1], %2, %3;\n\t"
        "selp.u32 %0, 1, 0, P;\n\t}\n"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity), "n"(BK_WAIT_HINT_NS)
        : "memory");
#else
    asm volatile(
        "{\n\t.reg .pred P;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 P, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, P;\n\t}\n"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
#endif
    return ok != 0;
}
// non-blocking probe (try_wait may suspend the thread for a system-dependent time before it returns false)
__device__ __forceinline__ bool mbar_test_wait(uint64_t *bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred P;\n\t"
        "mbarrier.test_wait.parity.shared::cta.b64 P, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, P;\n\t}\n"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    while (!mbar_try_wait(bar, parity)) {}
}

// ---------------------------------------------------------------- CTA pair (cluster of 2) helpers
// In a 2-CTA cluster the shared::cluster address of a CTA's own shared memory differs from its peer's only in
// bit 24 (CUTLASS: Sm100MmaPeerBitMask); clearing it addresses the same offset in the even ("leader") CTA.
constexpr uint32_t kPeerBitMask = 0xFEFFFFFFu;
__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// arrive (+ expect_tx) on a barrier given by its shared::cluster address (own or peer CTA).  Default .release.cta
// semantics as in CUTLASS' ClusterBarrier: an explicit .release.cluster compiles to MEMBAR.ALL + ERRBAR, which
// drains the issuing thread's outstanding TMA loads and serialises the operand ring (measured: 2x slower).
__device__ __forceinline__ void mbar_arrive_expect_tx_cluster(uint32_t bar_addr, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cluster.b64 _, [%0], %1;" ::"r"(bar_addr), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t bar_addr) {
    asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(bar_addr) : "memory");
}

// shared::cluster address of `local_addr` (a shared::cta address of this CTA) in the CTA of rank `rank`
__device__ __forceinline__ uint32_t mapa_cluster(uint32_t local_addr, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local_addr), "r"(rank));
    return r;
}
// 8-byte store into another CTA's shared memory that completes 8 transaction bytes on an mbarrier of the SAME
// remote CTA: the waiter of that barrier phase sees the data, no fence on either side
__device__ __forceinline__ void st_async_f32x2(uint32_t remote_addr, float a, float b, uint32_t remote_bar) {
    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v2.f32 [%0], {%1, %2}, [%3];"
                 ::"r"(remote_addr), "f"(a), "f"(b), "r"(remote_bar) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx_noelect(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive_n(uint64_t *bar, uint32_t n) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(n) : "memory");
}

// ------------------------------------------------------------------------------------------ TMA
__device__ __forceinline__ void tma_prefetch_desc(const void *tmap) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(tmap) : "memory");
}
// 2-D tiled load: coordinates {c0 = inner (contiguous) index, c1 = row index}
__device__ __forceinline__ void tma_load_2d(void *smem_dst, const void *tmap, uint64_t *bar, int32_t c0, int32_t c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(smem_u32(smem_dst)), "l"(tmap), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
        : "memory");
}
// the same load issued by either CTA of a pair: data lands in the issuing CTA's smem, complete_tx goes to the
// barrier at shared::cluster address bar_addr (normally the leader CTA's)
__device__ __forceinline__ void tma_load_2d_pair(void *smem_dst, const void *tmap, uint32_t bar_addr, int32_t c0, int32_t c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(smem_u32(smem_dst)), "l"(tmap), "r"(bar_addr), "r"(c0), "r"(c1)
        : "memory");
}
// 1-D bulk copy global -> shared (bytes multiple of 16, both 16 B aligned)
__device__ __forceinline__ void bulk_load_1d(void *smem_dst, const void *gsrc, uint32_t bytes, uint64_t *bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
        ::"r"(smem_u32(smem_dst)), "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
        : "memory");
}
// 2-D tiled store shared -> global (bulk async-group completion); rows/columns outside the tensor are clipped
__device__ __forceinline__ void tma_store_2d(const void *tmap, const void *smem_src, int32_t c0, int32_t c1) {
    asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
                 ::"l"(tmap), "r"(smem_u32(smem_src)), "r"(c0), "r"(c1)
                 : "memory");
}
__device__ __forceinline__ void bulk_commit_group() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
// all but the newest kPending bulk groups of this thread have finished READING their shared-memory source
template <int kPending>
__device__ __forceinline__ void bulk_wait_group_read() { asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(kPending) : "memory"); }
template <int kPending>
__device__ __forceinline__ void bulk_wait_group() { asm volatile("cp.async.bulk.wait_group %0;" ::"n"(kPending) : "memory"); }
// named barrier among `count` threads (count % 32 == 0); id 0 is __syncthreads
__device__ __forceinline__ void named_bar_sync(uint32_t id, uint32_t count) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(count) : "memory"); }
// generic-proxy writes to smem -> visible to the async proxy (UMMA / TMA reads)
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// -------------------------------------------------------------------------------------- tcgen05
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// whole warp; writes the TMEM base address into *smem_dst.  kCG = 2: the same warp of BOTH CTAs of the pair calls it.
template <uint32_t kCols, int kCG = 1>
__device__ __forceinline__ void tmem_alloc(uint32_t *smem_dst) {
    if (kCG == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_dst)), "n"(kCols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    } else {
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_dst)), "n"(kCols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
}
template <uint32_t kCols, int kCG = 1>
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr) {
    if (kCG == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "n"(kCols) : "memory");
    else asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "n"(kCols) : "memory");
}

// Shared-memory matrix descriptor for a K-major operand tile stored as rows of 128 B (64 halves)
// with the 128-byte swizzle (what TMA SWIZZLE_128B writes): 8-row groups are 1024 B apart (SBO),
// LBO is unused for swizzled K-major layouts.  Field layout: cute/arch/mma_sm100_desc.hpp
// (SmemDescriptor): [0,14) addr>>4, [16,30) LBO>>4, [32,46) SBO>>4, [46,48) version=1, [61,64) layout=2.
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t saddr) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr >> 4) & 0x3FFF);
    d |= (uint64_t)1 << 16;
    d |= (uint64_t)(1024 >> 4) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)2 << 61;
    return d;
}
// Instruction descriptor, kind::f16: A,B = F16 (format 0), D = F32 (c_format 1), both K-major.
// [4,6) c_format, [7,10) a_format, [10,13) b_format, 15 a_major, 16 b_major, [17,23) N>>3, [24,29) M>>4
__host__ __device__ constexpr uint32_t umma_idesc_f16(uint32_t M, uint32_t N) {
    return (1u << 4) | ((N >> 3) << 17) | ((M >> 4) << 24);
}
// D[tmem] (+)= A[smem] . B[smem]^T ; one thread issues for the CTA
__device__ __forceinline__ void umma_f16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}\n"
        ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// CTA-pair MMA (issued by the leader CTA only): D is 256 x N, rows 0-127 in the leader's TMEM and 128-255 in the
// peer's; each CTA's smem holds its own 128 rows of A and its own N/2 rows of B at the descriptor addresses
__device__ __forceinline__ void umma_f16_pair(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}\n"
        ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// pair commit: arrives on the barrier at this smem offset in BOTH CTAs when the MMAs issued so far complete
__device__ __forceinline__ void umma_commit_pair(uint64_t *bar) {
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(smem_u32(bar)), "h"((uint16_t)3) : "memory");
}
// all previously issued MMAs of this thread -> arrive on an mbarrier when they complete
__device__ __forceinline__ void umma_commit(uint64_t *bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// TMEM -> registers: lane i of the warp reads TMEM lane (taddr.lane + i), 32 consecutive columns
__device__ __forceinline__ void tmem_ld_32x32(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
          "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
          "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
}
// registers -> TMEM: lane i of the warp writes TMEM lane (taddr.lane + i), 32 consecutive columns
__device__ __forceinline__ void tmem_st_32x32(uint32_t taddr, const uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
        "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
        "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};"
        ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]),
          "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]),
          "r"(r[16]), "r"(r[17]), "r"(r[18]), "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]),
          "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]), "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31])
        : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// ------------------------------------------------------- warp-collective ("elected") issue variants
// Called by ALL lanes of a warp in uniform control flow; the instruction is issued once, by the lane
// elect.sync picks.  Unlike `if (lane == 0) op();` this keeps the enclosing loop warp-uniform, so ptxas keeps
// descriptors / addresses / loop counters in uniform registers and emits back-to-back UTCHMMA / UTMALDG
// instead of a per-instruction ELECT + R2UR "waterfall" loop (measured: ~700 -> ~250 cycles per K block).
#define BK_ELECT_PROLOGUE "{\n\t.reg .pred e;\n\telect.sync _|e, 0xffffffff;\n\t"
__device__ __forceinline__ void mbar_expect_tx_e(uint64_t *bar, uint32_t bytes) {
    asm volatile(BK_ELECT_PROLOGUE "@e mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n\t}" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx_cluster_e(uint32_t bar_addr, uint32_t bytes) {
    asm volatile(BK_ELECT_PROLOGUE "@e mbarrier.arrive.expect_tx.shared::cluster.b64 _, [%0], %1;\n\t}" ::"r"(bar_addr), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tma_load_2d_e(void *smem_dst, const void *tmap, uint64_t *bar, int32_t c0, int32_t c1) {
    asm volatile(BK_ELECT_PROLOGUE
                 "@e cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];\n\t}"
                 ::"r"(smem_u32(smem_dst)), "l"(tmap), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
                 : "memory");
}
__device__ __forceinline__ void tma_load_2d_pair_e(void *smem_dst, const void *tmap, uint32_t bar_addr, int32_t c0, int32_t c1) {
    asm volatile(BK_ELECT_PROLOGUE
                 "@e cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];\n\t}"
                 ::"r"(smem_u32(smem_dst)), "l"(tmap), "r"(bar_addr), "r"(c0), "r"(c1)
                 : "memory");
}
__device__ __forceinline__ void bulk_load_1d_e(void *smem_dst, const void *gsrc, uint32_t bytes, uint64_t *bar) {
    asm volatile(BK_ELECT_PROLOGUE "@e cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n\t}"
                 ::"r"(smem_u32(smem_dst)), "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
template <int kCG>
__device__ __forceinline__ void umma_f16_e(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    if (kCG == 1)
        asm volatile(BK_ELECT_PROLOGUE ".reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                     "@e tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                     ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
    else
        asm volatile(BK_ELECT_PROLOGUE ".reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                     "@e tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                     ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
// pair commit inside a larger cluster: mask = the two CTAs of this pair
__device__ __forceinline__ void umma_commit_pair_mask_e(uint64_t *bar, uint16_t mask) {
    asm volatile(BK_ELECT_PROLOGUE "@e tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;\n\t}"
                 ::"r"(smem_u32(bar)), "h"(mask) : "memory");
}
template <int kCG>
__device__ __forceinline__ void umma_commit_e(uint64_t *bar) {
    if (kCG == 1)
        asm volatile(BK_ELECT_PROLOGUE "@e tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n\t}" ::"r"(smem_u32(bar)) : "memory");
    else
        asm volatile(BK_ELECT_PROLOGUE "@e tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;\n\t}"
                     ::"r"(smem_u32(bar)), "h"((uint16_t)3) : "memory");
}

// ------------------------------------------------------------------------------ numerics helpers
// round to IEEE half (RNE) and widen back: what F16C _cvtss_sh/_cvtsh_ss do (ggml-impl.h:85-101)
__device__ __forceinline__ float round_f16(float x) { return __half2float(__float2half_rn(x)); }

// tanh-approximation GELU evaluated the way the reference's f16 table does (ggml.c:1418-1441,
// 2193-2201): input rounded to f16, fp32 evaluation, output rounded to f16.
__device__ __forceinline__ float gelu_ref(float x) {
    const float u = round_f16(x);
    // every product/sum rounded separately (no FMA contraction), left-to-right as the C expression
    const float inner = __fmul_rn(__fmul_rn(0.79788456080286535587989211986876f, u),
                                  __fadd_rn(1.0f, __fmul_rn(__fmul_rn(0.044715f, u), u)));
    const float g = __fmul_rn(__fmul_rn(0.5f, u), __fadd_rn(1.0f, tanhf(inner)));
    return round_f16(g);
}

// The same function for the GEMM epilogue without a tanhf call: tanh(u) = 2*sigmoid(2u) - 1 with
// sigmoid(2u) = 1 / (1 + 2^(-2u*log2(e))), u = sqrt(2/pi) x (1 + 0.044715 x^2), from ex2.approx / rcp.approx
// (accurate to ~2^-22).  The reference forms fl(1 + fl(tanhf(u))) in fp32, which for x < -3 cancels down to
// a multiple of 2^-24; rounding t = fl(2r - 1) to fp32 first and then adding 1 reproduces that cancellation,
// so the f16 result equals the table's except on a handful of near-ties (tests/test_gpu_kernels.py counts
// them over all 63488 finite f16 inputs).
__device__ __forceinline__ float gelu_fast(float x) {
    const float u = round_f16(x);
    const float z = u * fmaf(-0.1029432395800235f, u * u, -2.302208198144325f);  // -2u*log2(e)
    float e, r;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(z));
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(1.0f + e));
    const float t = fmaf(2.0f, r, -1.0f);  // tanh(u) rounded to fp32
    return round_f16(__fmul_rn(__fmul_rn(0.5f, u), __fadd_rn(1.0f, t)));
}

// Two elements at a time with Blackwell's packed fp32 instructions (fma/add/mul.f32x2, sm_100+): each lane of
// a packed op is the same IEEE operation as the scalar form above, so the results are bit-identical to
// gelu_fast; the epilogue of the FFN-up GEMM is issue-slot bound and this halves its FP32 instruction count.
// Returns the fp32 value BEFORE the final f16 rounding (the caller's f16 pack performs it once).
__device__ __forceinline__ float2 gelu_fast2_unrounded(float2 x) {
    const float2 u = __half22float2(__floats2half2_rn(x.x, x.y));
    const float2 w = __ffma2_rn(make_float2(-0.1029432395800235f, -0.1029432395800235f), __fmul2_rn(u, u),
                                make_float2(-2.302208198144325f, -2.302208198144325f));
    const float2 z = __fmul2_rn(u, w);  // -2u*log2(e)
    float2 e, r;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e.x) : "f"(z.x));
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e.y) : "f"(z.y));
    const float2 d = __fadd2_rn(make_float2(1.0f, 1.0f), e);
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r.x) : "f"(d.x));
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r.y) : "f"(d.y));
    const float2 t = __ffma2_rn(make_float2(2.0f, 2.0f), r, make_float2(-1.0f, -1.0f));  // tanh(u) rounded to fp32
    return __fmul2_rn(__fmul2_rn(make_float2(0.5f, 0.5f), u), __fadd2_rn(make_float2(1.0f, 1.0f), t));
}

// Q8_0 / Q8_1 activation rounding of one 32-element block given its amax (quantize_row_q8_0 /
// quantize_row_q8_1, AVX2 branches, ggml-quants.c:702-747, 944-985): d = amax/127 (Q8_0 stores it
// as f16, Q8_1 keeps f32), q = rne(x * (127/amax)); the value the Q4 dot product then sees is d*q.
// mode 1 = Q8_0, 2 = Q8_1.
__device__ __forceinline__ void q8_scales(float amax, int mode, float &id, float &dq) {
    const float d = amax / 127.f;
    id = (amax != 0.0f) ? 127.f / amax : 0.0f;
    dq = (mode == 1) ? round_f16(d) : d;
}
// round-to-nearest-even of x*id (|x*id| <= 127) by adding and subtracting 1.5 * 2^23: two full-rate FADDs instead of the quarter-rate
// FRND; the product is rounded first (__fmul_rn: no FMA contraction), as _mm256_mul_ps before _mm256_round_ps in the reference
__device__ __forceinline__ float q8_round(float x, float id, float dq) {
    return __fmul_rn(dq, __fadd_rn(__fadd_rn(__fmul_rn(x, id), 12582912.f), -12582912.f));
}
// exact (float)q for 0 <= q < 2^23 without the quarter-rate I2F: q ORed into the mantissa of 2^23, minus 2^23 (+ bias)
__device__ __forceinline__ float small_int_to_float(int q, float minus) { return __fadd_rn(__int_as_float(0x4B000000 | q), -8388608.f - minus); }
// packed form: rne(x*id) via the 1.5*2^23 trick (exact for |x*id| <= 2^22; here <= 127), all on the FMA pipe
__device__ __forceinline__ float2 q8_round2(float2 x, float id, float dq) {
    const float2 magic = make_float2(12582912.f, 12582912.f);
    // scalar products: ptxas contracts mul.rn.f32x2 + add.rn.f32x2 into FFMA2 (seen in SASS), which would skip the
    // fp32 rounding of x*id that _mm256_mul_ps performs before _mm256_round_ps in the reference
    const float2 m = make_float2(__fmul_rn(x.x, id), __fmul_rn(x.y, id));
    const float2 q = __fadd2_rn(__fadd2_rn(m, magic), make_float2(-12582912.f, -12582912.f));
    return __fmul2_rn(make_float2(dq, dq), q);
}

__device__ __forceinline__ uint32_t pack_h2(float a, float b) {
    __half2 h = __floats2half2_rn(a, b);
    return *reinterpret_cast<uint32_t *>(&h);
}
__device__ __forceinline__ float2 unpack_h2(uint32_t u) {
    __half2 h = *reinterpret_cast<__half2 *>(&u);
    return __half22float2(h);
}

}  // namespace bk
