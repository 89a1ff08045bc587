// hm_ptx.cuh — inline-PTX wrappers for sm_100a: mbarrier, TMA (cp.async.bulk.tensor),
// tcgen05 (alloc / mma / commit / ld / fences) and the UMMA descriptors.
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include "hm_act.h"

namespace hm {

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "elect.sync _|p, 0xffffffff;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(pred));
    return pred != 0;
}

// ---- mbarrier -------------------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void fence_barrier_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async_smem() {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
                 "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
// non-blocking probe of a phase (try_wait may suspend the thread for a while; this never does): issue it early,
// consume the result late, fall back to mbar_wait when it says "not yet"
__device__ __forceinline__ uint32_t mbar_test_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    while (!mbar_try_wait(bar, parity)) {
    }
}
// for long waits of many warps (an epilogue waiting for a whole mainloop): the pollers stay off the issue slots and the
// shared-memory port the MMA issuer and the tensor pipe are using
__device__ __forceinline__ void mbar_wait_backoff(uint64_t* bar, uint32_t parity, unsigned ns) {
    while (!mbar_try_wait(bar, parity)) __nanosleep(ns);
}

// ---- TMA ------------------------------------------------------------------------------
__device__ __forceinline__ void tma_prefetch_desc(const void* tmap) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(tmap) : "memory");
}
__device__ __forceinline__ void tma_load_2d(void* smem_dst, const void* tmap, uint64_t* bar,
                                            int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%3, %4}], [%2];" ::"r"(smem_u32(smem_dst)),
        "l"(tmap), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
        : "memory");
}
__device__ __forceinline__ void tma_load_4d(void* smem_dst, const void* tmap, uint64_t* bar,
                                            int c0, int c1, int c2, int c3) {
    asm volatile(
        "cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(smem_u32(smem_dst)),
        "l"(tmap), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
        : "memory");
}

__device__ __forceinline__ void tma_load_3d(void* smem_dst, const void* tmap, uint64_t* bar,
                                            int c0, int c1, int c2) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(smem_u32(smem_dst)),
        "l"(tmap), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2)
        : "memory");
}
// shared -> global tile store (bulk async group)
__device__ __forceinline__ void tma_store_3d(const void* tmap, const void* smem_src, int c0, int c1,
                                             int c2) {
    asm volatile(
        "cp.async.bulk.tensor.3d.global.shared::cta.tile.bulk_group [%0, {%2, %3, %4}], [%1];" ::"l"(tmap),
        "r"(smem_u32(smem_src)), "r"(c0), "r"(c1), "r"(c2)
        : "memory");
}
__device__ __forceinline__ void bulk_commit_group() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_group_read0() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_group0() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

// 1-D bulk copy global -> shared (size multiple of 16, 16-byte aligned), completes on mbarrier
__device__ __forceinline__ void bulk_load_1d(void* smem_dst, const void* gsrc, uint32_t bytes,
                                             uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
            "r"(smem_u32(smem_dst)), "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
        : "memory");
}
__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// ---- tcgen05 --------------------------------------------------------------------------
__device__ __forceinline__ void tmem_alloc(uint32_t* smem_dst, uint32_t ncols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                     smem_u32(smem_dst)),
                 "r"(ncols)
                 : "memory");
}
__device__ __forceinline__ void tmem_relinquish() {
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols)
                 : "memory");
}
__device__ __forceinline__ void tc_fence_before() {
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_after() {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}
// D[tmem] (+)= A[smem desc] * B[smem desc]; kind::f16 covers bf16/fp16 inputs, fp32 accumulate.
__device__ __forceinline__ void umma_f16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc,
                                         uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
        "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// Same, descriptors given as (low word, high word): the high words (LBO / SBO / layout) are
// loop constants and the low word is base + compile-time offset, which keeps the whole operand
// chain in uniform registers (one UIADD3 per descriptor instead of 64-bit vector adds + R2UR).
__device__ __forceinline__ void umma_f16_lh(uint32_t d_tmem, uint32_t a_lo, uint32_t a_hi,
                                            uint32_t b_lo, uint32_t b_hi, uint32_t idesc,
                                            uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\t"
        "setp.ne.b32 p, %6, 0;\n\t"
        "mov.b64 da, {%1, %2};\n\t"
        "mov.b64 db, {%3, %4};\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %5, p;\n\t}" ::"r"(d_tmem),
        "r"(a_lo), "r"(a_hi), "r"(b_lo), "r"(b_hi), "r"(idesc), "r"(accumulate)
        : "memory");
}
// EIGHT MMAs (two 16 KB A tiles x four K16 steps) and the two per-tile commits as ONE asm block: the descriptor
// arithmetic sits between the MMAs instead of in front of the burst.  The tensor pipe queues only two or three MMAs
// behind the running one (tools/probes/probe_issue.cu: a few extra instructions per burst already cost 8-10 % once
// they are hoisted in front of it), so everything the issuing thread does between two bursts — barrier wait,
// ~40 R2UR/UMOV of descriptor set-up — has to fit into ~150 cycles or the pipe starves; inside the block each MMA's
// set-up runs while the queue is still full.  Descriptor low words (16-byte units): first operand of tile j, step k =
// a_lo + j*AJ + k*AK; second = b_lo + j*BJ + k*BK.  acc_first: accumulate flag of the very first MMA (the others accumulate).
// MID_COMMIT = false: no commit after the first tile (callers with one barrier per 32 KB stage).  The descriptor HIGH
// words and the instruction descriptor are compile-time immediates (HI_A / HI_B / IDESC): they cost a UMOV, not a
// vector register + R2UR each.
constexpr uint32_t UMMA_DESC_HI_SW128 = (1024u >> 4) | (1u << 14) | (2u << 29);   // SBO 1024 B, version 1, SWIZZLE_128B
template <int AJ, int AK, int BJ, int BK, uint32_t IDESC, bool MID_COMMIT = true>
__device__ __forceinline__ void umma_f16_burst8(uint32_t d_tmem, uint32_t a_lo, uint32_t b_lo, uint32_t acc_first,
                                                uint64_t* bar_tile0, uint64_t* bar_tile1, uint32_t leader) {
#define HM_MMA_STEP(AOFF, BOFF)                                                   \
    "add.u32 ta, %1, " #AOFF ";\n\t add.u32 tb, %2, " #BOFF ";\n\t"              \
    "mov.b64 da, {ta, %22};\n\t mov.b64 db, {tb, %22};\n\t"                       \
    "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %23, q;\n\t"
    // called by the elected lane only, inside a warp-uniform `if (leader)`: predicating every instruction on a per-lane
    // predicate instead makes ptxas carry the descriptors in vector registers (two R2UR per descriptor)
    (void)leader;
    asm volatile(
        "{\n\t.reg .pred p, q;\n\t.reg .b64 da, db;\n\t.reg .b32 ta, tb;\n\t"
        "setp.ne.b32 p, %3, 0;\n\t setp.eq.b32 q, %0, %0;\n\t"
        "mov.b64 da, {%1, %22};\n\t mov.b64 db, {%2, %22};\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %23, p;\n\t"
        HM_MMA_STEP(%8, %9) HM_MMA_STEP(%10, %11) HM_MMA_STEP(%12, %13)
        "setp.ne.b32 p, %7, 0;\n\t"
        "@p tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%4];\n\t"
        HM_MMA_STEP(%14, %15) HM_MMA_STEP(%16, %17) HM_MMA_STEP(%18, %19) HM_MMA_STEP(%20, %21)
        "tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%5];\n\t}"
        ::"r"(d_tmem), "r"(a_lo), "r"(b_lo), "r"(acc_first),
          "r"(smem_u32(bar_tile0)), "r"(smem_u32(bar_tile1)), "r"(leader), "r"(MID_COMMIT ? 1u : 0u),
          "n"(AK), "n"(BK), "n"(2 * AK), "n"(2 * BK), "n"(3 * AK), "n"(3 * BK),
          "n"(AJ), "n"(BJ), "n"(AJ + AK), "n"(BJ + BK), "n"(AJ + 2 * AK), "n"(BJ + 2 * BK), "n"(AJ + 3 * AK), "n"(BJ + 3 * BK),
          "n"(UMMA_DESC_HI_SW128), "n"(IDESC)
        : "memory");
#undef HM_MMA_STEP
}
// The same burst with a run-time K-step count nk (1..8): steps k >= nk are skipped by a guard predicate (the last
// 128-channel pair of a layer whose channel count is not a multiple of 128); both commits are always issued, so the
// barrier protocol does not depend on nk.
template <int AJ, int AK, int BJ, int BK, uint32_t IDESC>
__device__ __forceinline__ void umma_f16_burst8n(uint32_t d_tmem, uint32_t a_lo, uint32_t b_lo, uint32_t acc_first,
                                                 uint32_t nk, uint64_t* bar_tile0, uint64_t* bar_tile1) {
#define HM_MMA_STEPG(K, AOFF, BOFF)                                               \
    "setp.gt.u32 g, %6, " #K ";\n\t"                                              \
    "add.u32 ta, %1, " #AOFF ";\n\t add.u32 tb, %2, " #BOFF ";\n\t"              \
    "mov.b64 da, {ta, %21};\n\t mov.b64 db, {tb, %21};\n\t"                       \
    "@g tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %22, q;\n\t"
    asm volatile(
        "{\n\t.reg .pred p, q, g;\n\t.reg .b64 da, db;\n\t.reg .b32 ta, tb;\n\t"
        "setp.ne.b32 p, %3, 0;\n\t setp.eq.b32 q, %0, %0;\n\t"
        "mov.b64 da, {%1, %21};\n\t mov.b64 db, {%2, %21};\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %22, p;\n\t"
        HM_MMA_STEPG(1, %7, %8) HM_MMA_STEPG(2, %9, %10) HM_MMA_STEPG(3, %11, %12)
        "tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%4];\n\t"
        HM_MMA_STEPG(4, %13, %14) HM_MMA_STEPG(5, %15, %16) HM_MMA_STEPG(6, %17, %18) HM_MMA_STEPG(7, %19, %20)
        "tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%5];\n\t}"
        ::"r"(d_tmem), "r"(a_lo), "r"(b_lo), "r"(acc_first),
          "r"(smem_u32(bar_tile0)), "r"(smem_u32(bar_tile1)), "r"(nk),
          "n"(AK), "n"(BK), "n"(2 * AK), "n"(2 * BK), "n"(3 * AK), "n"(3 * BK),
          "n"(AJ), "n"(BJ), "n"(AJ + AK), "n"(BJ + BK), "n"(AJ + 2 * AK), "n"(BJ + 2 * BK), "n"(AJ + 3 * AK), "n"(BJ + 3 * BK),
          "n"(UMMA_DESC_HI_SW128), "n"(IDESC)
        : "memory");
#undef HM_MMA_STEPG
}
// Arrive on an mbarrier when all previously issued tcgen05.mma of this thread completed.
// (implies tcgen05.fence::before_thread_sync)
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
    asm volatile(
        "tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(
            smem_u32(bar))
        : "memory");
}
// the same, issued only where `pred` != 0 (no divergent branch)
__device__ __forceinline__ void umma_commit_if(uint64_t* bar, uint32_t pred) {
    asm volatile(
        "{\n\t.reg .pred l;\n\t setp.ne.b32 l, %1, 0;\n\t"
        "@l tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n\t}" ::"r"(smem_u32(bar)),
        "r"(pred)
        : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() {
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
// 32 lanes x 32 columns of 32-bit: thread t of the warp receives row (lane base + t),
// columns [col, col+32).
__device__ __forceinline__ void tmem_ld_32x32b_x2(uint32_t taddr, uint32_t& r0, uint32_t& r1) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x2.b32 {%0, %1}, [%2];" : "=r"(r0), "=r"(r1) : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld_32x32b_x4(uint32_t taddr, uint32_t (&r)[4]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld_32x32b_x8(uint32_t taddr, uint32_t (&r)[8]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                 : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld_32x32b_x16(uint32_t taddr, uint32_t (&r)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld_32x32b_x32(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]),
          "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]),
          "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]),
          "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]),
          "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
}

// registers -> TMEM, same 32x32b shape: thread t of the warp writes row (lane base + t), columns [col, col+32)
__device__ __forceinline__ void tmem_st_32x32b_x32(uint32_t taddr, const uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
        "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
        "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};" ::"r"(taddr),
        "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
        "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]), "r"(r[18]),
        "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]),
        "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31])
        : "memory");
}
__device__ __forceinline__ void tmem_st_wait() {
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
}

// ---- UMMA descriptors -------------------------------------------------------------------
// Shared-memory matrix descriptor, K-major operand tile stored as rows of 128 bytes
// (64 bf16) with the 128-byte swizzle TMA applies (CU_TENSOR_MAP_SWIZZLE_128B):
//   bits [0,14)  start address >> 4
//   bits [16,30) leading byte offset >> 4   (ignored for swizzled K-major; 1)
//   bits [32,46) stride byte offset >> 4    (8 rows * 128 B = 1024 B between 8-row groups)
//   bits [46,48) descriptor version = 1 on sm_100
//   bits [61,64) layout type: 2 = SWIZZLE_128B
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t smem_addr) {
    uint64_t d = 0;
    d |= static_cast<uint64_t>((smem_addr & 0x3FFFFu) >> 4);
    d |= static_cast<uint64_t>(1) << 16;
    d |= static_cast<uint64_t>(1024 >> 4) << 32;
    d |= static_cast<uint64_t>(1) << 46;
    d |= static_cast<uint64_t>(2) << 61;
    return d;
}
// MN-major operand tile with the 128-byte swizzle (canonical layout, in 16-byte units,
// ((8,n),(8,k)):((1,LBO),(8,SBO)) — cute/atom/mma_traits_sm100.hpp): a row of 64 MN-contiguous
// 16-bit elements (128 B) per K index, 8 K rows per 1024-byte group; `lbo` = bytes between
// 64-element MN atoms, `sbo` = bytes between 8-row K groups.
__device__ __forceinline__ uint64_t umma_desc_mn_sw128(uint32_t smem_addr, uint32_t lbo, uint32_t sbo) {
    uint64_t d = 0;
    d |= static_cast<uint64_t>((smem_addr & 0x3FFFFu) >> 4);
    d |= static_cast<uint64_t>((lbo >> 4) & 0x3FFFu) << 16;
    d |= static_cast<uint64_t>((sbo >> 4) & 0x3FFFu) << 32;
    d |= static_cast<uint64_t>(1) << 46;
    d |= static_cast<uint64_t>(2) << 61;
    return d;
}
// Instruction descriptor, kind::f16: D = fp32, A = B = 16-bit (fmt 0 = F16, 1 = BF16), both
// K-major, dense.
//   bits [4,6) c_format = 1 (F32); [7,10) a_format; [10,13) b_format
//   bit 15 a_major = 0 (K); bit 16 b_major = 0 (K); [17,23) N >> 3; [24,29) M >> 4
__host__ __device__ constexpr uint32_t umma_idesc_16(uint32_t M, uint32_t N, uint32_t fmt) {
    return (1u << 4) | (fmt << 7) | (fmt << 10) | ((N >> 3) << 17) | ((M >> 4) << 24);
}
// same with A K-major and B MN-major (bit 16 b_major = 1)
__host__ __device__ constexpr uint32_t umma_idesc_16_bmn(uint32_t M, uint32_t N, uint32_t fmt) {
    return umma_idesc_16(M, N, fmt) | (1u << 16);
}

}  // namespace hm
