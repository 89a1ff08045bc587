// hm_trunk.cuh — a run of consecutive RISEv3 bottleneck blocks as ONE kernel (SURVEY.md §8a
// rows F2a-F2d, F3; reference builder_util.py:445-483, :49-80):
//
//   per block:  x' = eca ? x * hardsigmoid(Wc . mean(x) + bc) : x
//               y = relu(pw1(x') + b1);  z = relu(dw_kxk(y) + bd);  x <- x' + pw2(z) + b2
//
// Everything is channel-major ("transposed"): output channels are the MMA M dimension (TMEM
// lanes), the 128 board squares of the CTA's two positions are the N dimension (TMEM columns):
//
//   D1[128 op-channels x 128 sq]  = W1[chunk rows, 384] * X[384 x 128 sq]
//   D2[384 channels    x 128 sq] += W2[:, chunk cols]   * Z[128 op-channels x 128 sq]
//
// so a compute thread that reads its TMEM lane owns ONE channel: the depthwise k x k convolution,
// both ReLUs, the residual and the ECA channel sum are private register arithmetic — no
// shared-memory exchange, no shuffles.  The expanded tensor (up to 1152 channels) never leaves
// the SM.
//
// THE RESIDUAL STREAM IS FP32 AND LIVES IN TENSOR MEMORY.  D2 is never cleared between blocks: it
// is initialised with the fp32 image of the block-0 input, every pw2 MMA of every block
// accumulates on top of it, and the pw2 biases are carried as a per-block CUMULATIVE bias vector
// that is added when D2 is read.  The 16-bit copy X in shared memory is only the B operand of the
// next pw1 (rounded once from the fp32 stream, errors do not accumulate over the 15 blocks), the
// block epilogue reads no residual from shared memory, and an ECA gate rescales the fp32 stream
// in place (tcgen05.ld -> scale -> tcgen05.st) and re-bases the cumulative bias.
//
// Shared memory: X  [6][2 pos][64 ch][64 sq] 16-bit, MN-major 128-byte-swizzled = B operand of pw1
//                Z  [2][2 pos][64 ch][64 sq]                                     = B operand of pw2
//                ring of 6 x 16 KB weight tiles [128 rows x 64 K] K-major        = A operands
// TMEM (512 columns): D2 = 3 x 128 columns (fp32 residual stream + pw2 accumulation), D1 = 128.
// Warps (640 threads): 0-15 = compute, 16 = TMA producer, 17 = MMA issuer, 18-19 idle (they complete the control
// warpgroup that hands registers to the compute warpgroups with setmaxnreg).  A compute thread is
// (TMEM lane quarter q = warp & 3 -> channel r = 32q + lane, position h, board half t): sixteen
// compute warps = four per sub-partition, so the depthwise FFMA2 chains, the TMEM loads and the
// conversions of different warps overlap instead of running phase-locked (two warps per
// sub-partition left the FMA pipe at ~47 % on the 5x5 blocks).
#pragma once
#include <cuda.h>
#include <type_traits>
#include "hm_ptx.cuh"

namespace hm {

constexpr int TR_MAX_BLOCKS = 16;
constexpr int TR_PAR_FLOATS = 27 * 128;  // per 128-channel chunk: b1 | dw bias | 25 taps, each [128]

struct TrunkBlock {
    int chunk0;    // first global chunk (row block of W1cat / W2cat, entry of par)
    int n_chunks;  // ceil(cop / 128)
    int ksize;     // 3 or 5
    int bias2_off; // offset into bias2 of this block's CUMULATIVE pw2 bias (since the last re-base), floats
    int eca_chunk0;   // ECA on the block input: first of 3 row chunks of Wc in W1cat, or -1
    int eca_bias_off; // offset of the 384 ECA biases in bias2
};
struct TrunkParams {
    int n_blocks;
    TrunkBlock blk[TR_MAX_BLOCKS];
    const float* par;    // [chunk][TR_PAR_FLOATS]
    const float* bias2;  // cumulative pw2 biases and ECA biases, [...][384]
    unsigned long long* trace;  // -DHM_BRINGUP builds only: CTA 0 logs (tag << 48 | clock) per role
    int dbg;                    // -DHM_BRINGUP builds only, traced launches only (results are wrong when set): timing
                                // experiments — 1 skip the depthwise math, 2 skip the TMEM reads of D1, 4 skip the Z
                                // stores, 8 no weight loads (MMAs run on whatever the ring holds)
};
#ifdef HM_BRINGUP
#define TR_DBG(bit) (dbg & (bit))
#else
#define TR_DBG(bit) false
#endif

// timeline tracing (tools/trace_trunk.py): roles 0 = producer, 1 = MMA issuer, 2 = compute warp 2.
// Compiled in only with -DHM_BRINGUP; the default build carries no tracing or debug switches.
constexpr int TR_TRACE_CAP = 2048;
#ifdef HM_BRINGUP
#define TR_MARK(role, tag)                                                                       \
    do {                                                                                         \
        if (tr_trace != nullptr && trace_n < TR_TRACE_CAP)                                       \
            tr_trace[(role) * TR_TRACE_CAP + trace_n++] =                                        \
                (static_cast<unsigned long long>(tag) << 48) | (clock64() & 0xffffffffffffull);  \
    } while (0)
#else
#define TR_MARK(role, tag) do { } while (0)
#endif

constexpr int TR_COMPUTE_WARPS = 16;
constexpr int TR_COMPUTE = TR_COMPUTE_WARPS * 32;
constexpr int TR_PRODUCER_WARP = TR_COMPUTE_WARPS, TR_MMA_WARP = TR_COMPUTE_WARPS + 1;
constexpr int TR_THREADS = TR_COMPUTE + 128;   // + one warpgroup: producer, MMA issuer, two idle warps
// Weight ring: 6 slots of 16 KB = exactly one "group" (the 6 W1 tiles of a chunk's pw1, or the 6
// W2 tiles of its pw2), so slot indices are compile-time inside the fully unrolled group loops and
// the barrier parity simply toggles per group.  One "full" barrier guards TWO slots (32 KB,
// 8 MMAs), "empty" barriers are per slot: a thread pays ~40 cycles per MMA issue or successful
// wait, ~200 per 16 KB TMA issue and ~100 per barrier round trip (tools/probes/probe_issue.cu,
// probe_mcast.cu), so both single-thread roles keep up with the 256-cycle tile.
constexpr int TR_SLOTS = 6;
constexpr int TR_PAIRS = 3;
constexpr int TR_SLOT_BYTES = 16384;
constexpr int TR_X_OFF = 0;
constexpr int TR_Z_OFF = 6 * 16384;
constexpr int TR_RING_OFF = TR_Z_OFF + 32768;
constexpr int TR_BAR_OFF = TR_RING_OFF + TR_SLOTS * TR_SLOT_BYTES;
constexpr int TR_SMEM = TR_BAR_OFF + 256;

// X_READY[m] / D2_FULL[m] are per 128-channel output block m: the block epilogue of block m
// (channels m*128..) overlaps the pw2 MMAs of blocks m+1.. and the next bottleneck block's first
// pw1 starts on X blocks 2m, 2m+1 as soon as epilogue part m has rewritten them.
enum {
    TB_X_FULL = 0, TB_D1_FULL, TB_D1_EMPTY, TB_Z_FULL, TB_Z_EMPTY,
    TB_X_READY,                               // 3
    TB_D2_FULL = TB_X_READY + 3,              // 3
    TB_PAIR_FULL = TB_D2_FULL + 3,            // TR_PAIRS
    TB_TILE_EMPTY = TB_PAIR_FULL + TR_PAIRS,  // TR_SLOTS: per 16 KB tile, so its reload starts after ITS 4 MMAs
    TB_ECA_DONE = TB_TILE_EMPTY + TR_SLOTS,   // the gate pre-activations of all 384 channels are in TMEM (D1)
    TB_ECA_READY = TB_ECA_DONE + 1,           // the channel means of the block input are in shared memory
    TB_COUNT = TB_ECA_READY + 1
};

// k-th completion (k = 0, 1, ...) of a "full"-type barrier
__device__ __forceinline__ void tr_wait_full(uint64_t* bar, uint32_t k) { mbar_wait(bar, k & 1); }
// k-th reuse of a resource: free for k = 0, then needs the (k-1)-th completion of its "empty" barrier
__device__ __forceinline__ void tr_wait_empty(uint64_t* bar, uint32_t k) { mbar_wait(bar, (k & 1) ^ 1); }

// Depthwise k x k for FOUR output rows of one channel's board, inputs in registers: local input row
// l = 0..NR-1 (NR = 4 + KS/2), local output rows 0..3, out[lo] = sum_ky w[ky] * in[lo + ky - H] with
// rows above local row 0 off the board.  The thread owning the TOP half of the board (t = 0) loads
// image rows 0.. in ascending order; the thread owning the BOTTOM half loads image rows 7, 6, ... (so
// its local rows ascend away from ITS board edge too) and uses the taps flipped in ky: one code path.
// Packed fp32 FMAs (FFMA2, fma.rn.f32x2): two horizontally adjacent squares per instruction, identical
// fp32 products.  Every operand is an ALIGNED input pair A[a] = (in[2a], in[2a+1]) (no register
// shuffling): taps with an even horizontal offset accumulate into the aligned output pairs
// E[j] = (o[2j], o[2j+1]), taps with an odd offset into the odd-aligned pairs O[k] = (o[2k-1], o[2k]);
// one scalar add per square joins them.
template <int KS>
__device__ __forceinline__ void dw_row(int lo, const float (&y)[(4 + KS / 2) * 8], const float (&w)[KS * KS], float bd,
                                       uint32_t (&out)[4]) {
    constexpr int H = KS / 2;
    float2 E[4], O[5];
#pragma unroll
    for (int j = 0; j < 4; ++j) E[j] = make_float2(bd, bd);
#pragma unroll
    for (int j = 0; j < 5; ++j) O[j] = make_float2(0.0f, 0.0f);
#pragma unroll
    for (int ky = 0; ky < KS; ++ky) {
        const int li = lo + ky - H;
        if (li < 0) continue;           // off the board (li <= 3 + H < NR always)
#pragma unroll
        for (int kx = 0; kx < KS; ++kx) {
            const int dx = kx - H;
            const float2 ww = make_float2(w[ky * KS + kx], w[ky * KS + kx]);
#pragma unroll
            for (int a = 0; a < 4; ++a) {
                const float2 in = make_float2(y[li * 8 + 2 * a], y[li * 8 + 2 * a + 1]);
                if ((dx & 1) == 0) {
                    const int j = a - dx / 2;          // o[x] += w * in[x + dx]
                    if (j < 0 || j > 3) continue;
                    E[j] = __ffma2_rn(in, ww, E[j]);
                } else {
                    const int k = a + (1 - dx) / 2;    // pair (o[2a-dx], o[2a+1-dx]) = O[k]
                    O[k] = __ffma2_rn(in, ww, O[k]);
                }
            }
        }
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) out[j] = pack_act2_relu(E[j].x + O[j].y, E[j].y + O[j + 1].x);
}

// One 128-channel chunk on a compute thread (channel r, position h, board half t):
//   D1 rows of its half (+ halo) -> relu(. + b1) -> depthwise -> relu(. + bd) -> 16-bit -> Z.
// `par` = this chunk's parameters; they are requested BEFORE the D1 wait so that their L2/L1 latency
// hides behind it.
template <int KS>
__device__ __forceinline__ void trunk_chunk(const float* __restrict__ par, int r, int t, uint32_t tm_row0,
                                            int row_step, uint64_t* bars, uint32_t d1_n, uint32_t z_n,
                                            uint8_t* zrow, int key, int lane,
                                            unsigned long long* tr_trace, int& trace_n, int dbg) {
    constexpr int H = KS / 2, NR = 4 + H;
    float w[KS * KS];
#pragma unroll
    for (int ky = 0; ky < KS; ++ky) {
        const int kys = t ? KS - 1 - ky : ky;          // bottom half: taps flipped vertically
#pragma unroll
        for (int kx = 0; kx < KS; ++kx) w[ky * KS + kx] = __ldg(par + 256 + (kys * KS + kx) * 128 + r);
    }
    const float b1 = __ldg(par + r), bd = __ldg(par + 128 + r);
    TR_MARK(2, 1);
    mbar_wait_backoff(&bars[TB_D1_FULL], d1_n & 1, 20);   // sixteen warps wait here ~1 k cycles per 3x3 chunk: off the issue slots
    tc_fence_after();
    TR_MARK(2, 2);
    float y[NR * 8];
    {
        uint32_t a[NR][8];
        if (TR_DBG(2)) {
#pragma unroll
            for (int l = 0; l < NR; ++l)
#pragma unroll
                for (int j = 0; j < 8; ++j) a[l][j] = __float_as_uint(b1);
        } else {
#pragma unroll
            for (int l = 0; l < NR; ++l) tmem_ld_32x32b_x8(tm_row0 + static_cast<uint32_t>(l * row_step), a[l]);
        }
        tmem_ld_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&bars[TB_D1_EMPTY]);  // D1 is free as soon as it is in registers
        const float2 bb = make_float2(b1, b1);
#pragma unroll
        for (int l = 0; l < NR; ++l)
#pragma unroll
            for (int j = 0; j < 4; ++j) {   // packed bias add (FADD2), scalar max
                const float2 u = __fadd2_rn(make_float2(__uint_as_float(a[l][2 * j]), __uint_as_float(a[l][2 * j + 1])), bb);
                y[l * 8 + 2 * j] = fmaxf(u.x, 0.0f);
                y[l * 8 + 2 * j + 1] = fmaxf(u.y, 0.0f);
            }
    }
    TR_MARK(2, 3);
    auto store_row = [&](int lo, const uint32_t (&o)[4]) {
        const int yo = t ? 7 - lo : lo;     // image row of local output row lo = 16-byte piece of the Z row
        if (TR_DBG(4) && o[0] != 0x12345678u) return;
        *reinterpret_cast<uint4*>(zrow + ((yo ^ key) << 4)) = make_uint4(o[0], o[1], o[2], o[3]);
    };
    if (KS == 5) {
        // The 5x5 chunks are the compute-bound ones: by the time two of the four output rows are done, pw2 of the
        // previous chunk has long released Z, so the rows are stored as they are finished and only two rows of packed
        // results are ever live (the register pressure of holding all four spills — to L2, see the kernel comment).
        uint32_t r0[4], r1[4];
        dw_row<KS>(0, y, w, bd, r0);
        dw_row<KS>(1, y, w, bd, r1);
        TR_MARK(2, 4);
        tr_wait_empty(&bars[TB_Z_EMPTY], z_n);  // pw2 of the previous chunk has consumed Z
        TR_MARK(2, 5);
        store_row(0, r0);
        store_row(1, r1);
        dw_row<KS>(2, y, w, bd, r0);
        store_row(2, r0);
        dw_row<KS>(3, y, w, bd, r1);
        store_row(3, r1);
    } else {
        uint32_t packed[4][4];
#pragma unroll
        for (int lo = 0; lo < 4; ++lo) dw_row<KS>(lo, y, w, bd, packed[lo]);
        TR_MARK(2, 4);
        tr_wait_empty(&bars[TB_Z_EMPTY], z_n);  // pw2 of the previous chunk has consumed Z
        TR_MARK(2, 5);
#pragma unroll
        for (int lo = 0; lo < 4; ++lo) store_row(lo, packed[lo]);
    }
    fence_proxy_async_smem();
    __syncwarp();
    if (lane == 0) mbar_arrive(&bars[TB_Z_FULL]);
    TR_MARK(2, 6);
}

__global__ void __launch_bounds__(TR_THREADS, 1)
trunk_kernel(const __grid_constant__ TrunkParams p, const __grid_constant__ CUtensorMap tmX,
             const __grid_constant__ CUtensorMap tmW1, const __grid_constant__ CUtensorMap tmW2) {
    extern __shared__ __align__(1024) uint8_t tr_smem[];
    uint8_t* smem = tr_smem;
    if (smem_u32(smem) & 1023u) __trap();  // swizzle atoms need 1024-byte alignment
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + TR_BAR_OFF);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + TB_COUNT);
    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int tile = blockIdx.x;
#ifdef HM_BRINGUP
    const int dbg = p.dbg;
#else
    const int dbg = 0;
#endif
    (void)dbg;
#ifdef HM_BRINGUP
    unsigned long long* const tr_trace = (blockIdx.x == 0 && (lane == 0)) ? p.trace : nullptr;
#else
    unsigned long long* const tr_trace = nullptr;
#endif

    if (threadIdx.x == 0) {
        tma_prefetch_desc(&tmX);
        tma_prefetch_desc(&tmW1);
        tma_prefetch_desc(&tmW2);
        for (int i = 0; i < TB_COUNT; ++i) {
            uint32_t count = 1;
            if ((i >= TB_X_READY && i < TB_X_READY + 3) || i == TB_D1_EMPTY || i == TB_Z_FULL || i == TB_ECA_READY)
                count = TR_COMPUTE_WARPS;
            mbar_init(&bars[i], count);
        }
        fence_barrier_init();
    }
    if (warp == TR_MMA_WARP) {
        tmem_alloc(tmem_slot, 512);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const uint32_t tm_d2 = tmem_base;
    const uint32_t tm_d1 = tmem_base + 384u;

    // The two single-thread roles run with ALL 32 lanes in uniform control flow (waits, loop
    // counters, descriptor arithmetic stay warp-uniform -> uniform registers, no per-lane
    // election loops around every UTMALDG / UTCHMMA); only the issuing instructions themselves
    // are predicated on one elected lane.
    // Register budget: 640 threads x 96 at launch.  Every spill goes to L2 — with 224 KB of shared memory in use there
    // is next to no L1 left — so local-memory traffic is kept out of the hot loops (no dynamically indexed register
    // arrays: that alone was worth 5.6 % of the kernel).  Handing registers from the control warpgroup to the compute
    // warpgroups (setmaxnreg 64 / 104) was tried: ptxas then spills the MMA issuer and gives the depthwise nothing.
#ifdef HM_TRUNK_SETMAXNREG
    if (warp >= TR_COMPUTE_WARPS) asm volatile("setmaxnreg.dec.sync.aligned.u32 64;" ::: "memory");
    else asm volatile("setmaxnreg.inc.sync.aligned.u32 104;" ::: "memory");
#endif
    if (warp == TR_PRODUCER_WARP) {
        // ================================ TMA producer ================================
        const bool leader = elect_one();
        if (leader) {
            mbar_arrive_expect_tx(&bars[TB_X_FULL], 6 * 16384);
#pragma unroll
            for (int kb = 0; kb < 6; ++kb)
                tma_load_3d(smem + TR_X_OFF + kb * 16384, &tmX, &bars[TB_X_FULL], 0, kb * 64, tile * 2);
        }
        __syncwarp();
        // one group = 6 tiles = one lap over the ring; `par` is the empty-barrier parity of the lap
        uint32_t par = 1;
        int trace_n = 0;
        auto group_w1 = [&](int g) {   // W1 rows [g*128, +128) x K blocks 0..5
            TR_MARK(0, 1);
#pragma unroll
            for (int b2 = 0; b2 < TR_PAIRS; ++b2) {
#pragma unroll
                for (int j = 0; j < 2; ++j) {
                    mbar_wait(&bars[TB_TILE_EMPTY + 2 * b2 + j], par);
                    if (j == 0) TR_MARK(0, 10 + b2);
                    if (leader) {
                        if (j == 0) mbar_arrive_expect_tx(&bars[TB_PAIR_FULL + b2], 2 * TR_SLOT_BYTES);
                        tma_load_2d(smem + TR_RING_OFF + (2 * b2 + j) * TR_SLOT_BYTES, &tmW1,
                                    &bars[TB_PAIR_FULL + b2], (2 * b2 + j) * 64, g * 128);
                    }
                    __syncwarp();
                }
            }
            par ^= 1;
        };
        auto group_w2 = [&](int g) {   // W2 of chunk g: tile s = m*2 + kh -> rows g*384 + m*128, K half kh
            TR_MARK(0, 2);
#pragma unroll
            for (int b2 = 0; b2 < TR_PAIRS; ++b2) {
#pragma unroll
                for (int j = 0; j < 2; ++j) {   // pair b2 = output block m = b2, K halves kh = j
                    mbar_wait(&bars[TB_TILE_EMPTY + 2 * b2 + j], par);
                    if (j == 0) TR_MARK(0, 20 + b2);
                    if (leader) {
                        if (j == 0) mbar_arrive_expect_tx(&bars[TB_PAIR_FULL + b2], 2 * TR_SLOT_BYTES);
                        tma_load_2d(smem + TR_RING_OFF + (2 * b2 + j) * TR_SLOT_BYTES, &tmW2,
                                    &bars[TB_PAIR_FULL + b2], j * 64, g * 384 + b2 * 128);
                    }
                    __syncwarp();
                }
            }
            par ^= 1;
        };
        for (int b = 0; b < (TR_DBG(8) ? 0 : p.n_blocks); ++b) {
            const int g0 = p.blk[b].chunk0, nch = p.blk[b].n_chunks;
            if (b > 0 && p.blk[b].eca_chunk0 >= 0)   // in-kernel ECA: Wc streams through the ring like W1
                for (int m = 0; m < 3; ++m) group_w1(p.blk[b].eca_chunk0 + m);
            group_w1(g0);
            for (int c = 0; c < nch; ++c) {
                if (c + 1 < nch) group_w1(g0 + c + 1);
                group_w2(g0 + c);
            }
        }
    } else if (warp == TR_MMA_WARP) {
        // ================================ MMA issuer ================================
        // Issue discipline (tools/probes/probe_issue.cu): the tensor pipe queues only two or three MMAs behind the one
        // it runs, so whatever this thread does BETWEEN two bursts of MMAs must fit into ~150 cycles.  Hence (1) a
        // burst = 8 MMAs + the two tile commits in one asm block with the descriptor arithmetic between the MMAs
        // (umma_f16_burst8), and (2) the full-barrier of the NEXT burst is probed (non-blocking test_wait) BEFORE this
        // burst is issued and only consumed after it: in steady state the weight ring runs ahead and the wait costs
        // nothing; the blocking wait is the fallback.
        const bool leader = elect_one();
        constexpr uint32_t idesc = umma_idesc_16_bmn(128, 128, UMMA_FMT_16);
        uint32_t d1_n = 0, z_n = 0, par = 0;
        int trace_n = 0;
        // descriptors differ only in the start-address field (16-byte units): hoist the bases
        const uint64_t a0d = umma_desc_sw128(smem_u32(smem + TR_RING_OFF));
        const uint64_t xd0d = umma_desc_mn_sw128(smem_u32(smem + TR_X_OFF), 8192, 1024);
        const uint64_t zd0d = umma_desc_mn_sw128(smem_u32(smem + TR_Z_OFF), 8192, 1024);
        const uint32_t a0 = static_cast<uint32_t>(a0d);   // the high words of all four are UMMA_DESC_HI_SW128
        const uint32_t xd0 = static_cast<uint32_t>(xd0d), zd0 = static_cast<uint32_t>(zd0d);
        const uint64_t zmd = umma_desc_sw128(smem_u32(smem + TR_Z_OFF));   // ECA means: K-major, 16 rows per K block
        const uint32_t zm0 = static_cast<uint32_t>(zmd);
        // `have` = the full-barrier of ring pair 0 of the current lap is already known complete
        uint32_t have = 0;
        // one lap over the ring = three bursts; kind 0 = pw1 (D1 = W1 . X), 1 = pw2 (D2[m] += W2 . Z), 2 = ECA (G[m] = Wc . means)
        auto lap = [&](auto kind_tag, int xblock, int m_eca, bool last_chunk) {
            constexpr int KIND = decltype(kind_tag)::value;
#pragma unroll
            for (int b2 = 0; b2 < TR_PAIRS; ++b2) {
                if (KIND == 0 && xblock >= 0) tr_wait_full(&bars[TB_X_READY + b2], xblock);  // X blocks 2*b2, 2*b2+1 ready
                if (b2 == 0 && !have && !TR_DBG(8)) mbar_wait(&bars[TB_PAIR_FULL], par);   // b2 > 0: waited below
                tc_fence_after();
                TR_MARK(1, (KIND == 1 ? 20 : 10) + b2);
                // probe the NEXT burst's barrier now, look at the answer after this burst
                const int nb = b2 + 1 == TR_PAIRS ? 0 : b2 + 1;
                const uint32_t npar = b2 + 1 == TR_PAIRS ? par ^ 1 : par;
                const uint32_t next_ok = mbar_test_wait(&bars[TB_PAIR_FULL + nb], npar);
                const uint32_t ld = 1u;
                if (!leader) {
                } else if (KIND == 0) {
                    umma_f16_burst8<1024, 2, 1024, 128, idesc>(tm_d1, a0 + (2 * b2) * 1024, xd0 + (2 * b2) * 1024,
                                               b2 ? 1u : 0u, &bars[TB_TILE_EMPTY + 2 * b2], &bars[TB_TILE_EMPTY + 2 * b2 + 1], ld);
                    if (b2 == TR_PAIRS - 1) umma_commit_if(&bars[TB_D1_FULL], ld);
                } else if (KIND == 1) {   // always accumulates: D2 is the running fp32 residual stream
                    umma_f16_burst8<1024, 2, 1024, 128, idesc>(tm_d2 + b2 * 128, a0 + (2 * b2) * 1024, zd0, 1u,
                                               &bars[TB_TILE_EMPTY + 2 * b2], &bars[TB_TILE_EMPTY + 2 * b2 + 1], ld);
                    if (last_chunk) umma_commit_if(&bars[TB_D2_FULL + b2], ld);  // output block m = b2 complete
                    if (b2 == TR_PAIRS - 1) umma_commit_if(&bars[TB_Z_EMPTY], ld);
                } else {
                    constexpr uint32_t idesc_eca = umma_idesc_16(128, 16, UMMA_FMT_16);
                    umma_f16_burst8<1024, 2, 128, 2, idesc_eca>(tm_d1 + m_eca * 16, a0 + (2 * b2) * 1024, zm0 + (2 * b2) * 128,
                                            b2 ? 1u : 0u, &bars[TB_TILE_EMPTY + 2 * b2], &bars[TB_TILE_EMPTY + 2 * b2 + 1], ld);
                    if (m_eca == 2 && b2 == TR_PAIRS - 1) umma_commit_if(&bars[TB_ECA_DONE], ld);
                }
                __syncwarp();
                if (b2 + 1 < TR_PAIRS) {
                    if (!next_ok && !TR_DBG(8)) mbar_wait(&bars[TB_PAIR_FULL + nb], npar);
                } else {
                    have = next_ok;
                }
            }
            par ^= 1;
        };
        using K0 = std::integral_constant<int, 0>;
        using K1 = std::integral_constant<int, 1>;
        using K2 = std::integral_constant<int, 2>;
        auto pw1 = [&](int xblock) {   // xblock >= 0: first pw1 of bottleneck block `xblock`
            TR_MARK(1, 1);
            tr_wait_empty(&bars[TB_D1_EMPTY], d1_n);
            tc_fence_after();
            TR_MARK(1, 2);
            lap(K0{}, xblock, 0, false);
            ++d1_n;
            TR_MARK(1, 3);
        };
        uint32_t eca_n = 0;
        for (int b = 0; b < p.n_blocks; ++b) {
            const int nch = p.blk[b].n_chunks;
            if (b > 0 && p.blk[b].eca_chunk0 >= 0) {
                // ECA matvec on the tensor core: G[m] (128 out channels x N = 16) = Wc tiles (ring) x channel
                // means (fp16 hi/lo split, K-major, parked in the Z buffer by the compute warps: columns
                // 0-3 = hi parts of (half 0 pos 0, half 0 pos 1, half 1 pos 0, half 1 pos 1), 4-7 = lo parts);
                // accumulators in D1 columns [16m, 16m+16) — D1 is idle between blocks (the last chunk's
                // pre-activations were drained long ago), D2 holds the live residual stream
                tr_wait_full(&bars[TB_ECA_READY], eca_n);
                tr_wait_empty(&bars[TB_D1_EMPTY], d1_n);
                tc_fence_after();
                for (int m = 0; m < 3; ++m) lap(K2{}, -1, m, false);
                ++eca_n;
            }
            pw1(b);
            for (int c = 0; c < nch; ++c) {
                if (c + 1 < nch) pw1(-1);
                TR_MARK(1, 4);
                tr_wait_full(&bars[TB_Z_FULL], z_n);
                tc_fence_after();
                TR_MARK(1, 5);
                lap(K1{}, -1, 0, c + 1 == nch);
                ++z_n;
                TR_MARK(1, 6);
            }
        }
    } else if (warp < TR_COMPUTE_WARPS) {
        // ================================ compute warps ================================
        const int cw = warp;
        const int q = warp & 3;             // TMEM lane quarter this warp may access
        const int h = (cw >> 2) & 1;        // position of the tile handled by this thread
        const int t = cw >> 3;              // board half: 0 = ranks 0-3 (squares 0-31), 1 = ranks 4-7
        const int r = q * 32 + lane;        // channel within a 128-channel chunk / M block
        const uint32_t lane_off = static_cast<uint32_t>(q * 32) << 16;
        const int key = r & 7;
        // this thread's 128-byte row inside a [2 pos][64 ch][64 sq] block
        const uint32_t row_off = (r >> 6) * 16384 + h * 8192 + (r & 63) * 128;
        // its 32 columns of a 128-column accumulator block, and its four 16-byte pieces of a row
        const uint32_t col_half = static_cast<uint32_t>(h * 64 + t * 32);
        const int piece0 = t * 4;

        tr_wait_full(&bars[TB_X_FULL], 0);   // the tile's block input has landed
        // D2 <- fp32 image of the input: the residual stream starts here and stays in TMEM
#pragma unroll 1
        for (int m = 0; m < 3; ++m) {
            const uint8_t* xrow = smem + TR_X_OFF + m * 32768 + row_off;
            uint32_t v[32];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const uint4 u = *reinterpret_cast<const uint4*>(xrow + (((piece0 + j) ^ key) << 4));
                const uint32_t uu[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const float2 f = unpack_act2(uu[e]);
                    v[j * 8 + 2 * e] = __float_as_uint(f.x);
                    v[j * 8 + 2 * e + 1] = __float_as_uint(f.y);
                }
            }
            tmem_st_32x32b_x32(tm_d2 + lane_off + m * 128 + col_half, v);
        }
        tmem_st_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) {
            mbar_arrive(&bars[TB_X_READY + 0]);
            mbar_arrive(&bars[TB_X_READY + 1]);
            mbar_arrive(&bars[TB_X_READY + 2]);
        }

        uint32_t d1_n = 0, z_n = 0, eca_cnt = 0;
        int trace_n = 0;
        unsigned long long* const ctrace = warp == 0 ? tr_trace : nullptr;
        // this thread's D1 rows: local row l of its half = image row (t ? 7 - l : l) of position h
        const uint32_t tm_row0 = tm_d1 + lane_off + static_cast<uint32_t>(h * 64 + t * 56);
        const int row_step = t ? -8 : 8;
        uint8_t* const zrow = smem + TR_Z_OFF + row_off;
        for (int b = 0; b < p.n_blocks; ++b) {
            const int g0 = p.blk[b].chunk0, nch = p.blk[b].n_chunks, ks = p.blk[b].ksize;
            if (b > 0 && p.blk[b].eca_chunk0 >= 0) {
                // ---- ECA on the block input (builder_util.py:49-80, 481-483), inside the kernel:
                // gate[o] = hardsigmoid(Wc[o,:] . mean + b[o]).  The matvec ran on the tensor core (MMA warp);
                // here the pre-activation of this thread's channels comes back from TMEM, and the fp32
                // residual stream is rescaled IN TMEM (x' = g * (D2 + cumulative bias), which also re-bases
                // the cumulative bias to zero) together with its 16-bit operand copy in shared memory.
                const float* eb = p.bias2 + p.blk[b].eca_bias_off;
                const float* cbp = p.bias2 + p.blk[b - 1].bias2_off;
                float ebias[3], cb[3];
#pragma unroll
                for (int m = 0; m < 3; ++m) { ebias[m] = __ldg(eb + m * 128 + r); cb[m] = __ldg(cbp + m * 128 + r); }
                { unsigned long long* tr_trace = ctrace; TR_MARK(2, 30); }
                tr_wait_full(&bars[TB_ECA_DONE], eca_cnt);
                tc_fence_after();
                { unsigned long long* tr_trace = ctrace; TR_MARK(2, 31); }
                float gate[3];
#pragma unroll
                for (int m = 0; m < 3; ++m) {
                    uint32_t gv[8];
                    tmem_ld_32x32b_x8(tm_d1 + lane_off + m * 16, gv);
                    tmem_ld_wait();
                    // columns h, 2+h, 4+h, 6+h (selected without indexing the register array dynamically)
                    const float acc = (__uint_as_float(h ? gv[1] : gv[0]) + __uint_as_float(h ? gv[3] : gv[2])) +
                                      (__uint_as_float(h ? gv[5] : gv[4]) + __uint_as_float(h ? gv[7] : gv[6]));
                    // torch.nn.Hardsigmoid: relu6(x + 3) / 6
                    gate[m] = fminf(fmaxf((acc + ebias[m]) * (1.0f / 6.0f) + 0.5f, 0.0f), 1.0f);
                }
                // every thread has read ALL its gates before any X_READY below can complete (it needs all 16
                // warps), so the first pw1 of this block may overwrite D1 as soon as X_READY[0] fires
#pragma unroll
                for (int m = 0; m < 3; ++m) {
                    uint32_t a[32];
                    const uint32_t taddr = tm_d2 + lane_off + m * 128 + col_half;
                    tmem_ld_32x32b_x32(taddr, a);
                    tmem_ld_wait();
                    const float g = gate[m], c0 = cb[m];
                    uint8_t* xrow = smem + TR_X_OFF + m * 32768 + row_off;
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        uint32_t ou[4];
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            const int i = j * 8 + 2 * e;
                            const float v0 = (__uint_as_float(a[i]) + c0) * g, v1 = (__uint_as_float(a[i + 1]) + c0) * g;
                            a[i] = __float_as_uint(v0);
                            a[i + 1] = __float_as_uint(v1);
                            ou[e] = pack_act2(v0, v1);
                        }
                        *reinterpret_cast<uint4*>(xrow + (((piece0 + j) ^ key) << 4)) = make_uint4(ou[0], ou[1], ou[2], ou[3]);
                    }
                    tmem_st_32x32b_x32(taddr, a);
                    tmem_st_wait();
                    tc_fence_before();
                    fence_proxy_async_smem();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&bars[TB_X_READY + m]);
                }
                { unsigned long long* tr_trace = ctrace; TR_MARK(2, 35); }
                ++eca_cnt;
            }
            {
                const float* par0 = p.par + static_cast<size_t>(g0) * TR_PAR_FLOATS;
                if (ks == 3) {
#pragma unroll 1
                    for (int c = 0; c < nch; ++c, ++d1_n, ++z_n)
                        trunk_chunk<3>(par0 + static_cast<size_t>(c) * TR_PAR_FLOATS, r, t, tm_row0, row_step, bars, d1_n, z_n,
                                       zrow, key, lane, ctrace, trace_n, dbg);
                } else {
#pragma unroll 1
                    for (int c = 0; c < nch; ++c, ++d1_n, ++z_n)
                        trunk_chunk<5>(par0 + static_cast<size_t>(c) * TR_PAR_FLOATS, r, t, tm_row0, row_step, bars, d1_n, z_n,
                                       zrow, key, lane, ctrace, trace_n, dbg);
                }
            }
            // ---- block epilogue: the new 16-bit operand copy X <- D2 + cumulative bias (D2 itself stays) ----
            const bool last = (b + 1 == p.n_blocks);
            const float* b2 = p.bias2 + p.blk[b].bias2_off;
            const bool want_means = !last && p.blk[b + 1].eca_chunk0 >= 0;  // next block gates its input
            float sums3[3];
#pragma unroll
            for (int m = 0; m < 3; ++m) {
                const float bias = __ldg(b2 + m * 128 + r);
                uint8_t* xrow = smem + TR_X_OFF + m * 32768 + row_off;
                { unsigned long long* tr_trace = ctrace; TR_MARK(2, 7); }
                tr_wait_full(&bars[TB_D2_FULL + m], b);
                tc_fence_after();
                { unsigned long long* tr_trace = ctrace; TR_MARK(2, 8); }
                uint32_t a[32];
                tmem_ld_32x32b_x32(tm_d2 + lane_off + m * 128 + col_half, a);
                tmem_ld_wait();
                const float2 bb = make_float2(bias, bias);
                float s = 0.0f, s1 = 0.0f;
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    uint32_t ou[4];
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        const int i = j * 8 + 2 * e;
                        const float2 o = __fadd2_rn(make_float2(__uint_as_float(a[i]), __uint_as_float(a[i + 1])), bb);
                        ou[e] = pack_act2(o.x, o.y);
                        s += o.x;       // ECA squeeze of the next block: mean of the fp32 stream
                        s1 += o.y;
                    }
                    *reinterpret_cast<uint4*>(xrow + (((piece0 + j) ^ key) << 4)) = make_uint4(ou[0], ou[1], ou[2], ou[3]);
                }
                sums3[m] = s + s1;
                tc_fence_before();
                fence_proxy_async_smem();
                __syncwarp();
                if (!last && !want_means && lane == 0) mbar_arrive(&bars[TB_X_READY + m]);
                { unsigned long long* tr_trace = ctrace; TR_MARK(2, 9); }
            }
            if (want_means) {
                // every pw2 of this block has completed (D2_FULL[2]), so Z is idle: park the channel means there
                // as the K-major fp16 B operand of the ECA matvec: per 64-channel K block 16 rows of 128 bytes,
                // row n = column n of the N = 16 MMA.  This thread contributes the partial mean of ITS board half
                // as an fp16 hi/lo pair (rows 2t+h and 4+2t+h): the tensor core sums halves and hi + lo, so the
                // gate sees the mean of the fp32 stream to ~22 bits.  Rows 8..15 are zero.
                uint8_t* zm = smem + TR_Z_OFF;
                const int n_hi = 2 * t + h, n_lo = 4 + n_hi;
#pragma unroll
                for (int m = 0; m < 3; ++m) {
                    const int ch = m * 128 + r, kb = ch >> 6, k = ch & 63;
                    const float mean = sums3[m] * (1.0f / 64.0f);
                    const act_t hi = float_to_act(mean);
                    const act_t lo = float_to_act(mean - act_to_float(hi));
                    uint8_t* base = zm + kb * 2048 + (k & 7) * 2;
                    *reinterpret_cast<act_t*>(base + n_hi * 128 + (((k >> 3) ^ n_hi) << 4)) = hi;
                    *reinterpret_cast<act_t*>(base + n_lo * 128 + (((k >> 3) ^ n_lo) << 4)) = lo;
                }
                for (int i = threadIdx.x; i < 6 * 8 * 8; i += TR_COMPUTE) {
                    const int kb = i >> 6, rem = i & 63;
                    *reinterpret_cast<uint4*>(zm + kb * 2048 + (8 + (rem >> 3)) * 128 + ((rem & 7) << 4)) = make_uint4(0, 0, 0, 0);
                }
                fence_proxy_async_smem();
                __syncwarp();
                if (lane == 0) mbar_arrive(&bars[TB_ECA_READY]);
            }
        }
        // ---- run output: shared-memory tile -> global (TMA store) ----
        named_bar_sync(1, TR_COMPUTE);
        if (threadIdx.x == 0) {
            for (int kb = 0; kb < 6; ++kb) tma_store_3d(&tmX, smem + TR_X_OFF + kb * 16384, 0, kb * 64, tile * 2);
            bulk_commit_group();
            bulk_wait_group0();
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == TR_MMA_WARP) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

}  // namespace hm
