// hm_trunk.cuh — a run of consecutive RISEv3 bottleneck blocks as ONE kernel (SURVEY.md §8a
// rows F2a-F2d; reference builder_util.py:445-483):
//
//   per block:  y = relu(pw1(x) + b1);  z = relu(dw_kxk(y) + bd);  x <- x + pw2(z) + b2
//
// Everything is channel-major ("transposed"): output channels are the MMA M dimension (TMEM
// lanes), the 128 board squares of the CTA's two positions are the N dimension (TMEM columns):
//
//   D1[128 op-channels x 128 sq]  = W1[chunk rows, 384] * X[384 x 128 sq]
//   D2[384 channels    x 128 sq] += W2[:, chunk cols]   * Z[128 op-channels x 128 sq]
//
// so a compute thread that reads its TMEM lane owns ONE channel and all 64 squares of one
// position: the depthwise k x k convolution, both ReLUs, the residual add and the ECA channel
// sum are private register arithmetic — no shared-memory exchange, no shuffles.  The expanded
// tensor (up to 1152 channels) never leaves the SM, and the trunk activation X stays resident
// in shared memory across ALL blocks of the run (updated in place by the block epilogue): per
// run it is read from HBM once and written once.
//
// Shared memory: X  [6][2 pos][64 ch][64 sq] 16-bit, MN-major 128-byte-swizzled = B operand of pw1
//                Z  [2][2 pos][64 ch][64 sq]                                     = B operand of pw2
//                ring of 6 x 16 KB weight tiles [128 rows x 64 K] K-major        = A operands
// TMEM (512 columns): D2 = 3 x 128 columns, D1 = 128 columns.
// Warps (320 threads): 0 = TMA producer, 1 = MMA issuer, 2-9 = compute (lane quarter = warp & 3,
// position = (warp - 2) >> 2).
#pragma once
#include <cuda.h>
#include "hm_ptx.cuh"

namespace hm {

constexpr int TR_MAX_BLOCKS = 16;
constexpr int TR_PAR_FLOATS = 27 * 128;  // per 128-channel chunk: b1 | dw bias | 25 taps, each [128]

struct TrunkBlock {
    int chunk0;    // first global chunk (row block of W1cat / W2cat, entry of par)
    int n_chunks;  // ceil(cop / 128)
    int ksize;     // 3 or 5
    int bias2_off; // offset into bias2 (floats)
    int eca_chunk0;   // ECA on the block input: first of 3 row chunks of Wc in W1cat, or -1
    int eca_bias_off; // offset of the 384 ECA biases in bias2
};
struct TrunkParams {
    int n_blocks;
    TrunkBlock blk[TR_MAX_BLOCKS];
    const float* par;    // [chunk][TR_PAR_FLOATS]
    const float* bias2;  // [...][384]
    unsigned long long* trace;  // bring-up aid: CTA 0 logs (tag << 48 | clock) per role, or nullptr
    int dbg;                    // bring-up aid (timing experiments only, results are wrong when set):
                                // 1 skip depthwise math, 2 skip TMEM drain, 4 skip Z store, 8 / 16 skip the
                                // math only on sub-partitions 0,1 / 2,3, 64 scalar-FFMA depthwise (results
                                // stay right), 128 weight ring + MMAs only (compute warps idle)
};

// timeline tracing (tools/trace_trunk.py): roles 0 = producer, 1 = MMA issuer, 2 = compute warp 2
constexpr int TR_TRACE_CAP = 2048;
#define TR_MARK(role, tag)                                                                       \
    do {                                                                                         \
        if (tr_trace != nullptr && trace_n < TR_TRACE_CAP)                                       \
            tr_trace[(role) * TR_TRACE_CAP + trace_n++] =                                        \
                (static_cast<unsigned long long>(tag) << 48) | (clock64() & 0xffffffffffffull);  \
    } while (0)

constexpr int TR_THREADS = 320;
constexpr int TR_COMPUTE = 256;
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
    TB_ECA_DONE = TB_TILE_EMPTY + TR_SLOTS,   // 3: gate pre-activations of output block m are in TMEM
    TB_ECA_READY = TB_ECA_DONE + 3,           // the channel means of the block input are in shared memory
    TB_COUNT = TB_ECA_READY + 1
};

// k-th completion (k = 0, 1, ...) of a "full"-type barrier
__device__ __forceinline__ void tr_wait_full(uint64_t* bar, uint32_t k) { mbar_wait(bar, k & 1); }
// k-th reuse of a resource: free for k = 0, then needs the (k-1)-th completion of its "empty" barrier
__device__ __forceinline__ void tr_wait_empty(uint64_t* bar, uint32_t k) { mbar_wait(bar, (k & 1) ^ 1); }

// depthwise k x k on one channel of one board held in registers: out[s] = relu(bd + sum w*y)
template <int KS>
__device__ __forceinline__ void dw_board(const float (&y)[64], const float (&w)[KS * KS], float bd,
                                         uint32_t (&packed)[32]) {
    constexpr int H = KS / 2;
#pragma unroll
    for (int yo = 0; yo < 8; ++yo) {
        float o[8];
#pragma unroll
        for (int xo = 0; xo < 8; ++xo) {
            float acc = bd;
#pragma unroll
            for (int ky = 0; ky < KS; ++ky) {
                const int yy = yo + ky - H;
                if (yy < 0 || yy > 7) continue;
#pragma unroll
                for (int kx = 0; kx < KS; ++kx) {
                    const int xx = xo + kx - H;
                    if (xx < 0 || xx > 7) continue;
                    acc = fmaf(y[yy * 8 + xx], w[ky * KS + kx], acc);
                }
            }
            o[xo] = acc;
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) packed[yo * 4 + j] = pack_act2_relu(o[2 * j], o[2 * j + 1]);
    }
}

// The same with packed fp32 FMAs (FFMA2, fma.rn.f32x2): two horizontally adjacent squares per
// instruction, identical fp32 products.  Every operand is an ALIGNED input pair
// A[a] = (in[2a], in[2a+1]) (no register shuffling): taps with an even horizontal offset
// accumulate into the aligned output pairs E[j] = (o[2j], o[2j+1]), taps with an odd offset into
// the odd-aligned pairs O[k] = (o[2k-1], o[2k]); one scalar add per square joins them.  This
// cuts the issue slots of the depthwise by ~1/3 (3x3) to ~40 % (5x5) — the FMA-pipe time is the
// same — which is what the TMA-producer and MMA-issuer warps on the same sub-partitions compete
// for with the compute warps.
template <int KS>
__device__ __forceinline__ void dw_board2(const float (&y)[64], const float (&w)[KS * KS], float bd,
                                          uint32_t (&packed)[32]) {
    constexpr int H = KS / 2;
#pragma unroll
    for (int yo = 0; yo < 8; ++yo) {
        float2 E[4], O[5];
#pragma unroll
        for (int j = 0; j < 4; ++j) E[j] = make_float2(bd, bd);
#pragma unroll
        for (int j = 0; j < 5; ++j) O[j] = make_float2(0.0f, 0.0f);
#pragma unroll
        for (int ky = 0; ky < KS; ++ky) {
            const int yy = yo + ky - H;
            if (yy < 0 || yy > 7) continue;
#pragma unroll
            for (int kx = 0; kx < KS; ++kx) {
                const int dx = kx - H;
                const float2 ww = make_float2(w[ky * KS + kx], w[ky * KS + kx]);
#pragma unroll
                for (int a = 0; a < 4; ++a) {
                    const float2 in = make_float2(y[yy * 8 + 2 * a], y[yy * 8 + 2 * a + 1]);
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
        for (int j = 0; j < 4; ++j)
            packed[yo * 4 + j] = pack_act2_relu(E[j].x + O[j].y, E[j].y + O[j + 1].x);
    }
}

// Per-thread parameters of one 128-channel chunk: pw1 bias, depthwise bias, k*k depthwise taps.
template <int KS>
struct ChunkPar {
    float w[KS * KS];
    float b1, bd;
    __device__ __forceinline__ void load(const float* __restrict__ par, int r) {
#pragma unroll
        for (int t = 0; t < KS * KS; ++t) w[t] = __ldg(par + 256 + t * 128 + r);
        b1 = __ldg(par + r);
        bd = __ldg(par + 128 + r);
    }
};

// One chunk on the compute warps.  `cur` holds this chunk's parameters (already in registers);
// the next chunk's parameters are requested while this chunk's depthwise math runs, so their
// L2 latency never sits between "D1 is full" and the first use.
template <int KS>
__device__ __forceinline__ void trunk_chunk(ChunkPar<KS>& cur, const float* __restrict__ par_next,
                                            int r, uint32_t tm_d1, uint64_t* bars, uint32_t d1_n,
                                            uint32_t z_n, uint8_t* zrow, int lane,
                                            unsigned long long* tr_trace, int& trace_n, int dbg,
                                            long long* prof) {
    float y[64];
    TR_MARK(2, 1);
    tr_wait_full(&bars[TB_D1_FULL], d1_n);
    tc_fence_after();
    TR_MARK(2, 2);
    if (dbg & 2) {
#pragma unroll
        for (int j = 0; j < 64; ++j) y[j] = cur.b1;
    } else {
        uint32_t a0[32], a1[32];
        tmem_ld_32x32b_x32(tm_d1, a0);
        tmem_ld_wait();
        tmem_ld_32x32b_x32(tm_d1 + 32, a1);   // streams out while the first half is prepared
        const float2 bb = make_float2(cur.b1, cur.b1);
#pragma unroll
        for (int j = 0; j < 16; ++j) {   // packed bias add (FADD2), scalar max
            const float2 u = __fadd2_rn(make_float2(__uint_as_float(a0[2 * j]), __uint_as_float(a0[2 * j + 1])), bb);
            y[2 * j] = fmaxf(u.x, 0.0f);
            y[2 * j + 1] = fmaxf(u.y, 0.0f);
        }
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 16; ++j) {
            const float2 v = __fadd2_rn(make_float2(__uint_as_float(a1[2 * j]), __uint_as_float(a1[2 * j + 1])), bb);
            y[32 + 2 * j] = fmaxf(v.x, 0.0f);
            y[32 + 2 * j + 1] = fmaxf(v.y, 0.0f);
        }
    }
    tc_fence_before();
    __syncwarp();
    if (lane == 0) mbar_arrive(&bars[TB_D1_EMPTY]);  // D1 is free as soon as it is in registers
    TR_MARK(2, 3);
    ChunkPar<KS> nxt;
    if (par_next != nullptr) nxt.load(par_next, r);
    uint32_t packed[32];
    if ((dbg & 1) || ((dbg & 8) && ((threadIdx.x >> 5) & 3) < 2) || ((dbg & 16) && ((threadIdx.x >> 5) & 3) >= 2)) {
#pragma unroll
        for (int j = 0; j < 32; ++j) packed[j] = __float_as_uint(y[2 * j]);
    } else {
        const long long t0 = prof ? clock64() : 0;
        if (dbg & 64) dw_board<KS>(y, cur.w, cur.bd, packed);   // scalar FFMA variant kept for A/B timing
        else dw_board2<KS>(y, cur.w, cur.bd, packed);
        if (prof) *prof += clock64() - t0;
    }
    TR_MARK(2, 4);
    tr_wait_empty(&bars[TB_Z_EMPTY], z_n);  // pw2 of the previous chunk has consumed Z
    TR_MARK(2, 5);
    const int key = r & 7;
    if (!(dbg & 4)) {
#pragma unroll
        for (int j = 0; j < 8; ++j)
            *reinterpret_cast<uint4*>(zrow + ((j ^ key) << 4)) =
                make_uint4(packed[4 * j], packed[4 * j + 1], packed[4 * j + 2], packed[4 * j + 3]);
        fence_proxy_async_smem();
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(&bars[TB_Z_FULL]);
    TR_MARK(2, 6);
    if (par_next != nullptr) cur = nxt;
}

// all chunks of one bottleneck block
template <int KS>
__device__ __forceinline__ void trunk_block_chunks(const float* __restrict__ par0, int nch, int r,
                                                   uint32_t tm_d1, uint64_t* bars, uint32_t& d1_n,
                                                   uint32_t& z_n, uint8_t* zrow, int lane,
                                                   unsigned long long* tr_trace, int& trace_n, int dbg,
                                                   long long* prof) {
    ChunkPar<KS> cur;
    cur.load(par0, r);
    for (int c = 0; c < nch; ++c) {
        const float* nxt = (c + 1 < nch) ? par0 + static_cast<size_t>(c + 1) * TR_PAR_FLOATS : nullptr;
        trunk_chunk<KS>(cur, nxt, r, tm_d1, bars, d1_n, z_n, zrow, lane, tr_trace, trace_n, dbg, prof);
        ++d1_n;
        ++z_n;
    }
}

// effective-clock probe: CTA 0 and the last CTA record (SM cycles, nanoseconds) of their lifetime
// (debug view "trunk_clk"; tools/trunk_clock.py)
__device__ float g_trunk_clk[16];   // [0..3] lifetimes; [4..] CTA 0 phase marks (cycles since kernel entry)
__device__ __forceinline__ unsigned long long global_timer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
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
    unsigned long long* const tr_trace = (blockIdx.x == 0 && (lane == 0)) ? p.trace : nullptr;
    const long long clk_base = clock64();
#define TR_CLK(i) do { if (blockIdx.x == 0 && lane == 0) g_trunk_clk[4 + (i)] = static_cast<float>(clock64() - clk_base); } while (0)
    const bool clk_probe = threadIdx.x == 0 && (blockIdx.x == 0 || blockIdx.x == gridDim.x - 1);
    const long long clk_c0 = clk_probe ? clock64() : 0;
    const unsigned long long clk_t0 = clk_probe ? global_timer_ns() : 0;

    if (threadIdx.x == 0) {
        tma_prefetch_desc(&tmX);
        tma_prefetch_desc(&tmW1);
        tma_prefetch_desc(&tmW2);
        for (int i = 0; i < TB_COUNT; ++i) {
            uint32_t count = 1;
            if ((i >= TB_X_READY && i < TB_X_READY + 3) || i == TB_D1_EMPTY || i == TB_Z_FULL || i == TB_ECA_READY)
                count = 8;
            mbar_init(&bars[i], count);
        }
        fence_barrier_init();
    }
    if (warp == 1) {
        tmem_alloc(tmem_slot, 512);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    if (warp == 0) TR_CLK(0);   // prologue done
    const uint32_t tm_d2 = tmem_base;
    const uint32_t tm_d1 = tmem_base + 384u;

    // The two single-thread roles run with ALL 32 lanes in uniform control flow (waits, loop
    // counters, descriptor arithmetic stay warp-uniform -> uniform registers, no per-lane
    // election loops around every UTMALDG / UTCHMMA); only the issuing instructions themselves
    // are predicated on one elected lane.
    if (warp == 0) {
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
        for (int b = 0; b < ((p.dbg & 256) ? 0 : p.n_blocks); ++b) {
            const int g0 = p.blk[b].chunk0, nch = p.blk[b].n_chunks;
            if (b > 0 && p.blk[b].eca_chunk0 >= 0)   // in-kernel ECA: Wc streams through the ring like W1
                for (int m = 0; m < 3; ++m) group_w1(p.blk[b].eca_chunk0 + m);
            group_w1(g0);
            for (int c = 0; c < nch; ++c) {
                if (c + 1 < nch) group_w1(g0 + c + 1);
                group_w2(g0 + c);
            }
        }
    } else if (warp == 1) {
        // ================================ MMA issuer ================================
        const bool leader = elect_one();
        constexpr uint32_t idesc = umma_idesc_16_bmn(128, 128, UMMA_FMT_16);
        uint32_t d1_n = 0, z_n = 0, par = 0;
        int trace_n = 0;
        // descriptors differ only in the start-address field (16-byte units): hoist the bases
        const uint64_t a0d = umma_desc_sw128(smem_u32(smem + TR_RING_OFF));
        const uint64_t xd0d = umma_desc_mn_sw128(smem_u32(smem + TR_X_OFF), 8192, 1024);
        const uint64_t zd0d = umma_desc_mn_sw128(smem_u32(smem + TR_Z_OFF), 8192, 1024);
        const uint32_t a0 = static_cast<uint32_t>(a0d), a_hi = static_cast<uint32_t>(a0d >> 32);
        const uint32_t xd0 = static_cast<uint32_t>(xd0d), zd0 = static_cast<uint32_t>(zd0d);
        const uint32_t b_hi = static_cast<uint32_t>(xd0d >> 32);
        const uint64_t zmd = umma_desc_sw128(smem_u32(smem + TR_Z_OFF));   // ECA means: K-major, 16 rows per K block
        const uint32_t zm0 = static_cast<uint32_t>(zmd), zm_hi = static_cast<uint32_t>(zmd >> 32);
        auto pw1 = [&](int xblock) {   // xblock >= 0: first pw1 of bottleneck block `xblock`
            TR_MARK(1, 1);
            tr_wait_empty(&bars[TB_D1_EMPTY], d1_n);
            tc_fence_after();
            TR_MARK(1, 2);
#pragma unroll
            for (int b2 = 0; b2 < TR_PAIRS; ++b2) {
                if (xblock >= 0) tr_wait_full(&bars[TB_X_READY + b2], xblock);  // X blocks 2*b2, 2*b2+1 ready
                if (!(p.dbg & 256)) mbar_wait(&bars[TB_PAIR_FULL + b2], par);
                tc_fence_after();
                TR_MARK(1, 10 + b2);
                if (leader) {
#pragma unroll
                    for (int j = 0; j < 2; ++j) {
                        const int kb = 2 * b2 + j;  // ring slot = K block of X
#pragma unroll
                        for (int k = 0; k < 4; ++k)
                            umma_f16_lh(tm_d1, a0 + (kb * 1024 + 2 * k), a_hi, xd0 + (kb * 1024 + k * 128), b_hi,
                                        idesc, (kb | k) ? 1u : 0u);
                        umma_commit(&bars[TB_TILE_EMPTY + kb]);
                    }
                    if (b2 == TR_PAIRS - 1) umma_commit(&bars[TB_D1_FULL]);
                }
                __syncwarp();
            }
            par ^= 1;
            ++d1_n;
            TR_MARK(1, 3);
        };
        uint32_t eca_n = 0;
        TR_CLK(1);   // issuer enters its loop
        for (int b = 0; b < p.n_blocks; ++b) {
            if (b == 1) TR_CLK(6);   // block 0 issued
            const int nch = p.blk[b].n_chunks;
            if (b > 0 && p.blk[b].eca_chunk0 >= 0) {
                // ECA matvec on the tensor core: G[m] (128 out channels x N = 16, columns 0/1 = the
                // two positions) = Wc tiles (ring) x channel means (fp16, K-major, parked in the Z
                // buffer by the compute warps); accumulators in the idle D2 columns
                constexpr uint32_t idesc_eca = umma_idesc_16(128, 16, UMMA_FMT_16);
                if (!(p.dbg & 128)) tr_wait_full(&bars[TB_ECA_READY], eca_n);
                tc_fence_after();
                for (int m = 0; m < 3; ++m) {
#pragma unroll
                    for (int b2 = 0; b2 < TR_PAIRS; ++b2) {
                        if (!(p.dbg & 256)) mbar_wait(&bars[TB_PAIR_FULL + b2], par);
                        tc_fence_after();
                        if (leader) {
#pragma unroll
                            for (int j = 0; j < 2; ++j) {
                                const int kb = 2 * b2 + j;
#pragma unroll
                                for (int k = 0; k < 4; ++k)
                                    umma_f16_lh(tm_d2 + m * 16, a0 + (kb * 1024 + 2 * k), a_hi, zm0 + (kb * 128 + 2 * k), zm_hi,
                                                idesc_eca, (kb | k) ? 1u : 0u);
                                umma_commit(&bars[TB_TILE_EMPTY + kb]);
                            }
                            if (b2 == TR_PAIRS - 1) umma_commit(&bars[TB_ECA_DONE + m]);
                        }
                        __syncwarp();
                    }
                    par ^= 1;
                }
                ++eca_n;
            }
            pw1(b);
            for (int c = 0; c < nch; ++c) {
                if (c + 1 < nch) pw1(-1);
                TR_MARK(1, 4);
                tr_wait_full(&bars[TB_Z_FULL], z_n);
                tc_fence_after();
                TR_MARK(1, 5);
                const uint32_t acc0 = c > 0 ? 1u : 0u;
                const bool last_chunk = c + 1 == nch;
#pragma unroll
                for (int m = 0; m < TR_PAIRS; ++m) {   // pair = output channel block m
                    if (!(p.dbg & 256)) mbar_wait(&bars[TB_PAIR_FULL + m], par);
                    tc_fence_after();
                    TR_MARK(1, 20 + m);
                    if (leader) {
#pragma unroll
                        for (int kh = 0; kh < 2; ++kh) {
#pragma unroll
                            for (int k = 0; k < 4; ++k)
                                umma_f16_lh(tm_d2 + m * 128, a0 + ((2 * m + kh) * 1024 + 2 * k), a_hi,
                                            zd0 + (kh * 1024 + k * 128), b_hi, idesc, (kh | k) ? 1u : acc0);
                            umma_commit(&bars[TB_TILE_EMPTY + 2 * m + kh]);
                        }
                        if (last_chunk) umma_commit(&bars[TB_D2_FULL + m]);  // output block m complete
                        if (m == TR_PAIRS - 1) umma_commit(&bars[TB_Z_EMPTY]);
                    }
                    __syncwarp();
                }
                par ^= 1;
                ++z_n;
                TR_MARK(1, 6);
            }
        }
        TR_CLK(2);   // last MMA issued
    } else {
        // ================================ compute warps ================================
        const int q = warp & 3;             // TMEM lane quarter this warp may access
        const int h = (warp - 2) >> 2;      // position of the tile handled by this warp
        const int r = q * 32 + lane;        // channel within a 128-channel chunk / M block
        const uint32_t lane_off = static_cast<uint32_t>(q * 32) << 16;
        const int key = r & 7;
        // this thread's 128-byte row inside a [2 pos][64 ch][64 sq] block
        const uint32_t row_off = (r >> 6) * 16384 + h * 8192 + (r & 63) * 128;

        tr_wait_full(&bars[TB_X_FULL], 0);   // the tile's block input has landed
        if (warp == 2) TR_CLK(5);   // X landed
        if (lane == 0) {
            mbar_arrive(&bars[TB_X_READY + 0]);
            mbar_arrive(&bars[TB_X_READY + 1]);
            mbar_arrive(&bars[TB_X_READY + 2]);
        }

        uint32_t d1_n = 0, z_n = 0, eca_cnt = 0;
        int trace_n = 0;
        unsigned long long* const ctrace = warp == 2 ? tr_trace : nullptr;
        long long math_cycles = 0;   // per-warp total of the depthwise math (tracing only)
        long long* const prof = (blockIdx.x == 0 && p.trace != nullptr) ? &math_cycles : nullptr;
        for (int b = 0; b < p.n_blocks; ++b) {
            const int g0 = p.blk[b].chunk0, nch = p.blk[b].n_chunks, ks = p.blk[b].ksize;
            if (b > 0 && p.blk[b].eca_chunk0 >= 0) {
                // ---- ECA on the block input (builder_util.py:49-80, 481-483), inside the kernel:
                // gate[o] = hardsigmoid(Wc[o,:] . mean + b[o]).  The matvec runs on the tensor core
                // (MMA warp: Wc tiles from the weight ring x the fp16 means this block's predecessor
                // parked in the Z buffer, N = 16 with columns 0/1 = the two positions); here the
                // pre-activation of this thread's channel comes back from TMEM and scales its X row.
                const float* eb = p.bias2 + p.blk[b].eca_bias_off;
#pragma unroll 1
                for (int m = 0; m < 3; ++m) {
                    const float ebias = __ldg(eb + m * 128 + r);
                    { unsigned long long* tr_trace = ctrace; TR_MARK(2, 30); }
                    tr_wait_full(&bars[TB_ECA_DONE + m], eca_cnt);
                    tc_fence_after();
                    { unsigned long long* tr_trace = ctrace; TR_MARK(2, 31 + m); }
                    uint32_t g0, g1;
                    tmem_ld_32x32b_x2(tm_d2 + lane_off + m * 16, g0, g1);
                    tmem_ld_wait();
                    tc_fence_before();
                    const float acc = __uint_as_float(h == 0 ? g0 : g1);
                    // torch.nn.Hardsigmoid: relu6(x + 3) / 6
                    const float g = fminf(fmaxf((acc + ebias) * (1.0f / 6.0f) + 0.5f, 0.0f), 1.0f);
                    uint8_t* xrow = smem + TR_X_OFF + m * 32768 + row_off;
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        uint4* px = reinterpret_cast<uint4*>(xrow + ((j ^ key) << 4));
                        uint4 v = *px;
                        float2 a = unpack_act2(v.x), bb = unpack_act2(v.y), c = unpack_act2(v.z), d = unpack_act2(v.w);
                        v.x = pack_act2(a.x * g, a.y * g);
                        v.y = pack_act2(bb.x * g, bb.y * g);
                        v.z = pack_act2(c.x * g, c.y * g);
                        v.w = pack_act2(d.x * g, d.y * g);
                        *px = v;
                    }
                    fence_proxy_async_smem();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&bars[TB_X_READY + m]);
                    { unsigned long long* tr_trace = ctrace; TR_MARK(2, 35); }
                }
                ++eca_cnt;
            }
            {
                const float* par0 = p.par + static_cast<size_t>(g0) * TR_PAR_FLOATS;
                uint8_t* zrow = smem + TR_Z_OFF + row_off;
                if (ks == 3)
                    trunk_block_chunks<3>(par0, nch, r, tm_d1 + lane_off + h * 64, bars, d1_n, z_n, zrow, lane, ctrace, trace_n, p.dbg, prof);
                else
                    trunk_block_chunks<5>(par0, nch, r, tm_d1 + lane_off + h * 64, bars, d1_n, z_n, zrow, lane, ctrace, trace_n, p.dbg, prof);
            }
            // ---- block epilogue: x <- x + D2 + b2, in place in shared memory ----
            const bool last = (b + 1 == p.n_blocks);
            const float* b2 = p.bias2 + p.blk[b].bias2_off;
            const bool want_means = !last && p.blk[b + 1].eca_chunk0 >= 0;  // next block gates its input
            float sums3[3];
#pragma unroll
            for (int m = 0; m < 3; ++m) {
                const float bias = __ldg(b2 + m * 128 + r);
                uint8_t* xrow = smem + TR_X_OFF + m * 32768 + row_off;
                uint4 res[8];   // residual row: independent of D2, fetched before the wait
#pragma unroll
                for (int j = 0; j < 8; ++j) res[j] = *reinterpret_cast<const uint4*>(xrow + ((j ^ key) << 4));
                { unsigned long long* tr_trace = ctrace; TR_MARK(2, 7); }
                tr_wait_full(&bars[TB_D2_FULL + m], b);
                tc_fence_after();
                { unsigned long long* tr_trace = ctrace; TR_MARK(2, 8); }
                // the second half of the accumulator row streams out of TMEM while the first is processed
                uint32_t a0[32], a1[32];
                tmem_ld_32x32b_x32(tm_d2 + lane_off + m * 128 + h * 64, a0);
                tmem_ld_wait();
                tmem_ld_32x32b_x32(tm_d2 + lane_off + m * 128 + h * 64 + 32, a1);
                const float2 bb = make_float2(bias, bias);
                float s = 0.0f, s1 = 0.0f;
                auto piece = [&](int j, const uint32_t (&a)[32], int base) {
                    const uint4 v = res[j];
                    const uint32_t ru[4] = {v.x, v.y, v.z, v.w};
                    uint32_t ou[4];
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        const int i = j * 8 + e * 2 - base;
                        const float2 acc = __fadd2_rn(make_float2(__uint_as_float(a[i]), __uint_as_float(a[i + 1])), bb);
                        const float2 o = __fadd2_rn(acc, unpack_act2(ru[e]));
                        ou[e] = pack_act2(o.x, o.y);
                        if (want_means) {  // ECA squeeze of the next block: sum of what was actually stored
                            const float2 rd = unpack_act2(ou[e]);
                            s += rd.x;
                            s1 += rd.y;
                        }
                    }
                    *reinterpret_cast<uint4*>(xrow + ((j ^ key) << 4)) = make_uint4(ou[0], ou[1], ou[2], ou[3]);
                };
#pragma unroll
                for (int j = 0; j < 4; ++j) piece(j, a0, 0);
                tmem_ld_wait();
#pragma unroll
                for (int j = 4; j < 8; ++j) piece(j, a1, 32);
                s += s1;
                sums3[m] = s;
                tc_fence_before();
                fence_proxy_async_smem();
                __syncwarp();
                if (!last && !want_means && lane == 0) mbar_arrive(&bars[TB_X_READY + m]);
                { unsigned long long* tr_trace = ctrace; TR_MARK(2, 9); }
            }
            if (want_means) {
                // every pw2 of this block has completed (D2_FULL[2]), so Z is idle: park the channel means
                // there as the K-major fp16 B operand of the ECA matvec: per 64-channel K block 16 rows
                // of 128 bytes (row 0 / 1 = position 0 / 1, rows 2..15 zero), 128-byte swizzle
                uint8_t* zm = smem + TR_Z_OFF;
#pragma unroll
                for (int m = 0; m < 3; ++m) {
                    const int ch = m * 128 + r, kb = ch >> 6, k = ch & 63;
                    *reinterpret_cast<act_t*>(zm + kb * 2048 + h * 128 + (((k >> 3) ^ h) << 4) + (k & 7) * 2) =
                        float_to_act(sums3[m] * (1.0f / 64.0f));
                }
                for (int i = threadIdx.x - 64; i < 6 * 14 * 8; i += TR_COMPUTE) {
                    const int kb = i / 112, rem = i - kb * 112;
                    *reinterpret_cast<uint4*>(zm + kb * 2048 + (2 + (rem >> 3)) * 128 + ((rem & 7) << 4)) = make_uint4(0, 0, 0, 0);
                }
                fence_proxy_async_smem();
                __syncwarp();
                if (lane == 0) mbar_arrive(&bars[TB_ECA_READY]);
            }
        }
        if (prof != nullptr && lane == 0)   // tail of role 2's region: tag 100 + warp
            p.trace[3 * TR_TRACE_CAP - 16 + warp] =
                (static_cast<unsigned long long>(100 + warp) << 48) | (static_cast<unsigned long long>(math_cycles) & 0xffffffffffffull);
        // ---- run output: shared-memory tile -> global (TMA store) ----
        if (warp == 2) TR_CLK(3);   // last block epilogue done
        named_bar_sync(1, TR_COMPUTE);
        if (threadIdx.x == 64) {
            for (int kb = 0; kb < 6; ++kb) tma_store_3d(&tmX, smem + TR_X_OFF + kb * 16384, 0, kb * 64, tile * 2);
            bulk_commit_group();
            bulk_wait_group0();
            TR_CLK(4);   // output stored
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
    if (clk_probe) {
        const int o = blockIdx.x == 0 ? 0 : 2;
        g_trunk_clk[o] = static_cast<float>(clock64() - clk_c0);
        g_trunk_clk[o + 1] = static_cast<float>(global_timer_ns() - clk_t0);
    }
}

}  // namespace hm
