// hm_pgemm.cuh — tcgen05 + TMA implicit-GEMM 3x3 convolution for NARROW outputs (the policy
// projections: 384 -> 2 x 73 logits maps, rise_mobile_v3.py:38-57; per-board 192 -> 73,
// rise_mobile_v3.py:336-347).  Square-major formulation: the 128 board squares of two positions are
// the MMA M dimension, the output channels the N dimension,
//
//   D[(pos, sq), o] = sum_{tap, c} X[pos, c, sq + shift(tap)] * W[o, tap*Cpad + c]
//
// so 146 (73) output channels cost ONE M128 x N160 (N80) MMA per K step instead of the two (one)
// mostly empty 128-row M blocks the channel-major kernel (hm_cgemm.cuh) spends on them.
//
// A = activation block [2 positions][64 channels][64 squares], MN-major (squares contiguous), the
//     same 3-D TMA box hm_cgemm.cuh uses as its B operand; vertical tap = square shift by 8*dy with
//     TMA zero fill, horizontal tap resolved in the epilogue (see hm_cgemm.cuh).
// B = W tile [Npad output channels x 64 K], K-major, 128-byte swizzle (2-D TMA box of Npad rows).
// D = three accumulators P_dx (128 square lanes x Npad fp32 columns each, 3*Npad <= 480 columns).
//     TMEM lane = square, so the horizontal shift out[y][x] = P_-1[y][x-1] + P_0[y][x] + P_+1[y][x+1]
//     is a warp shuffle by one lane (a warp holds four whole ranks), and for one output channel the 32
//     lanes of a warp store 32 consecutive squares: coalesced 128-byte (fp32) rows of the NCHW output.
//
// CTA = 2 positions, all output channels; 192 threads: warp 0 weight producer, warp 1 MMA issuer + TMEM
// owner, warps 2-5 epilogue (warp 2 first loads the activation stages).
#pragma once
#include <cuda.h>
#include "hm_ptx.cuh"

namespace hm {

struct PgemmParams {
    int n_valid;        // cout: channels actually stored
    int npad;           // cout rounded up to 16 (MMA N), <= 160
    int kblocks;        // Cpad / 64 per tap (even)
    int ksteps_last;    // valid 16-channel K steps in the last 128-channel K pair
    const float* bias;  // [cout] or nullptr
    void* out;          // [pos][out_channels][64], act_t or fp32
    int out_channels;   // channel count of the output tensor
    int relu;
    int out_f32;
};

constexpr int PG_THREADS = 192;
constexpr int PG_WSTAGES = 4;
constexpr int PG_XSTAGES = 2;
constexpr int PG_XSTAGE_BYTES = 32768;
constexpr int pg_wstage_bytes(int npad) { return 2 * npad * 128; }   // two K blocks of [npad rows x 128 B]
constexpr int pg_smem_bytes(int npad) { return PG_WSTAGES * pg_wstage_bytes(npad) + PG_XSTAGES * PG_XSTAGE_BYTES + 256 + 1024; }

__global__ void __launch_bounds__(PG_THREADS)
pgemm_kernel(const PgemmParams p, const __grid_constant__ CUtensorMap tmW, const __grid_constant__ CUtensorMap tmX) {
    extern __shared__ uint8_t pg_smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(pg_smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
    const int wbytes = 2 * p.npad * 128;
    uint8_t* smem_x = smem + PG_WSTAGES * wbytes;
    uint64_t* full_w = reinterpret_cast<uint64_t*>(smem_x + PG_XSTAGES * PG_XSTAGE_BYTES);
    uint64_t* empty_w = full_w + PG_WSTAGES;
    uint64_t* full_x = empty_w + PG_WSTAGES;
    uint64_t* empty_x = full_x + PG_XSTAGES;
    uint64_t* accum_bar = empty_x + PG_XSTAGES;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(accum_bar + 1);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int tile = blockIdx.x;

    if (threadIdx.x == 0) {
        tma_prefetch_desc(&tmW);
        tma_prefetch_desc(&tmX);
        for (int s = 0; s < PG_WSTAGES; ++s) { mbar_init(&full_w[s], 1); mbar_init(&empty_w[s], 1); }
        for (int s = 0; s < PG_XSTAGES; ++s) { mbar_init(&full_x[s], 1); mbar_init(&empty_x[s], 1); }
        mbar_init(accum_bar, 1);
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
    const int kpairs = p.kblocks >> 1;

    if (warp == 0) {
        // weight producer (the activation stages are loaded by warp 2 before it turns to the epilogue: one
        // thread issuing all eight 16-20 KB boxes of a group needs ~2300 cycles against 1920 cycles of MMAs)
        const bool leader = elect_one();
        int ws = 0;
        uint32_t wphase = 1;
        const int tap_cols = p.kblocks * 64;
        for (int ky = 0; ky < 3; ++ky) {
            for (int kp = 0; kp < kpairs; ++kp) {
                const bool upper = kp + 1 < kpairs || p.ksteps_last > 4;
                for (int kx = 0; kx < 3; ++kx) {
                    const int kcol = (ky * 3 + kx) * tap_cols + kp * 128;
                    mbar_wait(&empty_w[ws], wphase);
                    if (leader) {
                        uint8_t* dst = smem + ws * wbytes;
                        mbar_arrive_expect_tx(&full_w[ws], upper ? wbytes : wbytes / 2);
                        tma_load_2d(dst, &tmW, &full_w[ws], kcol, 0);
                        if (upper) tma_load_2d(dst + (wbytes >> 1), &tmW, &full_w[ws], kcol + 64, 0);
                    }
                    __syncwarp();
                    if (++ws == PG_WSTAGES) { ws = 0; wphase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // issue discipline as in hm_trunk.cuh / hm_cgemm.cuh: fused 8-MMA bursts, next weight barrier probed early
        const bool leader = elect_one();
        const uint32_t ld = leader ? 1u : 0u;
        // A = activations MN-major (bit 15), B = weights K-major
        const uint32_t idesc = umma_idesc_16(128, static_cast<uint32_t>(p.npad), UMMA_FMT_16) | (1u << 15);
        const uint64_t w0 = umma_desc_sw128(smem_u32(smem));
        const uint64_t x0 = umma_desc_mn_sw128(smem_u32(smem_x), 8192, 1024);
        const uint32_t w_lo = static_cast<uint32_t>(w0), x_lo = static_cast<uint32_t>(x0);
        const uint32_t wunits = static_cast<uint32_t>(wbytes) >> 4;   // stage size in 16-byte descriptor units
        int ws = 0, xs = 0;
        uint32_t wphase = 0, xphase = 0, have_w = 0;
        for (int ky = 0; ky < 3; ++ky) {
            for (int kp = 0; kp < kpairs; ++kp) {
                mbar_wait(&full_x[xs], xphase);
                const uint32_t xo = xs * (PG_XSTAGE_BYTES / 16);
                const int nk = kp + 1 < kpairs ? 8 : p.ksteps_last;
                for (int kx = 0; kx < 3; ++kx) {
                    const uint32_t d = tmem_base + kx * p.npad;
                    if (!have_w) mbar_wait(&full_w[ws], wphase);
                    tc_fence_after();
                    const int nws = ws + 1 == PG_WSTAGES ? 0 : ws + 1;
                    const uint32_t nphase = ws + 1 == PG_WSTAGES ? wphase ^ 1 : wphase;
                    const uint32_t next_ok = mbar_test_wait(&full_w[nws], nphase);
                    const uint32_t wo = ws * wunits;
                    const uint32_t acc0 = (ky | kp) ? 1u : 0u;
                    if (nk == 8 && p.npad == 160) {        // second K block of a stage: 160 rows x 128 B = 1280 units on
                        if (leader) umma_f16_burst8<1024, 128, 1280, 2, umma_idesc_16(128, 160, UMMA_FMT_16) | (1u << 15), false>(d, x_lo + xo, w_lo + wo, acc0, &empty_w[ws], &empty_w[ws], ld);
                    } else if (nk == 8 && p.npad == 80) {
                        if (leader) umma_f16_burst8<1024, 128, 640, 2, umma_idesc_16(128, 80, UMMA_FMT_16) | (1u << 15), false>(d, x_lo + xo, w_lo + wo, acc0, &empty_w[ws], &empty_w[ws], ld);
                    } else if (leader) {
#pragma unroll
                        for (int j = 0; j < 2; ++j) {
#pragma unroll
                            for (int k = 0; k < 4; ++k)
                                if (j * 4 + k < nk)
                                    umma_f16(d, x0 + (xo + j * 1024 + k * 128), w0 + (wo + j * (wunits >> 1) + 2 * k), idesc,
                                             (j | k) ? 1u : acc0);
                        }
                        umma_commit(&empty_w[ws]);
                    }
                    if (kx == 2) umma_commit_if(&empty_x[xs], ld);
                    __syncwarp();
                    have_w = next_ok;
                    ws = nws;
                    wphase = nphase;
                }
                if (++xs == PG_XSTAGES) { xs = 0; xphase ^= 1; }
            }
        }
        if (leader) umma_commit(accum_bar);
        __syncwarp();
    } else {
        // ===== epilogue: thread = one square of one position, all output channels =====
        const int q = warp & 3;
        const int m = q * 32 + lane;                 // TMEM lane = row of D
        const int pos = tile * 2 + (m >> 6);
        const int sq = m & 63;
        const bool has_left = (sq & 7) != 0, has_right = (sq & 7) != 7;
        if (warp == 2) {   // activation producer
            const bool leader = elect_one();
            int xs = 0;
            uint32_t xphase = 1;
            const int pos0 = tile * 2;
            for (int ky = 0; ky < 3; ++ky) {
                const int sq0 = 8 * (ky - 1);
                for (int kp = 0; kp < kpairs; ++kp) {
                    const bool upper = kp + 1 < kpairs || p.ksteps_last > 4;
                    mbar_wait(&empty_x[xs], xphase);
                    if (leader) {
                        uint8_t* dst = smem_x + xs * PG_XSTAGE_BYTES;
                        mbar_arrive_expect_tx(&full_x[xs], upper ? PG_XSTAGE_BYTES : PG_XSTAGE_BYTES / 2);
                        tma_load_3d(dst, &tmX, &full_x[xs], sq0, kp * 128, pos0);
                        if (upper) tma_load_3d(dst + 16384, &tmX, &full_x[xs], sq0, kp * 128 + 64, pos0);
                    }
                    __syncwarp();
                    if (++xs == PG_XSTAGES) { xs = 0; xphase ^= 1; }
                }
            }
        }
        mbar_wait(accum_bar, 0);
        tc_fence_after();
        const uint32_t t_row = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
        const size_t out_base = static_cast<size_t>(pos) * p.out_channels * 64 + sq;
#pragma unroll 1
        for (int c0 = 0; c0 < p.npad; c0 += 16) {
            uint32_t l[16], c[16], r[16];
            tmem_ld_32x32b_x16(t_row + c0, l);                 // P_-1
            tmem_ld_32x32b_x16(t_row + p.npad + c0, c);        // P_0
            tmem_ld_32x32b_x16(t_row + 2 * p.npad + c0, r);    // P_+1
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 16; ++j) {
                const float fl = __shfl_up_sync(0xffffffffu, __uint_as_float(l[j]), 1);     // square x-1 (same rank)
                const float fr = __shfl_down_sync(0xffffffffu, __uint_as_float(r[j]), 1);   // square x+1
                const int ch = c0 + j;
                if (ch >= p.n_valid) continue;    // warp-uniform
                float v = __uint_as_float(c[j]) + (has_left ? fl : 0.0f) + (has_right ? fr : 0.0f);
                if (p.bias != nullptr) v += __ldg(p.bias + ch);
                if (p.relu == 1) v = fmaxf(v, 0.0f);
                const size_t o = out_base + static_cast<size_t>(ch) * 64;
                if (p.out_f32) static_cast<float*>(p.out)[o] = v;
                else static_cast<act_t*>(p.out)[o] = float_to_act(v);
            }
        }
        tc_fence_before();
    }
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

// ---------------------------------------------------------------------------------------------------------------
// pconv3_kernel<NPAD> — the same convolution driven like cconv3_kernel (hm_cconv.cuh): ONE persistent CTA per SM walks
// the position tiles; the MMA issuer is the trunk's fully unrolled burst loop (ring slots and barriers are compile-time
// offsets of uniform registers; with run-time ring indices a burst of 8 MMAs took ~1.4x its tensor time); weights
// stream through six [NPAD rows x 64 K] tiles with pair-full / tile-empty barriers, activations through three 32 KB
// stages loaded by a warp of their own (it has to keep loading while the epilogue of the previous tile runs);
// sixteen epilogue warps (four per TMEM lane quarter) share the 16-column chunks of the three accumulators.
// Warps: 0 weight producer, 1 MMA issuer + TMEM owner, 2 activation producer, 3 idle, 4-19 epilogue.
// ---------------------------------------------------------------------------------------------------------------
constexpr int PC_EPI_WARPS = 16;
constexpr int PC_THREADS = 128 + PC_EPI_WARPS * 32;
constexpr int PC_XSTAGES = 3;
__host__ __device__ constexpr int pc_tile_bytes(int npad) { return npad * 128; }
__host__ __device__ constexpr int pc_x_off(int npad) { return 6 * pc_tile_bytes(npad); }
__host__ __device__ constexpr int pc_bar_off(int npad) { return pc_x_off(npad) + PC_XSTAGES * PG_XSTAGE_BYTES; }
__host__ __device__ constexpr int pc_smem_bytes(int npad) { return pc_bar_off(npad) + 256 + 1024; }
enum {
    PB_X_FULL = 0, PB_X_EMPTY = PB_X_FULL + PC_XSTAGES, PB_PAIR_FULL = PB_X_EMPTY + PC_XSTAGES,
    PB_TILE_EMPTY = PB_PAIR_FULL + 3, PB_ACCUM = PB_TILE_EMPTY + 6, PB_DRAINED = PB_ACCUM + 1, PB_COUNT = PB_DRAINED + 1
};

template <int NPAD>
__global__ void __launch_bounds__(PC_THREADS, 1)
pconv3_kernel(const PgemmParams p, const int n_tiles, const __grid_constant__ CUtensorMap tmW,
              const __grid_constant__ CUtensorMap tmX) {
    static_assert(NPAD % 16 == 0 && NPAD <= 160, "three accumulators of NPAD columns must fit 512 TMEM columns");
    extern __shared__ uint8_t pc_smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(pc_smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
    constexpr int TILE = pc_tile_bytes(NPAD);
    uint8_t* smem_x = smem + pc_x_off(NPAD);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + pc_bar_off(NPAD));
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + PB_COUNT);
    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        tma_prefetch_desc(&tmW);
        tma_prefetch_desc(&tmX);
        for (int i = 0; i < PB_COUNT; ++i) mbar_init(&bars[i], i == PB_DRAINED ? PC_EPI_WARPS : 1);
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
    const int kpairs = p.kblocks >> 1;

    if (warp == 0) {
        // ===== weight producer =====
        const bool leader = elect_one();
        const int tap_cols = p.kblocks * 64;
        uint32_t wk = 0;
        for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
            int ky = 0, kp = 0;
            for (int step = 0; step < 3 * kpairs; ++step, ++wk) {
                const bool upper = kp + 1 < kpairs || p.ksteps_last > 4;
#pragma unroll
                for (int kx = 0; kx < 3; ++kx) {
                    const int kcol = (ky * 3 + kx) * tap_cols + kp * 128;
                    mbar_wait(&bars[PB_TILE_EMPTY + 2 * kx], (wk & 1) ^ 1);
                    if (leader) {
                        mbar_arrive_expect_tx(&bars[PB_PAIR_FULL + kx], upper ? 2 * TILE : TILE);
                        tma_load_2d(smem + (2 * kx) * TILE, &tmW, &bars[PB_PAIR_FULL + kx], kcol, 0);
                    }
                    __syncwarp();
                    mbar_wait(&bars[PB_TILE_EMPTY + 2 * kx + 1], (wk & 1) ^ 1);
                    if (leader && upper) tma_load_2d(smem + (2 * kx + 1) * TILE, &tmW, &bars[PB_PAIR_FULL + kx], kcol + 64, 0);
                    __syncwarp();
                }
                if (++kp == kpairs) { kp = 0; ++ky; }
            }
        }
    } else if (warp == 2) {
        // ===== activation producer =====
        const bool leader = elect_one();
        uint32_t xk = 0;
        for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
            int ky = 0, kp = 0;
            for (int s3 = 0; s3 < kpairs; ++s3, ++xk) {
#pragma unroll
                for (int u = 0; u < PC_XSTAGES; ++u) {
                    const bool upper = kp + 1 < kpairs || p.ksteps_last > 4;
                    mbar_wait(&bars[PB_X_EMPTY + u], (xk & 1) ^ 1);
                    if (leader) {
                        uint8_t* dst = smem_x + u * PG_XSTAGE_BYTES;
                        mbar_arrive_expect_tx(&bars[PB_X_FULL + u], upper ? PG_XSTAGE_BYTES : PG_XSTAGE_BYTES / 2);
                        tma_load_3d(dst, &tmX, &bars[PB_X_FULL + u], 8 * (ky - 1), kp * 128, tile * 2);
                        if (upper) tma_load_3d(dst + 16384, &tmX, &bars[PB_X_FULL + u], 8 * (ky - 1), kp * 128 + 64, tile * 2);
                    }
                    __syncwarp();
                    if (++kp == kpairs) { kp = 0; ++ky; }
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer: A = activations MN-major (idesc bit 15), B = weights K-major =====
        const bool leader = elect_one();
        constexpr uint32_t idesc = umma_idesc_16(128, NPAD, UMMA_FMT_16) | (1u << 15);
        const uint32_t w_lo = static_cast<uint32_t>(umma_desc_sw128(smem_u32(smem)));
        const uint32_t x_lo = static_cast<uint32_t>(umma_desc_mn_sw128(smem_u32(smem_x), 8192, 1024));
        uint32_t xpar = 0, wpar = 0, have = 0, it = 0;
        for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++it) {
            if (it > 0) {   // the epilogue warps have read the previous tile's accumulators
                mbar_wait(&bars[PB_DRAINED], (it - 1) & 1);
                tc_fence_after();
            }
            int kp = 0;
            uint32_t acc0 = 0;
            for (int s3 = 0; s3 < kpairs; ++s3) {
#pragma unroll
                for (int u = 0; u < PC_XSTAGES; ++u) {
                    const uint32_t nk = kp + 1 < kpairs ? 8u : static_cast<uint32_t>(p.ksteps_last);
                    mbar_wait(&bars[PB_X_FULL + u], xpar);
#pragma unroll
                    for (int kx = 0; kx < 3; ++kx) {
                        if (kx == 0 && !have) mbar_wait(&bars[PB_PAIR_FULL], wpar);
                        tc_fence_after();
                        const int nb = kx == 2 ? 0 : kx + 1;
                        const uint32_t npar = kx == 2 ? wpar ^ 1 : wpar;
                        const uint32_t next_ok = mbar_test_wait(&bars[PB_PAIR_FULL + nb], npar);
                        if (leader) {
                            umma_f16_burst8n<1024, 128, TILE / 16, 2, idesc>(tmem_base + kx * NPAD, x_lo + u * (PG_XSTAGE_BYTES / 16),
                                                                             w_lo + (2 * kx) * (TILE / 16), acc0, nk,
                                                                             &bars[PB_TILE_EMPTY + 2 * kx], &bars[PB_TILE_EMPTY + 2 * kx + 1]);
                            if (kx == 2) umma_commit(&bars[PB_X_EMPTY + u]);
                        }
                        __syncwarp();
                        if (kx < 2) {
                            if (!next_ok) mbar_wait(&bars[PB_PAIR_FULL + nb], npar);
                        } else {
                            have = next_ok;
                        }
                    }
                    wpar ^= 1;
                    acc0 = 1;
                    if (++kp == kpairs) kp = 0;
                }
                xpar ^= 1;
            }
            if (leader) umma_commit(&bars[PB_ACCUM]);
            __syncwarp();
        }
    } else if (warp >= 4) {
        // ===== epilogue: thread = one square of one position; warp = (lane quarter q, chunk group g) =====
        const int q = warp & 3;
        const int g = (warp - 4) >> 2;
        const int m = q * 32 + lane;                 // TMEM lane = row of D
        const int sq = m & 63;
        const bool has_left = (sq & 7) != 0, has_right = (sq & 7) != 7;
        const uint32_t t_row = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
        constexpr int CHUNKS = NPAD / 16;
        uint32_t it = 0;
        for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++it) {
            const int pos = tile * 2 + (m >> 6);
            const size_t out_base = static_cast<size_t>(pos) * p.out_channels * 64 + sq;
            mbar_wait_backoff(&bars[PB_ACCUM], it & 1, 64);
            tc_fence_after();
#pragma unroll 1
            for (int ci = g; ci < CHUNKS; ci += 4) {
                const int c0 = ci * 16;
                uint32_t l[16], c[16], r[16];
                tmem_ld_32x32b_x16(t_row + c0, l);                 // P_-1
                tmem_ld_32x32b_x16(t_row + NPAD + c0, c);          // P_0
                tmem_ld_32x32b_x16(t_row + 2 * NPAD + c0, r);      // P_+1
                tmem_ld_wait();
                if (ci + 4 >= CHUNKS) {   // this warp's last TMEM read of the tile
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&bars[PB_DRAINED]);
                }
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                    const float fl = __shfl_up_sync(0xffffffffu, __uint_as_float(l[j]), 1);     // square x-1 (same rank)
                    const float fr = __shfl_down_sync(0xffffffffu, __uint_as_float(r[j]), 1);   // square x+1
                    const int ch = c0 + j;
                    if (ch >= p.n_valid) continue;    // warp-uniform
                    float v = __uint_as_float(c[j]) + (has_left ? fl : 0.0f) + (has_right ? fr : 0.0f);
                    if (p.bias != nullptr) v += __ldg(p.bias + ch);
                    if (p.relu == 1) v = fmaxf(v, 0.0f);
                    const size_t o = out_base + static_cast<size_t>(ch) * 64;
                    if (p.out_f32) static_cast<float*>(p.out)[o] = v;
                    else static_cast<act_t*>(p.out)[o] = float_to_act(v);
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

}  // namespace hm
