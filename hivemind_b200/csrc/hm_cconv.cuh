// hm_cconv.cuh — persistent tcgen05 + TMA implicit-GEMM 3x3 convolution on NCHW activations (stem, policy convs;
// reference builder_util.py:154-178, 206-243, rise_mobile_v3.py:38-57).
//
// Same formulation, operand layouts, descriptors and epilogue arithmetic as cgemm_kernel (hm_cgemm.cuh): output
// channels = MMA M, the 128 squares of two positions = N, a vertical tap is a TMA square-coordinate shift, the
// horizontal taps are three TMEM accumulators P_dx combined in the epilogue,
// out[y][x] = P_-1[y][x-1] + P_0[y][x] + P_+1[y][x+1].  What differs is how the work is driven:
//   * ONE CTA per SM walks the (position tile, 128-row output block) work items, blockIdx.x, + gridDim.x, ...: barriers,
//     rings and TMEM are set up once and the next item's loads run under this item's drain (the one-CTA-per-item
//     kernel paid ~4 us of start-up and pipeline fill around 1.5-7 us of MMAs, six times per SM);
//   * the MMA issuer is the trunk's (hm_trunk.cuh): fully unrolled, ring slots and barrier addresses compile-time
//     offsets of uniform registers, 8 MMAs + commits per asm block, the next burst's barrier probed before the burst.
//     With run-time ring indices the operands of a burst came out of vector registers through R2UR and a burst took
//     ~705 cycles against 512 (tools/trace_cgemm.py);
//   * weights stream through six 16 KB tiles (pair-full / tile-empty barriers, a tile is refilled as soon as ITS four
//     MMAs retired), activations through three 32 KB stages, each serving the three horizontal taps;
//   * sixteen epilogue warps, four per TMEM lane quarter, one 32-column chunk each: the drain of an item takes
//     ~1.2 k cycles instead of ~3.5 k.
#pragma once
#include <cuda.h>
#include "hm_cgemm.cuh"

namespace hm {

// -DHM_BRINGUP builds: CTA 0 of the launch with CgemmParams::trace set logs (tag << 48 | SM clock) per role into
// g_cc_trace (roles: 0 producer, 1 MMA issuer, 2 first epilogue warp); hm_debug_cg_trace / tools/trace_cgemm.py
constexpr int CC_TRACE_CAP = 512;
#ifdef HM_BRINGUP
__device__ unsigned long long g_cc_trace[3 * CC_TRACE_CAP];
#define CC_MARK(role, tag)                                                                                  \
    do {                                                                                                    \
        if (p.trace && blockIdx.x == 0 && trace_n < CC_TRACE_CAP && (threadIdx.x & 31) == 0)                \
            g_cc_trace[(role) * CC_TRACE_CAP + trace_n++] =                                                 \
                (static_cast<unsigned long long>(tag) << 48) | (clock64() & 0xffffffffffffull);            \
    } while (0)
#else
#define CC_MARK(role, tag) do { } while (0)
#endif

constexpr int CC_EPI_WARPS = 16;
constexpr int CC_THREADS = 64 + CC_EPI_WARPS * 32;
constexpr int CC_WTILES = 6;                 // 16 KB weight tiles [128 rows x 64 K]; pair kx = tiles 2kx, 2kx+1
constexpr int CC_XSTAGES = 3;                // 32 KB activation stages [2 x 64 channels][2 positions][64 squares]
constexpr int CC_X_OFF = CC_WTILES * 16384;
constexpr int CC_OUT_OFF = CC_X_OFF + CC_XSTAGES * CG_STAGE_BYTES;   // 32 KB output staging tile (16-bit outputs)
constexpr int CC_BAR_OFF = CC_OUT_OFF + CG_STAGE_BYTES;
constexpr int CC_SMEM = CC_BAR_OFF + 256 + 1024;
enum {
    CB_X_FULL = 0,                           // 3
    CB_X_EMPTY = CB_X_FULL + CC_XSTAGES,     // 3
    CB_PAIR_FULL = CB_X_EMPTY + CC_XSTAGES,  // 3
    CB_TILE_EMPTY = CB_PAIR_FULL + 3,        // 6
    CB_ACCUM = CB_TILE_EMPTY + CC_WTILES,    // the three accumulators of an item are complete
    CB_DRAINED = CB_ACCUM + 1,               // ... and have been read by all epilogue warps
    CB_COUNT = CB_DRAINED + 1
};

// PACKED (layers with 64 < cin <= 80, i.e. the stem's 74 planes): a tap costs 80 K values, not 128.  Per horizontal
// tap kx the K dimension is four 64-row tiles, T0-T2 = channels 0-63 at ky = 0, 1, 2 and T3 = channels 64-79 of the
// three ky stacked (48 rows), i.e. TWO steps per item — (T0 | T1) with 8 K steps and (T2 | T3) with 7 — instead of
// three steps that each load 128 K values of which 80 are real: 246 KB instead of 384 KB through the ~64 B/clk
// per-SM ingest for 45 MMAs.  The weight matrix is packed to match ([n_rows][kx][256], hm_net.cu pack_conv_tap_packed),
// T3 is gathered by six 2 KB boxes (tmX16: 16 channels x 1 position).
template <bool PACKED>
__global__ void __launch_bounds__(CC_THREADS, 1)
cconv3_kernel(const CgemmParams p, const __grid_constant__ CUtensorMap tmW, const __grid_constant__ CUtensorMap tmX,
              const __grid_constant__ CUtensorMap tmX16, const __grid_constant__ CUtensorMap tmO) {
    extern __shared__ uint8_t cc_smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(cc_smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + CC_BAR_OFF);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + CB_COUNT);
    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
#ifdef HM_BRINGUP
    int trace_n = 0;
#endif
    if (threadIdx.x == 0) {
        tma_prefetch_desc(&tmW);
        tma_prefetch_desc(&tmX);
        for (int i = 0; i < CB_COUNT; ++i) mbar_init(&bars[i], i == CB_DRAINED ? CC_EPI_WARPS : 1);
        fence_barrier_init();
    }
    if (warp == 1) {
        tmem_alloc(tmem_slot, 512u);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const int kpairs = p.kblocks >> 1;   // 128-channel K pairs per tap; an item is 3 * kpairs steps of 3 bursts

    if (warp == 0) {
        // ===== TMA producer =====
        const bool leader = elect_one();
        const int tap_cols = p.kblocks * 64;   // K columns of one tap in the weight matrix
        uint32_t xk = 0, wk = 0;               // uses so far of an X stage / of a weight pair
        for (int item = blockIdx.x; item < p.n_items; item += gridDim.x) {
            const int tile = item / p.m_blocks;
            const int row0 = (item - tile * p.m_blocks) * 128, pos0 = tile * 2;
            CC_MARK(0, 1);
            if constexpr (PACKED) {
#pragma unroll
                for (int u = 0; u < 2; ++u) {
                    mbar_wait(&bars[CB_X_EMPTY + u], (xk & 1) ^ 1);
                    if (leader) {
                        uint8_t* dst = smem + CC_X_OFF + u * CG_STAGE_BYTES;
                        if (u == 0) {
                            mbar_arrive_expect_tx(&bars[CB_X_FULL], CG_STAGE_BYTES);
                            tma_load_3d(dst, &tmX, &bars[CB_X_FULL], -8, 0, pos0);
                            tma_load_3d(dst + 16384, &tmX, &bars[CB_X_FULL], 0, 0, pos0);
                        } else {
                            mbar_arrive_expect_tx(&bars[CB_X_FULL + 1], 16384 + 6 * 2048);
                            tma_load_3d(dst, &tmX, &bars[CB_X_FULL + 1], 8, 0, pos0);
#pragma unroll
                            for (int ky = 0; ky < 3; ++ky)
#pragma unroll
                                for (int pp = 0; pp < 2; ++pp)
                                    tma_load_3d(dst + 16384 + pp * 8192 + ky * 2048, &tmX16, &bars[CB_X_FULL + 1], 8 * (ky - 1), 64, pos0 + pp);
                        }
                    }
                    __syncwarp();
#pragma unroll
                    for (int kx = 0; kx < 3; ++kx) {
                        const int kcol = kx * 256 + u * 128;
                        mbar_wait(&bars[CB_TILE_EMPTY + 2 * kx], (wk & 1) ^ 1);
                        if (leader) {
                            mbar_arrive_expect_tx(&bars[CB_PAIR_FULL + kx], CG_STAGE_BYTES);
                            tma_load_2d(smem + (2 * kx) * 16384, &tmW, &bars[CB_PAIR_FULL + kx], kcol, row0);
                        }
                        __syncwarp();
                        mbar_wait(&bars[CB_TILE_EMPTY + 2 * kx + 1], (wk & 1) ^ 1);
                        if (leader) tma_load_2d(smem + (2 * kx + 1) * 16384, &tmW, &bars[CB_PAIR_FULL + kx], kcol + 64, row0);
                        __syncwarp();
                    }
                    ++wk;
                }
                ++xk;
                continue;
            }
            int ky = 0, kp = 0;
            for (int s3 = 0; s3 < kpairs; ++s3, ++xk) {
#pragma unroll
                for (int u = 0; u < CC_XSTAGES; ++u) {
                    // the upper 64 channels of the last pair may be pure padding: not loaded (and not multiplied)
                    const bool upper = kp + 1 < kpairs || p.ksteps_last > 4;
                    const uint32_t bytes = upper ? CG_STAGE_BYTES : CG_STAGE_BYTES / 2;
                    mbar_wait(&bars[CB_X_EMPTY + u], (xk & 1) ^ 1);
                    if (leader) {
                        uint8_t* dst = smem + CC_X_OFF + u * CG_STAGE_BYTES;
                        mbar_arrive_expect_tx(&bars[CB_X_FULL + u], bytes);
                        tma_load_3d(dst, &tmX, &bars[CB_X_FULL + u], 8 * (ky - 1), kp * 128, pos0);
                        if (upper) tma_load_3d(dst + 16384, &tmX, &bars[CB_X_FULL + u], 8 * (ky - 1), kp * 128 + 64, pos0);
                    }
                    __syncwarp();
#pragma unroll
                    for (int kx = 0; kx < 3; ++kx) {
                        const int kcol = (ky * 3 + kx) * tap_cols + kp * 128;
                        mbar_wait(&bars[CB_TILE_EMPTY + 2 * kx], (wk & 1) ^ 1);
                        if (leader) {
                            mbar_arrive_expect_tx(&bars[CB_PAIR_FULL + kx], bytes);
                            tma_load_2d(smem + (2 * kx) * 16384, &tmW, &bars[CB_PAIR_FULL + kx], kcol, row0);
                        }
                        __syncwarp();
                        mbar_wait(&bars[CB_TILE_EMPTY + 2 * kx + 1], (wk & 1) ^ 1);
                        if (leader && upper) tma_load_2d(smem + (2 * kx + 1) * 16384, &tmW, &bars[CB_PAIR_FULL + kx], kcol + 64, row0);
                        __syncwarp();
                    }
                    ++wk;
                    if (++kp == kpairs) { kp = 0; ++ky; }
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer (see hm_trunk.cuh for the issue discipline) =====
        const bool leader = elect_one();
        constexpr uint32_t idesc = umma_idesc_16_bmn(128, 128, UMMA_FMT_16);
        const uint32_t a_lo = static_cast<uint32_t>(umma_desc_sw128(smem_u32(smem)));
        const uint32_t b_lo = static_cast<uint32_t>(umma_desc_mn_sw128(smem_u32(smem + CC_X_OFF), 8192, 1024));
        uint32_t xpar = 0, wpar = 0, have = 0, it = 0;
        for (int item = blockIdx.x; item < p.n_items; item += gridDim.x, ++it) {
            CC_MARK(1, 1);
            if (it > 0) {   // the epilogue warps have read the previous item's accumulators
                mbar_wait(&bars[CB_DRAINED], (it - 1) & 1);
                tc_fence_after();
            }
            CC_MARK(1, 2);
            int kp = 0;
            uint32_t acc0 = 0;   // 0 on the first step of the item: the bursts overwrite the accumulators
            for (int s3 = 0; s3 < (PACKED ? 1 : kpairs); ++s3) {
#pragma unroll
                for (int u = 0; u < (PACKED ? 2 : CC_XSTAGES); ++u) {
                    const uint32_t nk = PACKED ? (u == 0 ? 8u : 7u) : kp + 1 < kpairs ? 8u : static_cast<uint32_t>(p.ksteps_last);
                    mbar_wait(&bars[CB_X_FULL + u], xpar);
                    CC_MARK(1, 3);
#pragma unroll
                    for (int kx = 0; kx < 3; ++kx) {
                        if (kx == 0 && !have) mbar_wait(&bars[CB_PAIR_FULL], wpar);
                        tc_fence_after();
                        const int nb = kx == 2 ? 0 : kx + 1;
                        const uint32_t npar = kx == 2 ? wpar ^ 1 : wpar;
                        const uint32_t next_ok = mbar_test_wait(&bars[CB_PAIR_FULL + nb], npar);
                        if (leader) {
                            umma_f16_burst8n<1024, 2, 1024, 128, idesc>(tmem_base + kx * 128, a_lo + (2 * kx) * 1024,
                                                                        b_lo + u * (CG_STAGE_BYTES / 16), acc0, nk,
                                                                        &bars[CB_TILE_EMPTY + 2 * kx], &bars[CB_TILE_EMPTY + 2 * kx + 1]);
                            if (kx == 2) umma_commit(&bars[CB_X_EMPTY + u]);
                        }
                        __syncwarp();
                        if (kx < 2) {
                            if (!next_ok) mbar_wait(&bars[CB_PAIR_FULL + nb], npar);
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
            if (leader) umma_commit(&bars[CB_ACCUM]);
            __syncwarp();
            CC_MARK(1, 4);
        }
    } else {
        // ===== epilogue: warp = (TMEM lane quarter q, 32-column chunk c); thread = one output channel, half a board =====
        const int q = warp & 3;                  // a warp may only touch TMEM lanes 32 * (warp % 4) ..
        const int c = (warp - 2) >> 2;           // chunk: position c >> 1, squares 32 * (c & 1) ..
        const uint32_t t_row = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + c * 32;
        // 16-bit outputs whose valid channels end on a 64-channel box boundary leave through a staging tile in the input
        // boxes' swizzled layout and two TMA stores (direct 16-byte stores put the 32 lanes of an instruction in 32
        // different 128-byte lines, ~8 B/clk per SM: the stem's items were bound by them)
        const bool staged = !p.out_f32 && (p.n_valid & 63) == 0;
        const bool storer = warp == 2 && lane == 0;
        const int rl = q * 32 + lane;
        uint8_t* my_row = smem + CC_OUT_OFF + (rl >> 6) * 16384 + (c >> 1) * 8192 + (rl & 63) * 128;
        uint32_t it = 0;
        for (int item = blockIdx.x; item < p.n_items; item += gridDim.x, ++it) {
            const int tile = item / p.m_blocks;
            const int ch = (item - tile * p.m_blocks) * 128 + q * 32 + lane;
            const bool valid = ch < p.n_valid;
            const float bias = (valid && p.bias != nullptr) ? __ldg(p.bias + ch) : 0.0f;
            const size_t row = (static_cast<size_t>(tile * 2 + (c >> 1)) * p.out_channels + ch) * 64 + (c & 1) * 32;
            if (warp == 2) CC_MARK(2, 1);
            mbar_wait_backoff(&bars[CB_ACCUM], it & 1, 64);
            tc_fence_after();
            if (warp == 2) CC_MARK(2, 2);
            float v[32];
            {
                uint32_t a[32];
                tmem_ld_32x32b_x32(t_row + 128, a);   // centre column (dx = 0)
                tmem_ld_wait();
#pragma unroll
                for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(a[j]) + bias;
                tmem_ld_32x32b_x32(t_row, a);         // P_-1
                tmem_ld_wait();
#pragma unroll
                for (int j = 0; j < 32; ++j)
                    if ((j & 7) != 0) v[j] += __uint_as_float(a[j - 1]);
                tmem_ld_32x32b_x32(t_row + 256, a);   // P_+1
                tmem_ld_wait();
#pragma unroll
                for (int j = 0; j < 32; ++j)
                    if ((j & 7) != 7) v[j] += __uint_as_float(a[j + 1]);
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&bars[CB_DRAINED]);   // this warp's part is in registers
            if (warp == 2) CC_MARK(2, 3);
            // (residual adds and GELU are cgemm_kernel's: the host routes no such layer here)
            if (p.relu == 1) {
#pragma unroll
                for (int j = 0; j < 32; ++j) v[j] = fmaxf(v[j], 0.0f);
            }
            if (p.out_f32) {
                if (valid) {
                    float4* o = reinterpret_cast<float4*>(static_cast<float*>(p.out) + row);
#pragma unroll
                    for (int j = 0; j < 8; ++j) o[j] = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
                }
            } else {
                uint4 w[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    w[j].x = pack_act2(v[8 * j + 0], v[8 * j + 1]);
                    w[j].y = pack_act2(v[8 * j + 2], v[8 * j + 3]);
                    w[j].z = pack_act2(v[8 * j + 4], v[8 * j + 5]);
                    w[j].w = pack_act2(v[8 * j + 6], v[8 * j + 7]);
                }
                if (!staged) {
                    if (valid) {
                        uint4* o = reinterpret_cast<uint4*>(static_cast<act_t*>(p.out) + row);
#pragma unroll
                        for (int j = 0; j < 4; ++j) o[j] = w[j];
                    }
                } else {   // kernel-uniform: all sixteen warps take the two barriers
                    if (storer) bulk_wait_group_read0();   // the previous item's stores have read the staging tile
                    named_bar_sync(1, CC_EPI_WARPS * 32);
                    if (valid) {
#pragma unroll
                        for (int j = 0; j < 4; ++j)
                            *reinterpret_cast<uint4*>(my_row + ((((c & 1) * 4 + j) ^ (rl & 7)) << 4)) = w[j];
                    }
                    fence_proxy_async_smem();
                    named_bar_sync(2, CC_EPI_WARPS * 32);
                    if (storer) {
                        const int r0 = (item - tile * p.m_blocks) * 128;
                        if (r0 < p.n_valid) tma_store_3d(&tmO, smem + CC_OUT_OFF, 0, r0, tile * 2);
                        if (r0 + 64 < p.n_valid) tma_store_3d(&tmO, smem + CC_OUT_OFF + 16384, 0, r0 + 64, tile * 2);
                        bulk_commit_group();
                    }
                }
            }
            if (warp == 2) CC_MARK(2, 4);
        }
        if (storer) bulk_wait_group0();
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512u);
    }
}

}  // namespace hm
