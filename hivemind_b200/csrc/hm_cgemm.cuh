// hm_cgemm.cuh — tcgen05 + TMA implicit-GEMM convolution on NCHW activations, used for every
// dense layer outside the residual trunk (stem 3x3, value 1x1,
// policy 3x3 convs; reference builder_util.py:154-178, 246-326, rise_mobile_v3.py:38-57).
//
// Channel-major ("transposed") formulation: output channels are the MMA M dimension, board
// squares the N dimension,
//
//   D[o, (pos, sq)] = act( sum_{tap, c} W[o, tap*Cpad + c] * X[pos, c, sq + shift(tap)] + bias[o] )
//
// A = W tile [128 out-channels x 64 K], K-major, 128-byte swizzle (TMA 2-D box).
// B = activation block [2 positions][64 channels][64 squares], MN-major (squares contiguous),
//     loaded by ONE 3-D TMA box (square, channel, position).  A vertical tap offset dy is a
//     shift of the square coordinate by 8*dy: squares shifted off the board and channels beyond
//     C are zero-filled by the TMA unit, so neither padding nor an im2col buffer exists.  (TMA
//     cannot shift the innermost dimension by one element — 16-byte granularity — so the
//     horizontal offset dx is NOT applied on the way in.)
// D = for a 3x3 convolution THREE accumulators P_dx (128 channel lanes x 128 fp32 columns each),
//     one per horizontal offset: P_dx[o,(y,x)] = sum_{dy,c} W[o;dy,dx;c] * X[c; y+dy, x].  The
//     epilogue thread owns one output channel and whole board ranks in registers, so it finishes
//     the convolution with out[y][x] = P_-1[y][x-1] + P_0[y][x] + P_+1[y][x+1] (zero beyond the
//     files a/h) — a register shift.  Its 64 squares of a position are one contiguous 128-byte
//     row of the NCHW output, written with 16-byte stores.
//
// CTA = (2 positions, 128 output channels); 192 threads: warp 0 TMA producer, warp 1 MMA
// issuer + TMEM owner, warps 2-5 epilogue.
#pragma once
#include <cuda.h>
#include "hm_ptx.cuh"

namespace hm {

struct CgemmParams {
    int m_blocks;       // ceil(cout / 128)
    int n_valid;        // cout: channels actually stored
    int ksize;          // 1 or 3
    int kblocks;        // Cpad / 64 per tap (even)
    const float* bias;  // [cout] or nullptr
    void* out;          // [pos][out_channels][64], act_t or fp32
    int out_channels;   // channel count of the output tensor
    int relu;           // activation: 0 none, 1 ReLU, 2 GELU (erf form, nn.GELU default)
    int out_f32;
    const act_t* residual;  // optional [pos][out_channels][64], added before the activation
    int ksteps_last;    // valid 16-channel K steps (1..8) in the last 128-channel K pair: the rest is zero padding
    int stages;         // ring depth (1..CG_STAGES); shallow-K 1x1 layers use 1 so that 3 CTAs share an SM
    int n_items;        // cconv3_kernel (hm_cconv.cuh): (position tile, 128-row output block) work items
    int trace;          // -DHM_BRINGUP builds only: CTA 0 of a cconv3_kernel launch logs its timeline
    // lgemm_kernel only (hm_lgemm.cuh)
    int kp_in2;              // K pairs kp >= kp_in2 are read from the second input tensor (channel (kp - kp_in2) * 128 ..); -1: none
    const float* post_add;   // [out_channels][64] fp32, added AFTER the activation (token embeddings), or nullptr
};

constexpr int CG_THREADS = 192;
// Shared memory: a ring of WEIGHT stages (two 16 KB A tiles = 128 K values of 128 output
// channels) and a ring of ACTIVATION stages (two 16 KB B blocks = 128 channels x 2 positions).
// For a 3x3 convolution one activation stage serves the three horizontal taps kx (they differ
// only in the weights and in the accumulator, the horizontal shift happens in the epilogue), so
// the activation tile crosses L2->SM once per (ky, 128 channels) instead of three times.
// 8 MMAs (512 tensor-pipe cycles) per weight-barrier round trip: one thread issues a 16 KB
// UTMALDG in ~200 cycles and pays ~100 per barrier round (tools/probes/probe_mcast.cu).
constexpr int CG_WSTAGES = 4;
constexpr int CG_XSTAGES = 2;
constexpr int CG_STAGE_BYTES = 32768;
constexpr int CG_STAGES = 3;   // legacy name: default ring depth handed to CgemmParams::stages
constexpr int cg_smem_bytes(int wstages, int xstages) { return (wstages + xstages) * CG_STAGE_BYTES + 256 + 1024; }
constexpr int CG_SMEM = cg_smem_bytes(CG_WSTAGES, CG_XSTAGES);

__global__ void __launch_bounds__(CG_THREADS)
cgemm_kernel(const CgemmParams p, const __grid_constant__ CUtensorMap tmW,
             const __grid_constant__ CUtensorMap tmX) {
    extern __shared__ uint8_t cg_smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(cg_smem_raw) + 1023) &
                                               ~static_cast<uintptr_t>(1023));
    // stages == 1: shallow-K 1x1 layers run with one W and one X stage so that 3 CTAs share an SM
    const int n_w = p.stages == 1 ? 1 : CG_WSTAGES;
    const int n_x = p.stages == 1 ? 1 : CG_XSTAGES;
    uint8_t* smem_x = smem + n_w * CG_STAGE_BYTES;
    uint64_t* full_w = reinterpret_cast<uint64_t*>(smem_x + n_x * CG_STAGE_BYTES);
    uint64_t* empty_w = full_w + CG_WSTAGES;
    uint64_t* full_x = empty_w + CG_WSTAGES;
    uint64_t* empty_x = full_x + CG_XSTAGES;
    uint64_t* accum_bar = empty_x + CG_XSTAGES;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(accum_bar + 1);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int tile = blockIdx.x / p.m_blocks;
    const int mb = blockIdx.x - tile * p.m_blocks;

    if (threadIdx.x == 0) {
        tma_prefetch_desc(&tmW);
        tma_prefetch_desc(&tmX);
        for (int s = 0; s < CG_WSTAGES; ++s) {
            mbar_init(&full_w[s], 1);
            mbar_init(&empty_w[s], 1);
        }
        for (int s = 0; s < CG_XSTAGES; ++s) {
            mbar_init(&full_x[s], 1);
            mbar_init(&empty_x[s], 1);
        }
        mbar_init(accum_bar, 1);
        fence_barrier_init();
    }
    const uint32_t tmem_cols = p.ksize == 3 ? 512u : 128u;
    if (warp == 1) {
        tmem_alloc(tmem_slot, tmem_cols);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    const int kpairs = p.kblocks >> 1;
    // both single-thread roles run warp-uniform; only the issuing instructions are predicated
    if (warp == 0) {
        const bool leader = elect_one();
        int ws = 0, xs = 0;
        uint32_t wphase = 1, xphase = 1;
        const int row0 = mb * 128, pos0 = tile * 2;
        const int tap_cols = p.kblocks * 64;            // K columns of one tap in the weight matrix
        for (int ky = 0; ky < p.ksize; ++ky) {
            const int sq0 = 8 * (ky - (p.ksize >> 1));   // vertical tap offset as a square shift
            for (int kp = 0; kp < kpairs; ++kp) {
                // the upper 64 channels of the last pair may be pure padding: neither loaded nor multiplied
                const bool upper = kp + 1 < kpairs || p.ksteps_last > 4;
                mbar_wait(&empty_x[xs], xphase);
                if (leader) {
                    uint8_t* dst = smem_x + xs * CG_STAGE_BYTES;
                    mbar_arrive_expect_tx(&full_x[xs], upper ? CG_STAGE_BYTES : CG_STAGE_BYTES / 2);
                    tma_load_3d(dst, &tmX, &full_x[xs], sq0, kp * 128, pos0);
                    if (upper) tma_load_3d(dst + 16384, &tmX, &full_x[xs], sq0, kp * 128 + 64, pos0);
                }
                __syncwarp();
                if (++xs == n_x) { xs = 0; xphase ^= 1; }
                for (int kx = 0; kx < p.ksize; ++kx) {
                    const int kcol = (ky * p.ksize + kx) * tap_cols + kp * 128;
                    mbar_wait(&empty_w[ws], wphase);
                    if (leader) {
                        uint8_t* dst = smem + ws * CG_STAGE_BYTES;
                        mbar_arrive_expect_tx(&full_w[ws], upper ? CG_STAGE_BYTES : CG_STAGE_BYTES / 2);
                        tma_load_2d(dst, &tmW, &full_w[ws], kcol, row0);
                        if (upper) tma_load_2d(dst + 16384, &tmW, &full_w[ws], kcol + 64, row0);
                    }
                    __syncwarp();
                    if (++ws == n_w) { ws = 0; wphase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // Issue discipline as in hm_trunk.cuh: the tensor pipe queues only two or three MMAs, so the work between two
        // bursts must be short — 8 MMAs + commit as one asm block (umma_f16_burst8), the NEXT weight stage's barrier
        // probed (test_wait) before the burst and consumed after it.
        const bool leader = elect_one();
        const uint32_t ld = leader ? 1u : 0u;
        constexpr uint32_t idesc = umma_idesc_16_bmn(128, 128, UMMA_FMT_16);
        const uint64_t a0 = umma_desc_sw128(smem_u32(smem));
        const uint64_t b0 = umma_desc_mn_sw128(smem_u32(smem_x), 8192, 1024);
        const uint32_t a_lo = static_cast<uint32_t>(a0), b_lo = static_cast<uint32_t>(b0);
        int ws = 0, xs = 0;
        uint32_t wphase = 0, xphase = 0;
        uint32_t have_w = 0;   // full_w[ws] already known complete
        for (int ky = 0; ky < p.ksize; ++ky) {
            for (int kp = 0; kp < kpairs; ++kp) {
                mbar_wait(&full_x[xs], xphase);
                const uint32_t xo = xs * (CG_STAGE_BYTES / 16);
                const int nk = kp + 1 < kpairs ? 8 : p.ksteps_last;   // K steps that hold real channels
                for (int kx = 0; kx < p.ksize; ++kx) {       // accumulator = horizontal offset kx
                    const uint32_t d = tmem_base + kx * 128;
                    if (!have_w) mbar_wait(&full_w[ws], wphase);
                    tc_fence_after();
                    const int nws = ws + 1 == n_w ? 0 : ws + 1;
                    const uint32_t nphase = ws + 1 == n_w ? wphase ^ 1 : wphase;
                    const uint32_t next_ok = mbar_test_wait(&full_w[nws], nphase);
                    const uint32_t wo = ws * (CG_STAGE_BYTES / 16);
                    const uint32_t acc0 = (ky | kp) ? 1u : 0u;  // first tile of this accumulator
                    if (nk == 8) {
                        if (leader) umma_f16_burst8<1024, 2, 1024, 128, idesc, false>(d, a_lo + wo, b_lo + xo, acc0, &empty_w[ws], &empty_w[ws], ld);
                    } else if (leader) {
#pragma unroll
                        for (int j = 0; j < 2; ++j) {
#pragma unroll
                            for (int k = 0; k < 4; ++k)
                                if (j * 4 + k < nk)
                                    umma_f16(d, a0 + (wo + j * 1024 + 2 * k), b0 + (xo + j * 1024 + k * 128), idesc,
                                             (j | k) ? 1u : acc0);
                        }
                        umma_commit(&empty_w[ws]);
                    }
                    if (kx == p.ksize - 1) umma_commit_if(&empty_x[xs], ld);
                    __syncwarp();
                    have_w = next_ok;
                    ws = nws;
                    wphase = nphase;
                }
                if (++xs == n_x) { xs = 0; xphase ^= 1; }
            }
        }
        if (leader) umma_commit(accum_bar);
        __syncwarp();
    } else {
        // ===== epilogue: thread = one output channel, both positions =====
        const int q = warp & 3;
        const int ch = mb * 128 + q * 32 + lane;
        const bool valid = ch < p.n_valid;
        const float bias = (valid && p.bias != nullptr) ? __ldg(p.bias + ch) : 0.0f;
        mbar_wait(accum_bar, 0);
        tc_fence_after();
        const uint32_t t_row = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
#pragma unroll 1
        for (int c0 = 0; c0 < 128; c0 += 32) {
            float v[32];
            {
                uint32_t acc[32];
                tmem_ld_32x32b_x32(t_row + (p.ksize == 3 ? 128 : 0) + c0, acc);  // centre column (dx = 0)
                tmem_ld_wait();
#pragma unroll
                for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(acc[j]) + bias;
            }
            if (p.ksize == 3) {  // CTA-uniform
                uint32_t l[32], r[32];
                tmem_ld_32x32b_x32(t_row + c0, l);        // P_-1
                tmem_ld_32x32b_x32(t_row + 256 + c0, r);  // P_+1
                tmem_ld_wait();
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                    if ((j & 7) != 0) v[j] += __uint_as_float(l[j - 1]);
                    if ((j & 7) != 7) v[j] += __uint_as_float(r[j + 1]);
                }
            }
            if (!valid) continue;
            const int pos = tile * 2 + (c0 >> 6);
            const size_t row = (static_cast<size_t>(pos) * p.out_channels + ch) * 64 + (c0 & 63);
            if (p.residual != nullptr) {
                const uint4* rr = reinterpret_cast<const uint4*>(p.residual + row);
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const uint4 u = __ldg(rr + j);
                    const uint32_t w4[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
                    for (int t = 0; t < 4; ++t) {
                        const float2 f = unpack_act2(w4[t]);
                        v[8 * j + 2 * t] += f.x;
                        v[8 * j + 2 * t + 1] += f.y;
                    }
                }
            }
            if (p.relu == 1) {
#pragma unroll
                for (int j = 0; j < 32; ++j) v[j] = fmaxf(v[j], 0.0f);
            } else if (p.relu == 2) {
#pragma unroll
                for (int j = 0; j < 32; ++j) v[j] = 0.5f * v[j] * (1.0f + erff(v[j] * 0.70710678118654752f));
            }
            if (p.out_f32) {
                float4* o = reinterpret_cast<float4*>(static_cast<float*>(p.out) + row);
#pragma unroll
                for (int j = 0; j < 8; ++j) o[j] = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
            } else {
                uint4* o = reinterpret_cast<uint4*>(static_cast<act_t*>(p.out) + row);
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    uint4 w;
                    w.x = pack_act2(v[8 * j + 0], v[8 * j + 1]);
                    w.y = pack_act2(v[8 * j + 2], v[8 * j + 3]);
                    w.z = pack_act2(v[8 * j + 4], v[8 * j + 5]);
                    w.w = pack_act2(v[8 * j + 6], v[8 * j + 7]);
                    o[j] = w;
                }
            }
        }
        tc_fence_before();
    }
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, tmem_cols);
    }
}

}  // namespace hm
