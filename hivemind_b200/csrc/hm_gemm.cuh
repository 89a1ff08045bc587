// hm_gemm.cuh — tcgen05 + TMA implicit-GEMM for every dense contraction of RISEv3
// (reference layers: builder_util.py:154-178,
// 465-472, 246-326 and rise_mobile_v3.py:38-57).
//
//   out[m, n] = act( sum_{tap, c} A[pos(m), y(m)+dy(tap), x(m)+dx(tap), c] * W[n, tap*Cpad + c]
//                    + bias[n] ) (+ residual[m, n])
//
// A is an NHWC 16-bit activation [B, 8, 8, C] seen through a 4-D TMA tensor map
// (dims C, x, y, pos); a k x k convolution is `taps = k*k` shifted box loads whose
// out-of-board parts TMA zero-fills, i.e. the conv zero padding costs nothing and no
// im2col buffer exists.  W is 16-bit [N, taps*Cpad] (K-major), bias fp32 (BatchNorm folded).
//
// CTA = 128 output rows (two positions) x BN columns, 192 threads:
//   warp 0     TMA producer (one elected lane)
//   warp 1     TMEM allocator + tcgen05.mma issuer (one elected lane)
//   warps 2-5  epilogue: tcgen05.ld -> bias/ReLU/residual -> global stores
// SMEM ring of STAGES x (A 128x64 + W BNx64) bf16 tiles, 128-byte swizzle,
// full/empty mbarriers; accumulator 128 lanes x BN fp32 columns in TMEM.
#pragma once
#include <cuda.h>
#include "hm_ptx.cuh"

namespace hm {

struct GemmParams {
    int M;             // rows = positions * 64 (multiple of 128)
    int n_valid;       // columns to store (multiple of 8)
    int ksize;         // 1, 3 or 5
    int kchunks;       // 64-wide K chunks per tap = Cpad / 64
    const float* bias; // [n_valid] or nullptr
    const act_t* residual; // [M, ldr] or nullptr
    int ldr;
    void* out;         // act_t or fp32 [M, ldc]
    int ldc;
    int relu;
    int out_f32;
    // optional second output: per-position channel sums of the stored values (fp32
    // [positions, n_valid], atomically accumulated) for the next block's ECA squeeze
    float* chan_sum;
};

constexpr int GEMM_BM = 128;
constexpr int GEMM_BK = 64;
constexpr int GEMM_THREADS = 192;

template <int BN, int STAGES>
struct GemmSmem {
    static constexpr int A_BYTES = GEMM_BM * GEMM_BK * 2;
    static constexpr int B_BYTES = BN * GEMM_BK * 2;
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int TILE_BYTES = STAGES * STAGE_BYTES;
    // + barriers (2*STAGES+1) * 8, tmem ptr, bias tile
    static constexpr int AUX_BYTES = 256 + BN * 4;
    static constexpr int TOTAL = TILE_BYTES + AUX_BYTES + 1024;  // 1024 = alignment slack
};

template <int BN, int STAGES>
__global__ void __launch_bounds__(GEMM_THREADS)
gemm_tcgen05_kernel(const GemmParams p, const __grid_constant__ CUtensorMap tmA,
                    const __grid_constant__ CUtensorMap tmW) {
    using S = GemmSmem<BN, STAGES>;
    constexpr uint32_t TMEM_COLS = BN <= 32 ? 32 : BN <= 64 ? 64 : BN <= 128 ? 128 : 256;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                               ~static_cast<uintptr_t>(1023));
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + S::TILE_BYTES);
    uint64_t* empty_bar = full_bar + STAGES;
    uint64_t* accum_bar = empty_bar + STAGES;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(accum_bar + 1);
    float* bias_s = reinterpret_cast<float*>(smem + S::TILE_BYTES + 256);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int n_tiles = (p.n_valid + BN - 1) / BN;
    const int m_tile = blockIdx.x / n_tiles;
    const int n_tile = blockIdx.x % n_tiles;
    const int m0 = m_tile * GEMM_BM;
    const int n0 = n_tile * BN;
    const int taps = p.ksize * p.ksize;
    const int k_iters = taps * p.kchunks;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmA);
        tma_prefetch_desc(&tmW);
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        mbar_init(accum_bar, 1);
        fence_barrier_init();
    }
    if (warp == 1) {
        tmem_alloc(tmem_slot, TMEM_COLS);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===== TMA producer =====
        if (lane == 0) {
            const int pos0 = m0 >> 6;
            const int half = p.ksize >> 1;
            int stage = 0;
            uint32_t phase = 0;
            for (int it = 0; it < k_iters; ++it) {
                const int tap = it / p.kchunks;
                const int kc = it - tap * p.kchunks;
                const int dy = tap / p.ksize - half;
                const int dx = tap % p.ksize - half;
                mbar_wait(&empty_bar[stage], phase ^ 1);
                uint8_t* a_dst = smem + stage * S::STAGE_BYTES;
                uint8_t* b_dst = a_dst + S::A_BYTES;
                mbar_arrive_expect_tx(&full_bar[stage], S::STAGE_BYTES);
                tma_load_4d(a_dst, &tmA, &full_bar[stage], kc * GEMM_BK, dx, dy, pos0);
                tma_load_2d(b_dst, &tmW, &full_bar[stage], it * GEMM_BK, n0);
                if (++stage == STAGES) { stage = 0; phase ^= 1; }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer =====
        if (lane == 0) {
            constexpr uint32_t idesc = umma_idesc_16(GEMM_BM, BN, UMMA_FMT_16);
            int stage = 0;
            uint32_t phase = 0;
            for (int it = 0; it < k_iters; ++it) {
                mbar_wait(&full_bar[stage], phase);
                tc_fence_after();
                const uint32_t a_addr = smem_u32(smem + stage * S::STAGE_BYTES);
                const uint64_t a_desc = umma_desc_sw128(a_addr);
                const uint64_t b_desc = umma_desc_sw128(a_addr + S::A_BYTES);
#pragma unroll
                for (int k = 0; k < GEMM_BK / 16; ++k) {
                    // +32 bytes (16 bf16) along K inside the 128-byte swizzle row: +2 in
                    // the descriptor's 16-byte address units
                    umma_f16(tmem_base, a_desc + 2 * k, b_desc + 2 * k, idesc,
                             (it > 0 || k > 0) ? 1u : 0u);
                }
                umma_commit(&empty_bar[stage]);  // frees the smem slot when the MMAs retire
                if (++stage == STAGES) { stage = 0; phase ^= 1; }
            }
            umma_commit(accum_bar);  // accumulator complete
        }
    } else {
        // ===== epilogue (warps 2..5) =====
        const int et = threadIdx.x - 64;  // 0..127
        for (int i = et; i < BN; i += 128) {
            const int n = n0 + i;
            bias_s[i] = (p.bias != nullptr && n < p.n_valid) ? __ldg(p.bias + n) : 0.0f;
        }
        asm volatile("bar.sync 1, 128;" ::: "memory");
        mbar_wait(accum_bar, 0);
        tc_fence_after();
        const int lane_base = (warp & 3) * 32;  // TMEM lanes this warp may touch
        const int row = m0 + lane_base + lane;
        const uint32_t t_row = tmem_base + (static_cast<uint32_t>(lane_base) << 16);
#pragma unroll 1
        for (int c0 = 0; c0 < BN; c0 += 32) {
            const int ncols = min(32, p.n_valid - (n0 + c0));  // CTA-uniform
            if (ncols <= 0) break;
            uint32_t acc[32];
            tmem_ld_32x32b_x32(t_row + c0, acc);
            tmem_ld_wait();
            float v[32];
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(acc[j]) + bias_s[c0 + j];
            if (p.relu) {
#pragma unroll
                for (int j = 0; j < 32; ++j) v[j] = fmaxf(v[j], 0.0f);
            }
            if (p.residual != nullptr) {
                const uint4* r = reinterpret_cast<const uint4*>(
                    p.residual + static_cast<size_t>(row) * p.ldr + n0 + c0);
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    if (q * 8 < ncols) {
                        const uint4 rv = __ldg(r + q);
                        const float2 f0 = unpack_act2(rv.x), f1 = unpack_act2(rv.y),
                                     f2 = unpack_act2(rv.z), f3 = unpack_act2(rv.w);
                        v[q * 8 + 0] += f0.x; v[q * 8 + 1] += f0.y;
                        v[q * 8 + 2] += f1.x; v[q * 8 + 3] += f1.y;
                        v[q * 8 + 4] += f2.x; v[q * 8 + 5] += f2.y;
                        v[q * 8 + 6] += f3.x; v[q * 8 + 7] += f3.y;
                    }
                }
            }
            if (p.out_f32) {
                float4* o = reinterpret_cast<float4*>(static_cast<float*>(p.out) +
                                                      static_cast<size_t>(row) * p.ldc + n0 + c0);
#pragma unroll
                for (int q = 0; q < 8; ++q)
                    if (q * 4 < ncols)
                        o[q] = make_float4(v[q * 4], v[q * 4 + 1], v[q * 4 + 2], v[q * 4 + 3]);
            } else {
                uint4* o = reinterpret_cast<uint4*>(static_cast<act_t*>(p.out) +
                                                    static_cast<size_t>(row) * p.ldc + n0 + c0);
#pragma unroll
                for (int q = 0; q < 4; ++q)
                    if (q * 8 < ncols) {
                        uint4 w;
                        w.x = pack_act2(v[q * 8 + 0], v[q * 8 + 1]);
                        w.y = pack_act2(v[q * 8 + 2], v[q * 8 + 3]);
                        w.z = pack_act2(v[q * 8 + 4], v[q * 8 + 5]);
                        w.w = pack_act2(v[q * 8 + 6], v[q * 8 + 7]);
                        o[q] = w;
                    }
            }
            if (p.chan_sum != nullptr) {
                // ECA squeeze of the NEXT block (builder_util.py:77-78): per-position channel
                // sums of what was just stored (bf16-rounded, to match what ECA would read).
                // Lanes = 32 consecutive squares of one position: reduce over lanes.
                const int pos = row >> 6;
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                    float s = p.out_f32 ? v[j] : act_to_float(float_to_act(v[j]));
                    s += __shfl_xor_sync(0xffffffffu, s, 16);
                    s += __shfl_xor_sync(0xffffffffu, s, 8);
                    s += __shfl_xor_sync(0xffffffffu, s, 4);
                    s += __shfl_xor_sync(0xffffffffu, s, 2);
                    s += __shfl_xor_sync(0xffffffffu, s, 1);
                    if (lane == 0 && j < ncols)
                        atomicAdd(p.chan_sum + static_cast<size_t>(pos) * p.n_valid + n0 + c0 + j, s);
                }
            }
        }
        tc_fence_before();
    }
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, TMEM_COLS);
    }
}

}  // namespace hm
