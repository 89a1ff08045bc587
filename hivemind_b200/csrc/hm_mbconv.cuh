// hm_mbconv.cuh — one RISEv3 bottleneck block as ONE persistent kernel (SURVEY.md §8a rows
// F2a-F2d and, through a precomputed gate, F3; reference builder_util.py:445-483):
//
//   x' = x * gate                       (ECA blocks only; gate from eca_gate_kernel)
//   y  = relu(pw1(x') + b1)             tcgen05, accumulator chunk in TMEM
//   z  = relu(dw_kxk(y) + bd)           CUDA cores on a shared-memory tile (half2 FMA)
//   out = x' + pw2(z) + b2              tcgen05, 128 x 384 fp32 accumulator in TMEM
//
// The expanded tensor (up to 1152 channels) never leaves the SM: the operating channels are
// processed in chunks of 64 — pw1 chunk -> TMEM -> registers -> smem Y -> depthwise -> smem Z
// (128-byte-swizzled K-major = A operand of pw2) -> accumulate into the 384-wide output tile.
// A CTA owns 128 rows = two whole positions, so the 8x8 board with its zero halo is local.
//
// Warp roles (320 threads): warp 0 = TMA producer, warp 1 = MMA issuer, warps 2-9 = compute
// (TMEM->Y epilogue, depthwise, final epilogue).  TMEM: D2 cols [0,384), D1[2] cols [384,512).
#pragma once
#include <cuda.h>
#include "hm_ptx.cuh"

namespace hm {

struct MbconvParams {
    int n_tiles;            // positions / 2
    int n_chunks;           // cop_pad / 64
    const uint8_t* par;     // per-chunk parameter blobs, stride par_stride bytes:
                            //   [bias1 64 x f32][dw bias 64 x half][dw w  taps x 64 half]
    int par_stride;
    const float* bias2;     // [384]
    const float* gate;      // [positions][384] fp32 or nullptr (ECA blocks)
    const act_t* x;         // block input [positions*64][384] (residual source)
    act_t* out;             // block output [positions*64][384]
};

constexpr int MB_THREADS = 320;
constexpr int MB_COMPUTE = 256;
constexpr int MB_C = 384;
constexpr int MB_SLOTS = 4;
constexpr int MB_SLOT_BYTES = 16384;
constexpr int MB_PAR_BUFS = 3;

template <int KS>
struct MbconvSmem {
    static constexpr int PAR_BYTES = 384 + KS * KS * 128;        // per chunk
    static constexpr int X_OFF = 0;                               // 6 x 16 KB
    static constexpr int RING_OFF = X_OFF + 6 * 16384;            // 4 x 16 KB
    static constexpr int Y_OFF = RING_OFF + MB_SLOTS * MB_SLOT_BYTES;
    static constexpr int Z_OFF = Y_OFF + 16384;                   // 2 x 16 KB
    static constexpr int PAR_OFF = Z_OFF + 2 * 16384;
    static constexpr int B2_OFF = PAR_OFF + MB_PAR_BUFS * ((PAR_BYTES + 127) / 128 * 128);
    static constexpr int GATE_OFF = B2_OFF + MB_C * 4;            // 2 tiles x 2 pos x 384 f32
    static constexpr int BAR_OFF = GATE_OFF + 2 * 2 * MB_C * 4;
    static constexpr int TOTAL = BAR_OFF + 512;
};

// barrier indices inside the barrier block
enum {
    BAR_X_FULL = 0, BAR_X_EMPTY = 1, BAR_XS_READY = 2, BAR_D2_FULL = 3, BAR_D2_EMPTY = 4,
    BAR_SLOT_FULL = 5,                       // 4
    BAR_SLOT_EMPTY = BAR_SLOT_FULL + MB_SLOTS,   // 4
    BAR_PAR_FULL = BAR_SLOT_EMPTY + MB_SLOTS,    // 3
    BAR_PAR_EMPTY = BAR_PAR_FULL + MB_PAR_BUFS,  // 3
    BAR_D1_FULL = BAR_PAR_EMPTY + MB_PAR_BUFS,   // 2
    BAR_D1_EMPTY = BAR_D1_FULL + 2,              // 2
    BAR_Z_FULL = BAR_D1_EMPTY + 2,               // 2
    BAR_Z_EMPTY = BAR_Z_FULL + 2,                // 2
    BAR_COUNT = BAR_Z_EMPTY + 2
};

// wait for the k-th completion (k = 0, 1, ...) of a "full"-type barrier
__device__ __forceinline__ void wait_full(uint64_t* bar, uint32_t k) { mbar_wait(bar, k & 1); }
// wait until the k-th reuse of a resource is allowed: passes immediately for k = 0, then
// needs the (k-1)-th completion of the "empty"-type barrier
__device__ __forceinline__ void wait_empty(uint64_t* bar, uint32_t k) { mbar_wait(bar, (k & 1) ^ 1); }

template <int KS>
__device__ void mbconv_compute(const MbconvParams& p, uint8_t* smem, uint64_t* bars, uint32_t tm_d2,
                               uint32_t tm_d1a, uint32_t tm_d1b);

template <int KS>
__global__ void __launch_bounds__(MB_THREADS, 1)
mbconv_kernel(const MbconvParams p, const __grid_constant__ CUtensorMap tmX,
              const __grid_constant__ CUtensorMap tmW1, const __grid_constant__ CUtensorMap tmW2) {
    using S = MbconvSmem<KS>;
    extern __shared__ __align__(1024) uint8_t mb_smem_raw[];
    uint8_t* smem = mb_smem_raw;
    if (smem_u32(smem) & 1023u) __trap();  // 128-byte swizzle atoms need 1024-byte alignment
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + S::BAR_OFF);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + BAR_COUNT);
    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const bool has_gate = p.gate != nullptr;
    const int nch = p.n_chunks;

    if (threadIdx.x == 0) {
        tma_prefetch_desc(&tmX);
        tma_prefetch_desc(&tmW1);
        tma_prefetch_desc(&tmW2);
        for (int i = 0; i < BAR_COUNT; ++i) {
            uint32_t count = 1;
            if (i == BAR_XS_READY || i == BAR_D2_EMPTY) count = 8;
            if (i >= BAR_PAR_EMPTY && i < BAR_PAR_EMPTY + MB_PAR_BUFS) count = 8;
            if (i >= BAR_D1_EMPTY && i < BAR_D1_EMPTY + 2) count = 8;
            if (i >= BAR_Z_FULL && i < BAR_Z_FULL + 2) count = 8;
            mbar_init(&bars[i], count);
        }
        fence_barrier_init();
    }
    if (warp == 1) {
        tmem_alloc(tmem_slot, 512);
        tmem_relinquish();
    }
    // bias2 is tile-invariant
    for (int i = threadIdx.x; i < MB_C; i += MB_THREADS)
        reinterpret_cast<float*>(smem + S::B2_OFF)[i] = __ldg(p.bias2 + i);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const uint32_t tm_d2 = tmem_base;
    const uint32_t tm_d1[2] = {tmem_base + 384u, tmem_base + 448u};

    if (warp == 0) {
        // ================================ TMA producer ================================
        if (lane == 0) {
            uint32_t slot_n = 0;   // ring uses so far
            uint32_t par_n = 0;    // parameter-buffer uses so far
            uint32_t tile_n = 0;
            auto load_w1 = [&](int c) {
                // parameters of chunk c
                const uint32_t pb = par_n % MB_PAR_BUFS;
                wait_empty(&bars[BAR_PAR_EMPTY + pb], par_n / MB_PAR_BUFS);
                mbar_arrive_expect_tx(&bars[BAR_PAR_FULL + pb], S::PAR_BYTES);
                bulk_load_1d(smem + S::PAR_OFF + pb * ((S::PAR_BYTES + 127) / 128 * 128),
                             p.par + static_cast<size_t>(c) * p.par_stride, S::PAR_BYTES,
                             &bars[BAR_PAR_FULL + pb]);
                ++par_n;
                // W1 rows [c*64, +64) x k [0,384): 3 slots of 2 x [64 rows x 64 k]
                for (int s = 0; s < 3; ++s) {
                    const uint32_t sl = slot_n % MB_SLOTS;
                    wait_empty(&bars[BAR_SLOT_EMPTY + sl], slot_n / MB_SLOTS);
                    uint8_t* dst = smem + S::RING_OFF + sl * MB_SLOT_BYTES;
                    mbar_arrive_expect_tx(&bars[BAR_SLOT_FULL + sl], MB_SLOT_BYTES);
                    tma_load_2d(dst, &tmW1, &bars[BAR_SLOT_FULL + sl], s * 128, c * 64);
                    tma_load_2d(dst + 8192, &tmW1, &bars[BAR_SLOT_FULL + sl], s * 128 + 64, c * 64);
                    ++slot_n;
                }
            };
            auto load_w2 = [&](int c) {
                // W2 rows [s*128, +128) x k [c*64, +64)
                for (int s = 0; s < 3; ++s) {
                    const uint32_t sl = slot_n % MB_SLOTS;
                    wait_empty(&bars[BAR_SLOT_EMPTY + sl], slot_n / MB_SLOTS);
                    uint8_t* dst = smem + S::RING_OFF + sl * MB_SLOT_BYTES;
                    mbar_arrive_expect_tx(&bars[BAR_SLOT_FULL + sl], MB_SLOT_BYTES);
                    tma_load_2d(dst, &tmW2, &bars[BAR_SLOT_FULL + sl], c * 64, s * 128);
                    ++slot_n;
                }
            };
            for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x, ++tile_n) {
                wait_empty(&bars[BAR_X_EMPTY], tile_n);
                mbar_arrive_expect_tx(&bars[BAR_X_FULL], 6 * 16384);
                for (int j = 0; j < 6; ++j)
                    tma_load_4d(smem + S::X_OFF + j * 16384, &tmX, &bars[BAR_X_FULL], j * 64, 0, 0, tile * 2);
                load_w1(0);
                for (int c = 0; c < nch; ++c) {
                    if (c + 1 < nch) load_w1(c + 1);
                    load_w2(c);
                }
            }
        }
    } else if (warp == 1) {
        // ================================ MMA issuer ================================
        if (lane == 0) {
            constexpr uint32_t idesc64 = umma_idesc_16(128, 64, UMMA_FMT_16);
            constexpr uint32_t idesc128 = umma_idesc_16(128, 128, UMMA_FMT_16);
            uint32_t slot_n = 0, tile_n = 0;
            uint32_t d1_n[2] = {0, 0}, z_n[2] = {0, 0};
            const uint32_t x_addr = smem_u32(smem + S::X_OFF);
            auto pw1 = [&](int c) {
                const int b = c & 1;
                wait_empty(&bars[BAR_D1_EMPTY + b], d1_n[b]);
                tc_fence_after();
                for (int s = 0; s < 3; ++s) {
                    const uint32_t sl = slot_n % MB_SLOTS;
                    wait_full(&bars[BAR_SLOT_FULL + sl], slot_n / MB_SLOTS);
                    tc_fence_after();
                    const uint32_t w_addr = smem_u32(smem + S::RING_OFF + sl * MB_SLOT_BYTES);
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        const uint64_t a_desc = umma_desc_sw128(x_addr + (s * 2 + h) * 16384);
                        const uint64_t b_desc = umma_desc_sw128(w_addr + h * 8192);
#pragma unroll
                        for (int k = 0; k < 4; ++k)
                            umma_f16(tm_d1[b], a_desc + 2 * k, b_desc + 2 * k, idesc64,
                                     (s | h | k) ? 1u : 0u);
                    }
                    umma_commit(&bars[BAR_SLOT_EMPTY + sl]);
                    ++slot_n;
                }
                umma_commit(&bars[BAR_D1_FULL + b]);
                ++d1_n[b];
            };
            for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x, ++tile_n) {
                wait_full(&bars[has_gate ? BAR_XS_READY : BAR_X_FULL], tile_n);
                tc_fence_after();
                pw1(0);
                for (int c = 0; c < nch; ++c) {
                    if (c + 1 < nch) pw1(c + 1);
                    if (c + 1 >= nch) umma_commit(&bars[BAR_X_EMPTY]);  // all pw1 of the tile issued
                    const int b = c & 1;
                    if (c == 0) {  // D2 of the previous tile must have been drained
                        wait_empty(&bars[BAR_D2_EMPTY], tile_n);
                    }
                    wait_full(&bars[BAR_Z_FULL + b], z_n[b]);
                    tc_fence_after();
                    const uint32_t z_addr = smem_u32(smem + S::Z_OFF + b * 16384);
                    for (int s = 0; s < 3; ++s) {
                        const uint32_t sl = slot_n % MB_SLOTS;
                        wait_full(&bars[BAR_SLOT_FULL + sl], slot_n / MB_SLOTS);
                        tc_fence_after();
                        const uint64_t a_desc = umma_desc_sw128(z_addr);
                        const uint64_t b_desc = umma_desc_sw128(smem_u32(smem + S::RING_OFF + sl * MB_SLOT_BYTES));
#pragma unroll
                        for (int k = 0; k < 4; ++k)
                            umma_f16(tm_d2 + s * 128, a_desc + 2 * k, b_desc + 2 * k, idesc128,
                                     (c | k) ? 1u : 0u);
                        umma_commit(&bars[BAR_SLOT_EMPTY + sl]);
                        ++slot_n;
                    }
                    umma_commit(&bars[BAR_Z_EMPTY + b]);
                    ++z_n[b];
                }
                umma_commit(&bars[BAR_D2_FULL]);
            }
        }
    } else {
        mbconv_compute<KS>(p, smem, bars, tm_d2, tm_d1[0], tm_d1[1]);
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

}  // namespace hm

#include "hm_mbconv_compute.cuh"
