// hm_mlp.cuh — the MLP of a cross-board attention block as ONE kernel (rise_mobile_v3.py:270-278:
// x + fc2(GELU(fc1(LayerNorm(x))))), for channel-major tokens [n2][192][64]:
//
//   H[384 x 128]  = W1[384 x 192] . LN[192 x 128 tokens]          (phase 1: three 128-row accumulators)
//   Hs            = GELU(H + b1), 16-bit, in SHARED MEMORY as the B operand of phase 2
//   O[192 x 128]  = W2[192 x 384] . Hs                            (phase 2: rows 0-127 and 128-191)
//   out           = O + b2 + x                                     (staged in shared memory, TMA store)
//
// As two lgemm launches the hidden tensor (n2 x 384 x 64 x 2 B = 100 MB at B = 1024) was written to HBM and read back:
// fc1 0.064 ms + fc2 0.062 ms per layer, both bound by the bytes they can keep in flight per SM.  Here it never leaves
// the SM.  One persistent CTA per SM walks the 128-token tiles; both weight matrices (336 KB) stream through a
// 5-tile ring per tile, from L2.
//
// Shared memory: Hs 6 boxes x 16 KB ([2 positions][64 channels][64 squares], 128-byte swizzled, = the layout of an
// input box; boxes 0-2 double as the output staging tile), X 3 boxes (the LayerNorm-ed tokens of the tile), weight
// ring 5 x 16 KB ([128 rows x 64 K] K-major).  TMEM: H in columns 0-383; O rows 0-127 in columns 384-511 (accumulated
// box by box as epilogue 1 writes Hs), O rows 128-191 in columns 256-383 (H block 2, drained by then: that pass runs
// after all of Hs is there).  The next tile's phase 1 starts on H blocks 0 and 1 while this tile's output is drained.
// Warps: 0 TMA producer, 1 MMA issuer + TMEM owner, 2-17 epilogue (TMEM lane quarter q = warp % 4, column chunk c).
#pragma once
#include <cuda.h>
#include "hm_lgemm.cuh"

namespace hm {

struct MlpParams {
    int trace;               // -DHM_BRINGUP builds only: CTA 0 logs its timeline (CC_MARK, tools/trace_cgemm.py)
    int n_tiles;             // n2 / 2
    const float* bias1;      // [384]
    const float* bias2;      // [192]
    const act_t* residual;   // x: [n2][192][64]
};

constexpr int ML_D = 192, ML_H = 384;
constexpr int ML_EPI_WARPS = 16;
constexpr int ML_THREADS = 64 + ML_EPI_WARPS * 32;
constexpr int ML_WSLOTS = 5;
constexpr int ML_HS_OFF = 0;
constexpr int ML_X_OFF = 6 * 16384;
constexpr int ML_RING_OFF = ML_X_OFF + 3 * 16384;
constexpr int ML_BAR_OFF = ML_RING_OFF + ML_WSLOTS * 16384;
constexpr int ML_SMEM = ML_BAR_OFF + 256 + 1024;
enum {
    MB_X_FULL = 0, MB_X_EMPTY,
    MB_W_FULL,                            // ML_WSLOTS
    MB_W_EMPTY = MB_W_FULL + ML_WSLOTS,   // ML_WSLOTS
    MB_H_FULL = MB_W_EMPTY + ML_WSLOTS,   // 3: accumulator block m of phase 1 complete
    MB_HS_READY = MB_H_FULL + 3,          // 3: Hs boxes 2m, 2m+1 written (and H block m drained), 16 arrivals
    MB_O_FULL = MB_HS_READY + 3,          // phase 2 complete
    MB_O_DRAINED,                         // both O blocks read by the epilogue warps, 16 arrivals
    MB_COUNT
};

// gelu_erf_fast (hm_lgemm.cuh) on two values per instruction (fma.rn.f32x2): the same operations in the same order,
// bit-identical results, about half the issue slots — epilogue 1 is the longest serial piece of a tile.
__device__ __forceinline__ float2 gelu_erf_fast2(float2 x) {
    const float2 z = make_float2(fabsf(x.x) * 0.70710678118654752f, fabsf(x.y) * 0.70710678118654752f);
    auto k2 = [](float k) { return make_float2(k, k); };
    float2 p = __ffma2_rn(k2(0.0000430638f), z, k2(0.0002765672f));
    p = __ffma2_rn(p, z, k2(0.0001520143f));
    p = __ffma2_rn(p, z, k2(0.0092705272f));
    p = __ffma2_rn(p, z, k2(0.0422820123f));
    p = __ffma2_rn(p, z, k2(0.0705230784f));
    p = __ffma2_rn(p, z, k2(1.0f));
    p = __fmul2_rn(p, p); p = __fmul2_rn(p, p); p = __fmul2_rn(p, p); p = __fmul2_rn(p, p);
    float2 r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r.x) : "f"(p.x));
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r.y) : "f"(p.y));
    const float2 t = make_float2(1.0f + copysignf(1.0f - r.x, x.x), 1.0f + copysignf(1.0f - r.y, x.y));
    return __fmul2_rn(__fmul2_rn(k2(0.5f), x), t);
}

__global__ void __launch_bounds__(ML_THREADS, 1)
mlp_kernel(const MlpParams p, const __grid_constant__ CUtensorMap tmW1, const __grid_constant__ CUtensorMap tmW2,
           const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmO) {
    extern __shared__ uint8_t ml_smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(ml_smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + ML_BAR_OFF);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + MB_COUNT);
    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
#ifdef HM_BRINGUP
    int trace_n = 0;
#endif
    if (threadIdx.x == 0) {
        tma_prefetch_desc(&tmW1);
        tma_prefetch_desc(&tmW2);
        tma_prefetch_desc(&tmX);
        tma_prefetch_desc(&tmO);
        for (int i = 0; i < MB_COUNT; ++i)
            mbar_init(&bars[i], ((i >= MB_HS_READY && i < MB_HS_READY + 3) || i == MB_O_DRAINED) ? ML_EPI_WARPS : 1);
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

    if (warp == 0) {
        // ===== TMA producer: X boxes of the tile, then its 9 + 12 weight tiles in the issuer's order =====
        const bool leader = elect_one();
        uint32_t wn = 0, it = 0;   // weight tiles issued so far; tiles of this CTA so far
        auto w_tile = [&](const CUtensorMap* tm, int kcol, int row) {
            const uint32_t slot = wn % ML_WSLOTS, use = wn / ML_WSLOTS;
            mbar_wait(&bars[MB_W_EMPTY + slot], (use & 1) ^ 1);
            if (leader) {
                mbar_arrive_expect_tx(&bars[MB_W_FULL + slot], 16384);
                tma_load_2d(smem + ML_RING_OFF + slot * 16384, tm, &bars[MB_W_FULL + slot], kcol, row);
            }
            __syncwarp();
            ++wn;
        };
        auto x_tile = [&](int tile, uint32_t n) {   // the n-th X load of this CTA: after phase 1 of the (n-1)-th tile
            mbar_wait(&bars[MB_X_EMPTY], (n & 1) ^ 1);
            if (leader) {
                mbar_arrive_expect_tx(&bars[MB_X_FULL], 3 * 16384);
#pragma unroll
                for (int kb = 0; kb < 3; ++kb) tma_load_3d(smem + ML_X_OFF + kb * 16384, &tmX, &bars[MB_X_FULL], 0, kb * 64, tile * 2);
            }
            __syncwarp();
        };
        if (static_cast<int>(blockIdx.x) < p.n_tiles) x_tile(blockIdx.x, 0);
        for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x, ++it) {
            for (int m = 0; m < 3; ++m)
                for (int k = 0; k < 3; ++k) w_tile(&tmW1, k * 64, m * 128);
            for (int mo = 0; mo < 2; ++mo)
                for (int k = 0; k < 6; ++k) {
                    w_tile(&tmW2, k * 64, mo * 128);
                    // the next tile's X: phase 1 of this tile is over by the time the ring has taken ML_WSLOTS tiles of
                    // phase 2 (their slots were freed by its last MMAs), so this wait does not hold up the weight stream
                    if (mo == 0 && k == ML_WSLOTS - 1) {
                        const int next = tile + static_cast<int>(gridDim.x);
                        if (next < p.n_tiles) x_tile(next, it + 1);
                    }
                }
        }
    } else if (warp == 1) {
        // ===== MMA issuer =====
        const bool leader = elect_one();
        constexpr uint32_t idesc = umma_idesc_16_bmn(128, 128, UMMA_FMT_16);
        const uint64_t a0 = umma_desc_sw128(smem_u32(smem + ML_RING_OFF));
        const uint64_t bx0 = umma_desc_mn_sw128(smem_u32(smem + ML_X_OFF), 8192, 1024);
        const uint64_t bh0 = umma_desc_mn_sw128(smem_u32(smem + ML_HS_OFF), 8192, 1024);
        uint32_t wn = 0, it = 0;
        // four K16 steps of one weight tile against the 64-channel box `b_box` of a B operand
        auto tile_mma = [&](uint32_t d, uint64_t b_box, uint32_t acc_first) {
            const uint32_t slot = wn % ML_WSLOTS, use = wn / ML_WSLOTS;
            mbar_wait(&bars[MB_W_FULL + slot], use & 1);
            tc_fence_after();
            if (leader) {
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    umma_f16(d, a0 + (slot * 1024 + 2 * k), b_box + k * 128, idesc, k ? 1u : acc_first);
                umma_commit(&bars[MB_W_EMPTY + slot]);
            }
            __syncwarp();
            ++wn;
        };
        for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x, ++it) {
            CC_MARK(1, 1);
            mbar_wait(&bars[MB_X_FULL], it & 1);
            tc_fence_after();
            CC_MARK(1, 2);
            // phase 1: H block m = W1[m] . X; blocks 0 and 1 are free as soon as the previous tile's epilogue 1 is done
            // (waited for in its phase 2), block 2 holds the previous tile's second output block until that is drained
            for (int m = 0; m < 3; ++m) {
                if (m == 2 && it > 0) {
                    mbar_wait(&bars[MB_O_DRAINED], (it - 1) & 1);
                    tc_fence_after();
                }
                for (int k = 0; k < 3; ++k) tile_mma(tmem_base + m * 128, bx0 + k * 1024, k ? 1u : 0u);
                if (leader) umma_commit(&bars[MB_H_FULL + m]);
                __syncwarp();
            }
            if (leader) umma_commit(&bars[MB_X_EMPTY]);
            __syncwarp();
            CC_MARK(1, 3);
            // phase 2: O rows 0-127 -> columns 384.., box k of Hs as soon as it is written; then rows 128-191 -> columns
            // 256.. (H block 2: drained once HS_READY[2] has been seen)
            for (int k = 0; k < 6; ++k) {
                if ((k & 1) == 0) {
                    mbar_wait(&bars[MB_HS_READY + (k >> 1)], it & 1);
                    tc_fence_after();
                    CC_MARK(1, 10 + (k >> 1));
                }
                tile_mma(tmem_base + 384, bh0 + k * 1024, k ? 1u : 0u);
            }
            for (int k = 0; k < 6; ++k) tile_mma(tmem_base + 256, bh0 + k * 1024, k ? 1u : 0u);
            if (leader) umma_commit(&bars[MB_O_FULL]);
            __syncwarp();
            CC_MARK(1, 4);
        }
    } else {
        // ===== epilogue warps: thread = one channel row (TMEM lane), 32 squares (column chunk c) =====
        const int q = warp & 3;
        const int c = (warp - 2) >> 2;           // chunk: position c >> 1, squares 32 * (c & 1) ..
        const int rl = q * 32 + lane;            // row within a 128-row accumulator block
        const bool storer = warp == 2 && lane == 0;
        const uint32_t t_lane = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + c * 32;
        // this thread's 64-byte half row inside box (rl >> 6) of a 128-channel block: position c >> 1, row rl & 63
        const uint32_t row_off = (rl >> 6) * 16384 + (c >> 1) * 8192 + (rl & 63) * 128;
        auto put_row = [&](uint8_t* block, const float (&v)[32]) {   // 16-bit, 128-byte swizzle (conflict-free)
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                uint4 w;
                w.x = pack_act2(v[8 * j + 0], v[8 * j + 1]);
                w.y = pack_act2(v[8 * j + 2], v[8 * j + 3]);
                w.z = pack_act2(v[8 * j + 4], v[8 * j + 5]);
                w.w = pack_act2(v[8 * j + 6], v[8 * j + 7]);
                *reinterpret_cast<uint4*>(block + row_off + ((((c & 1) * 4 + j) ^ (rl & 7)) << 4)) = w;
            }
        };
        uint32_t it = 0;
        for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x, ++it) {
            // ---- epilogue 1: Hs = GELU(H + b1) ----
#pragma unroll 1
            for (int m = 0; m < 3; ++m) {
                const float b1 = __ldg(p.bias1 + m * 128 + rl);
                mbar_wait(&bars[MB_H_FULL + m], it & 1);
                tc_fence_after();
                if (warp == 2) CC_MARK(2, 10 + m);
                uint32_t acc[32];
                tmem_ld_32x32b_x32(t_lane + m * 128, acc);
                tmem_ld_wait();
                float v[32];
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                    const float2 g = gelu_erf_fast2(make_float2(__uint_as_float(acc[2 * j]) + b1, __uint_as_float(acc[2 * j + 1]) + b1));
                    v[2 * j] = g.x;
                    v[2 * j + 1] = g.y;
                }
                if (m == 0 && it > 0) {   // boxes 0-2 of Hs were the previous tile's output staging: its stores have read them
                    if (storer) bulk_wait_group_read0();
                    named_bar_sync(1, ML_EPI_WARPS * 32);
                }
                put_row(smem + ML_HS_OFF + m * 32768, v);
                fence_proxy_async_smem();   // generic-proxy writes -> visible to the tensor core's operand reads
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&bars[MB_HS_READY + m]);
                if (warp == 2) CC_MARK(2, 20 + m);
            }
            // ---- epilogue 2: out = O + b2 + x ----
            const int pos = tile * 2 + (c >> 1);
            uint4 res[8];   // residual rows of channel rl (block 0) and 128 + rl (block 1, quarters 0-1), requested before the wait
            {
                const uint4* r0 = reinterpret_cast<const uint4*>(p.residual + (static_cast<size_t>(pos) * ML_D + rl) * 64 + (c & 1) * 32);
#pragma unroll
                for (int j = 0; j < 4; ++j) res[j] = __ldg(r0 + j);
                if (q < 2) {
                    const uint4* r1 = r0 + 128 * 64 * 2 / 16;
#pragma unroll
                    for (int j = 0; j < 4; ++j) res[4 + j] = __ldg(r1 + j);
                }
            }
            const float b2v[2] = {__ldg(p.bias2 + rl), q < 2 ? __ldg(p.bias2 + 128 + rl) : 0.0f};
            mbar_wait(&bars[MB_O_FULL], it & 1);
            tc_fence_after();
            if (warp == 2) CC_MARK(2, 2);
#pragma unroll
            for (int mo = 0; mo < 2; ++mo) {
                if (mo == 1 && q >= 2) break;   // rows 192-255 of the second block are padding
                uint32_t acc[32];
                tmem_ld_32x32b_x32(t_lane + (mo == 0 ? 384 : 256), acc);
                tmem_ld_wait();
                const float b2 = b2v[mo];
                float v[32];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const uint4 u = res[4 * mo + j];
                    const uint32_t w4[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
                    for (int t = 0; t < 4; ++t) {
                        const float2 f = unpack_act2(w4[t]);
                        v[8 * j + 2 * t] = __uint_as_float(acc[8 * j + 2 * t]) + b2 + f.x;
                        v[8 * j + 2 * t + 1] = __uint_as_float(acc[8 * j + 2 * t + 1]) + b2 + f.y;
                    }
                }
                put_row(smem + ML_HS_OFF + mo * 32768, v);   // staging: output channels mo * 128 + rl -> boxes 2 mo + (rl >> 6)
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&bars[MB_O_DRAINED]);
            fence_proxy_async_smem();
            named_bar_sync(2, ML_EPI_WARPS * 32);
            if (storer) {
#pragma unroll
                for (int b = 0; b < 3; ++b) tma_store_3d(&tmO, smem + ML_HS_OFF + b * 16384, 0, b * 64, tile * 2);
                bulk_commit_group();
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
