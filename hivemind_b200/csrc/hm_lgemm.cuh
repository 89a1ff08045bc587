// hm_lgemm.cuh — persistent tcgen05 GEMM for the shallow-K 1x1 layers (crossboard token projections,
// q / k|v / out-proj / fc1 / fc2 Linears, value fusion: K = 192..384, rise_mobile_v3.py:242-278,301-329).
//
// cgemm_kernel gives such a layer one CTA per (2 positions, 128 output channels): 12-24 MMAs (0.4-0.8 us of
// tensor work) behind ~4 us of CTA start-up, two exposed TMA round trips and an epilogue, i.e. the layer is
// bound by per-CTA latency (14-25 % of the tensor roofline).  Here a CTA stays on its SM:
//   * the 128 x K weight block of its output-channel block is loaded ONCE and stays resident (<= 96 KB);
//   * it walks position pairs t = blockIdx.x, += gridDim.x; activations stream through a 4-5-stage ring;
//   * TWO TMEM accumulators (2 x 128 columns): the MMAs of tile i+1 run while the 4 epilogue warps drain tile i.
// Same operand layouts, descriptors, padding rules and epilogue semantics as cgemm_kernel (hm_cgemm.cuh).
#pragma once
#include "hm_cgemm.cuh"
#include "hm_cconv.cuh"   // CC_MARK (bring-up timeline)

namespace hm {

constexpr int LG_XSTAGES = 5;     // barrier slots; a launch uses n_x = 7 - kpairs of them (224 KB with the resident weights)
constexpr int LG_EPI_GROUPS = 4;   // epilogue warps per TMEM lane quarter: each drains one 32-column chunk of a tile
constexpr int LG_THREADS = 64 + 4 * LG_EPI_GROUPS * 32;   // producer + MMA + 16 epilogue warps: with ONE CTA per SM
                                                          // the epilogue (GELU, residual, stores) needs the threads
                                                          // that three co-resident cgemm CTAs used to provide
constexpr int LG_MAX_KPAIRS = 4;   // K <= 512 (the folded token projection: 384 trunk channels + the 74 planes)
// one 32 KB stage is the OUTPUT staging tile (16-bit outputs): [2 boxes of 64 channels][2 positions][64 squares], 128-byte
// swizzled like the input boxes, written by the epilogue warps and stored by TMA.  (Each thread owns one channel's
// 64-byte half row, so direct stores put the 32 lanes of an instruction in 32 different 128-byte lines: ~8 B/clk per SM,
// 2.2 TB/s over the chip — the Linears were bound by their own epilogue stores, not by HBM.)
__host__ __device__ constexpr int lg_xstages(int kpairs) { return 6 - kpairs < LG_XSTAGES ? 6 - kpairs : LG_XSTAGES; }
constexpr int lg_smem_bytes(int kpairs) { return (kpairs + lg_xstages(kpairs) + 1) * CG_STAGE_BYTES + 256 + 1024; }

// erf by Abramowitz & Stegun 7.1.28, erf(z) = 1 - (1 + a1 z + ... + a6 z^6)^-16 (|error| <= 3e-7 on erf, far below
// the 16-bit output rounding): 10 FMA-pipe instructions + ONE MUFU instead of erff's ~40 instructions — with one
// CTA per SM the exact-erf GELU epilogue of fc1 was issue-bound (73 % issue-active, 2.8 us per tile against
// 0.4 us of MMAs).  (7.1.26, with an exp and a reciprocal, was MUFU-bound and no faster.)
__device__ __forceinline__ float gelu_erf_fast(float x) {
    const float z = fabsf(x) * 0.70710678118654752f;
    float p = fmaf(0.0000430638f, z, 0.0002765672f);
    p = fmaf(p, z, 0.0001520143f);
    p = fmaf(p, z, 0.0092705272f);
    p = fmaf(p, z, 0.0422820123f);
    p = fmaf(p, z, 0.0705230784f);
    p = fmaf(p, z, 1.0f);
    p *= p; p *= p; p *= p; p *= p;                 // ^16; overflows to +inf for large |x|: erf -> 1
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(p));   // one MUFU; rcp(+inf) = 0
    const float e = 1.0f - r;                       // erf(|x| / sqrt 2)
    return 0.5f * x * (1.0f + copysignf(e, x));
}

__global__ void __launch_bounds__(LG_THREADS)
lgemm_kernel(const CgemmParams p, const int n_tiles, const __grid_constant__ CUtensorMap tmW,
             const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmO,
             const __grid_constant__ CUtensorMap tmX2) {
    extern __shared__ uint8_t lg_smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(lg_smem_raw) + 1023) &
                                               ~static_cast<uintptr_t>(1023));
    const int kpairs = p.kblocks >> 1;
    uint8_t* smem_x = smem + kpairs * CG_STAGE_BYTES;                      // after the resident weights
    const int n_x = lg_xstages(kpairs);
    uint8_t* smem_o = smem_x + n_x * CG_STAGE_BYTES;                       // output staging tile
    uint64_t* w_full = reinterpret_cast<uint64_t*>(smem_o + CG_STAGE_BYTES);
    uint64_t* full_x = w_full + 1;
    uint64_t* empty_x = full_x + LG_XSTAGES;
    uint64_t* acc_full = empty_x + LG_XSTAGES;    // [2]
    uint64_t* acc_empty = acc_full + 2;           // [2]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 2);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int mb = blockIdx.y;
#ifdef HM_BRINGUP
    int trace_n = blockIdx.y == 0 ? 0 : CC_TRACE_CAP;   // only the CTA (0, 0) logs
#endif

    if (threadIdx.x == 0) {
        tma_prefetch_desc(&tmW);
        tma_prefetch_desc(&tmX);
        mbar_init(w_full, 1);
        for (int s = 0; s < LG_XSTAGES; ++s) {
            mbar_init(&full_x[s], 1);
            mbar_init(&empty_x[s], 1);
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(&acc_full[b], 1);
            mbar_init(&acc_empty[b], 4 * LG_EPI_GROUPS);   // one arrival per epilogue warp
        }
        fence_barrier_init();
    }
    if (warp == 1) {
        tmem_alloc(tmem_slot, 256u);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    // the upper 64 channels of the last K pair may be pure padding: neither loaded nor multiplied
    const bool upper_last = p.ksteps_last > 4;

    if (warp == 0) {
        // ===== TMA producer =====
        const bool leader = elect_one();
        if (leader) {
            const uint32_t bytes = static_cast<uint32_t>(kpairs) * CG_STAGE_BYTES - (upper_last ? 0u : CG_STAGE_BYTES / 2);
            mbar_arrive_expect_tx(w_full, bytes);
            for (int kp = 0; kp < kpairs; ++kp) {
                uint8_t* dst = smem + kp * CG_STAGE_BYTES;
                tma_load_2d(dst, &tmW, w_full, kp * 128, mb * 128);
                if (kp + 1 < kpairs || upper_last) tma_load_2d(dst + 16384, &tmW, w_full, kp * 128 + 64, mb * 128);
            }
        }
        __syncwarp();
        int xs = 0;
        uint32_t xphase = 1;
        for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
            for (int kp = 0; kp < kpairs; ++kp) {
                const bool upper = kp + 1 < kpairs || upper_last;
                mbar_wait(&empty_x[xs], xphase);
                if (leader) {
                    uint8_t* dst = smem_x + xs * CG_STAGE_BYTES;
                    mbar_arrive_expect_tx(&full_x[xs], upper ? CG_STAGE_BYTES : CG_STAGE_BYTES / 2);
                    if (p.kp_in2 >= 0 && kp >= p.kp_in2) {   // second input tensor (warp-uniform)
                        const int c0 = (kp - p.kp_in2) * 128;
                        tma_load_3d(dst, &tmX2, &full_x[xs], 0, c0, tile * 2);
                        if (upper) tma_load_3d(dst + 16384, &tmX2, &full_x[xs], 0, c0 + 64, tile * 2);
                    } else {
                        tma_load_3d(dst, &tmX, &full_x[xs], 0, kp * 128, tile * 2);
                        if (upper) tma_load_3d(dst + 16384, &tmX, &full_x[xs], 0, kp * 128 + 64, tile * 2);
                    }
                }
                __syncwarp();
                if (++xs == n_x) { xs = 0; xphase ^= 1; }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer =====
        const bool leader = elect_one();
        constexpr uint32_t idesc = umma_idesc_16_bmn(128, 128, UMMA_FMT_16);
        const uint64_t a0 = umma_desc_sw128(smem_u32(smem));
        const uint64_t b0 = umma_desc_mn_sw128(smem_u32(smem_x), 8192, 1024);
        mbar_wait(w_full, 0);
        int xs = 0, it = 0;
        uint32_t xphase = 0;
        for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++it) {
            const int buf = it & 1;
            const uint32_t use = static_cast<uint32_t>(it >> 1);
            CC_MARK(1, 1);
            mbar_wait(&acc_empty[buf], (use & 1u) ^ 1u);   // passes at once for the first use of a buffer
            tc_fence_after();
            CC_MARK(1, 2);
            const uint32_t d = tmem_base + buf * 128;
            for (int kp = 0; kp < kpairs; ++kp) {
                mbar_wait(&full_x[xs], xphase);
                tc_fence_after();
                if (leader) {
                    const uint32_t xo = xs * (CG_STAGE_BYTES / 16), wo = kp * (CG_STAGE_BYTES / 16);
                    const int nk = kp + 1 < kpairs ? 8 : p.ksteps_last;
#pragma unroll
                    for (int j = 0; j < 2; ++j) {
#pragma unroll
                        for (int k = 0; k < 4; ++k)
                            if (j * 4 + k < nk)
                                umma_f16(d, a0 + (wo + j * 1024 + 2 * k), b0 + (xo + j * 1024 + k * 128), idesc,
                                         (kp | j | k) ? 1u : 0u);
                    }
                    umma_commit(&empty_x[xs]);
                    if (kp == kpairs - 1) umma_commit(&acc_full[buf]);
                }
                __syncwarp();
                if (++xs == n_x) { xs = 0; xphase ^= 1; }
            }
            CC_MARK(1, 4);
        }
    } else {
        // ===== epilogue: warp = (TMEM lane quarter q, 32-column chunk c); thread = one output channel =====
        const int q = warp & 3;                 // a warp may only touch TMEM lanes 32*(warp % 4) ..
        const int c = (warp - 2) >> 2;          // chunk: position c >> 1, squares 32*(c & 1) ..
        const int chl = q * 32 + lane;          // channel within the 128-row output block
        const int ch = mb * 128 + chl;
        const bool valid = ch < p.n_valid;
        const float bias = (valid && p.bias != nullptr) ? __ldg(p.bias + ch) : 0.0f;
        // 16-bit outputs whose valid channels end on a 64-channel box boundary go through the staging tile + TMA store
        const bool staged = !p.out_f32 && (p.n_valid & 63) == 0;
        const bool storer = warp == 2 && lane == 0;
        uint8_t* my_row = smem_o + (chl >> 6) * 16384 + (c >> 1) * 8192 + (chl & 63) * 128;
        int it = 0;
        for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++it) {
            const int buf = it & 1;
            const int pos = tile * 2 + (c >> 1);
            const size_t row = (static_cast<size_t>(pos) * p.out_channels + ch) * 64 + (c & 1) * 32;
            uint4 res[4];   // the residual row is requested BEFORE the wait for the accumulator
            if (valid && p.residual != nullptr) {
                const uint4* rr = reinterpret_cast<const uint4*>(p.residual + row);
#pragma unroll
                for (int j = 0; j < 4; ++j) res[j] = __ldg(rr + j);
            }
            if (warp == 2) CC_MARK(2, 1);
            mbar_wait(&acc_full[buf], static_cast<uint32_t>(it >> 1) & 1u);
            tc_fence_after();
            if (warp == 2) CC_MARK(2, 2);
            const uint32_t t_row = tmem_base + buf * 128 + (static_cast<uint32_t>(q * 32) << 16);
            uint32_t acc[32];
            tmem_ld_32x32b_x32(t_row + c * 32, acc);
            tmem_ld_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&acc_empty[buf]);   // this warp's part is in registers
            float v[32];
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(acc[j]) + bias;
            if (valid && p.residual != nullptr) {
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const uint4 u = res[j];
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
                for (int j = 0; j < 32; ++j) v[j] = gelu_erf_fast(v[j]);
            }
            if (valid && p.post_add != nullptr) {
                const float4* e = reinterpret_cast<const float4*>(p.post_add + static_cast<size_t>(ch) * 64 + (c & 1) * 32);
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const float4 f = __ldg(e + j);
                    v[4 * j] += f.x; v[4 * j + 1] += f.y; v[4 * j + 2] += f.z; v[4 * j + 3] += f.w;
                }
            }
            if (p.out_f32) {
                if (valid) {
                    float4* o = reinterpret_cast<float4*>(static_cast<float*>(p.out) + row);
#pragma unroll
                    for (int j = 0; j < 8; ++j) o[j] = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
                }
                continue;
            }
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
                continue;
            }
            if (warp == 2) CC_MARK(2, 3);
            if (storer) bulk_wait_group_read0();          // the previous tile's stores have read the staging tile
            named_bar_sync(1, 32 * 4 * LG_EPI_GROUPS);
            if (warp == 2) CC_MARK(2, 5);
            if (valid) {
#pragma unroll
                for (int j = 0; j < 4; ++j)               // 16-byte piece (c & 1) * 4 + j of the row, 128-byte swizzle
                    *reinterpret_cast<uint4*>(my_row + ((((c & 1) * 4 + j) ^ (chl & 7)) << 4)) = w[j];
            }
            fence_proxy_async_smem();
            named_bar_sync(2, 32 * 4 * LG_EPI_GROUPS);
            if (storer) {
                if (mb * 128 < p.n_valid) tma_store_3d(&tmO, smem_o, 0, mb * 128, tile * 2);
                if (mb * 128 + 64 < p.n_valid) tma_store_3d(&tmO, smem_o + 16384, 0, mb * 128 + 64, tile * 2);
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
        tmem_dealloc(tmem_base, 256u);
    }
}

}  // namespace hm
