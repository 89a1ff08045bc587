// hm_mbconv_compute.cuh — the eight compute warps of mbconv_kernel (see hm_mbconv.cuh):
// TMEM -> Y epilogue of pw1, depthwise k x k on the shared-memory tile, ECA scaling of the
// input tile, and the final epilogue (bias + residual -> global).
#pragma once

namespace hm {

#ifdef HM_ACT_BF16
using act2_t = __nv_bfloat162;
#else
using act2_t = __half2;
#endif

__device__ __forceinline__ act2_t as_act2(uint32_t u) { return *reinterpret_cast<act2_t*>(&u); }
__device__ __forceinline__ uint32_t as_u32(act2_t v) { return *reinterpret_cast<uint32_t*>(&v); }

template <int KS>
__device__ void mbconv_compute(const MbconvParams& p, uint8_t* smem, uint64_t* bars, uint32_t tm_d2,
                               uint32_t tm_d1a, uint32_t tm_d1b) {
    using S = MbconvSmem<KS>;
    constexpr int PAR_STRIDE = (S::PAR_BYTES + 127) / 128 * 128;
    constexpr int H = KS / 2;
    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int ct = threadIdx.x - 64;        // 0..255
    const int quarter = warp & 3;           // TMEM lane quarter this warp may access
    const int colhalf = (warp - 2) >> 2;    // warps 2-5: 0, warps 6-9: 1
    const int trow = quarter * 32 + lane;   // tile row owned in TMEM loads
    const uint32_t lane_off = static_cast<uint32_t>(quarter * 32) << 16;
    const bool has_gate = p.gate != nullptr;
    const int nch = p.n_chunks;
    float* b2s = reinterpret_cast<float*>(smem + S::B2_OFF);
    float* gates = reinterpret_cast<float*>(smem + S::GATE_OFF);

    uint32_t tile_n = 0, par_n = 0;
    uint32_t d1_n[2] = {0, 0}, z_n[2] = {0, 0};

    // x' = x * gate on the freshly landed input tile (A operand of pw1), in place
    auto scale_x = [&](int tile, uint32_t tn) {
        float* gs = gates + (tn & 1) * 2 * MB_C;
        for (int i = ct; i < 2 * MB_C; i += MB_COMPUTE) gs[i] = __ldg(p.gate + static_cast<size_t>(tile) * 2 * MB_C + i);
        wait_full(&bars[BAR_X_FULL], tn);
        named_bar_sync(3, MB_COMPUTE);
#pragma unroll 4
        for (int i = 0; i < 24; ++i) {
            const int idx = ct + i * MB_COMPUTE;
            const int q = idx & 7, r = (idx >> 3) & 127, j = idx >> 10;
            const int ch = j * 64 + ((q ^ (r & 7)) << 3);
            uint4* px = reinterpret_cast<uint4*>(smem + S::X_OFF + j * 16384 + r * 128 + q * 16);
            const float* g = gs + (r >> 6) * MB_C + ch;
            uint4 v = *px;
            float2 a = unpack_act2(v.x), b = unpack_act2(v.y), c = unpack_act2(v.z), d = unpack_act2(v.w);
            v.x = pack_act2(a.x * g[0], a.y * g[1]);
            v.y = pack_act2(b.x * g[2], b.y * g[3]);
            v.z = pack_act2(c.x * g[4], c.y * g[5]);
            v.w = pack_act2(d.x * g[6], d.y * g[7]);
            *px = v;
        }
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) mbar_arrive(&bars[BAR_XS_READY]);
    };

    if (has_gate && static_cast<int>(blockIdx.x) < p.n_tiles) scale_x(blockIdx.x, 0);

    for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x, ++tile_n) {
        for (int c = 0; c < nch; ++c) {
            const int b = c & 1;
            const uint32_t pb = par_n % MB_PAR_BUFS;
            const uint8_t* par = smem + S::PAR_OFF + pb * PAR_STRIDE;
            wait_full(&bars[BAR_PAR_FULL + pb], par_n / MB_PAR_BUFS);
            // ---- pw1 epilogue: D1 chunk -> bias + ReLU -> Y (swizzled [row][64 ch]) ----
            wait_full(&bars[BAR_D1_FULL + b], d1_n[b]);
            tc_fence_after();
            {
                uint32_t acc[32];
                tmem_ld_32x32b_x32((b ? tm_d1b : tm_d1a) + lane_off + colhalf * 32, acc);
                tmem_ld_wait();
                const float* b1 = reinterpret_cast<const float*>(par) + colhalf * 32;
                uint8_t* yrow = smem + S::Y_OFF + trow * 128;
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    uint4 o;
                    const int j = i * 8;
                    o.x = pack_act2(fmaxf(__uint_as_float(acc[j + 0]) + b1[j + 0], 0.f), fmaxf(__uint_as_float(acc[j + 1]) + b1[j + 1], 0.f));
                    o.y = pack_act2(fmaxf(__uint_as_float(acc[j + 2]) + b1[j + 2], 0.f), fmaxf(__uint_as_float(acc[j + 3]) + b1[j + 3], 0.f));
                    o.z = pack_act2(fmaxf(__uint_as_float(acc[j + 4]) + b1[j + 4], 0.f), fmaxf(__uint_as_float(acc[j + 5]) + b1[j + 5], 0.f));
                    o.w = pack_act2(fmaxf(__uint_as_float(acc[j + 6]) + b1[j + 6], 0.f), fmaxf(__uint_as_float(acc[j + 7]) + b1[j + 7], 0.f));
                    const int u = colhalf * 4 + i;
                    *reinterpret_cast<uint4*>(yrow + ((u ^ (trow & 7)) << 4)) = o;
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&bars[BAR_D1_EMPTY + b]);
            ++d1_n[b];
            named_bar_sync(1, MB_COMPUTE);  // Y complete

            // ---- depthwise k x k + bias + ReLU: Y -> Z[b] (A operand of pw2) ----
            wait_empty(&bars[BAR_Z_EMPTY + b], z_n[b]);
            {
                const int g = ct & 7;
                const uint4 bd = *reinterpret_cast<const uint4*>(par + 256 + g * 16);
                const uint8_t* wbase = par + 384 + g * 16;
                uint8_t* zbuf = smem + S::Z_OFF + b * 16384;
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const int r = (ct >> 3) + 32 * i;
                    const int pbase = r & 64, y = (r >> 3) & 7, x = r & 7;
                    act2_t a0 = as_act2(bd.x), a1 = as_act2(bd.y), a2 = as_act2(bd.z), a3 = as_act2(bd.w);
#pragma unroll
                    for (int ky = 0; ky < KS; ++ky) {
                        const int yy = y + ky - H;
                        if (yy < 0 || yy > 7) continue;
#pragma unroll
                        for (int kx = 0; kx < KS; ++kx) {
                            const int xx = x + kx - H;
                            if (xx < 0 || xx > 7) continue;
                            const int nr = pbase + yy * 8 + xx;
                            const uint4 v = *reinterpret_cast<const uint4*>(smem + S::Y_OFF + nr * 128 + ((g ^ (nr & 7)) << 4));
                            const uint4 w = *reinterpret_cast<const uint4*>(wbase + (ky * KS + kx) * 128);
                            a0 = __hfma2(as_act2(v.x), as_act2(w.x), a0);
                            a1 = __hfma2(as_act2(v.y), as_act2(w.y), a1);
                            a2 = __hfma2(as_act2(v.z), as_act2(w.z), a2);
                            a3 = __hfma2(as_act2(v.w), as_act2(w.w), a3);
                        }
                    }
                    const act2_t zero = as_act2(0u);
                    uint4 o;
                    o.x = as_u32(__hmax2(a0, zero));
                    o.y = as_u32(__hmax2(a1, zero));
                    o.z = as_u32(__hmax2(a2, zero));
                    o.w = as_u32(__hmax2(a3, zero));
                    *reinterpret_cast<uint4*>(zbuf + r * 128 + ((g ^ (r & 7)) << 4)) = o;
                }
            }
            fence_proxy_async_smem();  // generic-proxy writes of Z -> visible to the tensor core
            __syncwarp();
            if (lane == 0) {
                mbar_arrive(&bars[BAR_Z_FULL + b]);
                mbar_arrive(&bars[BAR_PAR_EMPTY + pb]);
            }
            ++z_n[b];
            ++par_n;
            named_bar_sync(2, MB_COMPUTE);  // all reads of Y done before the next chunk overwrites it
        }

        // the next tile's input is already in flight: scale it before draining this tile
        const int next = tile + gridDim.x;
        if (has_gate && next < p.n_tiles) scale_x(next, tile_n + 1);

        // ---- final epilogue: D2 + bias2 + residual -> out ----
        wait_full(&bars[BAR_D2_FULL], tile_n);
        tc_fence_after();
        {
            const size_t grow = static_cast<size_t>(tile) * 128 + trow;
            const act_t* xr = p.x + grow * MB_C;
            act_t* orow = p.out + grow * MB_C;
            const float* gs = gates + (tile_n & 1) * 2 * MB_C + (trow >> 6) * MB_C;
#pragma unroll 1
            for (int cc = 0; cc < 6; ++cc) {
                const int col0 = colhalf * 192 + cc * 32;
                uint32_t acc[32];
                tmem_ld_32x32b_x32(tm_d2 + lane_off + col0, acc);
                uint4 rv[4];
#pragma unroll
                for (int q = 0; q < 4; ++q) rv[q] = __ldg(reinterpret_cast<const uint4*>(xr + col0) + q);
                tmem_ld_wait();
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const uint32_t ru[4] = {rv[q].x, rv[q].y, rv[q].z, rv[q].w};
                    uint32_t ou[4];
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        const int j = q * 8 + e * 2;
                        float2 r = unpack_act2(ru[e]);
                        if (has_gate) {
                            r = unpack_act2(pack_act2(r.x * gs[col0 + j], r.y * gs[col0 + j + 1]));
                        }
                        ou[e] = pack_act2(__uint_as_float(acc[j]) + b2s[col0 + j] + r.x,
                                          __uint_as_float(acc[j + 1]) + b2s[col0 + j + 1] + r.y);
                    }
                    reinterpret_cast<uint4*>(orow + col0)[q] = make_uint4(ou[0], ou[1], ou[2], ou[3]);
                }
            }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&bars[BAR_D2_EMPTY]);
    }
}

}  // namespace hm
