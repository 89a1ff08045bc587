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
            // Thread = (position, file x, 4-channel group): it owns the 8 squares of one file.
            // Input rows slide through registers: every Y element is loaded once per kx and feeds
            // up to KS outputs; rank validity is resolved at compile time, file validity by a
            // select (no branches), so all loads of a row are in flight together.
            wait_empty(&bars[BAR_Z_EMPTY + b], z_n[b]);
            {
                const int cg = ct & 15, xcol = (ct >> 4) & 7, pbase = (ct >> 7) << 6;
                const uint32_t sub = (cg & 1) * 8, unit = cg >> 1;
                const uint2 bd = *reinterpret_cast<const uint2*>(par + 256 + cg * 8);
                uint2 w[KS * KS];
#pragma unroll
                for (int t = 0; t < KS * KS; ++t) w[t] = *reinterpret_cast<const uint2*>(par + 384 + t * 128 + cg * 8);
                act2_t acc[8][2];
#pragma unroll
                for (int y = 0; y < 8; ++y) { acc[y][0] = as_act2(bd.x); acc[y][1] = as_act2(bd.y); }
                uint32_t xoff[KS];
                bool xok[KS];
#pragma unroll
                for (int kx = 0; kx < KS; ++kx) {
                    const int xx = xcol + kx - H;
                    xok[kx] = static_cast<unsigned>(xx) < 8u;
                    const int xc = xok[kx] ? xx : xcol;
                    // row & 7 == file, so the swizzle key is the (clamped) file
                    xoff[kx] = static_cast<uint32_t>(xc) * 128u + ((unit ^ static_cast<uint32_t>(xc)) << 4) + sub;
                }
                const uint8_t* ybase = smem + S::Y_OFF + pbase * 128;
#pragma unroll
                for (int yy = 0; yy < 8; ++yy) {
                    uint2 v[KS];
#pragma unroll
                    for (int kx = 0; kx < KS; ++kx) {
                        v[kx] = *reinterpret_cast<const uint2*>(ybase + yy * 1024 + xoff[kx]);
                        if (!xok[kx]) v[kx] = make_uint2(0u, 0u);
                    }
#pragma unroll
                    for (int ky = 0; ky < KS; ++ky) {
                        const int y = yy - ky + H;
                        if (y < 0 || y > 7) continue;  // compile-time after unrolling
#pragma unroll
                        for (int kx = 0; kx < KS; ++kx) {
                            acc[y][0] = __hfma2(as_act2(v[kx].x), as_act2(w[ky * KS + kx].x), acc[y][0]);
                            acc[y][1] = __hfma2(as_act2(v[kx].y), as_act2(w[ky * KS + kx].y), acc[y][1]);
                        }
                    }
                }
                const act2_t zero = as_act2(0u);
                uint8_t* zbuf = smem + S::Z_OFF + b * 16384 + pbase * 128;
                const uint32_t zoff = static_cast<uint32_t>(xcol) * 128u + ((unit ^ static_cast<uint32_t>(xcol)) << 4) + sub;
#pragma unroll
                for (int y = 0; y < 8; ++y) {
                    uint2 o;
                    o.x = as_u32(__hmax2(acc[y][0], zero));
                    o.y = as_u32(__hmax2(acc[y][1], zero));
                    *reinterpret_cast<uint2*>(zbuf + y * 1024 + zoff) = o;
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
        {
            const size_t grow = static_cast<size_t>(tile) * 128 + trow;
            const act_t* xr = p.x + grow * MB_C + colhalf * 192;
            act_t* orow = p.out + grow * MB_C + colhalf * 192;
            const float* gs = gates + (tile_n & 1) * 2 * MB_C + (trow >> 6) * MB_C + colhalf * 192;
            const float* b2 = b2s + colhalf * 192;
            uint4 rv[2][4];
#pragma unroll
            for (int q = 0; q < 4; ++q) rv[0][q] = __ldg(reinterpret_cast<const uint4*>(xr) + q);
            wait_full(&bars[BAR_D2_FULL], tile_n);
            tc_fence_after();
#pragma unroll
            for (int cc = 0; cc < 6; ++cc) {
                const int col0 = cc * 32;
                uint32_t acc[32];
                tmem_ld_32x32b_x32(tm_d2 + lane_off + colhalf * 192 + col0, acc);
                if (cc + 1 < 6) {
#pragma unroll
                    for (int q = 0; q < 4; ++q) rv[(cc + 1) & 1][q] = __ldg(reinterpret_cast<const uint4*>(xr + col0 + 32) + q);
                }
                tmem_ld_wait();
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const uint4 rq = rv[cc & 1][q];
                    const uint32_t ru[4] = {rq.x, rq.y, rq.z, rq.w};
                    uint32_t ou[4];
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        const int j = q * 8 + e * 2;
                        float2 r = unpack_act2(ru[e]);
                        if (has_gate) r = unpack_act2(pack_act2(r.x * gs[col0 + j], r.y * gs[col0 + j + 1]));
                        ou[e] = pack_act2(__uint_as_float(acc[j]) + b2[col0 + j] + r.x,
                                          __uint_as_float(acc[j + 1]) + b2[col0 + j + 1] + r.y);
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
