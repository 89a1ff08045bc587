// hm_layers.cu — the small non-GEMM layer of RISEv3 outside the trunk kernel (SURVEY.md §8a row F4
// tail): the value-head tail.  (The ECA gates, row F3, live inside trunk_kernel, hm_trunk.cuh.)
// Activations are NCHW 16-bit: x[pos][ch][sq], sq = rank*8 + file (the reference's own layout).
#include "hm_kernels.h"
#include "hm_ptx.cuh"

namespace hm {

// ---------------------------------------------------------------------------------------
// Value-head tail (reference: builder_util.py:311-324): flatten [16,8,8] channel-major,
// wdl = Linear(1024,3), plys = sigmoid(Linear(1024,1)), value = softmax(wdl)[2]-softmax(wdl)[0].
// One warp per position over the 1024 features [16 ch][64 sq]; w4 = [1024][4] (3 wdl rows + plys).
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128)
value_tail_kernel(const act_t* __restrict__ feat, const float* __restrict__ w4,
                  const float* __restrict__ b4, int n, float* __restrict__ raw5,
                  __half* __restrict__ value_h, __half* __restrict__ wdl_h,
                  __half* __restrict__ ml_h) {
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (warp >= n) return;
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    // feat[pos] is the reference's flatten order already: i = c*64 + sq (builder_util.py:311)
    const uint4* f = reinterpret_cast<const uint4*>(feat + static_cast<size_t>(warp) * 1024);
#pragma unroll
    for (int t = 0; t < 4; ++t) {
        const int v8 = lane + 32 * t;
        const uint4 v = __ldg(f + v8);
        const uint32_t u[4] = {v.x, v.y, v.z, v.w};
        const float4* wp = reinterpret_cast<const float4*>(w4) + static_cast<size_t>(v8) * 8;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const float2 x = unpack_act2(u[q]);
            const float4 wa = __ldg(wp + 2 * q), wb = __ldg(wp + 2 * q + 1);
            acc[0] = fmaf(x.x, wa.x, acc[0]); acc[1] = fmaf(x.x, wa.y, acc[1]);
            acc[2] = fmaf(x.x, wa.z, acc[2]); acc[3] = fmaf(x.x, wa.w, acc[3]);
            acc[0] = fmaf(x.y, wb.x, acc[0]); acc[1] = fmaf(x.y, wb.y, acc[1]);
            acc[2] = fmaf(x.y, wb.z, acc[2]); acc[3] = fmaf(x.y, wb.w, acc[3]);
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[j] += __shfl_xor_sync(0xffffffffu, acc[j], o);
    }
    if (lane == 0) {
        const float l = acc[0] + b4[0], d = acc[1] + b4[1], w = acc[2] + b4[2];
        const float pl = acc[3] + b4[3];
        const float mx = fmaxf(l, fmaxf(d, w));
        const float el = expf(l - mx), ed = expf(d - mx), ew = expf(w - mx);
        const float inv = 1.0f / (el + ed + ew);
        const float value = ew * inv - el * inv;       // builder_util.py:322-324
        const float ml = 1.0f / (1.0f + expf(-pl));    // sigmoid, builder_util.py:287-290
        float* r = raw5 + static_cast<size_t>(warp) * 5;
        r[0] = value; r[1] = l; r[2] = d; r[3] = w; r[4] = ml;
        if (value_h) value_h[warp] = __float2half_rn(value);
        if (wdl_h) {
            wdl_h[warp * 3 + 0] = __float2half_rn(l);
            wdl_h[warp * 3 + 1] = __float2half_rn(d);
            wdl_h[warp * 3 + 2] = __float2half_rn(w);
        }
        if (ml_h) ml_h[warp] = __float2half_rn(ml);
    }
}

void launch_value_tail(const act_t* feat, const float* w4, const float* b4, int n,
                       float* raw5, __half* value_h, __half* wdl_h, __half* ml_h,
                       cudaStream_t s) {
    if (n <= 0) return;
    value_tail_kernel<<<(n + 3) / 4, 128, 0, s>>>(feat, w4, b4, n, raw5, value_h, wdl_h, ml_h);
}

}  // namespace hm
