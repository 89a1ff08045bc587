// hm_layers.cu — the small non-GEMM layers of RISEv3 outside the trunk kernel (SURVEY.md §8a row F4
// tail): the value-head tail in its three reference configurations (WDL + plys, + value MLP, tanh).  (The ECA gates, row F3, live inside trunk_kernel, hm_trunk.cuh.)
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
                  const float* __restrict__ b4, const float* __restrict__ mlp, int n, float* __restrict__ raw5,
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
        float value = ew * inv - el * inv;             // builder_util.py:322-324
        const float ml = 1.0f / (1.0f + expf(-pl));    // sigmoid, builder_util.py:287-290
        if (mlp != nullptr) {
            // use_mlp_wdl_ply (builder_util.py:292-296, :317-319): value = tanh(Linear(8,1)(relu(Linear(4,8)(cat(wdl, plys)))))
            // mlp = [8][4] first weight | [8] first bias | [8] second weight | second bias
            const float in4[4] = {l, d, w, ml};
            float o = mlp[48];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                float hsum = mlp[32 + j];
#pragma unroll
                for (int i = 0; i < 4; ++i) hsum = fmaf(mlp[j * 4 + i], in4[i], hsum);
                o = fmaf(mlp[40 + j], fmaxf(hsum, 0.0f), o);
            }
            value = tanhf(o);
        }
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

void launch_value_tail(const act_t* feat, const float* w4, const float* b4, const float* mlp, int n,
                       float* raw5, __half* value_h, __half* wdl_h, __half* ml_h,
                       cudaStream_t s) {
    if (n <= 0) return;
    value_tail_kernel<<<(n + 3) / 4, 128, 0, s>>>(feat, w4, b4, mlp, n, raw5, value_h, wdl_h, ml_h);
}

// ---------------------------------------------------------------------------------------
// Tanh value head (reference: builder_util.py:297-301, :326 — the value path of nets built WITHOUT the WDL / plys
// heads): value = tanh(Linear(512,1)(relu(Linear(1024,512)(flatten)))).  CTA = 8 positions x 256 threads; thread t
// owns hidden units t and t + 256 for all 8 positions: the 2 MB fp32 weight matrix (stored transposed,
// w1t[feature][hidden], so a warp reads 128 contiguous bytes per feature) is read once per 8 positions from L2,
// the features sit in shared memory as fp32 and are read as broadcasts.  There are no auxiliary heads: the
// evaluator reports NaN WDL logits — the reference's shaping then keeps the scalar value
// (searchthread.cc:581-583) — and moves_left = 0 (no discount, :609-617).
// ---------------------------------------------------------------------------------------
constexpr int VT_POS = 8;
__global__ void __launch_bounds__(256)
value_tanh_kernel(const act_t* __restrict__ feat, const float* __restrict__ w1t, const float* __restrict__ b1,
                  const float* __restrict__ w2, const float* __restrict__ b2, int n, float* __restrict__ raw5,
                  __half* __restrict__ value_h, __half* __restrict__ wdl_h, __half* __restrict__ ml_h) {
    __shared__ __align__(16) float sf[VT_POS][1024];
    __shared__ float red[VT_POS][8];
    const int p0 = blockIdx.x * VT_POS;
    const int np = min(VT_POS, n - p0);
    for (int i = threadIdx.x; i < VT_POS * 512; i += 256) {   // 2 features per iteration
        const int p = i >> 9, f2 = i & 511;
        float2 v = make_float2(0.0f, 0.0f);
        if (p < np) v = unpack_act2(reinterpret_cast<const uint32_t*>(feat + static_cast<size_t>(p0 + p) * 1024)[f2]);
        sf[p][2 * f2] = v.x;
        sf[p][2 * f2 + 1] = v.y;
    }
    __syncthreads();
    float acc0[VT_POS], acc1[VT_POS];
    const float bb0 = __ldg(b1 + threadIdx.x), bb1 = __ldg(b1 + 256 + threadIdx.x);
#pragma unroll
    for (int p = 0; p < VT_POS; ++p) { acc0[p] = bb0; acc1[p] = bb1; }
#pragma unroll 4
    for (int f = 0; f < 1024; ++f) {
        const float wa = __ldg(w1t + static_cast<size_t>(f) * 512 + threadIdx.x);
        const float wb = __ldg(w1t + static_cast<size_t>(f) * 512 + 256 + threadIdx.x);
#pragma unroll
        for (int p = 0; p < VT_POS; ++p) {
            const float x = sf[p][f];
            acc0[p] = fmaf(wa, x, acc0[p]);
            acc1[p] = fmaf(wb, x, acc1[p]);
        }
    }
    const float v0 = __ldg(w2 + threadIdx.x), v1 = __ldg(w2 + 256 + threadIdx.x);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
    for (int p = 0; p < VT_POS; ++p) {
        float t = fmaf(v0, fmaxf(acc0[p], 0.0f), v1 * fmaxf(acc1[p], 0.0f));
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
        if (lane == 0) red[p][warp] = t;
    }
    __syncthreads();
    if (threadIdx.x < np) {
        const int p = threadIdx.x;
        float t = b2[0];
#pragma unroll
        for (int w = 0; w < 8; ++w) t += red[p][w];
        const float value = tanhf(t);
        const float qnan = __int_as_float(0x7fc00000);
        float* r = raw5 + static_cast<size_t>(p0 + p) * 5;
        r[0] = value; r[1] = qnan; r[2] = qnan; r[3] = qnan; r[4] = 0.0f;
        if (value_h) value_h[p0 + p] = __float2half_rn(value);
        if (wdl_h) {
            wdl_h[(p0 + p) * 3 + 0] = __float2half_rn(qnan);
            wdl_h[(p0 + p) * 3 + 1] = __float2half_rn(qnan);
            wdl_h[(p0 + p) * 3 + 2] = __float2half_rn(qnan);
        }
        if (ml_h) ml_h[p0 + p] = __float2half_rn(0.0f);
    }
}

void launch_value_tanh(const act_t* feat, const float* w1t, const float* b1, const float* w2, const float* b2, int n,
                       float* raw5, __half* value_h, __half* wdl_h, __half* ml_h, cudaStream_t s) {
    if (n <= 0) return;
    value_tanh_kernel<<<(n + VT_POS - 1) / VT_POS, 256, 0, s>>>(feat, w1t, b1, w2, b2, n, raw5, value_h, wdl_h, ml_h);
}

}  // namespace hm
