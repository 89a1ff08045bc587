// hm_layers.cu — the small non-GEMM layers of RISEv3 (SURVEY.md §8a rows F3, F4 tail):
// ECA gate from channel sums, value-head tail.  Activations are NCHW 16-bit: x[pos][ch][sq],
// sq = rank*8 + file (the reference's own layout).
#include "hm_kernels.h"
#include "hm_ptx.cuh"

namespace hm {

// ---------------------------------------------------------------------------------------
// ECA gate (reference: builder_util.py:49-80).  Conv1d(C, C, k=5, pad=2) runs on a length-1
// sequence, so only the centre tap meets data: gate = hardsigmoid(Wc * mean_hw(x) + b).
// The channel sums over the 64 squares come from the epilogue of the trunk kernel that
// produced x (hm_trunk.cuh); the scaling x * gate happens on the shared-memory tile of the
// trunk kernel that consumes it.  Block = 8 positions x 128 output channels; thread = two
// adjacent channels x one quarter of the 384-long dot product (partials combined in a fixed order).
// ---------------------------------------------------------------------------------------
constexpr int ECA_C = 384;
constexpr int ECA_GP = 8;      // positions per CTA
constexpr int ECA_CB = 128;    // output channels per CTA
constexpr int ECA_KS = 8;      // k slices (threads = 64 channel pairs x 8 slices)
constexpr int ECA_THREADS = 64 * ECA_KS;
__global__ void __launch_bounds__(ECA_THREADS)
eca_gate_kernel(const float* __restrict__ sums, const act_t* __restrict__ wt,
                const float* __restrict__ bias, float* __restrict__ gate, int n) {
    __shared__ __align__(16) float mean_s[ECA_C][ECA_GP];   // [k][pos]
    __shared__ float part_s[ECA_KS][ECA_GP][ECA_CB];  // per-slice partial sums (combined in a fixed order)
    const int p0 = blockIdx.x * ECA_GP;
    const int c0 = blockIdx.y * ECA_CB;
    const int np = min(ECA_GP, n - p0);
    {
        constexpr int PER = ECA_C * ECA_GP / ECA_THREADS;  // independent loads, all in flight together
        float v[PER];
#pragma unroll
        for (int j = 0; j < PER; ++j) {
            const int i = threadIdx.x + j * ECA_THREADS;
            const int p = i / ECA_C, k = i - p * ECA_C;  // coalesced over k
            v[j] = p < np ? __ldg(sums + static_cast<size_t>(p0 + p) * ECA_C + k) : 0.0f;
        }
#pragma unroll
        for (int j = 0; j < PER; ++j) {
            const int i = threadIdx.x + j * ECA_THREADS;
            const int p = i / ECA_C, k = i - p * ECA_C;
            mean_s[k][p] = v[j] * (1.0f / 64.0f);  // AdaptiveAvgPool2d(1)
        }
    }
    __syncthreads();
    const int cp = threadIdx.x & 63, ks = threadIdx.x >> 6;
    const int c = c0 + 2 * cp;
    float a0[ECA_GP], a1[ECA_GP];
#pragma unroll
    for (int p = 0; p < ECA_GP; ++p) { a0[p] = 0.0f; a1[p] = 0.0f; }
    const int k0 = ks * (ECA_C / ECA_KS);
#pragma unroll 16
    for (int k = k0; k < k0 + ECA_C / ECA_KS; ++k) {
        const uint32_t w2 = __ldg(reinterpret_cast<const uint32_t*>(wt + static_cast<size_t>(k) * ECA_C + c));
        const float2 w = unpack_act2(w2);
        const float4 m0 = *reinterpret_cast<const float4*>(&mean_s[k][0]);
        const float4 m1 = *reinterpret_cast<const float4*>(&mean_s[k][4]);
        const float m[8] = {m0.x, m0.y, m0.z, m0.w, m1.x, m1.y, m1.z, m1.w};
#pragma unroll
        for (int p = 0; p < ECA_GP; ++p) { a0[p] = fmaf(w.x, m[p], a0[p]); a1[p] = fmaf(w.y, m[p], a1[p]); }
    }
#pragma unroll
    for (int p = 0; p < ECA_GP; ++p) {
        part_s[ks][p][2 * cp] = a0[p];
        part_s[ks][p][2 * cp + 1] = a1[p];
    }
    __syncthreads();
    for (int i = threadIdx.x; i < np * ECA_CB; i += ECA_THREADS) {
        const int p = i / ECA_CB, cc = i - p * ECA_CB;
        // torch.nn.Hardsigmoid: relu6(x + 3) / 6
        float acc = 0.0f;
#pragma unroll
        for (int q = 0; q < ECA_KS; ++q) acc += part_s[q][p][cc];
        gate[static_cast<size_t>(p0 + p) * ECA_C + c0 + cc] =
            fminf(fmaxf((acc + bias[c0 + cc]) * (1.0f / 6.0f) + 0.5f, 0.0f), 1.0f);
    }
}
void launch_eca_gate(const float* sums, const act_t* wt, const float* b, float* gate, int n,
                     cudaStream_t s) {
    if (n <= 0) return;
    eca_gate_kernel<<<dim3((n + ECA_GP - 1) / ECA_GP, ECA_C / ECA_CB), ECA_THREADS, 0, s>>>(sums, wt, b, gate, n);
}

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
