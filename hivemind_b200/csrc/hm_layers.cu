// hm_layers.cu — the bandwidth-bound layers of RISEv3 (SURVEY.md §8a rows F2b, F3, F4 tail,
// F6): depthwise k x k + BN + ReLU, ECA squeeze/excite, value-head tail, policy layout.
// Activations are NHWC 16-bit: x[pos][sq][C], sq = rank*8 + file.
#include "hm_kernels.h"
#include "hm_ptx.cuh"

namespace hm {

// ---------------------------------------------------------------------------------------
// Depthwise conv (reference: builder_util.py:468-470, Conv2d(groups=C) + BN + ReLU).
// Block = one position x 64 channels: the 64x64 16-bit tile (8 KB) is staged in shared memory
// once, every output reads its k*k neighbours from there; one work item = one square x 8
// channels (16-byte load/store).  Reads and writes each element of the operating tensor
// exactly once from/to global memory.
// ---------------------------------------------------------------------------------------
constexpr int DW_CH = 64;
constexpr int DW_THREADS = 256;

template <int KS>
__global__ void __launch_bounds__(DW_THREADS)
depthwise_kernel(const act_t* __restrict__ in, act_t* __restrict__ out,
                 const float* __restrict__ w, const float* __restrict__ bias, int C) {
    __shared__ __align__(16) act_t tile[64][DW_CH];      // [sq][ch]
    __shared__ __align__(16) float ws[KS * KS][DW_CH];
    __shared__ __align__(16) float bs[DW_CH];
    const int pos = blockIdx.x;
    const int c0 = blockIdx.y * DW_CH;
    const int cw = min(DW_CH, C - c0);  // multiple of 8
    const int groups = cw >> 3;
    const act_t* src = in + static_cast<size_t>(pos) * 64 * C + c0;
    for (int it = threadIdx.x; it < 64 * 8; it += DW_THREADS) {
        const int sq = it >> 3, g = it & 7;
        if (g < groups)
            *reinterpret_cast<uint4*>(&tile[sq][g * 8]) =
                __ldg(reinterpret_cast<const uint4*>(src + static_cast<size_t>(sq) * C + g * 8));
    }
    for (int it = threadIdx.x; it < KS * KS * DW_CH; it += DW_THREADS) {
        const int t = it / DW_CH, c = it - t * DW_CH;
        ws[t][c] = c < cw ? __ldg(w + static_cast<size_t>(t) * C + c0 + c) : 0.0f;
    }
    if (threadIdx.x < DW_CH) bs[threadIdx.x] = threadIdx.x < cw ? __ldg(bias + c0 + threadIdx.x) : 0.0f;
    __syncthreads();

    act_t* dst = out + static_cast<size_t>(pos) * 64 * C + c0;
    constexpr int H = KS / 2;
    for (int it = threadIdx.x; it < 64 * 8; it += DW_THREADS) {
        const int sq = it >> 3, g = it & 7;
        if (g >= groups) continue;
        const int y = sq >> 3, x = sq & 7;
        float acc[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] = bs[g * 8 + j];
#pragma unroll
        for (int ky = 0; ky < KS; ++ky) {
            const int yy = y + ky - H;
            if (yy < 0 || yy > 7) continue;
#pragma unroll
            for (int kx = 0; kx < KS; ++kx) {
                const int xx = x + kx - H;
                if (xx < 0 || xx > 7) continue;
                const uint4 v = *reinterpret_cast<const uint4*>(&tile[yy * 8 + xx][g * 8]);
                const float4 w0 = *reinterpret_cast<const float4*>(&ws[ky * KS + kx][g * 8]);
                const float4 w1 = *reinterpret_cast<const float4*>(&ws[ky * KS + kx][g * 8 + 4]);
                const float2 a = unpack_act2(v.x), b = unpack_act2(v.y),
                             c = unpack_act2(v.z), d = unpack_act2(v.w);
                acc[0] = fmaf(a.x, w0.x, acc[0]); acc[1] = fmaf(a.y, w0.y, acc[1]);
                acc[2] = fmaf(b.x, w0.z, acc[2]); acc[3] = fmaf(b.y, w0.w, acc[3]);
                acc[4] = fmaf(c.x, w1.x, acc[4]); acc[5] = fmaf(c.y, w1.y, acc[5]);
                acc[6] = fmaf(d.x, w1.z, acc[6]); acc[7] = fmaf(d.y, w1.w, acc[7]);
            }
        }
        uint4 o;
        o.x = pack_act2(fmaxf(acc[0], 0.f), fmaxf(acc[1], 0.f));
        o.y = pack_act2(fmaxf(acc[2], 0.f), fmaxf(acc[3], 0.f));
        o.z = pack_act2(fmaxf(acc[4], 0.f), fmaxf(acc[5], 0.f));
        o.w = pack_act2(fmaxf(acc[6], 0.f), fmaxf(acc[7], 0.f));
        *reinterpret_cast<uint4*>(dst + static_cast<size_t>(sq) * C + g * 8) = o;
    }
}

void launch_depthwise(const act_t* in, act_t* out, const float* w, const float* b,
                      int n, int C, int ksize, cudaStream_t s) {
    if (n <= 0) return;
    dim3 grid(n, (C + DW_CH - 1) / DW_CH);
    if (ksize == 3) depthwise_kernel<3><<<grid, DW_THREADS, 0, s>>>(in, out, w, b, C);
    else depthwise_kernel<5><<<grid, DW_THREADS, 0, s>>>(in, out, w, b, C);
}

// ---------------------------------------------------------------------------------------
// ECA (reference: builder_util.py:49-80).  Conv1d(C, C, k=5, pad=2) runs on a length-1
// sequence, so only the centre tap meets data: gate = hardsigmoid(Wc * mean_hw(x) + b),
// out = x * gate.  Block = ECA_P positions x 384 channels, thread = channel.
// ---------------------------------------------------------------------------------------
constexpr int ECA_C = 384;
constexpr int ECA_P = 4;

__global__ void __launch_bounds__(ECA_C)
eca_kernel(const act_t* __restrict__ in, act_t* __restrict__ out,
           const act_t* __restrict__ wt, const float* __restrict__ bias, int n) {
    __shared__ float mean_s[ECA_P][ECA_C];
    __shared__ float gate_s[ECA_P][ECA_C];
    const int p0 = blockIdx.x * ECA_P;
    const int np = min(ECA_P, n - p0);
    const int c = threadIdx.x;
    for (int p = 0; p < np; ++p) {
        const act_t* x = in + static_cast<size_t>(p0 + p) * 64 * ECA_C + c;
        float s = 0.0f;
#pragma unroll 8
        for (int sq = 0; sq < 64; ++sq) s += act_to_float(x[static_cast<size_t>(sq) * ECA_C]);
        mean_s[p][c] = s * (1.0f / 64.0f);  // AdaptiveAvgPool2d(1)
    }
    __syncthreads();
    float acc[ECA_P];
#pragma unroll
    for (int p = 0; p < ECA_P; ++p) acc[p] = 0.0f;
    for (int k = 0; k < ECA_C; ++k) {
        const float wv = act_to_float(wt[static_cast<size_t>(k) * ECA_C + c]);
#pragma unroll
        for (int p = 0; p < ECA_P; ++p) acc[p] = fmaf(wv, mean_s[p][k], acc[p]);
    }
    const float bv = bias[c];
#pragma unroll
    for (int p = 0; p < ECA_P; ++p) {
        // torch.nn.Hardsigmoid: relu6(x + 3) / 6
        gate_s[p][c] = fminf(fmaxf((acc[p] + bv) * (1.0f / 6.0f) + 0.5f, 0.0f), 1.0f);
    }
    __syncthreads();
    // scale: 16-byte vectors, 48 per square row
    for (int p = 0; p < np; ++p) {
        const uint4* xi = reinterpret_cast<const uint4*>(in + static_cast<size_t>(p0 + p) * 64 * ECA_C);
        uint4* xo = reinterpret_cast<uint4*>(out + static_cast<size_t>(p0 + p) * 64 * ECA_C);
        for (int i = threadIdx.x; i < 64 * (ECA_C / 8); i += ECA_C) {
            const int g = i % (ECA_C / 8);
            const uint4 v = xi[i];
            const float* gs = &gate_s[p][g * 8];
            float2 a = unpack_act2(v.x), b = unpack_act2(v.y), cc = unpack_act2(v.z),
                   d = unpack_act2(v.w);
            uint4 o;
            o.x = pack_act2(a.x * gs[0], a.y * gs[1]);
            o.y = pack_act2(b.x * gs[2], b.y * gs[3]);
            o.z = pack_act2(cc.x * gs[4], cc.y * gs[5]);
            o.w = pack_act2(d.x * gs[6], d.y * gs[7]);
            xo[i] = o;
        }
    }
}

// Gate only (fused path): gate[pos][c] = hardsigmoid(Wc * mean_hw(x[pos]) + b) as fp32; the
// scaling itself happens inside mbconv_kernel on the shared-memory input tile.
constexpr int ECA_GP = 2;
__global__ void __launch_bounds__(ECA_C)
eca_gate_kernel(const act_t* __restrict__ in, const act_t* __restrict__ wt,
                const float* __restrict__ bias, float* __restrict__ gate, int n) {
    __shared__ float mean_s[ECA_GP][ECA_C];
    const int p0 = blockIdx.x * ECA_GP;
    const int np = min(ECA_GP, n - p0);
    const int c = threadIdx.x;
    for (int p = 0; p < np; ++p) {
        const act_t* x = in + static_cast<size_t>(p0 + p) * 64 * ECA_C + c;
        float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
#pragma unroll 4
        for (int sq = 0; sq < 64; sq += 4) {
            s0 += act_to_float(x[static_cast<size_t>(sq) * ECA_C]);
            s1 += act_to_float(x[static_cast<size_t>(sq + 1) * ECA_C]);
            s2 += act_to_float(x[static_cast<size_t>(sq + 2) * ECA_C]);
            s3 += act_to_float(x[static_cast<size_t>(sq + 3) * ECA_C]);
        }
        mean_s[p][c] = ((s0 + s1) + (s2 + s3)) * (1.0f / 64.0f);
    }
    for (int p = np; p < ECA_GP; ++p) mean_s[p][c] = 0.f;
    __syncthreads();
    float acc[ECA_GP];
#pragma unroll
    for (int p = 0; p < ECA_GP; ++p) acc[p] = 0.0f;
#pragma unroll 8
    for (int k = 0; k < ECA_C; ++k) {
        const float wv = act_to_float(wt[static_cast<size_t>(k) * ECA_C + c]);
#pragma unroll
        for (int p = 0; p < ECA_GP; ++p) acc[p] = fmaf(wv, mean_s[p][k], acc[p]);
    }
    const float bv = bias[c];
    for (int p = 0; p < np; ++p)
        gate[static_cast<size_t>(p0 + p) * ECA_C + c] = fminf(fmaxf((acc[p] + bv) * (1.0f / 6.0f) + 0.5f, 0.0f), 1.0f);
}
void launch_eca_gate(const act_t* in, const act_t* wt, const float* b, float* gate, int n,
                     cudaStream_t s) {
    if (n <= 0) return;
    eca_gate_kernel<<<(n + ECA_GP - 1) / ECA_GP, ECA_C, 0, s>>>(in, wt, b, gate, n);
}

void launch_eca(const act_t* in, act_t* out, const act_t* wt,
                const float* b, int n, cudaStream_t s) {
    if (n <= 0) return;
    eca_kernel<<<(n + ECA_P - 1) / ECA_P, ECA_C, 0, s>>>(in, out, wt, b, n);
}

// ---------------------------------------------------------------------------------------
// Value-head tail (reference: builder_util.py:311-324): flatten [16,8,8] channel-major,
// wdl = Linear(1024,3), plys = sigmoid(Linear(1024,1)), value = softmax(wdl)[2]-softmax(wdl)[0].
// One warp per position; lane handles squares lane and lane+32.
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
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const int sq = lane + 32 * h;
        const uint4* f = reinterpret_cast<const uint4*>(feat + (static_cast<size_t>(warp) * 64 + sq) * 16);
        const uint4 v0 = __ldg(f), v1 = __ldg(f + 1);
        const uint32_t u[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
        const float4* wp = reinterpret_cast<const float4*>(w4 + static_cast<size_t>(sq) * 16 * 4);
#pragma unroll
        for (int q = 0; q < 8; ++q) {
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

// ---------------------------------------------------------------------------------------
// Dense policy maps in the reference layout (rise_mobile_v3.py:55 `.view(-1, 4672)`,
// index = ch*64 + sq) as fp16, from the NHWC fp32 logits [pos*64+sq][POLICY_LD].
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
policy_dense_half_kernel(const float* __restrict__ logits, int n, __half* __restrict__ pi_a,
                         __half* __restrict__ pi_b) {
    __shared__ float tile[64][146 + 1];
    const int p = blockIdx.x;
    const float* src = logits + static_cast<size_t>(p) * 64 * POLICY_LD;
    for (int i = threadIdx.x; i < 64 * POLICY_LD; i += blockDim.x) {
        const int sq = i / POLICY_LD, c = i - sq * POLICY_LD;
        if (c < 146) tile[sq][c] = src[i];
    }
    __syncthreads();
    __half* a = pi_a + static_cast<size_t>(p) * HM_NB_POLICY_VALUES;
    __half* b = pi_b + static_cast<size_t>(p) * HM_NB_POLICY_VALUES;
    for (int i = threadIdx.x; i < HM_NB_POLICY_VALUES; i += blockDim.x) {
        const int ch = i >> 6, sq = i & 63;
        a[i] = __float2half_rn(tile[sq][ch]);
        b[i] = __float2half_rn(tile[sq][73 + ch]);
    }
}

void launch_policy_dense_half(const float* logits, int n, __half* pi_a, __half* pi_b,
                              cudaStream_t s) {
    if (n <= 0) return;
    policy_dense_half_kernel<<<n, 256, 0, s>>>(logits, n, pi_a, pi_b);
}

}  // namespace hm
