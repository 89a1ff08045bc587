// hm_xboard.cu — the non-GEMM kernels of `crossboard-risev33` (SURVEY.md §8a stage X;
// reference src/architectures/rise_mobile_v3.py:242-412, class CrossBoardRiseV3).
//
// Tokens stay in the network's channel-major layout: the token matrix of a board is the
// transpose of its map, so `tok[pos][board][192][64]` (fp16) is at once the [64 tokens][192]
// residual stream of the attention blocks and — read as [pos][384][64] — the
// cat(board_a_map, board_b_map) the value fusion conv wants (:397-404).  Every Linear of the
// blocks is therefore a 1x1 convolution on cgemm_kernel (hm_cgemm.cuh); what is left here is
// the bandwidth work around them: token embedding, context tokens, LayerNorms, and the
// 64x71 attention core.
#include <cuda_fp16.h>
#include <math.h>
#include <stdlib.h>

#include "hm_kernels.h"

namespace hm {

constexpr int XD = 192;      // attention_dim
constexpr int XH = 6;        // heads
constexpr int XDH = 32;      // head dim
constexpr int XCTX = 7;      // context tokens
constexpr float LN_EPS = 1e-5f;

// ---- X1/X2: board tokens ---------------------------------------------------------------
// pre: fp32 [n][384][64] = W_trunk(board) * trunk for both boards stacked (cgemm, no bias).
// tok[pos][b][o][sq] = relu(pre + bias[b][o] + sum_c wp[b][o][c] * planes[pos][37b + c][sq])
//                      + pos_emb_t[o][sq] + board_emb[b][o]        (rise_mobile_v3.py:374-385;
//                      pos_emb_t = the position embedding transposed to channel-major at load)
// grid (ceil(n / TE_POS), 2), 256 threads = 4 channel groups x 64 squares; a CTA stages the board's
// 192 x 37 plane weights in shared memory once and walks TE_POS positions with them.
constexpr int TE_POS = 4;
__global__ void __launch_bounds__(256)
token_embed_kernel(const float* __restrict__ pre, const act_t* __restrict__ planes,
                   const float* __restrict__ wp, const float* __restrict__ bias,
                   const float* __restrict__ pos_emb, const float* __restrict__ board_emb,
                   act_t* __restrict__ tok, int n) {
    __shared__ __align__(16) float s_wp[XD * 40];   // rows padded 37 -> 40 for 16-byte broadcast reads
    __shared__ float s_add[XD];                     // bias (before the ReLU)
    __shared__ float s_emb[XD];                     // board embedding (after it)
    const int b = blockIdx.y;
    for (int i = threadIdx.x; i < XD * 40; i += 256) {
        const int o = i / 40, c = i - o * 40;
        s_wp[i] = c < 37 ? wp[(b * XD + o) * 37 + c] : 0.0f;
    }
    if (threadIdx.x < XD) {
        s_add[threadIdx.x] = bias[b * XD + threadIdx.x];
        s_emb[threadIdx.x] = board_emb[b * XD + threadIdx.x];
    }
    const int sq = threadIdx.x & 63, grp = threadIdx.x >> 6;
    __syncthreads();
    // two positions per step: one 16-byte weight broadcast feeds 8 FMAs instead of 4 (the loop was bound by the
    // shared-memory weight reads, one LDS.128 per 4 FMAs)
    for (int pp = 0; pp < TE_POS; pp += 2) {
        const int pos0 = blockIdx.x * TE_POS + pp, pos1 = pos0 + 1;
        if (pos0 >= n) break;
        const bool has1 = pos1 < n;
        float pl0[40], pl1[40];
#pragma unroll
        for (int c = 0; c < 37; ++c) {
            pl0[c] = act_to_float(planes[(static_cast<size_t>(pos0) * 74 + 37 * b + c) * 64 + sq]);
            pl1[c] = has1 ? act_to_float(planes[(static_cast<size_t>(pos1) * 74 + 37 * b + c) * 64 + sq]) : 0.0f;
        }
        pl0[37] = pl0[38] = pl0[39] = pl1[37] = pl1[38] = pl1[39] = 0.0f;
        const float* prow0 = pre + (static_cast<size_t>(pos0) * 384 + b * XD) * 64 + sq;
        const float* prow1 = prow0 + (has1 ? 384 * 64 : 0);
        act_t* trow0 = tok + (static_cast<size_t>(pos0) * 384 + b * XD) * 64 + sq;
        act_t* trow1 = trow0 + 384 * 64;
#pragma unroll 1
        for (int o0 = grp * 48; o0 < grp * 48 + 48; o0 += 4) {
            float v0[4], v1[4], e[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                v0[j] = prow0[(o0 + j) * 64];
                v1[j] = prow1[(o0 + j) * 64];
                e[j] = pos_emb[(o0 + j) * 64 + sq];
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int o = o0 + j;
                float a0 = v0[j] + s_add[o], a1 = v1[j] + s_add[o];
                const float4* w = reinterpret_cast<const float4*>(s_wp + o * 40);
#pragma unroll
                for (int c4 = 0; c4 < 10; ++c4) {
                    const float4 ww = w[c4];
                    a0 = fmaf(ww.x, pl0[4 * c4], a0);      a1 = fmaf(ww.x, pl1[4 * c4], a1);
                    a0 = fmaf(ww.y, pl0[4 * c4 + 1], a0);  a1 = fmaf(ww.y, pl1[4 * c4 + 1], a1);
                    a0 = fmaf(ww.z, pl0[4 * c4 + 2], a0);  a1 = fmaf(ww.z, pl1[4 * c4 + 2], a1);
                    a0 = fmaf(ww.w, pl0[4 * c4 + 3], a0);  a1 = fmaf(ww.w, pl1[4 * c4 + 3], a1);
                }
                trow0[o * 64] = float_to_act(fmaxf(a0, 0.0f) + e[j] + s_emb[o]);
                if (has1) trow1[o * 64] = float_to_act(fmaxf(a1, 0.0f) + e[j] + s_emb[o]);
            }
        }
    }
}

// ---- X3: context tokens ------------------------------------------------------------------
// 4 pocket tokens Linear(5->192) on the spatial means of planes 12-16, 17-21, 49-53, 54-58 and
// 3 scalar tokens Linear(1->192) on the means of planes 25, 62, 31, plus the role embedding
// (rise_mobile_v3.py:353-369).  ctx: fp32 [n][7][192].  grid n, 192 threads.
__global__ void __launch_bounds__(XD)
context_tokens_kernel(const act_t* __restrict__ planes, const float* __restrict__ pocket_w,
                      const float* __restrict__ pocket_b, const float* __restrict__ scalar_w,
                      const float* __restrict__ scalar_b, const float* __restrict__ role,
                      float* __restrict__ ctx) {
    __shared__ float s_mean[23];
    const int pos = blockIdx.x;
    if (threadIdx.x < 23) {
        const int t = threadIdx.x;
        const int plane = t < 5 ? 12 + t : t < 10 ? 17 + (t - 5) : t < 15 ? 49 + (t - 10) : t < 20 ? 54 + (t - 15)
                        : t == 20 ? 25 : t == 21 ? 62 : 31;
        const act_t* p = planes + (static_cast<size_t>(pos) * 74 + plane) * 64;
        float s = 0.0f;
        for (int i = 0; i < 64; ++i) s += act_to_float(p[i]);
        s_mean[t] = s * (1.0f / 64.0f);
    }
    __syncthreads();
    const int o = threadIdx.x;
    float* out = ctx + static_cast<size_t>(pos) * XCTX * XD;
#pragma unroll
    for (int t = 0; t < 4; ++t) {
        float v = pocket_b[o];
#pragma unroll
        for (int j = 0; j < 5; ++j) v = fmaf(pocket_w[o * 5 + j], s_mean[t * 5 + j], v);
        out[t * XD + o] = v + role[t * XD + o];
    }
#pragma unroll
    for (int t = 0; t < 3; ++t)
        out[(4 + t) * XD + o] = fmaf(scalar_w[o], s_mean[20 + t], scalar_b[o]) + role[(4 + t) * XD + o];
}

// ---- LayerNorm over the 192 channels of each token ----------------------------------------
// x: [n2][192][64] (n2 = positions x boards).  One statistic, up to two affine outputs
// (query_norm and memory_norm read the same tokens, rise_mobile_v3.py:263-268).
// grid n2, 256 threads = 32 channel groups x 8 square groups: a thread holds 6 channels x 8 squares and moves
// them as 16-byte pieces (a warp = 4 channel groups x 8 square groups reads/writes 4 whole 128-byte rows per
// instruction); the per-square statistics are reduced over the 32 channel groups through shared memory.
__global__ void __launch_bounds__(256)
token_layernorm_kernel(const act_t* __restrict__ x, const float* __restrict__ g1,
                       const float* __restrict__ b1, act_t* __restrict__ y1,
                       const float* __restrict__ g2, const float* __restrict__ b2,
                       act_t* __restrict__ y2) {
    __shared__ float s_part[32][65];   // [channel group][square] partial sums (padded: conflict-free columns)
    __shared__ float s_stat[2][64];    // mean, rstd per square
    const size_t base = static_cast<size_t>(blockIdx.x) * XD * 64;
    const int sg = threadIdx.x & 7, cg = threadIdx.x >> 3;   // squares 8*sg .., channels 6*cg ..
    float v[6][8];
#pragma unroll
    for (int i = 0; i < 6; ++i) {
        const uint4 u = __ldg(reinterpret_cast<const uint4*>(x + base + static_cast<size_t>(cg * 6 + i) * 64 + sg * 8));
        const float2 a = unpack_act2(u.x), b = unpack_act2(u.y), c = unpack_act2(u.z), d = unpack_act2(u.w);
        v[i][0] = a.x; v[i][1] = a.y; v[i][2] = b.x; v[i][3] = b.y;
        v[i][4] = c.x; v[i][5] = c.y; v[i][6] = d.x; v[i][7] = d.y;
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        float s = 0.0f;
#pragma unroll
        for (int i = 0; i < 6; ++i) s += v[i][k];
        s_part[cg][sg * 8 + k] = s;
    }
    __syncthreads();
    if (threadIdx.x < 64) {
        float s = 0.0f;
#pragma unroll
        for (int r = 0; r < 32; ++r) s += s_part[r][threadIdx.x];
        s_stat[0][threadIdx.x] = s * (1.0f / XD);
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const float mean = s_stat[0][sg * 8 + k];
        float q = 0.0f;
#pragma unroll
        for (int i = 0; i < 6; ++i) {
            v[i][k] -= mean;
            q = fmaf(v[i][k], v[i][k], q);
        }
        s_part[cg][sg * 8 + k] = q;
    }
    __syncthreads();
    if (threadIdx.x < 64) {
        float q = 0.0f;
#pragma unroll
        for (int r = 0; r < 32; ++r) q += s_part[r][threadIdx.x];
        s_stat[1][threadIdx.x] = rsqrtf(q * (1.0f / XD) + LN_EPS);
    }
    __syncthreads();
    float rstd[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) rstd[k] = s_stat[1][sg * 8 + k];
#pragma unroll
    for (int i = 0; i < 6; ++i) {
        const int c = cg * 6 + i;
        const size_t off = base + static_cast<size_t>(c) * 64 + sg * 8;
        {
            const float g = g1[c], b = b1[c];
            uint4 o;
            o.x = pack_act2(fmaf(v[i][0] * rstd[0], g, b), fmaf(v[i][1] * rstd[1], g, b));
            o.y = pack_act2(fmaf(v[i][2] * rstd[2], g, b), fmaf(v[i][3] * rstd[3], g, b));
            o.z = pack_act2(fmaf(v[i][4] * rstd[4], g, b), fmaf(v[i][5] * rstd[5], g, b));
            o.w = pack_act2(fmaf(v[i][6] * rstd[6], g, b), fmaf(v[i][7] * rstd[7], g, b));
            *reinterpret_cast<uint4*>(y1 + off) = o;
        }
        if (y2 != nullptr) {
            const float g = g2[c], b = b2[c];
            uint4 o;
            o.x = pack_act2(fmaf(v[i][0] * rstd[0], g, b), fmaf(v[i][1] * rstd[1], g, b));
            o.y = pack_act2(fmaf(v[i][2] * rstd[2], g, b), fmaf(v[i][3] * rstd[3], g, b));
            o.z = pack_act2(fmaf(v[i][4] * rstd[4], g, b), fmaf(v[i][5] * rstd[5], g, b));
            o.w = pack_act2(fmaf(v[i][6] * rstd[6], g, b), fmaf(v[i][7] * rstd[7], g, b));
            *reinterpret_cast<uint4*>(y2 + off) = o;
        }
    }
}

// ---- K/V of the 7 context tokens -----------------------------------------------------------
// mem = memory_norm(ctx); K = Wk mem + bk, V = Wv mem + bv (nn.MultiheadAttention in_proj rows
// 192..575).  wkv_t: fp32 [192 in][384 out] (K outputs 0..191, V outputs 192..383), bkv [384].
// ckv: fp32 [n][7][384].  grid n, 384 threads.
constexpr int CKV_POS = 4;   // positions per CTA: the 295 KB weight matrix is read once for all of them
__global__ void __launch_bounds__(384)
context_kv_kernel(const float* __restrict__ ctx, const float* __restrict__ g, const float* __restrict__ b,
                  const float* __restrict__ wkv_t, const float* __restrict__ bkv, float* __restrict__ ckv, int n) {
    constexpr int NT = CKV_POS * XCTX;   // 28 tokens
    __shared__ __align__(16) float s_n[NT][XD];
    __shared__ float s_stat[NT][2];
    const int pos0 = blockIdx.x * CKV_POS;
    const int n_tok = min(CKV_POS, n - pos0) * XCTX;
    const float* c = ctx + static_cast<size_t>(pos0) * XCTX * XD;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int t = warp; t < n_tok; t += 12) {
        float s = 0.0f;
        for (int i = lane; i < XD; i += 32) s += c[t * XD + i];
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        const float mean = s * (1.0f / XD);
        float q = 0.0f;
        for (int i = lane; i < XD; i += 32) { const float d = c[t * XD + i] - mean; q = fmaf(d, d, q); }
        for (int o = 16; o > 0; o >>= 1) q += __shfl_xor_sync(0xffffffffu, q, o);
        if (lane == 0) { s_stat[t][0] = mean; s_stat[t][1] = rsqrtf(q * (1.0f / XD) + LN_EPS); }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < NT * XD; i += 384) {
        const int t = i / XD, ch = i - t * XD;
        s_n[t][ch] = t < n_tok ? fmaf((c[i] - s_stat[t][0]) * s_stat[t][1], g[ch], b[ch]) : 0.0f;
    }
    __syncthreads();
    const int o = threadIdx.x;
    float acc[NT];
#pragma unroll
    for (int t = 0; t < NT; ++t) acc[t] = bkv[o];
    // four input channels per step: the normalised tokens come out of shared memory as one 16-byte
    // broadcast per (token, step) instead of four 4-byte ones (the loop was LDS-bound: one LDS per FMA)
    for (int i = 0; i < XD; i += 4) {
        const float w0 = wkv_t[i * 384 + o], w1 = wkv_t[(i + 1) * 384 + o];
        const float w2 = wkv_t[(i + 2) * 384 + o], w3 = wkv_t[(i + 3) * 384 + o];
#pragma unroll
        for (int t = 0; t < NT; ++t) {
            const float4 sv = *reinterpret_cast<const float4*>(&s_n[t][i]);
            acc[t] = fmaf(w0, sv.x, acc[t]);   // same order of accumulation as the one-channel loop
            acc[t] = fmaf(w1, sv.y, acc[t]);
            acc[t] = fmaf(w2, sv.z, acc[t]);
            acc[t] = fmaf(w3, sv.w, acc[t]);
        }
    }
#pragma unroll
    for (int t = 0; t < NT; ++t)
        if (t < n_tok) ckv[(static_cast<size_t>(pos0) * XCTX + t) * 384 + o] = acc[t];
}

// ---- X4: cross-board attention core --------------------------------------------------------
// For board i of a position: queries = its 64 tokens, memory = the OTHER board's 64 tokens + the
// 7 context tokens (rise_mobile_v3.py:262-271), 6 heads of 32 channels, softmax(QK^T/sqrt(32))V.
// q:  fp16 [n2][192][64]  (Wq * query_norm(tokens) + bq, channel-major)
// kv: fp16 [n2][384][64]  (rows 0..191 = K, 192..383 = V of memory_norm(tokens))
// ckv: fp32 [n][7][384]   (context K | V)
// att: fp16 [n2][192][64] (heads concatenated, before out_proj)
// grid n2 = positions x boards, 384 threads: thread = (head, query square).
constexpr int XATT_SMEM = (384 * 64 + XCTX * 384) * 4;

__global__ void __launch_bounds__(384)
cross_attention_kernel(const act_t* __restrict__ q, const act_t* __restrict__ kv,
                       const float* __restrict__ ckv, act_t* __restrict__ att) {
    extern __shared__ float xs[];
    float* s_kv = xs;                 // [384][64]: K rows then V rows of the other board
    float* s_c = xs + 384 * 64;       // [7][384]
    const int n2 = blockIdx.x, pos = n2 >> 1, other = n2 ^ 1;
    {
        // coalesced fill: consecutive threads convert consecutive 16-byte pieces (8 values)
        const uint4* src = reinterpret_cast<const uint4*>(kv + static_cast<size_t>(other) * 384 * 64);
        float4* dst = reinterpret_cast<float4*>(s_kv);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const int i = threadIdx.x + 384 * j;
            const uint4 u = __ldg(src + i);
            const float2 a = unpack_act2(u.x), b = unpack_act2(u.y), c = unpack_act2(u.z), d = unpack_act2(u.w);
            dst[2 * i] = make_float4(a.x, a.y, b.x, b.y);
            dst[2 * i + 1] = make_float4(c.x, c.y, d.x, d.y);
        }
        for (int i = threadIdx.x; i < XCTX * 384; i += 384) s_c[i] = ckv[static_cast<size_t>(pos) * XCTX * 384 + i];
    }
    const int h = threadIdx.x >> 6, sq = threadIdx.x & 63;
    float qv[XDH];
    const float scale = 0.17677669529663689f;  // 1/sqrt(32)
#pragma unroll
    for (int c = 0; c < XDH; ++c)
        qv[c] = act_to_float(q[(static_cast<size_t>(n2) * XD + h * XDH + c) * 64 + sq]) * scale;
    __syncthreads();

    float s[64 + XCTX];
#pragma unroll
    for (int k = 0; k < 64 + XCTX; ++k) s[k] = 0.0f;
    const float* kh = s_kv + h * XDH * 64;
#pragma unroll
    for (int c = 0; c < XDH; ++c) {
        const float4* row = reinterpret_cast<const float4*>(kh + c * 64);
#pragma unroll
        for (int k4 = 0; k4 < 16; ++k4) {
            const float4 kk = row[k4];
            s[4 * k4 + 0] = fmaf(qv[c], kk.x, s[4 * k4 + 0]);
            s[4 * k4 + 1] = fmaf(qv[c], kk.y, s[4 * k4 + 1]);
            s[4 * k4 + 2] = fmaf(qv[c], kk.z, s[4 * k4 + 2]);
            s[4 * k4 + 3] = fmaf(qv[c], kk.w, s[4 * k4 + 3]);
        }
#pragma unroll
        for (int t = 0; t < XCTX; ++t) s[64 + t] = fmaf(qv[c], s_c[t * 384 + h * XDH + c], s[64 + t]);
    }
    float mx = s[0];
#pragma unroll
    for (int k = 1; k < 64 + XCTX; ++k) mx = fmaxf(mx, s[k]);
    float sum = 0.0f;
#pragma unroll
    for (int k = 0; k < 64 + XCTX; ++k) {
        s[k] = __expf(s[k] - mx);
        sum += s[k];
    }
    const float inv = 1.0f / sum;
    const float* vh = s_kv + (XD + h * XDH) * 64;
#pragma unroll 2
    for (int c = 0; c < XDH; ++c) {
        const float4* row = reinterpret_cast<const float4*>(vh + c * 64);
        float a0 = 0.0f, a1 = 0.0f, a2 = 0.0f, a3 = 0.0f;
#pragma unroll
        for (int k4 = 0; k4 < 16; ++k4) {
            const float4 vv = row[k4];
            a0 = fmaf(s[4 * k4 + 0], vv.x, a0);
            a1 = fmaf(s[4 * k4 + 1], vv.y, a1);
            a2 = fmaf(s[4 * k4 + 2], vv.z, a2);
            a3 = fmaf(s[4 * k4 + 3], vv.w, a3);
        }
#pragma unroll
        for (int t = 0; t < XCTX; ++t) a0 = fmaf(s[64 + t], s_c[t * 384 + XD + h * XDH + c], a0);
        att[(static_cast<size_t>(n2) * XD + h * XDH + c) * 64 + sq] = float_to_act(((a0 + a1) + (a2 + a3)) * inv);
    }
}

// ---- X4 on the tensor cores (warp-level mma.sync m16n8k16, fp32 accumulate) --------------------------
// The same computation as cross_attention_kernel with the two small GEMMs of a head — S = Q K^T
// (64 x 80 x 32) and O = P V (64 x 32 x 80) — on mma.sync; the op is HBM-bound (q 24 KB + kv 48 KB in,
// 24 KB out per (position, board): 98 KB -> 0.03 ms at B = 1024), the tensor instructions only have to
// get the 3.5 MFLOP per CTA out of the way of the loads.  Everything stays in the channel-major token
// layout ([channel][square]): Q^T and K^T tiles are read with ldmatrix.trans, V tiles with plain
// ldmatrix, the probabilities go from the S accumulators straight into the A fragments of the PV
// product (fp16/bf16, values in [0,1]), and O is staged through the warp's own (dead) Q tile so the
// global store is 16-byte coalesced.  Keys 64..70 are the 7 context tokens (fp32 ckv rounded to the
// 16-bit type), 71..79 zero padding masked out of the softmax.
// CTA = (position, board), 12 warps: warp = (head, half of the 64 queries) = two m16 tiles.
constexpr int XKEYS = 80;                 // 64 + 7 context + 9 padding
constexpr int XKP = 88;                   // K/V row pitch in elements (176 B: ldmatrix conflict-free)
constexpr int XQP = 72;                   // Q row pitch (144 B)
constexpr int XATT2_SMEM = (384 * XKP + XD * XQP) * 2;

#ifdef HM_ACT_BF16
#define HM_MMA_16816 "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32"
#else
#define HM_MMA_16816 "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32"
#endif
__device__ __forceinline__ void mma16816(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile(HM_MMA_16816 " {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void ldsm_x4(uint32_t (&r)[4], const void* p) {
    const uint32_t a = static_cast<uint32_t>(__cvta_generic_to_shared(p));
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(a));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t (&r)[4], const void* p) {
    const uint32_t a = static_cast<uint32_t>(__cvta_generic_to_shared(p));
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(a));
}
__device__ __forceinline__ void cp_async16(void* dst, const void* src) {
    const uint32_t d = static_cast<uint32_t>(__cvta_generic_to_shared(dst));
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(src) : "memory");
}

__global__ void __launch_bounds__(384, 2)
cross_attention_mma_kernel(const act_t* __restrict__ q, const act_t* __restrict__ kv,
                           const float* __restrict__ ckv, act_t* __restrict__ att) {
    extern __shared__ __align__(16) unsigned char xraw[];
    act_t* s_kv = reinterpret_cast<act_t*>(xraw);   // [384][XKP]: K rows 0..191, V rows 192..383; cols = keys
    act_t* s_q = s_kv + 384 * XKP;                  // [192][XQP]: Q^T, later the O^T staging
    const int n2 = blockIdx.x, pos = n2 >> 1, other = n2 ^ 1;
    const int tid = threadIdx.x;
    {
        const act_t* gkv = kv + static_cast<size_t>(other) * 384 * 64;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const int i = tid + 384 * j, row = i >> 3, c = i & 7;
            cp_async16(s_kv + row * XKP + c * 8, gkv + row * 64 + c * 8);
        }
        const act_t* gq = q + static_cast<size_t>(n2) * XD * 64;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int i = tid + 384 * j, row = i >> 3, c = i & 7;
            cp_async16(s_q + row * XQP + c * 8, gq + row * 64 + c * 8);
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
        // context keys 64..70 and zero padding 71..79 of row `tid` (K row or V row: ckv holds K | V)
        const float* c = ckv + static_cast<size_t>(pos) * XCTX * 384 + tid;
        act_t* dst = s_kv + tid * XKP + 64;
#pragma unroll
        for (int t = 0; t < XCTX; ++t) dst[t] = float_to_act(c[t * 384]);
#pragma unroll
        for (int t = XCTX; t < XKEYS - 64; ++t) dst[t] = float_to_act(0.0f);
        asm volatile("cp.async.wait_group 0;" ::: "memory");
    }
    __syncthreads();

    const int warp = tid >> 5, lane = tid & 31;
    const int h = warp >> 1;
    const int g = lane >> 2, t4 = lane & 3;
    const int lm = lane >> 3, lr = lane & 7;          // ldmatrix: matrix index, row within it
    const act_t* kh = s_kv + (h * XDH) * XKP;         // K^T rows of this head: [dim][key]
    const act_t* vh = s_kv + (XD + h * XDH) * XKP;    // V^T rows: [dim][key]
    act_t* qh = s_q + (h * XDH) * XQP;                // Q^T rows: [dim][query]
    const float scale = 0.17677669529663689f;         // 1/sqrt(32)

#pragma unroll 1
    for (int mt = 0; mt < 2; ++mt) {
        const int m0 = ((warp & 1) * 2 + mt) * 16;
        // ---- S = Q K^T : 10 n-tiles of 8 keys, 2 k-steps of 16 dims ----
        float sacc[10][4];
#pragma unroll
        for (int n = 0; n < 10; ++n) sacc[n][0] = sacc[n][1] = sacc[n][2] = sacc[n][3] = 0.0f;
#pragma unroll
        for (int ks = 0; ks < 2; ++ks) {
            uint32_t a[4];
            // matrices: (m 0-7,k 0-7) (m 8-15,k 0-7) (m 0-7,k 8-15) (m 8-15,k 8-15); stored rows = k (dim)
            ldsm_x4_t(a, qh + (ks * 16 + (lm >> 1) * 8 + lr) * XQP + m0 + (lm & 1) * 8);
#pragma unroll
            for (int np = 0; np < 5; ++np) {
                uint32_t b[4];
                // matrices: (k 0-7,n 0-7) (k 8-15,n 0-7) (k 0-7,n 8-15) (k 8-15,n 8-15); stored rows = k (dim)
                ldsm_x4_t(b, kh + (ks * 16 + (lm & 1) * 8 + lr) * XKP + np * 16 + (lm >> 1) * 8);
                mma16816(sacc[2 * np], a, b[0], b[1]);
                mma16816(sacc[2 * np + 1], a, b[2], b[3]);
            }
        }
        // ---- softmax over the 71 real keys; rows g (regs 0,1) and g+8 (regs 2,3) ----
        if (t4 == 3) sacc[8][1] = sacc[8][3] = -INFINITY;   // key 71
#pragma unroll
        for (int r = 0; r < 4; ++r) sacc[9][r] = -INFINITY;  // keys 72..79
        float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
        for (int n = 0; n < 10; ++n) {
            mx0 = fmaxf(mx0, fmaxf(sacc[n][0], sacc[n][1]));
            mx1 = fmaxf(mx1, fmaxf(sacc[n][2], sacc[n][3]));
        }
        mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1));
        mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
        mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1));
        mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
        float sum0 = 0.0f, sum1 = 0.0f;
        uint32_t pfr[10][2];
#pragma unroll
        for (int n = 0; n < 10; ++n) {
            const float e0 = __expf((sacc[n][0] - mx0) * scale), e1 = __expf((sacc[n][1] - mx0) * scale);
            const float e2 = __expf((sacc[n][2] - mx1) * scale), e3 = __expf((sacc[n][3] - mx1) * scale);
            sum0 += e0 + e1;
            sum1 += e2 + e3;
            pfr[n][0] = pack_act2(e0, e1);
            pfr[n][1] = pack_act2(e2, e3);
        }
        sum0 += __shfl_xor_sync(0xffffffffu, sum0, 1);
        sum0 += __shfl_xor_sync(0xffffffffu, sum0, 2);
        sum1 += __shfl_xor_sync(0xffffffffu, sum1, 1);
        sum1 += __shfl_xor_sync(0xffffffffu, sum1, 2);
        // ---- O = P V : 4 n-tiles of 8 dims, 5 k-steps of 16 keys ----
        float oacc[4][4];
#pragma unroll
        for (int n = 0; n < 4; ++n) oacc[n][0] = oacc[n][1] = oacc[n][2] = oacc[n][3] = 0.0f;
#pragma unroll
        for (int ks = 0; ks < 5; ++ks) {
            const uint32_t a[4] = {pfr[2 * ks][0], pfr[2 * ks][1], pfr[2 * ks + 1][0], pfr[2 * ks + 1][1]};
#pragma unroll
            for (int np = 0; np < 2; ++np) {
                uint32_t b[4];
                // matrices: (n 0-7,k 0-7) (n 0-7,k 8-15) (n 8-15,k 0-7) (n 8-15,k 8-15); stored rows = n (dim)
                ldsm_x4(b, vh + (np * 16 + (lm >> 1) * 8 + lr) * XKP + ks * 16 + (lm & 1) * 8);
                mma16816(oacc[2 * np], a, b[0], b[1]);
                mma16816(oacc[2 * np + 1], a, b[2], b[3]);
            }
        }
        // ---- O^T into the warp's own Q tile (all its Q fragments are consumed) ----
        const float inv0 = 1.0f / sum0, inv1 = 1.0f / sum1;
        __syncwarp();
#pragma unroll
        for (int n = 0; n < 4; ++n) {
            act_t* o0 = qh + (n * 8 + 2 * t4) * XQP + m0 + g;
            o0[0] = float_to_act(oacc[n][0] * inv0);
            o0[XQP] = float_to_act(oacc[n][1] * inv0);
            o0[8] = float_to_act(oacc[n][2] * inv1);
            o0[XQP + 8] = float_to_act(oacc[n][3] * inv1);
        }
    }
    __syncthreads();
    {
        act_t* gout = att + static_cast<size_t>(n2) * XD * 64;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int i = tid + 384 * j, row = i >> 3, c = i & 7;
            *reinterpret_cast<uint4*>(gout + row * 64 + c * 8) = *reinterpret_cast<const uint4*>(s_q + row * XQP + c * 8);
        }
    }
}

// ---- launchers ---------------------------------------------------------------------------
void launch_token_embed(const float* pre, const act_t* planes, const float* wp, const float* bias,
                        const float* pos_emb, const float* board_emb, act_t* tok, int n, cudaStream_t s) {
    if (n > 0) token_embed_kernel<<<dim3((n + TE_POS - 1) / TE_POS, 2), 256, 0, s>>>(pre, planes, wp, bias, pos_emb, board_emb, tok, n);
}
void launch_context_tokens(const act_t* planes, const float* pocket_w, const float* pocket_b,
                           const float* scalar_w, const float* scalar_b, const float* role, float* ctx,
                           int n, cudaStream_t s) {
    if (n > 0) context_tokens_kernel<<<n, XD, 0, s>>>(planes, pocket_w, pocket_b, scalar_w, scalar_b, role, ctx);
}
void launch_token_layernorm(const act_t* x, const float* g1, const float* b1, act_t* y1, const float* g2,
                            const float* b2, act_t* y2, int n2, cudaStream_t s) {
    if (n2 > 0) token_layernorm_kernel<<<n2, 256, 0, s>>>(x, g1, b1, y1, g2, b2, y2);
}
void launch_context_kv(const float* ctx, const float* g, const float* b, const float* wkv_t,
                       const float* bkv, float* ckv, int n, cudaStream_t s) {
    if (n > 0) context_kv_kernel<<<(n + CKV_POS - 1) / CKV_POS, 384, 0, s>>>(ctx, g, b, wkv_t, bkv, ckv, n);
}
bool xboard_prepare() {
    return cudaFuncSetAttribute(cross_attention_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                XATT_SMEM) == cudaSuccess &&
           cudaFuncSetAttribute(cross_attention_mma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                XATT2_SMEM) == cudaSuccess;
}
void launch_cross_attention(const act_t* q, const act_t* kv, const float* ckv, act_t* att, int n,
                            cudaStream_t s) {
    if (n <= 0) return;
    // HM_XATT_FP32=1 selects the CUDA-core fp32 kernel (bring-up / numerical comparison)
    static const bool fp32_path = getenv("HM_XATT_FP32") != nullptr;
    if (fp32_path) cross_attention_kernel<<<2 * n, 384, XATT_SMEM, s>>>(q, kv, ckv, att);
    else cross_attention_mma_kernel<<<2 * n, 384, XATT2_SMEM, s>>>(q, kv, ckv, att);
}

}  // namespace hm
