// hm_policy.cu — stage D: policy index, legal-move gather + masked softmax, value shaping.
// Reference: engine/src/common/utils.h:127-243, engine/src/common/globals.cc:82-136,
// engine/src/search/searchthread.cc:569-619.
#include <math.h>
#include "hm_kernels.h"
#include "hm_ptx.cuh"

namespace hm {

// ---------------------------------------------------------------------------------------
// Policy index without tables.  The reference builds POLICY_TABLE_NORMAL / _DROP from the
// 4672 UCI labels (globals.cc:82-136); the label layout is regular (index = channel*64 +
// square), so the index is computed arithmetically:
//   ch 0 pass | ch 1 P@ (ranks 2..7) | ch 2..5 N@,R@,B@,Q@ | ch 6+7d+(k-1): ray d in
//   {N,NE,E,SE,S,SW,W,NW}, distance k | ch 62..69 knight jumps | ch 70..72 knight
//   under-promotion (capture-left, push, capture-right) from rank 7.
// Black to move mirrors ranks (sq ^ 56) first (globals.cc:36-52).
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ int label_index_normal(int from, int to) {
    const int df = (to & 7) - (from & 7), dr = (to >> 3) - (from >> 3);
    if (df == 0 && dr == 0) return -1;
    const int adf = abs(df), adr = abs(dr);
    if (df == 0 || dr == 0 || adf == adr) {
        const int k = max(adf, adr);
        const int sf = (df > 0) - (df < 0), sr = (dr > 0) - (dr < 0);
        // direction order N,NE,E,SE,S,SW,W,NW
        int d;
        if (sf == 0) d = sr > 0 ? 0 : 4;
        else if (sf > 0) d = sr > 0 ? 1 : (sr == 0 ? 2 : 3);
        else d = sr < 0 ? 5 : (sr == 0 ? 6 : 7);
        return (6 + 7 * d + (k - 1)) * 64 + from;
    }
    if (adf * adr == 2) {
        // (df,dr): (1,2),(2,1),(2,-1),(1,-2),(-1,-2),(-2,-1),(-2,1),(-1,2)
        int j;
        if (df == 1) j = dr == 2 ? 0 : 3;
        else if (df == 2) j = dr == 1 ? 1 : 2;
        else if (df == -1) j = dr == -2 ? 4 : 7;
        else j = dr == -1 ? 5 : 6;
        return (62 + j) * 64 + from;
    }
    return -1;
}

// get_fast_policy_index (utils.h:184-216).  Move bits (Fairy-Stockfish types.h:249-265,
// 735-798, SQUARE_BITS = 6): to [0:6], from [6:12], type [12:16] (PROMOTION 3, DROP 4),
// piece type [16:22] (P 1, N 2, B 3, R 4, Q 5).
__device__ __forceinline__ int policy_index(uint32_t move, int stm_black) {
    if (move == 0u) return 0;
    const int m = stm_black ? 56 : 0;
    const int to = static_cast<int>(move & 63u) ^ m;
    const int type = static_cast<int>((move >> 12) & 15u);
    const int pt = static_cast<int>((move >> 16) & 63u);
    if (type == 4) {
        if (pt < 1 || pt > 5) return -1;
        if (pt == 1 && ((to >> 3) == 0 || (to >> 3) == 7)) return -1;
        const int ch = pt == 1 ? 1 : pt == 2 ? 2 : pt == 3 ? 4 : pt == 4 ? 3 : 5;  // P,N,R,B,Q
        return ch * 64 + to;
    }
    const int from = static_cast<int>((move >> 6) & 63u) ^ m;
    if (type == 3) {
        if (pt == 2) {  // knight: "..n" labels exist only rank 7 -> rank 8, |df| <= 1
            const int df = (to & 7) - (from & 7);
            if ((from >> 3) != 6 || (to >> 3) != 7 || df < -1 || df > 1) return -1;
            return (71 + df) * 64 + from;
        }
        if (pt != 5) return -1;  // R/B under-promotions are not representable
    }
    return label_index_normal(from, to);
}

__global__ void policy_indices_kernel(const uint32_t* __restrict__ moves,
                                      const uint8_t* __restrict__ stm, int n,
                                      int32_t* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = policy_index(moves[i], stm[i]);
}
void launch_policy_indices(const uint32_t* moves, const uint8_t* stm_black, int n, int32_t* out,
                           cudaStream_t s) {
    if (n <= 0) return;
    policy_indices_kernel<<<(n + 255) / 256, 256, 0, s>>>(moves, stm_black, n, out);
}

// ---------------------------------------------------------------------------------------
// get_normalized_probability (utils.h:226-243) + normalize_logits (utils.h:127-167):
// one warp per (position, board).  exp() and the sum are evaluated in double so that the
// float priors agree with the reference's (float exp, double sum) to the last ulp or one.
// ---------------------------------------------------------------------------------------
struct LogitsPacked {
    const float* base;  // [pos][2*73][64]: the reference's ch*64+sq index per board, boards A|B stacked
    __device__ __forceinline__ float at(int pos, int board, int idx) const {
        return base[(static_cast<size_t>(pos) * 2 + board) * HM_NB_POLICY_VALUES + idx];
    }
};
struct LogitsDenseHalf {
    const __half* a;
    const __half* b;
    __device__ __forceinline__ float at(int pos, int board, int idx) const {
        const __half* p = board ? b : a;
        return __half2float(p[static_cast<size_t>(pos) * HM_NB_POLICY_VALUES + idx]);
    }
};

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ int warp_sum_i(int v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

constexpr int PRI_REG = 4;   // moves per lane held in registers by the short-list path

template <typename L>
__global__ void __launch_bounds__(128)
policy_priors_kernel(L logits, const hm_packed_pos* __restrict__ pos, int n_pos,
                     const uint32_t* __restrict__ moves, const uint32_t* __restrict__ offsets,
                     float* __restrict__ priors) {
    const int w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;  // (pos, board)
    const int lane = threadIdx.x & 31;
    if (w >= 2 * n_pos) return;
    const int p = w >> 1, board = w & 1;
    const uint32_t lo = offsets[w], hi = offsets[w + 1];
    const int n = static_cast<int>(hi - lo);
    if (n <= 0) return;
    const int stm_black = (pos[p].board[board].flags & HM_FLAG_STM_BLACK) ? 1 : 0;
    const float NEG_INF = -INFINITY;

    if (n <= 32 * PRI_REG) {
        // Lists of up to 128 moves (every list of the golden positions; bughouse lists beyond that are rare) stay in
        // registers: offsets | flags -> moves -> logits -> priors are the only dependent global round trips, the
        // loads of a round are all in flight together, and nothing is staged through `priors`.  The arithmetic is the
        // long-list path's, operation for operation.
        float l[PRI_REG];
        uint32_t mv[PRI_REG];
#pragma unroll
        for (int k = 0; k < PRI_REG; ++k) mv[k] = lane + 32 * k < n ? __ldg(moves + lo + lane + 32 * k) : 0u;
        float mx = NEG_INF;
        int finite = 0;
#pragma unroll
        for (int k = 0; k < PRI_REG; ++k) {
            l[k] = NEG_INF;
            if (lane + 32 * k < n) {
                const int idx = policy_index(mv[k], stm_black);
                if (idx >= 0) l[k] = logits.at(p, board, idx);
            }
        }
#pragma unroll
        for (int k = 0; k < PRI_REG; ++k)
            if (isfinite(l[k])) { mx = fmaxf(mx, l[k]); ++finite; }
        mx = warp_max(mx);
        if (!isfinite(mx)) {  // utils.h:136-141
            const float u = 1.0f / static_cast<float>(n);
#pragma unroll
            for (int k = 0; k < PRI_REG; ++k)
                if (lane + 32 * k < n) priors[lo + lane + 32 * k] = u;
            return;
        }
        double sum = 0.0;
        float e[PRI_REG];
#pragma unroll
        for (int k = 0; k < PRI_REG; ++k) {
            e[k] = 0.0f;
            if (isfinite(l[k])) {
                e[k] = static_cast<float>(exp(static_cast<double>(l[k] - mx)));
                sum += static_cast<double>(e[k]);
            }
        }
        // the long-list path adds lane-strided partial sums in the same order: i = lane, lane + 32, ...
        sum = warp_sum_d(sum);
        if (!isfinite(sum) || sum <= 0.0) {  // utils.h:151-161
            finite = warp_sum_i(finite);
            const float u = finite > 0 ? 1.0f / static_cast<float>(finite) : 0.0f;
#pragma unroll
            for (int k = 0; k < PRI_REG; ++k)
                if (lane + 32 * k < n) priors[lo + lane + 32 * k] = isfinite(l[k]) ? u : 0.0f;
            return;
        }
#pragma unroll
        for (int k = 0; k < PRI_REG; ++k)  // utils.h:163-165
            if (lane + 32 * k < n) priors[lo + lane + 32 * k] = static_cast<float>(static_cast<double>(e[k]) / sum);
        return;
    }

    // pass 1: gather, max over finite logits (utils.h:129-134)
    float mx = NEG_INF;
    int finite = 0;
    for (int i = lane; i < n; i += 32) {
        const int idx = policy_index(moves[lo + i], stm_black);
        const float l = idx >= 0 ? logits.at(p, board, idx) : NEG_INF;
        if (isfinite(l)) { mx = fmaxf(mx, l); ++finite; }
        priors[lo + i] = l;  // stage the gathered logit
    }
    mx = warp_max(mx);
    finite = warp_sum_i(finite);
    __syncwarp();
    if (!isfinite(mx)) {  // utils.h:136-141: no finite logit -> uniform over all
        const float u = 1.0f / static_cast<float>(n);
        for (int i = lane; i < n; i += 32) priors[lo + i] = u;
        return;
    }
    // pass 2: exp and double sum (utils.h:143-149)
    double sum = 0.0;
    for (int i = lane; i < n; i += 32) {
        const float l = priors[lo + i];
        float e = 0.0f;
        if (isfinite(l)) {
            e = static_cast<float>(exp(static_cast<double>(l - mx)));
            sum += static_cast<double>(e);
        }
        priors[lo + i] = e;
    }
    sum = warp_sum_d(sum);
    __syncwarp();
    if (!isfinite(sum) || sum <= 0.0) {  // utils.h:151-161 (unreachable once max is finite; kept)
        // the staged values are exps now; finiteness of the logit == (e > 0 or l == mx path);
        // recompute from the gather to stay faithful
        const float u = finite > 0 ? 1.0f / static_cast<float>(finite) : 0.0f;
        for (int i = lane; i < n; i += 32) {
            const int idx = policy_index(moves[lo + i], stm_black);
            const float l = idx >= 0 ? logits.at(p, board, idx) : NEG_INF;
            priors[lo + i] = isfinite(l) ? u : 0.0f;
        }
        return;
    }
    for (int i = lane; i < n; i += 32)  // utils.h:163-165
        priors[lo + i] = static_cast<float>(static_cast<double>(priors[lo + i]) / sum);
}

void launch_policy_priors_packed(const float* logits, const hm_packed_pos* pos, int n_pos,
                               const uint32_t* moves, const uint32_t* offsets, float* priors,
                               cudaStream_t s) {
    if (n_pos <= 0) return;
    const int warps = 2 * n_pos;
    policy_priors_kernel<LogitsPacked><<<(warps + 3) / 4, 128, 0, s>>>(LogitsPacked{logits}, pos, n_pos,
                                                                      moves, offsets, priors);
}
void launch_policy_priors_dense(const __half* pi_a, const __half* pi_b, const hm_packed_pos* pos,
                                int n_pos, const uint32_t* moves, const uint32_t* offsets,
                                float* priors, cudaStream_t s) {
    if (n_pos <= 0) return;
    const int warps = 2 * n_pos;
    policy_priors_kernel<LogitsDenseHalf><<<(warps + 3) / 4, 128, 0, s>>>(
        LogitsDenseHalf{pi_a, pi_b}, pos, n_pos, moves, offsets, priors);
}

// ---------------------------------------------------------------------------------------
// Value shaping (searchthread.cc:569-619)
// ---------------------------------------------------------------------------------------
// std::clamp semantics (NaN propagates, as in the reference)
__device__ __forceinline__ float clampf(float v, float lo, float hi) { return v < lo ? lo : (v > hi ? hi : v); }

__device__ __forceinline__ float shape_value(float value, float l, float d, float w, float ml,
                                             const hm_value_config& cfg) {
    const float scalar = isfinite(value) ? clampf(value, -1.0f, 1.0f) : 0.0f;
    float neural = scalar;
    if (cfg.enable_wdl_eval && isfinite(l) && isfinite(d) && isfinite(w)) {
        const float mx = fmaxf(l, fmaxf(d, w));
        const float el = expf(l - mx), ed = expf(d - mx), ew = expf(w - mx);
        const float s = el + ed + ew;
        if (isfinite(s) && s > 0.0f) {
            const float pl = el / s, pd = ed / s, pw = ew / s;
            const float wdl_value = pw - pl - cfg.draw_contempt * pd;
            const float wt = clampf(cfg.wdl_value_weight, 0.0f, 1.0f);
            neural = (1.0f - wt) * scalar + wt * wdl_value;
        }
    }
    if (cfg.moves_left_discount > 0.0f) {
        const float plies = clampf(ml, 0.0f, 1.0f);
        const float disc = clampf(cfg.moves_left_discount, 0.0f, 1.0f);
        neural *= 1.0f - disc * plies;
    }
    return clampf(neural, -1.0f, 1.0f);
}

// cfg lives in device memory (one struct per engine): hm_engine_set_value_config updates it with a copy,
// the captured graphs never have to be rebuilt
__global__ void shape_values_raw5_kernel(const float* __restrict__ raw5, int n, const hm_value_config* __restrict__ cfgp,
                                         float* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const hm_value_config cfg = *cfgp;
    const float* r = raw5 + static_cast<size_t>(i) * 5;
    out[i] = shape_value(r[0], r[1], r[2], r[3], r[4], cfg);
}
__global__ void shape_values_half_kernel(const __half* __restrict__ value,
                                         const __half* __restrict__ wdl,
                                         const __half* __restrict__ ml, int n, hm_value_config cfg,
                                         float* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    out[i] = shape_value(__half2float(value[i]), __half2float(wdl[3 * i]), __half2float(wdl[3 * i + 1]),
                         __half2float(wdl[3 * i + 2]), __half2float(ml[i]), cfg);
}
void launch_shape_values_raw5(const float* raw5, int n, const hm_value_config* cfg, float* out,
                              cudaStream_t s) {
    if (n <= 0) return;
    shape_values_raw5_kernel<<<(n + 127) / 128, 128, 0, s>>>(raw5, n, cfg, out);
}
void launch_shape_values_half(const __half* value, const __half* wdl, const __half* ml, int n,
                              hm_value_config cfg, float* out, cudaStream_t s) {
    if (n <= 0) return;
    shape_values_half_kernel<<<(n + 127) / 128, 128, 0, s>>>(value, wdl, ml, n, cfg, out);
}

}  // namespace hm
