// hm_planes.cu — stage P: packed bughouse positions -> input planes, on device.
// Mirrors board_to_planes_impl (reference engine/src/environment/planes.cc:213-257) value for
// value; the channel order is planes.cc:236-256 and is cross-checked bit-exactly against the
// reference-compiled oracle (tests/test_planes_gpu.py).
#include "hm_kernels.h"
#include "hm_ptx.cuh"

namespace hm {

// Value of channel `ch` (0..36) of one board at square `sq` (a1 = 0).
// first = colour listed first (team on board A, ~team on board B: planes.cc:99-100),
// flip = needs_flipping (planes.cc:90-93).  A flipped bitboard's bit sq is the original's
// bit sq^56 (flip_vertical = bswap64, utils.h:111-113).
__device__ __forceinline__ float plane_value(const hm_packed_board& b, int first, int flip,
                                             int tadv, int ch, int sq) {
    const int src = flip ? (sq ^ 56) : sq;
    if (ch < 12) {  // pieces, planes.cc:97-111
        const int ci = ch / 6, pt = ch - ci * 6;
        const int color = ci == 0 ? first : 1 - first;
        return static_cast<float>((b.pieces[color * 6 + pt] >> src) & 1ull);
    }
    if (ch < 22) {  // pockets, planes.cc:114-124: (float)count / 16.0f
        const int k = ch - 12, ci = k / 5, pt = k - ci * 5;
        const int color = ci == 0 ? first : 1 - first;
        return __fdiv_rn(static_cast<float>(b.pocket[color * 5 + pt]), 16.0f);
    }
    if (ch < 24) {  // promoted, planes.cc:127-138
        const int color = (ch == 22) ? first : 1 - first;
        return static_cast<float>((b.promoted[color] >> src) & 1ull);
    }
    switch (ch) {
        case 24: return (b.ep_square == src) ? 1.0f : 0.0f;                        // :141-151
        case 25: return (((b.flags & HM_FLAG_STM_BLACK) ? 1 : 0) == first) ? 1.0f : 0.0f;  // :154-159
        case 26: return 1.0f;                                                      // :242,253
        case 27: return (b.flags & (first ? HM_FLAG_BLACK_OO : HM_FLAG_WHITE_OO)) ? 1.0f : 0.0f;   // :162-176
        case 28: return (b.flags & (first ? HM_FLAG_BLACK_OOO : HM_FLAG_WHITE_OOO)) ? 1.0f : 0.0f;
        case 29: return (b.flags & (first ? HM_FLAG_WHITE_OO : HM_FLAG_BLACK_OO)) ? 1.0f : 0.0f;
        case 30: return (b.flags & (first ? HM_FLAG_WHITE_OOO : HM_FLAG_BLACK_OOO)) ? 1.0f : 0.0f;
        case 31: return tadv ? 1.0f : 0.0f;                                        // :244,255
        case 32: return (b.last_to < 64 && b.last_from == src) ? 1.0f : 0.0f;      // :183-198
        case 33: return (b.last_to == src) ? 1.0f : 0.0f;
        case 34: return __fdiv_rn(static_cast<float>(min(static_cast<int>(b.rule50), 50)), 50.0f);  // :203-206
        case 35: return b.repetitions >= 2 ? 1.0f : 0.0f;                          // :208-210
        default: return b.repetitions >= 3 ? 1.0f : 0.0f;
    }
}

constexpr int PL_POS_PER_BLOCK = 4;
constexpr int PL_THREADS = 256;

__device__ __forceinline__ void load_records(const hm_packed_pos* pos, int p0, int n,
                                             hm_packed_pos* s_pos) {
    // 4 records x 256 B = 64 x 16 B: one coalesced 16-byte load per thread
    const int np = min(PL_POS_PER_BLOCK, n - p0);
    const uint4* src = reinterpret_cast<const uint4*>(pos + p0);
    uint4* dst = reinterpret_cast<uint4*>(s_pos);
    for (int i = threadIdx.x; i < np * 16; i += blockDim.x) dst[i] = __ldg(src + i);
    __syncthreads();
}

// NCHW, dtype templated; one work item = 8 consecutive squares (one rank) of one channel.
template <typename T>
__device__ __forceinline__ T cvt_plane(float v);
template <> __device__ __forceinline__ float cvt_plane<float>(float v) { return v; }
template <> __device__ __forceinline__ __half cvt_plane<__half>(float v) { return __float2half_rn(v); }  // planes.cc:22-28
template <> __device__ __forceinline__ __nv_bfloat16 cvt_plane<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }
template <> __device__ __forceinline__ uint8_t cvt_plane<uint8_t>(float v) {
    // selfplay.cc:471-473: lround(clamp(x, 0, 1) * 255)
    return static_cast<uint8_t>(lroundf(fminf(fmaxf(v, 0.0f), 1.0f) * 255.0f));
}

template <typename T>
__global__ void __launch_bounds__(PL_THREADS)
planes_nchw_kernel(const hm_packed_pos* __restrict__ pos, int n, T* __restrict__ out) {
    __shared__ __align__(16) hm_packed_pos s_pos[PL_POS_PER_BLOCK];
    const int p0 = blockIdx.x * PL_POS_PER_BLOCK;
    load_records(pos, p0, n, s_pos);
    const int np = min(PL_POS_PER_BLOCK, n - p0);
    const int items = np * 74 * 8;
    for (int it = threadIdx.x; it < items; it += blockDim.x) {
        const int pl = it / (74 * 8);
        const int rem = it - pl * 74 * 8;
        const int c = rem >> 3, rank = rem & 7;
        const int bi = c >= 37 ? 1 : 0, ch = c - bi * 37;
        const hm_packed_pos& P = s_pos[pl];
        const int team_black = (P.board[0].flags & HM_FLAG_TEAM_BLACK) ? 1 : 0;
        const int tadv = (P.board[0].flags & HM_FLAG_TIME_ADVANTAGE) ? 1 : 0;
        const int first = bi == 0 ? team_black : 1 - team_black;
        const int flip = first;  // planes.cc:90-93: board A iff team BLACK, board B iff team WHITE
        __align__(16) T vals[8];
#pragma unroll
        for (int f = 0; f < 8; ++f)
            vals[f] = cvt_plane<T>(plane_value(P.board[bi], first, flip, tadv, ch, rank * 8 + f));
        T* dst = out + (static_cast<size_t>(p0 + pl) * HM_NB_INPUT_VALUES + c * 64 + rank * 8);
        if constexpr (sizeof(T) == 4) {
            reinterpret_cast<uint4*>(dst)[0] = reinterpret_cast<uint4*>(vals)[0];
            reinterpret_cast<uint4*>(dst)[1] = reinterpret_cast<uint4*>(vals)[1];
        } else if constexpr (sizeof(T) == 2) {
            reinterpret_cast<uint4*>(dst)[0] = reinterpret_cast<uint4*>(vals)[0];
        } else {
            reinterpret_cast<uint2*>(dst)[0] = reinterpret_cast<uint2*>(vals)[0];
        }
    }
}

// NHWC 16-bit [n][64][80]: one work item = 8 consecutive channels of one square (16-byte store).
__global__ void __launch_bounds__(PL_THREADS)
planes_nhwc_kernel(const hm_packed_pos* __restrict__ pos, int n, act_t* __restrict__ out) {
    __shared__ __align__(16) hm_packed_pos s_pos[PL_POS_PER_BLOCK];
    const int p0 = blockIdx.x * PL_POS_PER_BLOCK;
    load_records(pos, p0, n, s_pos);
    const int np = min(PL_POS_PER_BLOCK, n - p0);
    constexpr int G = PLANES_NHWC_C / 8;  // 10 groups of 8 channels
    const int items = np * 64 * G;
    for (int it = threadIdx.x; it < items; it += blockDim.x) {
        const int pl = it / (64 * G);
        const int rem = it - pl * 64 * G;
        const int sq = rem / G, g = rem - sq * G;
        const hm_packed_pos& P = s_pos[pl];
        const int team_black = (P.board[0].flags & HM_FLAG_TEAM_BLACK) ? 1 : 0;
        const int tadv = (P.board[0].flags & HM_FLAG_TIME_ADVANTAGE) ? 1 : 0;
        float v[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const int c = g * 8 + j;
            if (c < 74) {
                const int bi = c >= 37 ? 1 : 0, ch = c - bi * 37;
                const int first = bi == 0 ? team_black : 1 - team_black;
                v[j] = plane_value(P.board[bi], first, first, tadv, ch, sq);
            } else {
                v[j] = 0.0f;
            }
        }
        uint4 w;
        w.x = pack_act2(v[0], v[1]); w.y = pack_act2(v[2], v[3]);
        w.z = pack_act2(v[4], v[5]); w.w = pack_act2(v[6], v[7]);
        reinterpret_cast<uint4*>(out + (static_cast<size_t>(p0 + pl) * 64 + sq) * PLANES_NHWC_C)[g] = w;
    }
}

// dense fp16 NCHW [n][74][64] -> 16-bit NHWC [n][64][80], through a shared-memory transpose
__global__ void __launch_bounds__(256)
planes_f16_to_nhwc_kernel(const __half* __restrict__ in, int n, act_t* __restrict__ out) {
    __shared__ __half tile[74][64 + 2];
    const int p = blockIdx.x;
    if (p >= n) return;
    const __half* src = in + static_cast<size_t>(p) * HM_NB_INPUT_VALUES;
    for (int i = threadIdx.x; i < HM_NB_INPUT_VALUES; i += blockDim.x) tile[i >> 6][i & 63] = src[i];
    __syncthreads();
    constexpr int G = PLANES_NHWC_C / 8;
    for (int it = threadIdx.x; it < 64 * G; it += blockDim.x) {
        const int sq = it / G, g = it - sq * G;
        float v[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const int c = g * 8 + j;
            v[j] = c < 74 ? __half2float(tile[c][sq]) : 0.0f;
        }
        uint4 w;
        w.x = pack_act2(v[0], v[1]); w.y = pack_act2(v[2], v[3]);
        w.z = pack_act2(v[4], v[5]); w.w = pack_act2(v[6], v[7]);
        reinterpret_cast<uint4*>(out + (static_cast<size_t>(p) * 64 + sq) * PLANES_NHWC_C)[g] = w;
    }
}

void launch_planes_nchw(const hm_packed_pos* pos, int n, void* out, int dtype, cudaStream_t s) {
    if (n <= 0) return;
    const int grid = (n + PL_POS_PER_BLOCK - 1) / PL_POS_PER_BLOCK;
    switch (dtype) {
        case HM_DTYPE_F32: planes_nchw_kernel<float><<<grid, PL_THREADS, 0, s>>>(pos, n, static_cast<float*>(out)); break;
        case HM_DTYPE_F16: planes_nchw_kernel<__half><<<grid, PL_THREADS, 0, s>>>(pos, n, static_cast<__half*>(out)); break;
        case HM_DTYPE_BF16: planes_nchw_kernel<__nv_bfloat16><<<grid, PL_THREADS, 0, s>>>(pos, n, static_cast<__nv_bfloat16*>(out)); break;
        default: planes_nchw_kernel<uint8_t><<<grid, PL_THREADS, 0, s>>>(pos, n, static_cast<uint8_t*>(out)); break;
    }
}
void launch_planes_nhwc(const hm_packed_pos* pos, int n, act_t* out, cudaStream_t s) {
    if (n <= 0) return;
    planes_nhwc_kernel<<<(n + PL_POS_PER_BLOCK - 1) / PL_POS_PER_BLOCK, PL_THREADS, 0, s>>>(pos, n, out);
}
void launch_planes_f16_to_nhwc(const __half* in, int n, act_t* out, cudaStream_t s) {
    if (n <= 0) return;
    planes_f16_to_nhwc_kernel<<<n, 256, 0, s>>>(in, n, out);
}

}  // namespace hm
