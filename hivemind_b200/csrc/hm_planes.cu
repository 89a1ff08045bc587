// hm_planes.cu — stage P: packed bughouse positions -> input planes, on device.
// Mirrors board_to_planes_impl (reference engine/src/environment/planes.cc:213-257) value for
// value; the channel order is planes.cc:236-256 and is cross-checked bit-exactly against the
// reference-compiled oracle (tests/test_gpu_planes_policy.py).
#include <stddef.h>
#include "hm_kernels.h"
#include "hm_ptx.cuh"

namespace hm {

// Every plane is either a bitboard (1 where a bit is set) or a constant fill, so a channel of
// one board is fully described by (mask, value): plane[sq] = (mask >> sq) & 1 ? value : 0.
// first = colour listed first (team on board A, ~team on board B: planes.cc:99-100),
// flip = needs_flipping (planes.cc:90-93).  flip_vertical = bswap64 (utils.h:111-113), and
// orient_square = sq ^ 56 (planes.cc:178-180).
struct PlaneDesc {
    uint64_t mask;
    float value;
};
__device__ __forceinline__ uint64_t orient_bb(uint64_t bb, int flip) {
    return flip ? __byte_perm(static_cast<uint32_t>(bb >> 32), 0, 0x0123) |
                      (static_cast<uint64_t>(__byte_perm(static_cast<uint32_t>(bb), 0, 0x0123)) << 32)
                : bb;
}
__device__ __forceinline__ uint64_t square_bb(int sq, int flip) {
    return sq < 64 ? (1ull << (flip ? (sq ^ 56) : sq)) : 0ull;
}
__device__ __forceinline__ PlaneDesc plane_desc(const hm_packed_board& b, int first, int flip,
                                                int tadv, int ch) {
    const uint64_t ALL = ~0ull;
    if (ch < 12) {  // pieces, planes.cc:97-111
        const int ci = ch / 6, pt = ch - ci * 6;
        const int color = ci == 0 ? first : 1 - first;
        return {orient_bb(b.pieces[color * 6 + pt], flip), 1.0f};
    }
    if (ch < 22) {  // pockets, planes.cc:114-124: (float)count / 16.0f
        const int k = ch - 12, ci = k / 5, pt = k - ci * 5;
        const int color = ci == 0 ? first : 1 - first;
        return {ALL, __fdiv_rn(static_cast<float>(b.pocket[color * 5 + pt]), 16.0f)};
    }
    if (ch < 24) {  // promoted, planes.cc:127-138
        const int color = (ch == 22) ? first : 1 - first;
        return {orient_bb(b.promoted[color], flip), 1.0f};
    }
    const int f = b.flags;
    switch (ch) {
        case 24: return {square_bb(b.ep_square, flip), 1.0f};                                   // :141-151
        case 25: return {(((f & HM_FLAG_STM_BLACK) ? 1 : 0) == first) ? ALL : 0ull, 1.0f};      // :154-159
        case 26: return {ALL, 1.0f};                                                            // :242,253
        case 27: return {(f & (first ? HM_FLAG_BLACK_OO : HM_FLAG_WHITE_OO)) ? ALL : 0ull, 1.0f};   // :162-176
        case 28: return {(f & (first ? HM_FLAG_BLACK_OOO : HM_FLAG_WHITE_OOO)) ? ALL : 0ull, 1.0f};
        case 29: return {(f & (first ? HM_FLAG_WHITE_OO : HM_FLAG_BLACK_OO)) ? ALL : 0ull, 1.0f};
        case 30: return {(f & (first ? HM_FLAG_WHITE_OOO : HM_FLAG_BLACK_OOO)) ? ALL : 0ull, 1.0f};
        case 31: return {tadv ? ALL : 0ull, 1.0f};                                              // :244,255
        case 32: return {b.last_to < 64 ? square_bb(b.last_from, flip) : 0ull, 1.0f};           // :183-198 (drops: from = 65)
        case 33: return {square_bb(b.last_to, flip), 1.0f};
        case 34: return {ALL, __fdiv_rn(static_cast<float>(min(static_cast<int>(b.rule50), 50)), 50.0f)};  // :203-206
        case 35: return {b.repetitions >= 2 ? ALL : 0ull, 1.0f};                                // :208-210
        default: return {b.repetitions >= 3 ? ALL : 0ull, 1.0f};
    }
}

template <typename T>
__device__ __forceinline__ T cvt_plane(float v);
template <> __device__ __forceinline__ float cvt_plane<float>(float v) { return v; }
template <> __device__ __forceinline__ __half cvt_plane<__half>(float v) { return __float2half_rn(v); }  // planes.cc:22-28
template <> __device__ __forceinline__ __nv_bfloat16 cvt_plane<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }
template <> __device__ __forceinline__ uint8_t cvt_plane<uint8_t>(float v) {
    // selfplay.cc:471-473: lround(clamp(x, 0, 1) * 255)
    return static_cast<uint8_t>(lroundf(fminf(fmaxf(v, 0.0f), 1.0f) * 255.0f));
}

// NCHW [n][74][64].  A lane works out the (mask, value) of ONE (position, channel) plane — so the table
// logic runs once per plane — and the warp then writes its 32 planes cooperatively: in step k lane l
// stores 16-byte piece l % PIECES of plane k*(32/PIECES) + l / PIECES, fetching that plane's mask and
// value by shuffle, so every store instruction covers 512 CONSECUTIVE bytes (whole 128-byte lines).
// The 256-byte record is read through the read-only path (each record is shared by 74 lanes).
constexpr int PL_THREADS = 256;
template <typename T>
__global__ void __launch_bounds__(PL_THREADS)
planes_nchw_kernel(const hm_packed_pos* __restrict__ pos, int n, T* __restrict__ out) {
    constexpr int PER = 16 / sizeof(T);     // outputs per 16-byte store
    constexpr int PIECES = 64 / PER;        // stores per plane: 16 (f32), 8 (f16/bf16), 4 (u8)
    constexpr int PPS = 32 / PIECES;        // planes written per step
    const long long total = static_cast<long long>(n) * HM_NB_INPUT_CHANNELS;
    const long long item = static_cast<long long>(blockIdx.x) * PL_THREADS + threadIdx.x;
    const int lane = threadIdx.x & 31;
    uint64_t mask = 0ull;
    float value = 0.0f;
    if (item < total) {
        const int p = static_cast<int>(item / HM_NB_INPUT_CHANNELS);
        const int c = static_cast<int>(item - static_cast<long long>(p) * HM_NB_INPUT_CHANNELS);
        const int bi = c >= HM_NB_CHANNELS_PER_BOARD ? 1 : 0, ch = c - bi * HM_NB_CHANNELS_PER_BOARD;
        const int flags0 = reinterpret_cast<const uint8_t*>(pos + p)[offsetof(hm_packed_board, flags)];
        const int team_black = (flags0 & HM_FLAG_TEAM_BLACK) ? 1 : 0;
        const int tadv = (flags0 & HM_FLAG_TIME_ADVANTAGE) ? 1 : 0;
        const int first = bi == 0 ? team_black : 1 - team_black;
        // planes.cc:90-93: board A flipped iff team BLACK, board B iff team WHITE  (== first)
        const PlaneDesc d = plane_desc(pos[p].board[bi], first, first, tadv, ch);
        mask = d.mask;
        value = d.value;
    }
    const long long base = item - lane;     // first plane of this warp
    const T off = cvt_plane<T>(0.0f);
    const int j = lane % PIECES;
#pragma unroll
    for (int k = 0; k < PIECES; ++k) {
        const int src = k * PPS + lane / PIECES;
        const uint64_t m = __shfl_sync(0xffffffffu, mask, src);
        const T on = cvt_plane<T>(__shfl_sync(0xffffffffu, value, src));
        const uint32_t bits = static_cast<uint32_t>(m >> (j * PER));
        __align__(16) T v[PER];
#pragma unroll
        for (int e = 0; e < PER; ++e) v[e] = ((bits >> e) & 1u) ? on : off;
        if (base + src < total)
            reinterpret_cast<uint4*>(out + (base + src) * 64)[j] = *reinterpret_cast<const uint4*>(v);
    }
}

void launch_planes_nchw(const hm_packed_pos* pos, int n, void* out, int dtype, cudaStream_t s) {
    if (n <= 0) return;
    const unsigned grid = static_cast<unsigned>((static_cast<long long>(n) * HM_NB_INPUT_CHANNELS + PL_THREADS - 1) / PL_THREADS);
    switch (dtype) {
        case HM_DTYPE_F32: planes_nchw_kernel<float><<<grid, PL_THREADS, 0, s>>>(pos, n, static_cast<float*>(out)); break;
        case HM_DTYPE_F16: planes_nchw_kernel<__half><<<grid, PL_THREADS, 0, s>>>(pos, n, static_cast<__half*>(out)); break;
        case HM_DTYPE_BF16: planes_nchw_kernel<__nv_bfloat16><<<grid, PL_THREADS, 0, s>>>(pos, n, static_cast<__nv_bfloat16*>(out)); break;
        default: planes_nchw_kernel<uint8_t><<<grid, PL_THREADS, 0, s>>>(pos, n, static_cast<uint8_t*>(out)); break;
    }
}

}  // namespace hm
