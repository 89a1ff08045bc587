// hm_act.h — the 16-bit storage type of activations and weights.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <stdint.h>

namespace hm {

// ---- 16-bit activation / weight type -------------------------------------------------------
// fp16 by default: it is what the reference deploys (its TensorRT plan is all-fp16,
// engine.cc:331-354) and bf16 storage misses the parity bar on a 15-block trunk (DESIGN.md
// "Precision").  tcgen05 kind::f16 runs fp16 and bf16 at the same rate; -DHM_ACT_BF16 switches.
#ifdef HM_ACT_BF16
using act_t = __nv_bfloat16;
constexpr uint32_t UMMA_FMT_16 = 1;  // BF16
__device__ __forceinline__ uint32_t pack_act2(float lo, float hi) {
    __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ uint32_t pack_act2_relu(float lo, float hi) {
    return pack_act2(fmaxf(lo, 0.0f), fmaxf(hi, 0.0f));
}
__device__ __forceinline__ float2 unpack_act2(uint32_t u) {
    __nv_bfloat162 v = *reinterpret_cast<__nv_bfloat162*>(&u);
    return __bfloat1622float2(v);
}
__host__ __device__ __forceinline__ float act_to_float(act_t v) { return __bfloat162float(v); }
__host__ __device__ __forceinline__ act_t float_to_act(float v) { return __float2bfloat16(v); }
#else
using act_t = __half;
constexpr uint32_t UMMA_FMT_16 = 0;  // F16
__device__ __forceinline__ uint32_t pack_act2(float lo, float hi) {
    // one F2FP.SATFINITE: round to nearest, saturate to +-65504 instead of overflowing to inf
    uint32_t r;
    asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
    return r;
}
// max(x, 0) folded into the conversion (F2FP.RELU.SATFINITE)
__device__ __forceinline__ uint32_t pack_act2_relu(float lo, float hi) {
    uint32_t r;
    asm("cvt.rn.relu.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
    return r;
}
__device__ __forceinline__ float2 unpack_act2(uint32_t u) {
    __half2 v = *reinterpret_cast<__half2*>(&u);
    return __half22float2(v);
}
__host__ __device__ __forceinline__ float act_to_float(act_t v) { return __half2float(v); }
__host__ __device__ __forceinline__ act_t float_to_act(float v) { return __float2half_rn(v); }
#endif


#ifdef HM_ACT_BF16
#define HM_TMA_DTYPE CU_TENSOR_MAP_DATA_TYPE_BFLOAT16
#define HM_ACT_NAME "bf16"
#else
#define HM_TMA_DTYPE CU_TENSOR_MAP_DATA_TYPE_FLOAT16
#define HM_ACT_NAME "fp16"
#endif

}  // namespace hm
