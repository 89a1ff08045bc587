// hm_kernels.h — launchers of the bandwidth-bound kernels (planes, depthwise, ECA, heads,
// policy gather/softmax, value shaping).  All take the stream to launch on.
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include "../../include/hm_engine.h"
#include "hm_act.h"

namespace hm {

constexpr int PLANES_NHWC_C = 80;   // 74 input channels padded to a 16-byte-aligned row
constexpr int POLICY_LD = 160;      // 2 x 73 policy channels padded to a multiple of 32

// ---- stage P ------------------------------------------------------------------------
// packed positions -> [n,74,8,8] (NCHW) in dtype HM_DTYPE_* (planes.cc:213-265)
void launch_planes_nchw(const hm_packed_pos* pos, int n, void* out, int dtype, cudaStream_t s);
// packed positions -> [n,64,80] 16-bit NHWC (network input layout; channels 74..79 zero)
void launch_planes_nhwc(const hm_packed_pos* pos, int n, act_t* out, cudaStream_t s);
// dense fp16 NCHW planes -> 16-bit NHWC (reference-shaped entry, engine.cc:598-602 input)
void launch_planes_f16_to_nhwc(const __half* in, int n, act_t* out, cudaStream_t s);

// ---- stage F, bandwidth-bound layers ------------------------------------------------------
// depthwise k x k + folded BN + ReLU on [n,64,C] 16-bit (builder_util.py:468-470)
// w: [k*k][C] fp32 (tap-major), b: [C] fp32
void launch_depthwise(const act_t* in, act_t* out, const float* w, const float* b,
                      int n, int C, int ksize, cudaStream_t s);
// ECA (builder_util.py:49-80 with hard sigmoid): x[n,64,384] -> x * hardsigmoid(Wc*mean(x)+b)
// wt: [384 in][384 out] 16-bit (centre tap, transposed), b: [384] fp32.  in may equal out.
void launch_eca(const act_t* in, act_t* out, const act_t* wt,
                const float* b, int n, cudaStream_t s);
// ECA gate only: gate[n][384] fp32 (the fused block kernel applies it)
void launch_eca_gate(const act_t* in, const act_t* wt, const float* b, float* gate, int n,
                     cudaStream_t s);
// value head tail (builder_util.py:311-324): feat [n,64,16] 16-bit (after conv+BN+ReLU),
// w4: [64 sq][16 c][4] fp32 (3 wdl rows + plys row, from Linear weights indexed c*64+sq),
// b4: [4].  raw5: [n,5] fp32 {value, wdl_loss, wdl_draw, wdl_win, moves_left};
// half outputs may be null.
void launch_value_tail(const act_t* feat, const float* w4, const float* b4, int n,
                       float* raw5, __half* value_h, __half* wdl_h, __half* ml_h,
                       cudaStream_t s);
// policy logits [n*64, POLICY_LD] fp32 (col = board*73 + ch) -> fp16 [n,4672] x2, idx = ch*64+sq
void launch_policy_dense_half(const float* logits, int n, __half* pi_a, __half* pi_b,
                              cudaStream_t s);

// ---- stage D ------------------------------------------------------------------------
// get_normalized_probability for both boards of n_pos positions (utils.h:226-243) reading
// the NHWC fp32 logits; optional value shaping of raw5 (searchthread.cc:569-619).
void launch_policy_priors_nhwc(const float* logits, const hm_packed_pos* pos, int n_pos,
                               const uint32_t* moves, const uint32_t* offsets, float* priors,
                               cudaStream_t s);
// the same on dense fp16 [n,4672] logits per board
void launch_policy_priors_dense(const __half* pi_a, const __half* pi_b, const hm_packed_pos* pos,
                                int n_pos, const uint32_t* moves, const uint32_t* offsets,
                                float* priors, cudaStream_t s);
void launch_policy_indices(const uint32_t* moves, const uint8_t* stm_black, int n, int32_t* out,
                           cudaStream_t s);
void launch_shape_values_raw5(const float* raw5, int n, hm_value_config cfg, float* out,
                              cudaStream_t s);
void launch_shape_values_half(const __half* value, const __half* wdl, const __half* ml, int n,
                              hm_value_config cfg, float* out, cudaStream_t s);

}  // namespace hm
