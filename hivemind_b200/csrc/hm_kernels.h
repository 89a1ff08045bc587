// hm_kernels.h — launchers of the non-GEMM kernels (planes, value tail, crossboard token kernels,
// policy gather/softmax, value shaping).  All take the stream to launch on.
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include "../../include/hm_engine.h"
#include "hm_act.h"

namespace hm {

constexpr int POLICY_CH = 146;       // two 73-channel projections stacked (board A | board B)

// ---- stage P ------------------------------------------------------------------------
// packed positions -> [n,74,8,8] (NCHW, the network's input layout) in dtype HM_DTYPE_*
// (planes.cc:213-265)
void launch_planes_nchw(const hm_packed_pos* pos, int n, void* out, int dtype, cudaStream_t s);

// ---- stage F, non-GEMM layers ---------------------------------------------------------------
// value head tail (builder_util.py:311-324): feat [n,16,64] 16-bit (after conv+BN+ReLU),
// w4: [1024][4] fp32 (3 wdl rows + plys row of the Linear weights, feature index c*64+sq),
// b4: [4].  raw5: [n,5] fp32 {value, wdl_loss, wdl_draw, wdl_win, moves_left};
// half outputs may be null.
// mlp: null, or the 49 floats of the use_mlp_wdl_ply value MLP (builder_util.py:292-296): [8][4] | [8] | [8] | [1].
void launch_value_tail(const act_t* feat, const float* w4, const float* b4, const float* mlp, int n,
                       float* raw5, __half* value_h, __half* wdl_h, __half* ml_h,
                       cudaStream_t s);
// tanh value head (builder_util.py:297-301, :326): w1t fp32 [1024 features][512 hidden], b1 [512], w2 [512], b2 [1];
// raw5 = {value, NaN, NaN, NaN, 0}
void launch_value_tanh(const act_t* feat, const float* w1t, const float* b1, const float* w2, const float* b2, int n,
                       float* raw5, __half* value_h, __half* wdl_h, __half* ml_h, cudaStream_t s);

// ---- stage X (crossboard-risev33, rise_mobile_v3.py:242-412), hm_xboard.cu ------------------------
// tokens: fp16 [n][2 boards][192][64] channel-major; see the kernels for the formulas
void launch_token_embed(const float* pre, const act_t* planes, const float* wp, const float* bias,
                        const float* pos_emb, const float* board_emb, act_t* tok, int n, cudaStream_t s);
void launch_context_tokens(const act_t* planes, const float* pocket_w, const float* pocket_b,
                           const float* scalar_w, const float* scalar_b, const float* role, float* ctx,
                           int n, cudaStream_t s);
void launch_token_layernorm(const act_t* x, const float* g1, const float* b1, act_t* y1, const float* g2,
                            const float* b2, act_t* y2, int n2, cudaStream_t s);
void launch_context_kv(const float* ctx, const float* g, const float* b, const float* wkv_t,
                       const float* bkv, float* ckv, int n, cudaStream_t s);
void launch_cross_attention(const act_t* q, const act_t* kv, const float* ckv, act_t* att, int n,
                            cudaStream_t s);
bool xboard_prepare();

// ---- stage D ------------------------------------------------------------------------
// get_normalized_probability for both boards of n_pos positions (utils.h:226-243) reading
// the fp32 logits [n][146][64] (index (pos*2+board)*4672 + ch*64+sq).
void launch_policy_priors_packed(const float* logits, const hm_packed_pos* pos, int n_pos,
                               const uint32_t* moves, const uint32_t* offsets, float* priors,
                               cudaStream_t s);
// the same on dense fp16 [n,4672] logits per board
void launch_policy_priors_dense(const __half* pi_a, const __half* pi_b, const hm_packed_pos* pos,
                                int n_pos, const uint32_t* moves, const uint32_t* offsets,
                                float* priors, cudaStream_t s);
void launch_policy_indices(const uint32_t* moves, const uint8_t* stm_black, int n, int32_t* out,
                           cudaStream_t s);
// cfg: DEVICE pointer (the engine's value configuration, updated in place by hm_engine_set_value_config)
void launch_shape_values_raw5(const float* raw5, int n, const hm_value_config* cfg, float* out,
                              cudaStream_t s);
void launch_shape_values_half(const __half* value, const __half* wdl, const __half* ml, int n,
                              hm_value_config cfg, float* out, cudaStream_t s);

// ---- root policy (hm_root.cu): temperature, Dirichlet noise, joint-action sampling --------------
struct RootPolicyArgs {
    double temperature = 0.0;
    float alpha = 0.0f, epsilon = 0.0f;
    uint64_t seed = 0;
    const uint64_t* hash = nullptr;     // device [n_pos]
    const uint8_t* captures = nullptr;  // device, per move
    const double* uniforms = nullptr;   // device [n_pos][3]
};
// priors updated in place; noise = scratch of the same size as priors (required when the noise stage
// runs); picks [n_pos][2] / status [n_pos] may be null
void launch_root_policy(const hm_packed_pos* pos, int n_pos, const uint32_t* moves, const uint32_t* offsets,
                        float* priors, const RootPolicyArgs& a, float* noise, uint32_t* picks, int32_t* status,
                        cudaStream_t s);


}  // namespace hm
