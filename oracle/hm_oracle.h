/*
 * hm_oracle.h — TEST INFRASTRUCTURE.  CPU restatement (plain C) of the reference's
 * algorithms for stages P (planes) and D (policy index, masked softmax, value shaping).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
 * may load liboracle.so; the product (hivemind_b200/) never does.
 *
 * Parity status: PINNED.  tests/test_oracle_vs_reference.py checks every function here
 * against the reference itself (oracle/_ref/libhmref.so, compiled from /root/reference)
 * where that library exists, and tests/test_oracle_golden.py checks it against committed
 * golden vectors generated from the reference (tests/golden/, oracle/make_golden.py) and
 * against the reference's own known-answer tests (engine/tests/test_move_gen.cc:68-157,
 * 306-326, 643-767).
 */
#ifndef HM_ORACLE_H
#define HM_ORACLE_H
#include <stddef.h>
#include <stdint.h>
#include "../include/hm_engine.h"

#ifdef __cplusplus
extern "C" {
#endif

/* IEEE conversions, round-to-nearest-even (what __float2half_rn / __float2bfloat16_rn do). */
uint16_t hmo_float_to_half(float f);
float hmo_half_to_float(uint16_t h);
uint16_t hmo_float_to_bf16(float f);
float hmo_bf16_to_float(uint16_t h);

/* planes.cc:213-257 board_to_planes_impl<float>: out = [n][74][64] fp32. */
void hmo_encode_planes_f32(const hm_packed_pos* pos, size_t n, float* out);
/* planes.cc:22-28,263-265 (__half overload) */
void hmo_encode_planes_f16(const hm_packed_pos* pos, size_t n, uint16_t* out);
/* our extension: __float2bfloat16_rn of the fp32 planes */
void hmo_encode_planes_bf16(const hm_packed_pos* pos, size_t n, uint16_t* out);
/* tools/selfplay.cc:464-476: lround(clamp(x,0,1)*255) */
void hmo_encode_planes_u8(const hm_packed_pos* pos, size_t n, uint8_t* out);

/* globals.cc:82-136 init_policy_index — tables derived from the label layout
 * (constants.h:24 UCI_MOVES), without the label strings. Idempotent. */
void hmo_policy_init(void);
/* copies of POLICY_TABLE_NORMAL[2][64][64][2] and POLICY_TABLE_DROP[2][64][8] */
void hmo_policy_tables(int32_t* normal, int32_t* drop);
/* utils.h:184-216 get_fast_policy_index */
int32_t hmo_policy_index(uint32_t move, int stm_black);
/* utils.h:127-167 normalize_logits */
void hmo_normalize_logits(const float* logits, size_t n, float* out);
/* utils.h:226-243 get_normalized_probability<__half> / <float> */
void hmo_priors_f16(const uint16_t* policy, const uint32_t* moves, size_t n, int stm_black,
                    float* out);
void hmo_priors_f32(const float* policy, const uint32_t* moves, size_t n, int stm_black,
                    float* out);
/* searchthread.cc:569-619 value shaping; wdl may be NULL, moves_left may be NULL */
float hmo_shape_value(float value, const float* wdl, const float* moves_left,
                      const hm_value_config* cfg);
/* The same on fp16 bit patterns, as process_batch reads them (searchthread.cc:474-484). */
float hmo_shape_value_half(uint16_t value, const uint16_t* wdl, const uint16_t* moves_left,
                           const hm_value_config* cfg);

/* Batched: priors for n_pos positions given dense fp16 logits [n_pos][4672] per board and the
 * concatenated move lists (layout of hm_engine_enqueue_positions). */
void hmo_batch_priors_f16(const uint16_t* logits_a, const uint16_t* logits_b,
                          const hm_packed_pos* pos, size_t n_pos, const uint32_t* moves,
                          const uint32_t* move_offsets, float* priors_out);

/* ---- root policy stages ---------------------------------------------
 * Pinned against the reference through oracle/_ref (hmref_sample_joint runs the reference's own
 * JointActionCandidate with the real std::discrete_distribution / std::mt19937_64) and the committed
 * fixture tests/golden/root_policy.npz (oracle/make_golden_root.py). */
/* tools/selfplay.cc:167-185 apply_temperature; returns 0 if temperature <= 0 (the reference throws) */
int hmo_apply_temperature(float* p, size_t n, double temperature);
/* search/node.h:301-311: total = sum(noise) in float, p = (1-eps)*p + eps*noise/total (eps clamped) */
void hmo_dirichlet_mix(float* p, const float* noise, size_t n, float epsilon);
/* std::discrete_distribution<size_t>(p, p+n) drawn with the canonical double u, as libstdc++ does it
 * (bits/random.tcc _M_initialize / operator()); n < 2 returns 0 and uses no draw */
size_t hmo_discrete_pick_f32(const float* p, size_t n, double u);
size_t hmo_discrete_pick_f64(const double* w, size_t n, double u);
/* tools/selfplay.cc:301-376 sample_raw_policy_action after the priors are known: independent picks,
 * legality by environment/joint_action.h:14-110, resample over the legal pairs.  u = {uA, uB, uJoint}.
 * Returns HM_ROOT_OK / HM_ROOT_RESAMPLED / HM_ROOT_RESAMPLED_FORCED / HM_ROOT_NO_LEGAL_JOINT / HM_ROOT_EMPTY_LIST / HM_ROOT_BAD_WEIGHTS. */
int hmo_sample_joint(const hm_packed_pos* pos, const uint32_t* moves_a, const float* p_a, const uint8_t* cap_a,
                     size_t n_a, const uint32_t* moves_b, const float* p_b, const uint8_t* cap_b, size_t n_b,
                     const double* u, uint32_t* pick_a, uint32_t* pick_b);
/* std::generate_canonical<double,53> on std::mt19937_64(seed): n successive draws (restated MT19937-64) */
void hmo_mt19937_64_canonical(uint64_t seed, size_t n, double* out);

#ifdef __cplusplus
}
#endif
#endif
