/*
 * hm_engine.h — C ABI of the B200-native batched leaf evaluator for Hivemind.
 *
 * This is the drop-in boundary for the reference's nn/ inference engine
 * (reference: engine/src/nn/engine.h:43-132, class Engine).  Every entry point
 * below names the reference interface it replaces.  Plain pointers and sizes
 * only; no C++ or torch types cross this boundary.  All functions return
 * 1 on success and 0 on failure (the reference returns bool and never throws,
 * engine.cc:105-111); a failure message is retrievable with hm_last_error().
 *
 * Threading contract (reference: engine.h:72-76, engine.cc:579-588,655-664):
 * an engine owns `n_workers` independent execution slots (stream + buffers).
 * Distinct workers may be driven concurrently from distinct host threads;
 * each worker may have at most ONE request in flight (a second enqueue before
 * the matching sync returns 0).
 */
#ifndef HM_ENGINE_H
#define HM_ENGINE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define HM_NB_INPUT_CHANNELS 74          /* constants.h:12-13 */
#define HM_NB_CHANNELS_PER_BOARD 37
#define HM_NB_INPUT_VALUES 4736          /* 74*8*8, constants.h:17-19 */
#define HM_NB_POLICY_CHANNELS 73
#define HM_NB_POLICY_VALUES 4672         /* 73*8*8, constants.h:21-23 */

/* Output element types for hm_encode_planes. */
#define HM_DTYPE_F32 0
#define HM_DTYPE_F16 1
#define HM_DTYPE_BF16 2
#define HM_DTYPE_U8 3                    /* selfplay.cc:464-476 quantisation */

/* Model architectures (reference: src/architectures/rise_mobile_v3.py:820-888). */
#define HM_ARCH_RISEV33 0
#define HM_ARCH_CROSSBOARD_RISEV33 1

/*
 * Packed bughouse position: everything planes.cc reads through the Board
 * accessors (board.h:239-332), 128 bytes per board, 256 bytes per position.
 * Colours are absolute (0 = WHITE, 1 = BLACK), squares a1 = 0 .. h8 = 63.
 */
typedef struct hm_packed_board {
    uint64_t pieces[12];   /* [color*6 + (pt-1)], pt = P,N,B,R,Q,K  (board.h:259-261) */
    uint64_t promoted[2];  /* promoted_pieces & pieces(color)       (board.h:269-271,287-289) */
    uint8_t pocket[10];    /* [color*5 + (pt-1)], pt = P,N,B,R,Q    (board.h:239-241) */
    uint8_t ep_square;     /* 0..63, or 64 = none                   (board.h:278-280) */
    uint8_t flags;         /* bit0 side to move (1 = BLACK)         (board.h:248-250)
                              bit1 WHITE_OO bit2 WHITE_OOO bit3 BLACK_OO bit4 BLACK_OOO (board.h:306-308)
                              board 0 only: bit5 teamSide (1 = BLACK), bit6 hasTimeAdvantage */
    uint8_t last_from;     /* 0..63; 64 = no last move; 65 = last move was a drop (board.h:319-324) */
    uint8_t last_to;       /* 0..63; 64 = no last move */
    uint8_t rule50;        /* min(rule50_count, 255)                (board.h:315-317) */
    uint8_t repetitions;   /* min(repetition_count, 255)            (board.h:326-332) */
} hm_packed_board;

typedef struct hm_packed_pos {
    hm_packed_board board[2];
} hm_packed_pos;

#define HM_FLAG_STM_BLACK 0x01
#define HM_FLAG_WHITE_OO 0x02
#define HM_FLAG_WHITE_OOO 0x04
#define HM_FLAG_BLACK_OO 0x08
#define HM_FLAG_BLACK_OOO 0x10
#define HM_FLAG_TEAM_BLACK 0x20
#define HM_FLAG_TIME_ADVANTAGE 0x40
#define HM_SQ_NONE 64
#define HM_LAST_FROM_DROP 65

/* Value shaping parameters (search_params.h:122-128,148,275-295 RuntimeConfig). */
typedef struct hm_value_config {
    int enable_wdl_eval;        /* ENABLE_WDL_EVAL = true */
    float wdl_value_weight;     /* WDL_VALUE_WEIGHT = 0.25 */
    float moves_left_discount;  /* MOVES_LEFT_DISCOUNT = 0.005 */
    float draw_contempt;        /* DRAW_CONTEMPT = 0 */
} hm_value_config;

typedef struct hm_engine hm_engine;

/* ---- lifetime ------------------------------------------------------------------ */

/* Engine::Engine(int deviceId, int batchSize) — engine.h:53, engine.cc:186-189.
 * n_workers replaces the compile-time NUM_SEARCH_THREADS (search_params.h:30). */
hm_engine* hm_engine_create(int device, int batch, int n_workers);

/* Engine::~Engine — engine.cc:191-215. */
void hm_engine_destroy(hm_engine* e);

/* Engine::loadNetwork(onnxFile, engineFile) — engine.h:60, engine.cc:217-248.
 * `path` names a packed-weight file (.hvmw, see hivemind_b200/weights.py); a path
 * ending in .onnx is mapped to the sibling .hvmw. */
int hm_engine_load(hm_engine* e, const char* path);

/* Same, from a blob already in host memory (used for NCCL-broadcast refreshes). */
int hm_engine_load_memory(hm_engine* e, const void* blob, size_t bytes);

/* Engine::getBatchSize — engine.h:81. */
int hm_engine_batch_size(const hm_engine* e);
int hm_engine_num_workers(const hm_engine* e);
int hm_engine_arch(const hm_engine* e);

/* ---- reference-shaped inference (dense fp16 planes in, dense fp16 outputs out) ---- */

/* Engine::enqueueInferenceHalf(const __half* obs, size_t workerIndex) —
 * engine.h:74, engine.cc:573-650.  obs = [B,74,8,8] fp16 NCHW host memory
 * (pinned for full speed) that must stay unchanged until the matching sync. */
int hm_engine_enqueue_half(hm_engine* e, const uint16_t* obs, size_t worker);

/* Engine::synchronizeInferenceHalf(HalfInferenceOutputs&, size_t) —
 * engine.h:75-76, engine.cc:652-679.  Pointers are into engine-owned pinned host
 * buffers, valid until this worker's next enqueue.  value[B], pi_a[B*4672],
 * pi_b[B*4672] raw logits (index ch*64+sq), wdl[B*3] raw logits (loss,draw,win),
 * moves_left[B] in [0,1]; all fp16 bit patterns. */
int hm_engine_sync_half(hm_engine* e, size_t worker, const uint16_t** value,
                        const uint16_t** pi_a, const uint16_t** pi_b,
                        const uint16_t** wdl, const uint16_t** moves_left);

/* Engine::runInferenceHalf — engine.h:69-70 (enqueue + sync). */
int hm_engine_run_half(hm_engine* e, const uint16_t* obs, size_t worker,
                       const uint16_t** value, const uint16_t** pi_a,
                       const uint16_t** pi_b, const uint16_t** wdl,
                       const uint16_t** moves_left);

/* Engine::runInference(float* obs, float* value, float* piA, float* piB,
 * float* wdl, float* movesLeft, size_t workerIndex) — engine.h:66-67,
 * engine.cc:537-566: converts through fp16 both ways. */
int hm_engine_run_float(hm_engine* e, const float* obs, float* value, float* pi_a,
                        float* pi_b, float* wdl, float* moves_left, size_t worker);

/* ---- fused B200 path: packed positions in, priors + shaped values out ----------- */

/* Replaces, in one device pass: board_to_planes (planes.cc:213-265), the forward
 * (engine.cc:577-650), get_normalized_probability for both boards
 * (utils.h:226-243 as called at searchthread.cc:525-545) and the value shaping
 * (searchthread.cc:569-619).  `moves` holds Fairy-Stockfish Move bit patterns
 * (types.h:735-798), concatenated per (position, board): board A of position i
 * is moves[move_offsets[2i] .. move_offsets[2i+1]), board B is
 * [move_offsets[2i+1] .. move_offsets[2i+2]).  Lists are exactly what the
 * reference passes to get_normalized_probability (legal moves minus R/B
 * under-promotions, MOVE_NONE appended).  n_pos <= batch. */
int hm_engine_enqueue_positions(hm_engine* e, const hm_packed_pos* pos, size_t n_pos,
                                const uint32_t* moves, const uint32_t* move_offsets,
                                size_t worker);

/* priors: float per move, same offsets as `moves`; values: n_pos shaped values
 * (searchthread.cc:569-619); raw: n_pos*5 floats {value, wdl_loss, wdl_draw,
 * wdl_win, moves_left} before shaping.  Any out-pointer may be NULL. */
int hm_engine_sync_priors(hm_engine* e, size_t worker, const float** priors,
                          const float** values, const float** raw);

/* Runtime configuration of the value shaping (SearchParams::RuntimeConfig fields of search_params.h:275-295 read by
 * searchthread.cc:569-619).  Returns 0 (and changes nothing) while any worker has a request in flight. */
int hm_engine_set_value_config(hm_engine* e, const hm_value_config* cfg);

/* ---- root policy on the device: temperature, Dirichlet noise, joint-action sampling ---- */

/* Replaces the host loops that post-process root / opening priors in the reference:
 *   apply_temperature             tools/selfplay.cc:167-185 (raw-policy openings, :283-299)
 *   root Dirichlet noise          search/node.h:288-315 (Node::try_init_and_expand, depth 0)
 *   sample_raw_policy_action      tools/selfplay.cc:301-376, legality = JointActionCandidate
 *                                 (environment/joint_action.h:14-110)
 * Stages run in that order; each is skipped when its inputs are absent. */
typedef struct hm_root_policy {
    double temperature;            /* > 0: p_i = pow(max(0,p_i), 1/T) renormalised; 0 = skip; < 0 = error */
    float dirichlet_alpha;         /* RuntimeConfig::rootDirichletAlpha   (search_params.h:292) */
    float dirichlet_epsilon;       /* RuntimeConfig::rootDirichletEpsilon (search_params.h:293); noise runs when
                                      alpha > 0, epsilon > 0, position_hash != NULL and the list has > 1 move */
    uint64_t noise_seed;           /* RuntimeConfig::rootNoiseSeed; key = seed ^ position_hash[i] ^ board salt
                                      (node.h:295-296).  The variates are Gamma(alpha,1) as in the reference but come
                                      from a counter-based generator, not mt19937_64: same distribution, other stream */
    const uint64_t* position_hash; /* [n_pos] or NULL */
    const uint8_t* captures;       /* one flag per move (same offsets as `moves`): Board::is_capture, 0 for MOVE_NONE
                                      (selfplay.cc:337-347); NULL = no move captures */
    const double* uniforms;        /* [n_pos][3] or NULL (no sampling): the std::generate_canonical<double,53> draws
                                      std::discrete_distribution makes for board A, board B and the legal-pair
                                      resample.  A list of < 2 entries consumes no draw in the reference, nor does a
                                      resample over a single legal pair: see INTEGRATION.md for the host recipe that
                                      keeps a std::mt19937_64 in step */
} hm_root_policy;

#define HM_ROOT_OK 0
#define HM_ROOT_RESAMPLED 1          /* independent picks were an illegal sit; drawn again from the legal pairs
                                        (uniforms[3i+2] was consumed) */
#define HM_ROOT_RESAMPLED_FORCED 2   /* the same, but only ONE pair was legal: taken without a draw, as
                                        std::discrete_distribution of one weight does */
#define HM_ROOT_EMPTY_LIST (-1)      /* a board's move list was empty (the reference always appends MOVE_NONE) */
#define HM_ROOT_BAD_WEIGHTS (-2)     /* a prior list did not sum to a positive finite number */
#define HM_ROOT_NO_LEGAL_JOINT (-3)  /* selfplay.cc:369-371 throws "Raw policy produced no legal joint action" */

/* Stand-alone stage on HOST priors (same offsets as `moves`); priors are updated in place.
 * picks: [n_pos][2] indices into board A's and board B's move list; status: [n_pos] HM_ROOT_*;
 * noise_out: the raw gamma variates per move (for inspection).  picks/status/noise_out may be NULL. */
int hm_root_policy_apply(hm_engine* e, const hm_packed_pos* pos, size_t n_pos, const uint32_t* moves,
                         const uint32_t* move_offsets, float* priors, const hm_root_policy* rp,
                         float* noise_out, uint32_t* picks, int32_t* status);

/* Fused: hm_engine_enqueue_positions followed by the root stages on the device priors, one
 * stream, no host round trip between the forward and the sampler (selfplay.cc:240-260 runs a
 * whole fp32 runInference per opening ply for this). */
int hm_engine_enqueue_positions_root(hm_engine* e, const hm_packed_pos* pos, size_t n_pos,
                                     const uint32_t* moves, const uint32_t* move_offsets,
                                     const hm_root_policy* rp, size_t worker);
/* As hm_engine_sync_priors, plus picks [n_pos][2] and status [n_pos] (engine-owned pinned memory). */
int hm_engine_sync_root(hm_engine* e, size_t worker, const float** priors, const float** values,
                        const float** raw, const uint32_t** picks, const int32_t** status);

/* Allocation-free variant of hm_root_policy_apply.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  It calls no cudaMalloc / cudaFree / allocating library routine, copies nothing and does
 * not synchronise: it launches the root-policy kernel on `stream` (a cudaStream_t; NULL = the default stream) of
 * `device` and returns.
 *   - every buffer is the CALLER's DEVICE memory: d_pos [n_pos], d_moves [n_moves], d_move_offsets [2*n_pos+1]
 *     (non-decreasing, d_move_offsets[2*n_pos] == n_moves: not checked, they are not readable from the host here),
 *     d_priors [n_moves] (updated in place), d_picks [n_pos][2] / d_status [n_pos] (may be NULL), and the three
 *     pointers inside *rp (position_hash, captures, uniforms — DEVICE pointers here; the struct itself is host memory);
 *   - the kernel's scratch (one Gamma variate per move for the Dirichlet-noise stage) is a static device array of
 *     HM_ROOT_NOALLOC_MAX_MOVES floats (4 MiB) that every GPU which loads the library holds until the process exits;
 *     n_moves beyond that limit is refused with an error.  The first call on a device resolves the array's device
 *     address (cudaGetSymbolAddress): make it once outside stream capture;
 *   - all calls on a device share that array, which keeps the previous call's contents (nothing in it has to start at
 *     zero: the kernel writes every variate before it reads it).  Two calls whose kernels overlap on the GPU (different
 *     streams, no dependency) RACE on it: order them on one stream or with events.
 * Returns 1 / 0 (hm_last_error) like hm_root_policy_apply; results are bit-identical to that entry point's. */
#define HM_ROOT_NOALLOC_MAX_MOVES (1u << 20)
int hm_root_policy_apply_noalloc(int device, const hm_packed_pos* d_pos, size_t n_pos, const uint32_t* d_moves,
                                 const uint32_t* d_move_offsets, size_t n_moves, float* d_priors,
                                 const hm_root_policy* rp, uint32_t* d_picks, int32_t* d_status, void* stream);

/* ---- stand-alone stages (parity checks and benches) ---------------------------- */

/* board_to_planes for n packed positions -> [n,74,8,8] of `dtype` in HOST memory
 * (planes.cc:259-265; HM_DTYPE_U8 additionally applies selfplay.cc:464-476). */
int hm_encode_planes(hm_engine* e, const hm_packed_pos* pos, size_t n, void* out,
                     int dtype);

/* get_normalized_probability<__half> on host-provided dense logits
 * (utils.h:226-243): logits_a/b = [n_pos,4672] fp16. */
int hm_policy_priors_half(hm_engine* e, const uint16_t* logits_a,
                          const uint16_t* logits_b, const hm_packed_pos* pos,
                          size_t n_pos, const uint32_t* moves,
                          const uint32_t* move_offsets, float* priors_out);

/* get_fast_policy_index (utils.h:184-216) for n moves, computed on the device.
 * stm_black[i] is the side to move for moves[i]. */
int hm_policy_indices(hm_engine* e, const uint32_t* moves, const uint8_t* stm_black,
                      size_t n, int32_t* idx_out);

/* Value shaping (searchthread.cc:569-619) of n fp16 quintuples on the device. */
int hm_shape_values_half(hm_engine* e, const uint16_t* value, const uint16_t* wdl,
                         const uint16_t* moves_left, size_t n, float* out);

/* ---- device-resident entry points (used by bench.py's `value` leg) --------------- */

/* Runs `iters` complete fused passes (planes -> forward -> priors/values) on data the
 * previous hm_engine_enqueue_positions left resident on the device; returns elapsed
 * device milliseconds measured with CUDA events on the worker's stream, and the number
 * of kernel launches per pass. */
int hm_engine_bench_resident(hm_engine* e, size_t worker, int iters, float* ms_total,
                             int* launches_per_pass);

/* The same with `n_workers` execution slots in flight (passes issued round-robin on the
 * workers' streams, as concurrent search threads do, engine.cc:493-532); elapsed device time
 * between a fork event and a join event on worker 0's stream. */
int hm_engine_bench_resident_multi(hm_engine* e, int n_workers, int iters, float* ms_total,
                                   int* launches_per_pass);

/* Stand-alone, device-resident timing of the bandwidth stages at sizes where GB/s means something
 * (the plane encoder is launch-bound at B = 512).  planes: n positions (copied to the
 * device once) encoded `iters` times into a device buffer of `dtype`; priors: get_normalized_probability
 * for n_pos positions over synthetic fp32 logits already on the device.  Elapsed CUDA-event
 * milliseconds for all `iters` launches. */
int hm_bench_planes(hm_engine* e, const hm_packed_pos* pos, size_t n, int dtype, int iters, float* ms_total);
int hm_bench_priors(hm_engine* e, const hm_packed_pos* pos, size_t n_pos, const uint32_t* moves,
                    const uint32_t* move_offsets, int iters, float* ms_total);

/* End-to-end loop as a C/C++ caller drives it: `iters` requests of hm_engine_enqueue_positions /
 * hm_engine_sync_priors from the given HOST buffers, `slots` workers kept in flight round-robin
 * (searchthread.cc:661-708 keeps one request per worker in flight).  Wall-clock milliseconds
 * (std::chrono::steady_clock) from the first enqueue to the last sync, H2D and D2H included;
 * *checksum = sum of the shaped values of the last request. */
int hm_engine_bench_e2e(hm_engine* e, const hm_packed_pos* pos, size_t n_pos, const uint32_t* moves,
                        const uint32_t* move_offsets, int iters, int slots, double* ms_total, double* checksum);

/* Per-kernel timing of one pass with CUDA events between launches: fills up to `cap`
 * entries of (name, ms); returns the count in *n. */
typedef struct hm_kernel_time {
    char name[48];
    float ms;
    double flops;   /* algorithmic dense FLOPs of this launch (0 for bandwidth kernels) */
    double bytes;   /* algorithmic bytes of this launch */
} hm_kernel_time;
int hm_engine_profile_pass(hm_engine* e, size_t worker, hm_kernel_time* out, int cap,
                           int* n);

/* Debug: launch trunk run `run` of the layer program once with timeline tracing: CTA 0 logs
 * (tag << 48 | SM clock) events of its TMA producer, MMA issuer and first compute warp.
 * out = 3 roles x *cap entries (pass room for 3 x 2048); unused entries are 0. */
int hm_engine_trace_trunk(hm_engine* e, size_t worker, int run, unsigned long long* out, int* cap);

/* Debug: copy an intermediate activation of the last pass on `worker` to host as fp32.
 * name in {"planes","stem","block<i>","trunk","value_feat","policy_shared"}. */
int hm_engine_debug_fetch(hm_engine* e, size_t worker, const char* name, float* out,
                          size_t cap_floats, size_t* n_floats);

/* ---- misc ---------------------------------------------------------------------- */
const char* hm_last_error(void);
const char* hm_version(void);
/* pinned host memory helpers (searchthread.cc:153-174 uses cudaMallocHost). */
void* hm_alloc_pinned(size_t bytes);
void hm_free_pinned(void* p);

#ifdef __cplusplus
}
#endif
#endif /* HM_ENGINE_H */
