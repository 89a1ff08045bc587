// hm_net.h — host-side runtime of the leaf evaluator: packed weights, per-worker activation
// arenas, the layer program (a list of kernel launches) and its CUDA graphs.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <functional>
#include <map>
#include <memory>
#include <string>
#include <vector>
#include "../../include/hm_engine.h"
#include "hm_act.h"

namespace hm {

void set_error(const std::string& msg);
const char* get_error();
bool cuda_ok(cudaError_t e, const char* what);  // engine.cc:105-111 checkCuda semantics

// ---- .hvmw packed-weight container (written by hivemind_b200/weights.py) -----------------
struct HvmwTensor {
    std::string name;
    std::vector<uint32_t> dims;
    const float* data = nullptr;  // into the owning blob
    size_t numel() const { size_t n = 1; for (auto d : dims) n *= d; return n; }
};
struct HvmwBlob {
    std::vector<uint8_t> bytes;
    uint32_t arch = 0;
    std::map<std::string, HvmwTensor> tensors;
    bool parse();  // fills arch + tensors from bytes
    const HvmwTensor* find(const std::string& n) const {
        auto it = tensors.find(n);
        return it == tensors.end() ? nullptr : &it->second;
    }
};

// ---- device-resident network parameters ----------------------------------------------------
struct DenseLayer {            // one tcgen05 implicit-GEMM convolution (hm_cgemm.cuh)
    std::string name;
    int cin = 0, cout = 0, ksize = 1;
    int cpad = 0;              // cin rounded up to 64 (K columns per tap)
    int n_rows = 0;            // cout rounded up to 128 (rows of the padded W)
    act_t* w = nullptr;        // [n_rows][taps*cpad], K-major
    float* bias = nullptr;     // [cout] or null
    CUtensorMap tm_w;
    // 3x3 layers with 64 < cin <= 80 (the stem): second copy packed for cconv3_kernel<PACKED> (hm_cconv.cuh),
    // [n_rows][kx][256]: channels 0-63 of ky = 0, 1, 2, then channels 64-79 of the three ky, then 16 zero columns
    act_t* w_tap = nullptr;
    CUtensorMap tm_w_tap;
    double macs_per_pos() const { return 64.0 * cin * cout * ksize * ksize; }
};
struct Block {                 // one bottleneck block inside the trunk kernel (hm_trunk.cuh)
    int cop = 0, ksize = 3;
    bool eca = false;
    int chunk0 = 0, n_chunks = 0;  // 128-channel chunks in W1cat / W2cat / par
    int bias2_off = 0;             // offset into Net::trunk_bias2
    int eca_chunk0 = -1;           // row chunk of the ECA weight inside W1cat (in-kernel ECA)
    int eca_bias_off = 0;          // offset of the ECA bias inside Net::trunk_bias2
    double macs_per_pos() const { return 2.0 * 64.0 * 384.0 * cop; }
};
struct XLayer {                // one _CrossBoardAttentionBlock (rise_mobile_v3.py:242-278)
    DenseLayer wq, wkv, wo, fc1, fc2;   // 1x1 "convolutions" over channel-major tokens
    float *qn_g = nullptr, *qn_b = nullptr, *mn_g = nullptr, *mn_b = nullptr, *ln_g = nullptr, *ln_b = nullptr;
    float* wkv_t = nullptr;        // fp32 [192 in][384 out] for the 7 context tokens
    float* bkv = nullptr;          // fp32 [384]
};
struct Net {
    int arch = HM_ARCH_RISEV33;
    int channels = 384;
    DenseLayer stem;
    std::vector<Block> blocks;
    // trunk parameters, all blocks concatenated by 128-channel chunk
    act_t* w1cat = nullptr;       // [chunks*128][384]
    act_t* w2cat = nullptr;       // [chunks*384][128]
    float* trunk_par = nullptr;   // [chunks][27*128]: pw1 bias | dw bias | 25 dw taps
    float* trunk_bias2 = nullptr; // [blocks][384]
    int total_chunks = 0;
    CUtensorMap tm_w1, tm_w2;
    DenseLayer value_conv;       // 384 -> 16
    float* value_w4 = nullptr;   // [1024][4]
    float* value_b4 = nullptr;   // [4]
    float* value_mlp = nullptr;  // use_mlp_wdl_ply value MLP (49 floats, hm_kernels.h) or null
    // tanh value head (nets built without the WDL / plys heads, builder_util.py:297-301): present iff value_w4 is null
    float *value_f0t = nullptr, *value_f0b = nullptr, *value_f2w = nullptr, *value_f2b = nullptr;
    DenseLayer policy_shared;    // 3x3 384 -> 384
    DenseLayer policy_proj;      // 3x3 384 -> 146 (A: 0..72, B: 73..145)
    // risev33 with shared_policy_trunk=False (rise_mobile_v3.py:198-204): two _PolicyHead(384); the two 3x3
    // 384->384 first convs are ONE stacked 384->768 layer over the trunk, the 384->73 second convs are pol_c2[]
    bool policy_per_board = false;
    DenseLayer pol_c1x;
    // crossboard-risev33 only (arch 1)
    DenseLayer tok_proj;          // 1x1 384 -> 384: trunk part of both board token projections stacked
    float *tok_wp = nullptr, *tok_bias = nullptr;      // [2][192][37], [2][192]
    // the whole board-token projection as ONE GEMM (rise_mobile_v3.py:374-385): K = 384 trunk channels + the 74 input
    // planes (board 0's rows see planes 0-36, board 1's rows planes 37-73, zeros elsewhere), bias = tok_bias, ReLU,
    // then + tok_post[row][sq] = pos_emb_t[row % 192][sq] + board_emb[row]
    DenseLayer tok_fold;
    float* tok_post = nullptr;                         // [384][64]
    float *pos_emb = nullptr, *board_emb = nullptr, *ctx_role = nullptr;
    float *pocket_w = nullptr, *pocket_b = nullptr, *scalar_w = nullptr, *scalar_b = nullptr;
    std::vector<XLayer> xlayers;
    DenseLayer vfuse;             // 1x1 384 -> 384 on cat(board maps)
    DenseLayer pol_c1[2], pol_c2[2];   // per-board _PolicyHead(192): 3x3 192->192, 3x3 192->73
    std::vector<void*> allocations;
    size_t param_count = 0;
    ~Net();
};
std::unique_ptr<Net> build_net(const HvmwBlob& blob);

// ---- one launch of the layer program ------------------------------------------------------
struct Op {
    std::string name;
    std::string klass;   // "dense", "depthwise", "eca", "planes", "policy", "head"
    double flops = 0;    // algorithmic dense FLOPs (2*MAC) of this launch
    double bytes = 0;    // algorithmic bytes of this launch (bandwidth-bound kernels)
    std::function<void(cudaStream_t)> launch;
};

struct TrunkRun;  // hm_net.cu (parameters of one trunk_kernel launch, kept for tracing)
struct Worker {
    int batch = 0;        // positions per request (even)
    size_t move_cap = 0;
    cudaStream_t stream = nullptr;
    cudaStream_t side_stream = nullptr;   // captured only: the value-head branch of the graphs
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
    cudaStream_t ctx_stream = nullptr;    // captured only: the context-token branch (ctx.* ops) beside the trunk
    cudaEvent_t ev_fork2 = nullptr, ev_join2 = nullptr;
    bool pending = false;
    int pending_kind = 0;  // 1 = dense half request, 2 = packed positions request, 3 = positions + root stages
    size_t pending_npos = 0, pending_nmoves = 0;

    // device arena.  d_pos | d_offsets | d_moves are consecutive pieces of d_in, d_values | d_raw5 | d_priors
    // of d_out (pinned mirrors h_in / h_out): one copy each way per fused request
    uint8_t *d_in = nullptr, *d_out = nullptr, *h_in = nullptr, *h_out = nullptr;
    size_t in_fixed = 0, out_fixed = 0;   // bytes before the per-move part of each block
    hm_packed_pos* d_pos = nullptr;
    uint32_t* d_moves = nullptr;
    uint32_t* d_offsets = nullptr;
    float* d_priors = nullptr;
    float* d_values = nullptr;
    // root-policy stages (hm_root.cu): per-move capture flags / noise scratch, per-position hashes,
    // uniforms [B][3], picks [B][2], status [B]
    uint8_t* d_captures = nullptr;
    float* d_noise = nullptr;
    uint64_t* d_hash = nullptr;
    double* d_uniforms = nullptr;
    uint32_t* d_picks = nullptr;
    int32_t* d_status = nullptr;
    act_t* d_planes = nullptr;      // [B][74][64] network input (NCHW, the reference's obs layout)
    act_t* d_trunk = nullptr;       // [B][384][64], updated in place by the trunk kernels
    act_t* d_pol_shared = nullptr;  // [B][384][64]
    float* d_logits = nullptr;      // [B][146][64] fp32 (fused path)
    __half* d_pi = nullptr;         // [B][146][64] fp16 (reference-shaped path): A = ch 0..72, B = 73..145
    act_t* d_vfeat = nullptr;       // [B][16][64]
    float* d_raw5 = nullptr;
    __half *d_value_h = nullptr, *d_wdl_h = nullptr, *d_ml_h = nullptr;
    // crossboard-risev33 (arch 1): tokens [B][2][192][64] channel-major and their intermediates
    float* d_tokpre = nullptr;      // fp32 [B][384][64]
    act_t *d_tok[2] = {nullptr, nullptr}, *d_nq = nullptr, *d_nm = nullptr, *d_q = nullptr, *d_kv = nullptr,
          *d_att = nullptr, *d_x1 = nullptr, *d_h = nullptr, *d_fuse = nullptr;
    float *d_ctx = nullptr, *d_ckv = nullptr;
    // pinned host mirrors (engine-owned outputs, engine.h:45-51)
    hm_packed_pos* h_pos = nullptr;
    uint32_t* h_moves = nullptr;
    uint32_t* h_offsets = nullptr;
    float* h_priors = nullptr;
    float* h_values = nullptr;
    float* h_raw5 = nullptr;
    uint8_t* h_captures = nullptr;
    uint64_t* h_hash = nullptr;
    double* h_uniforms = nullptr;
    uint32_t* h_picks = nullptr;
    int32_t* h_status = nullptr;
    __half* h_obs = nullptr;        // staging of runInference's float -> half conversion
    __half *h_value = nullptr, *h_wdl = nullptr, *h_ml = nullptr, *h_pi_a = nullptr,
           *h_pi_b = nullptr;

    std::vector<Op> ops_trunk;    // stem .. policy trunk + value head (planes in -> features)
    std::vector<Op> ops_dense_in, ops_dense_out;   // reference-shaped pre/post
    std::vector<Op> ops_fused_in, ops_fused_out;   // packed-position pre/post
    cudaGraphExec_t graph_dense = nullptr, graph_fused = nullptr;
    std::map<std::string, std::pair<const void*, std::vector<int>>> debug_views;  // name -> (ptr, {rows, cols, is_f32})
    std::vector<void*> dev_allocs, host_allocs;
    std::vector<cudaEvent_t> events;
    std::vector<std::shared_ptr<TrunkRun>> trunk_runs;
};
// bring-up aid: launch trunk run `run` of `w` once with timeline tracing; out = 3 roles x cap
bool trace_trunk_run(Worker& w, int run, unsigned long long* out_host, int* cap);

struct Engine {
    int device = 0, batch = 0, n_workers = 0;
    std::unique_ptr<Net> net;
    std::vector<std::unique_ptr<Worker>> workers;
    hm_value_config vcfg{1, 0.25f, 0.005f, 0.0f};  // search_params.h:122-128,148
    hm_value_config* d_vcfg = nullptr;             // device copy read by the value.shape launch of every graph
    // scratch for the stand-alone stage entry points (guarded by a mutex)
    cudaStream_t util_stream = nullptr;
    void* util_mutex = nullptr;
};

bool engine_setup_workers(Engine* e);  // allocate arenas, build op lists, capture graphs
void engine_free_workers(Engine* e);
bool tensormap_init();                 // resolves cuTensorMapEncodeTiled

}  // namespace hm
