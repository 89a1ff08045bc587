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
struct DenseLayer {            // one tcgen05 GEMM layer (1x1 or kxk convolution)
    std::string name;
    int cin = 0, cout = 0, ksize = 1;
    int cpad = 0;              // cin rounded up to 64 (K columns per tap)
    int n_rows = 0;            // cout rounded up to 128 (rows of the padded W)
    int n_valid = 0;           // columns stored by the epilogue (multiple of 8)
    act_t* w = nullptr;  // [n_rows][taps*cpad]
    float* bias = nullptr;       // [n_valid] or null
    CUtensorMap tm_w;
    double macs_per_pos() const { return 64.0 * cin * cout * ksize * ksize; }
};
struct Block {
    int cop = 0, ksize = 3;
    bool eca = false;
    DenseLayer pw1, pw2;
    float* dw_w = nullptr;  // [k*k][cop]
    float* dw_b = nullptr;  // [cop]
    act_t* eca_wt = nullptr;  // [384 in][384 out]
    float* eca_b = nullptr;
    // fused-block (mbconv_kernel) parameter layout
    int cop_pad = 0;          // cop rounded up to 64
    act_t* mb_w1 = nullptr;   // [cop_pad][384]
    act_t* mb_w2 = nullptr;   // [384][cop_pad]
    uint8_t* mb_par = nullptr;  // per 64-channel chunk: bias1 f32[64] | dw bias half[64] | dw w half[taps][64]
    int mb_par_stride = 0;
    CUtensorMap tm_mw1, tm_mw2;
};
struct Net {
    int arch = HM_ARCH_RISEV33;
    int channels = 384;
    DenseLayer stem;
    std::vector<Block> blocks;
    DenseLayer value_conv;       // 384 -> 16
    float* value_w4 = nullptr;   // [64][16][4]
    float* value_b4 = nullptr;   // [4]
    DenseLayer policy_shared;    // 3x3 384 -> 384
    DenseLayer policy_proj;      // 3x3 384 -> 146 (A: 0..72, B: 73..145), padded to 160
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

struct Worker {
    int batch = 0;        // positions per request (even)
    size_t move_cap = 0;
    cudaStream_t stream = nullptr;
    bool pending = false;
    int pending_kind = 0;  // 1 = dense half request, 2 = packed positions request
    size_t pending_npos = 0, pending_nmoves = 0;

    // device arena
    hm_packed_pos* d_pos = nullptr;
    uint32_t* d_moves = nullptr;
    uint32_t* d_offsets = nullptr;
    float* d_priors = nullptr;
    float* d_values = nullptr;
    __half* d_obs = nullptr;
    act_t* d_planes = nullptr;
    act_t* d_trunk[2] = {nullptr, nullptr};
    act_t* d_exp[2] = {nullptr, nullptr};
    act_t* d_pol_shared = nullptr;
    float* d_logits = nullptr;
    act_t* d_vfeat = nullptr;
    float* d_raw5 = nullptr;
    float* d_gate = nullptr;
    __half *d_value_h = nullptr, *d_wdl_h = nullptr, *d_ml_h = nullptr, *d_pi_a = nullptr,
           *d_pi_b = nullptr;
    // pinned host mirrors (engine-owned outputs, engine.h:45-51)
    hm_packed_pos* h_pos = nullptr;
    uint32_t* h_moves = nullptr;
    uint32_t* h_offsets = nullptr;
    float* h_priors = nullptr;
    float* h_values = nullptr;
    float* h_raw5 = nullptr;
    __half* h_obs = nullptr;
    __half *h_value = nullptr, *h_wdl = nullptr, *h_ml = nullptr, *h_pi_a = nullptr,
           *h_pi_b = nullptr;

    std::vector<Op> ops_trunk;    // stem .. heads (planes NHWC in -> logits/raw5)
    std::vector<Op> ops_dense_in, ops_dense_out;   // reference-shaped pre/post
    std::vector<Op> ops_fused_in, ops_fused_out;   // packed-position pre/post
    cudaGraphExec_t graph_dense = nullptr, graph_fused = nullptr;
    std::map<std::string, std::pair<const void*, std::vector<int>>> debug_views;  // name -> (ptr, {rows, cols, is_f32})
    std::vector<void*> dev_allocs, host_allocs;
    std::vector<cudaEvent_t> events;
};

struct Engine {
    int device = 0, batch = 0, n_workers = 0;
    std::unique_ptr<Net> net;
    std::vector<std::unique_ptr<Worker>> workers;
    hm_value_config vcfg{1, 0.25f, 0.005f, 0.0f};  // search_params.h:122-128,148
    // scratch for the stand-alone stage entry points (guarded by a mutex)
    cudaStream_t util_stream = nullptr;
    void* util_mutex = nullptr;
};

bool engine_setup_workers(Engine* e);  // allocate arenas, build op lists, capture graphs
void engine_free_workers(Engine* e);
bool tensormap_init();                 // resolves cuTensorMapEncodeTiled

}  // namespace hm
