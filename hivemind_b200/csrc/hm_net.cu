// hm_net.cu — weights ingress (.hvmw), device parameter layout, per-worker layer program.
// Layer list follows the reference model (src/architectures/rise_mobile_v3.py:820-849
// get_rise_v33_model; builder_util.py:154-178 stem, :445-483 bottleneck block, :49-80 ECA,
// :246-326 value head; rise_mobile_v3.py:38-57 shared policy heads).
#include "hm_net.h"

#include <cstring>
#include <mutex>

#include "hm_gemm.cuh"
#include "hm_mbconv.cuh"
#include "hm_kernels.h"

namespace hm {

// ---- errors ---------------------------------------------------------------------------
static thread_local std::string g_error;
void set_error(const std::string& msg) {
    g_error = msg;
    fprintf(stderr, "[hm_engine] %s\n", msg.c_str());  // reference logs to std::cerr
}
const char* get_error() { return g_error.c_str(); }
bool cuda_ok(cudaError_t e, const char* what) {
    if (e == cudaSuccess) return true;
    set_error(std::string(what) + ": " + cudaGetErrorString(e));
    return false;
}

// ---- .hvmw ------------------------------------------------------------------------------
// header (64 B): "HVMW", u32 version, u32 arch, u32 n_tensors, u64 data_offset, zero pad
// entry (104 B): char name[64], u32 dtype (0 = f32), u32 ndim, u32 dims[4], u64 offset, u64 nbytes
bool HvmwBlob::parse() {
    if (bytes.size() < 64 || std::memcmp(bytes.data(), "HVMW", 4) != 0) {
        set_error("hvmw: bad magic");
        return false;
    }
    uint32_t version, n;
    std::memcpy(&version, bytes.data() + 4, 4);
    std::memcpy(&arch, bytes.data() + 8, 4);
    std::memcpy(&n, bytes.data() + 12, 4);
    if (version != 1) { set_error("hvmw: unsupported version"); return false; }
    if (bytes.size() < 64 + static_cast<size_t>(n) * 104) { set_error("hvmw: truncated table"); return false; }
    for (uint32_t i = 0; i < n; ++i) {
        const uint8_t* e = bytes.data() + 64 + static_cast<size_t>(i) * 104;
        char name[65];
        std::memcpy(name, e, 64);
        name[64] = 0;
        uint32_t dtype, ndim, dims[4];
        uint64_t off, nbytes;
        std::memcpy(&dtype, e + 64, 4);
        std::memcpy(&ndim, e + 68, 4);
        std::memcpy(dims, e + 72, 16);
        std::memcpy(&off, e + 88, 8);
        std::memcpy(&nbytes, e + 96, 8);
        if (dtype != 0 || ndim > 4 || off + nbytes > bytes.size() || (off & 3)) {
            set_error(std::string("hvmw: bad entry ") + name);
            return false;
        }
        HvmwTensor t;
        t.name = name;
        t.dims.assign(dims, dims + ndim);
        t.data = reinterpret_cast<const float*>(bytes.data() + off);
        if (t.numel() * 4 != nbytes) { set_error(std::string("hvmw: size mismatch ") + name); return false; }
        tensors[t.name] = t;
    }
    return true;
}

// ---- TMA tensor maps -------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn g_encode = nullptr;
bool tensormap_init() {
    static std::once_flag once;
    std::call_once(once, []() {
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            g_encode = reinterpret_cast<EncodeTiledFn>(fn);
    });
    if (!g_encode) set_error("cuTensorMapEncodeTiled unavailable (no CUDA driver?)");
    return g_encode != nullptr;
}
// W: bf16 [rows][kcols] K-major, box = 64 K-elements x box_rows rows, 128-byte swizzle.
static bool make_tm_weights(CUtensorMap* tm, const void* w, int rows, int kcols, int box_rows) {
    cuuint64_t dims[2] = {static_cast<cuuint64_t>(kcols), static_cast<cuuint64_t>(rows)};
    cuuint64_t strides[1] = {static_cast<cuuint64_t>(kcols) * 2};
    cuuint32_t box[2] = {64, static_cast<cuuint32_t>(box_rows)};
    cuuint32_t es[2] = {1, 1};
    CUresult r = g_encode(tm, HM_TMA_DTYPE, 2, const_cast<void*>(w), dims, strides,
                          box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                          CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled(W) failed: " + std::to_string(r)); return false; }
    return true;
}
// A: bf16 NHWC activation [n][8][8][C] as a 4-D map (C, x, y, pos); box = 64 channels x 8 x 8 x
// 2 positions = the 128-row GEMM tile.  Shifted boxes read zeros outside the board.
static bool make_tm_act(CUtensorMap* tm, const void* x, int n, int C) {
    cuuint64_t dims[4] = {static_cast<cuuint64_t>(C), 8, 8, static_cast<cuuint64_t>(n)};
    cuuint64_t strides[3] = {static_cast<cuuint64_t>(C) * 2, static_cast<cuuint64_t>(C) * 16,
                             static_cast<cuuint64_t>(C) * 128};
    cuuint32_t box[4] = {64, 8, 8, 2};
    cuuint32_t es[4] = {1, 1, 1, 1};
    CUresult r = g_encode(tm, HM_TMA_DTYPE, 4, const_cast<void*>(x), dims, strides,
                          box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                          CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled(A) failed: " + std::to_string(r)); return false; }
    return true;
}

// ---- parameters -------------------------------------------------------------------------
Net::~Net() {
    for (void* p : allocations) cudaFree(p);
}
template <typename T>
static T* upload(Net& net, const std::vector<T>& host) {
    void* d = nullptr;
    if (!cuda_ok(cudaMalloc(&d, host.size() * sizeof(T)), "cudaMalloc(weights)")) return nullptr;
    net.allocations.push_back(d);
    if (!cuda_ok(cudaMemcpy(d, host.data(), host.size() * sizeof(T), cudaMemcpyHostToDevice), "upload")) return nullptr;
    return static_cast<T*>(d);
}
static inline int round_up(int v, int m) { return (v + m - 1) / m * m; }

constexpr int GEMM_BN = 128;
constexpr int GEMM_STAGES = 3;

// conv weight [cout][cin][k][k] fp32 (+ optional bias [cout]) -> bf16 [n_rows][k*k*cpad],
// K index = tap*cpad + c with tap = ky*k + kx (cross-correlation: in[y+ky-k/2][x+kx-k/2]).
// `row_offset`/`dst` let two convs share one padded matrix (policy projections A|B).
static void pack_conv(std::vector<act_t>& dst, int ktot, int cpad, const float* w, int cout,
                      int cin, int k, int row_offset) {
    for (int o = 0; o < cout; ++o)
        for (int c = 0; c < cin; ++c)
            for (int t = 0; t < k * k; ++t)
                dst[static_cast<size_t>(row_offset + o) * ktot + t * cpad + c] =
                    float_to_act(w[(static_cast<size_t>(o) * cin + c) * k * k + t]);
}
static bool finish_dense(Net& net, DenseLayer& L, const std::vector<act_t>& w,
                         const std::vector<float>* bias) {
    L.w = upload(net, w);
    if (!L.w) return false;
    if (bias) {
        std::vector<float> b(L.n_valid, 0.0f);
        std::copy(bias->begin(), bias->end(), b.begin());
        L.bias = upload(net, b);
        if (!L.bias) return false;
    }
    return make_tm_weights(&L.tm_w, L.w, L.n_rows, L.ksize * L.ksize * L.cpad, GEMM_BN);
}
static bool make_dense(Net& net, DenseLayer& L, const std::string& name, const HvmwTensor* w,
                       const HvmwTensor* b, int cin_pad_to = 0) {
    if (!w || w->dims.size() < 2) { set_error("missing tensor " + name + ".w"); return false; }
    L.name = name;
    L.cout = w->dims[0];
    L.cin = w->dims[1];
    L.ksize = w->dims.size() == 4 ? w->dims[2] : 1;
    L.cpad = round_up(cin_pad_to > 0 ? cin_pad_to : L.cin, 64);
    L.n_rows = round_up(L.cout, GEMM_BN);
    L.n_valid = round_up(L.cout, 8);
    const int ktot = L.ksize * L.ksize * L.cpad;
    std::vector<act_t> packed(static_cast<size_t>(L.n_rows) * ktot, float_to_act(0.0f));
    pack_conv(packed, ktot, L.cpad, w->data, L.cout, L.cin, L.ksize, 0);
    std::vector<float> bias;
    if (b) bias.assign(b->data, b->data + b->numel());
    net.param_count += w->numel() + (b ? b->numel() : 0);
    return finish_dense(net, L, packed, b ? &bias : nullptr);
}

std::unique_ptr<Net> build_net(const HvmwBlob& blob) {
    if (!tensormap_init()) return nullptr;
    auto net = std::make_unique<Net>();
    net->arch = blob.arch;
    if (blob.arch != HM_ARCH_RISEV33) {
        set_error("architecture " + std::to_string(blob.arch) + " not supported by this build");
        return nullptr;
    }
    auto T = [&](const std::string& n) { return blob.find(n); };
    if (!make_dense(*net, net->stem, "stem", T("stem.w"), T("stem.b"), PLANES_NHWC_C)) return nullptr;
    net->channels = net->stem.cout;
    if (net->channels != 384) { set_error("trunk width must be 384"); return nullptr; }
    for (int i = 0;; ++i) {
        const std::string p = "b" + std::to_string(i) + ".";
        if (!T(p + "pw1.w")) break;
        Block blk;
        if (!make_dense(*net, blk.pw1, p + "pw1", T(p + "pw1.w"), T(p + "pw1.b"))) return nullptr;
        if (!make_dense(*net, blk.pw2, p + "pw2", T(p + "pw2.w"), T(p + "pw2.b"))) return nullptr;
        const HvmwTensor* dw = T(p + "dw.w");
        const HvmwTensor* db = T(p + "dw.b");
        if (!dw || !db || dw->dims.size() != 3) { set_error("missing " + p + "dw"); return nullptr; }
        blk.cop = dw->dims[0];
        blk.ksize = dw->dims[1];
        if (blk.cop % 8 != 0 || blk.cop != blk.pw1.cout || (blk.ksize != 3 && blk.ksize != 5)) {
            set_error("unsupported depthwise shape in " + p);
            return nullptr;
        }
        const int kk = blk.ksize * blk.ksize;
        std::vector<float> dww(static_cast<size_t>(kk) * blk.cop);
        for (int c = 0; c < blk.cop; ++c)
            for (int t = 0; t < kk; ++t) dww[static_cast<size_t>(t) * blk.cop + c] = dw->data[c * kk + t];
        blk.dw_w = upload(*net, dww);
        blk.dw_b = upload(*net, std::vector<float>(db->data, db->data + blk.cop));
        net->param_count += dw->numel() + db->numel();
        if (const HvmwTensor* ew = T(p + "eca.w")) {
            const HvmwTensor* eb = T(p + "eca.b");
            if (!eb || ew->dims.size() != 2 || ew->dims[0] != 384 || ew->dims[1] != 384) {
                set_error("bad ECA tensors in " + p);
                return nullptr;
            }
            blk.eca = true;
            std::vector<act_t> wt(384 * 384);
            for (int o = 0; o < 384; ++o)
                for (int k = 0; k < 384; ++k) wt[k * 384 + o] = float_to_act(ew->data[o * 384 + k]);
            blk.eca_wt = upload(*net, wt);
            blk.eca_b = upload(*net, std::vector<float>(eb->data, eb->data + 384));
            net->param_count += ew->numel() + eb->numel();
        }
        if (!blk.dw_w || !blk.dw_b) return nullptr;
        {   // ---- fused-block layout (hm_mbconv.cuh) ----
            const HvmwTensor *w1 = T(p + "pw1.w"), *b1 = T(p + "pw1.b"), *w2 = T(p + "pw2.w");
            blk.cop_pad = round_up(blk.cop, 64);
            const int cp = blk.cop_pad;
            std::vector<act_t> w1p(static_cast<size_t>(cp) * 384, float_to_act(0.0f));
            for (int o = 0; o < blk.cop; ++o)
                for (int k = 0; k < 384; ++k) w1p[static_cast<size_t>(o) * 384 + k] = float_to_act(w1->data[static_cast<size_t>(o) * 384 + k]);
            std::vector<act_t> w2p(static_cast<size_t>(384) * cp, float_to_act(0.0f));
            for (int o = 0; o < 384; ++o)
                for (int k = 0; k < blk.cop; ++k) w2p[static_cast<size_t>(o) * cp + k] = float_to_act(w2->data[static_cast<size_t>(o) * blk.cop + k]);
            const int par_bytes = 384 + kk * 128;
            blk.mb_par_stride = (par_bytes + 127) / 128 * 128;
            std::vector<uint8_t> par(static_cast<size_t>(cp / 64) * blk.mb_par_stride, 0);
            for (int ch = 0; ch < cp / 64; ++ch) {
                uint8_t* base = par.data() + static_cast<size_t>(ch) * blk.mb_par_stride;
                float* pb1 = reinterpret_cast<float*>(base);
                act_t* pbd = reinterpret_cast<act_t*>(base + 256);
                act_t* pw = reinterpret_cast<act_t*>(base + 384);
                for (int j = 0; j < 64; ++j) {
                    const int c = ch * 64 + j;
                    const bool ok = c < blk.cop;
                    pb1[j] = ok ? b1->data[c] : 0.0f;
                    pbd[j] = float_to_act(ok ? db->data[c] : 0.0f);
                    for (int t = 0; t < kk; ++t) pw[t * 64 + j] = float_to_act(ok ? dw->data[c * kk + t] : 0.0f);
                }
            }
            blk.mb_w1 = upload(*net, w1p);
            blk.mb_w2 = upload(*net, w2p);
            blk.mb_par = upload(*net, par);
            if (!blk.mb_w1 || !blk.mb_w2 || !blk.mb_par) return nullptr;
            if (!make_tm_weights(&blk.tm_mw1, blk.mb_w1, cp, 384, 64)) return nullptr;
            if (!make_tm_weights(&blk.tm_mw2, blk.mb_w2, 384, cp, 128)) return nullptr;
        }
        net->blocks.push_back(blk);
    }
    if (net->blocks.empty()) { set_error("no residual blocks in blob"); return nullptr; }
    // value head
    if (!make_dense(*net, net->value_conv, "value.conv", T("value.conv.w"), T("value.conv.b"))) return nullptr;
    {
        const HvmwTensor *ww = T("value.wdl.w"), *wb = T("value.wdl.b"), *pw = T("value.plys.w"),
                         *pb = T("value.plys.b");
        if (!ww || !wb || !pw || !pb || net->value_conv.cout != 16 || ww->numel() != 3 * 1024 || pw->numel() != 1024) {
            set_error("value head tensors missing (WDL + plys configuration required)");
            return nullptr;
        }
        std::vector<float> w4(64 * 16 * 4), b4(4);
        for (int sq = 0; sq < 64; ++sq)
            for (int c = 0; c < 16; ++c) {
                for (int o = 0; o < 3; ++o) w4[(sq * 16 + c) * 4 + o] = ww->data[o * 1024 + c * 64 + sq];
                w4[(sq * 16 + c) * 4 + 3] = pw->data[c * 64 + sq];
            }
        for (int o = 0; o < 3; ++o) b4[o] = wb->data[o];
        b4[3] = pb->data[0];
        net->value_w4 = upload(*net, w4);
        net->value_b4 = upload(*net, b4);
        net->param_count += ww->numel() + wb->numel() + pw->numel() + pb->numel();
    }
    // policy heads (shared trunk, two 73-channel projections packed side by side)
    if (!make_dense(*net, net->policy_shared, "policy.shared", T("policy.shared.w"), T("policy.shared.b"))) return nullptr;
    {
        const HvmwTensor *pa = T("policy.a.w"), *pb = T("policy.b.w");
        if (!pa || !pb || pa->dims.size() != 4 || pa->dims[0] != 73 || pb->dims[0] != 73) {
            set_error("policy projection tensors missing");
            return nullptr;
        }
        DenseLayer& L = net->policy_proj;
        L.name = "policy.proj";
        L.cout = 146;
        L.cin = pa->dims[1];
        L.ksize = pa->dims[2];
        L.cpad = round_up(L.cin, 64);
        L.n_rows = round_up(POLICY_LD, GEMM_BN);
        L.n_valid = POLICY_LD;
        const int ktot = L.ksize * L.ksize * L.cpad;
        std::vector<act_t> packed(static_cast<size_t>(L.n_rows) * ktot, float_to_act(0.0f));
        pack_conv(packed, ktot, L.cpad, pa->data, 73, L.cin, L.ksize, 0);
        pack_conv(packed, ktot, L.cpad, pb->data, 73, L.cin, L.ksize, 73);
        net->param_count += pa->numel() + pb->numel();
        if (!finish_dense(*net, L, packed, nullptr)) return nullptr;
    }
    return net;
}

// ---- per-worker arenas and layer program ------------------------------------------------
template <typename T>
static bool dev_alloc(Worker& w, T** p, size_t count) {
    void* d = nullptr;
    if (!cuda_ok(cudaMalloc(&d, count * sizeof(T)), "cudaMalloc(arena)")) return false;
    if (!cuda_ok(cudaMemset(d, 0, count * sizeof(T)), "cudaMemset(arena)")) return false;
    w.dev_allocs.push_back(d);
    *p = static_cast<T*>(d);
    return true;
}
template <typename T>
static bool host_alloc(Worker& w, T** p, size_t count) {
    void* h = nullptr;
    if (!cuda_ok(cudaMallocHost(&h, count * sizeof(T)), "cudaMallocHost")) return false;
    std::memset(h, 0, count * sizeof(T));
    w.host_allocs.push_back(h);
    *p = static_cast<T*>(h);
    return true;
}

static bool g_gemm_attr_set = false;
static bool gemm_prepare() {
    if (g_gemm_attr_set) return true;
    using S = GemmSmem<GEMM_BN, GEMM_STAGES>;
    if (!cuda_ok(cudaFuncSetAttribute(gemm_tcgen05_kernel<GEMM_BN, GEMM_STAGES>,
                                      cudaFuncAttributeMaxDynamicSharedMemorySize, S::TOTAL),
                 "cudaFuncSetAttribute(gemm)"))
        return false;
    if (!cuda_ok(cudaFuncSetAttribute(mbconv_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, MbconvSmem<3>::TOTAL), "cudaFuncSetAttribute(mbconv3)") ||
        !cuda_ok(cudaFuncSetAttribute(mbconv_kernel<5>, cudaFuncAttributeMaxDynamicSharedMemorySize, MbconvSmem<5>::TOTAL), "cudaFuncSetAttribute(mbconv5)"))
        return false;
    g_gemm_attr_set = true;
    return true;
}

// One fused bottleneck block as an Op.
static bool add_mbconv(std::vector<Op>& ops, const Block& b, int idx, const act_t* x, act_t* out,
                       const float* gate, int n_pos, int num_sms) {
    CUtensorMap tm_x;
    if (!make_tm_act(&tm_x, x, n_pos, 384)) return false;
    MbconvParams p{};
    p.n_tiles = n_pos / 2;
    p.n_chunks = b.cop_pad / 64;
    p.par = b.mb_par;
    p.par_stride = b.mb_par_stride;
    p.bias2 = b.pw2.bias;
    p.gate = gate;
    p.x = x;
    p.out = out;
    const int grid = std::min(p.n_tiles, num_sms);
    const CUtensorMap tw1 = b.tm_mw1, tw2 = b.tm_mw2;
    const int ks = b.ksize;
    Op op;
    op.name = "b" + std::to_string(idx) + ".mbconv" + std::to_string(ks) + (gate ? "+eca" : "");
    op.klass = "dense";
    op.flops = 2.0 * (b.pw1.macs_per_pos() + b.pw2.macs_per_pos()) * n_pos;
    op.bytes = 2.0 * n_pos * 64 * 384 * 2;
    op.launch = [p, tm_x, tw1, tw2, grid, ks](cudaStream_t s) {
        if (ks == 3)
            mbconv_kernel<3><<<grid, MB_THREADS, MbconvSmem<3>::TOTAL, s>>>(p, tm_x, tw1, tw2);
        else
            mbconv_kernel<5><<<grid, MB_THREADS, MbconvSmem<5>::TOTAL, s>>>(p, tm_x, tw1, tw2);
    };
    ops.push_back(std::move(op));
    return true;
}

// One dense layer as an Op: in = NHWC bf16 [n][64][cin_buf] -> out [n*64][ldc].
static bool add_dense(std::vector<Op>& ops, const DenseLayer& L, const act_t* in,
                      int cin_buf, int n_pos, void* out, int ldc, bool relu, bool out_f32,
                      const act_t* residual, int ldr) {
    CUtensorMap tm_a;
    if (!make_tm_act(&tm_a, in, n_pos, cin_buf)) return false;
    GemmParams p{};
    p.M = n_pos * 64;
    p.n_valid = L.n_valid;
    p.ksize = L.ksize;
    p.kchunks = L.cpad / 64;
    p.bias = L.bias;
    p.residual = residual;
    p.ldr = ldr;
    p.out = out;
    p.ldc = ldc;
    p.relu = relu ? 1 : 0;
    p.out_f32 = out_f32 ? 1 : 0;
    p.chan_sum = nullptr;
    const int n_tiles = (L.n_valid + GEMM_BN - 1) / GEMM_BN;
    const int grid = (p.M / GEMM_BM) * n_tiles;
    const CUtensorMap tm_w = L.tm_w;
    Op op;
    op.name = L.name;
    op.klass = "dense";
    op.flops = 2.0 * L.macs_per_pos() * n_pos;
    op.bytes = 0;
    op.launch = [p, tm_a, tm_w, grid](cudaStream_t s) {
        gemm_tcgen05_kernel<GEMM_BN, GEMM_STAGES>
            <<<grid, GEMM_THREADS, GemmSmem<GEMM_BN, GEMM_STAGES>::TOTAL, s>>>(p, tm_a, tm_w);
    };
    ops.push_back(std::move(op));
    return true;
}

static bool capture_graph(Worker& w, const std::vector<const std::vector<Op>*>& lists,
                          cudaGraphExec_t* exec) {
    cudaGraph_t graph = nullptr;
    if (!cuda_ok(cudaStreamBeginCapture(w.stream, cudaStreamCaptureModeThreadLocal), "BeginCapture")) return false;
    for (auto* l : lists)
        for (const Op& op : *l) op.launch(w.stream);
    if (!cuda_ok(cudaStreamEndCapture(w.stream, &graph), "EndCapture")) return false;
    const bool ok = cuda_ok(cudaGraphInstantiate(exec, graph, 0), "GraphInstantiate");
    cudaGraphDestroy(graph);
    return ok;
}

static bool setup_worker(Engine* e, Worker& w) {
    const Net& net = *e->net;
    const int B = w.batch;
    const size_t rows = static_cast<size_t>(B) * 64;
    int max_cop = 0;
    for (const Block& b : net.blocks) max_cop = std::max(max_cop, b.cop);
    w.move_cap = static_cast<size_t>(B) * 2 * 256;
    if (!cuda_ok(cudaStreamCreateWithFlags(&w.stream, cudaStreamNonBlocking), "cudaStreamCreate")) return false;
    bool ok = dev_alloc(w, &w.d_pos, B) && dev_alloc(w, &w.d_moves, w.move_cap) &&
              dev_alloc(w, &w.d_offsets, 2 * B + 1) && dev_alloc(w, &w.d_priors, w.move_cap) &&
              dev_alloc(w, &w.d_values, B) && dev_alloc(w, &w.d_obs, static_cast<size_t>(B) * HM_NB_INPUT_VALUES) &&
              dev_alloc(w, &w.d_planes, rows * PLANES_NHWC_C) && dev_alloc(w, &w.d_trunk[0], rows * 384) &&
              dev_alloc(w, &w.d_trunk[1], rows * 384) && dev_alloc(w, &w.d_exp[0], rows * max_cop) &&
              dev_alloc(w, &w.d_exp[1], rows * max_cop) && dev_alloc(w, &w.d_pol_shared, rows * 384) &&
              dev_alloc(w, &w.d_logits, rows * POLICY_LD) && dev_alloc(w, &w.d_vfeat, rows * 16) &&
              dev_alloc(w, &w.d_raw5, static_cast<size_t>(B) * 5) && dev_alloc(w, &w.d_gate, static_cast<size_t>(B) * 384) && dev_alloc(w, &w.d_value_h, B) &&
              dev_alloc(w, &w.d_wdl_h, static_cast<size_t>(B) * 3) && dev_alloc(w, &w.d_ml_h, B) &&
              dev_alloc(w, &w.d_pi_a, static_cast<size_t>(B) * HM_NB_POLICY_VALUES) &&
              dev_alloc(w, &w.d_pi_b, static_cast<size_t>(B) * HM_NB_POLICY_VALUES);
    ok = ok && host_alloc(w, &w.h_pos, B) && host_alloc(w, &w.h_moves, w.move_cap) &&
         host_alloc(w, &w.h_offsets, 2 * B + 1) && host_alloc(w, &w.h_priors, w.move_cap) &&
         host_alloc(w, &w.h_values, B) && host_alloc(w, &w.h_raw5, static_cast<size_t>(B) * 5) &&
         host_alloc(w, &w.h_obs, static_cast<size_t>(B) * HM_NB_INPUT_VALUES) &&
         host_alloc(w, &w.h_value, B) && host_alloc(w, &w.h_wdl, static_cast<size_t>(B) * 3) &&
         host_alloc(w, &w.h_ml, B) && host_alloc(w, &w.h_pi_a, static_cast<size_t>(B) * HM_NB_POLICY_VALUES) &&
         host_alloc(w, &w.h_pi_b, static_cast<size_t>(B) * HM_NB_POLICY_VALUES);
    if (!ok || !gemm_prepare()) return false;

    // ---- trunk: stem + blocks + heads ---------------------------------------------------
    auto& ops = w.ops_trunk;
    int cur = 0;
    if (!add_dense(ops, net.stem, w.d_planes, PLANES_NHWC_C, B, w.d_trunk[cur], 384, true, false, nullptr, 0)) return false;
    w.debug_views["stem"] = {w.d_trunk[cur], {static_cast<int>(rows), 384, 0}};
    const char* unfused_env = getenv("HM_UNFUSED");
    const bool fused = !(unfused_env && unfused_env[0] == '1');
    int num_sms = 148;
    cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, e->device);
    for (size_t i = 0; i < net.blocks.size(); ++i) {
        const Block& b = net.blocks[i];
        act_t* x = w.d_trunk[cur];
        act_t* y = w.d_trunk[cur ^ 1];
        if (fused) {
            const float* gate = nullptr;
            if (b.eca) {
                Op op;
                op.name = "b" + std::to_string(i) + ".eca_gate";
                op.klass = "eca";
                op.bytes = static_cast<double>(rows) * 384 * 2;
                const act_t* wt = b.eca_wt;
                const float* eb = b.eca_b;
                float* g = w.d_gate;
                op.launch = [x, wt, eb, g, B](cudaStream_t s) { launch_eca_gate(x, wt, eb, g, B, s); };
                ops.push_back(std::move(op));
                gate = w.d_gate;
            }
            if (!add_mbconv(ops, b, static_cast<int>(i), x, y, gate, B, num_sms)) return false;
            cur ^= 1;
            w.debug_views["block" + std::to_string(i)] = {w.d_trunk[cur], {static_cast<int>(rows), 384, 0}};
            continue;
        }
        if (b.eca) {
            Op op;
            op.name = "b" + std::to_string(i) + ".eca";
            op.klass = "eca";
            op.bytes = 2.0 * rows * 384 * 2;
            const act_t* wt = b.eca_wt;
            const float* eb = b.eca_b;
            op.launch = [x, wt, eb, B](cudaStream_t s) { launch_eca(x, x, wt, eb, B, s); };
            ops.push_back(std::move(op));
        }
        if (!add_dense(ops, b.pw1, x, 384, B, w.d_exp[0], b.cop, true, false, nullptr, 0)) return false;
        {
            Op op;
            op.name = "b" + std::to_string(i) + ".dw" + std::to_string(b.ksize);
            op.klass = "depthwise";
            op.bytes = 2.0 * rows * b.cop * 2;
            const act_t* in = w.d_exp[0];
            act_t* out = w.d_exp[1];
            const float *dw = b.dw_w, *db = b.dw_b;
            const int cop = b.cop, ks = b.ksize;
            op.launch = [in, out, dw, db, B, cop, ks](cudaStream_t s) { launch_depthwise(in, out, dw, db, B, cop, ks, s); };
            ops.push_back(std::move(op));
        }
        if (!add_dense(ops, b.pw2, w.d_exp[1], b.cop, B, y, 384, false, false, x, 384)) return false;
        cur ^= 1;
        w.debug_views["block" + std::to_string(i)] = {w.d_trunk[cur], {static_cast<int>(rows), 384, 0}};
    }
    // NOTE: trunk buffers ping-pong, so only the LAST writer of each buffer is inspectable
    // after a pass; "trunk" is the final activation.
    act_t* trunk = w.d_trunk[cur];
    w.debug_views["trunk"] = {trunk, {static_cast<int>(rows), 384, 0}};
    if (!add_dense(ops, net.value_conv, trunk, 384, B, w.d_vfeat, 16, true, false, nullptr, 0)) return false;
    w.debug_views["value_feat"] = {w.d_vfeat, {static_cast<int>(rows), 16, 0}};
    {
        Op op;
        op.name = "value.tail";
        op.klass = "head";
        op.bytes = static_cast<double>(rows) * 16 * 2;
        const act_t* f = w.d_vfeat;
        const float *w4 = net.value_w4, *b4 = net.value_b4;
        float* raw5 = w.d_raw5;
        __half *vh = w.d_value_h, *wh = w.d_wdl_h, *mh = w.d_ml_h;
        op.launch = [f, w4, b4, B, raw5, vh, wh, mh](cudaStream_t s) { launch_value_tail(f, w4, b4, B, raw5, vh, wh, mh, s); };
        ops.push_back(std::move(op));
    }
    if (!add_dense(ops, net.policy_shared, trunk, 384, B, w.d_pol_shared, 384, true, false, nullptr, 0)) return false;
    w.debug_views["policy_shared"] = {w.d_pol_shared, {static_cast<int>(rows), 384, 0}};
    if (!add_dense(ops, net.policy_proj, w.d_pol_shared, 384, B, w.d_logits, POLICY_LD, false, true, nullptr, 0)) return false;
    w.debug_views["logits"] = {w.d_logits, {static_cast<int>(rows), POLICY_LD, 1}};
    w.debug_views["planes"] = {w.d_planes, {static_cast<int>(rows), PLANES_NHWC_C, 0}};
    w.debug_views["exp_pw1"] = {w.d_exp[0], {static_cast<int>(rows), net.blocks.back().cop, 0}};
    w.debug_views["exp_dw"] = {w.d_exp[1], {static_cast<int>(rows), net.blocks.back().cop, 0}};

    // ---- reference-shaped pre/post ----------------------------------------------------------
    {
        Op op;
        op.name = "planes.f16_to_nhwc";
        op.klass = "planes";
        op.bytes = static_cast<double>(B) * (HM_NB_INPUT_VALUES * 2 + 64 * PLANES_NHWC_C * 2);
        const __half* in = w.d_obs;
        act_t* out = w.d_planes;
        op.launch = [in, out, B](cudaStream_t s) { launch_planes_f16_to_nhwc(in, B, out, s); };
        w.ops_dense_in.push_back(std::move(op));
    }
    {
        Op op;
        op.name = "policy.dense_half";
        op.klass = "policy";
        op.bytes = static_cast<double>(rows) * POLICY_LD * 4 + 2.0 * B * HM_NB_POLICY_VALUES * 2;
        const float* lg = w.d_logits;
        __half *pa = w.d_pi_a, *pb = w.d_pi_b;
        op.launch = [lg, B, pa, pb](cudaStream_t s) { launch_policy_dense_half(lg, B, pa, pb, s); };
        w.ops_dense_out.push_back(std::move(op));
    }
    // ---- fused pre/post ---------------------------------------------------------------------
    {
        Op op;
        op.name = "planes.encode_nhwc";
        op.klass = "planes";
        op.bytes = static_cast<double>(B) * (256 + 64 * PLANES_NHWC_C * 2);
        const hm_packed_pos* pp = w.d_pos;
        act_t* out = w.d_planes;
        op.launch = [pp, out, B](cudaStream_t s) { launch_planes_nhwc(pp, B, out, s); };
        w.ops_fused_in.push_back(std::move(op));
    }
    {
        Op op;
        op.name = "policy.priors";
        op.klass = "policy";
        op.bytes = static_cast<double>(B) * 740;  // SURVEY.md §8d per-position estimate
        const float* lg = w.d_logits;
        const hm_packed_pos* pp = w.d_pos;
        const uint32_t *mv = w.d_moves, *off = w.d_offsets;
        float* pri = w.d_priors;
        op.launch = [lg, pp, B, mv, off, pri](cudaStream_t s) { launch_policy_priors_nhwc(lg, pp, B, mv, off, pri, s); };
        w.ops_fused_out.push_back(std::move(op));
        Op ov;
        ov.name = "value.shape";
        ov.klass = "head";
        ov.bytes = static_cast<double>(B) * 24;
        const float* raw5 = w.d_raw5;
        float* vals = w.d_values;
        const hm_value_config* cfg = &e->vcfg;  // read at launch time: graphs are re-captured on change
        ov.launch = [raw5, B, cfg, vals](cudaStream_t s) { launch_shape_values_raw5(raw5, B, *cfg, vals, s); };
        w.ops_fused_out.push_back(std::move(ov));
    }
    if (!capture_graph(w, {&w.ops_dense_in, &w.ops_trunk, &w.ops_dense_out}, &w.graph_dense)) return false;
    if (!capture_graph(w, {&w.ops_fused_in, &w.ops_trunk, &w.ops_fused_out}, &w.graph_fused)) return false;
    return cuda_ok(cudaStreamSynchronize(w.stream), "worker init sync");
}

bool engine_setup_workers(Engine* e) {
    engine_free_workers(e);
    for (int i = 0; i < e->n_workers; ++i) {
        auto w = std::make_unique<Worker>();
        w->batch = (e->batch + 1) & ~1;
        if (!setup_worker(e, *w)) {
            e->workers.push_back(std::move(w));
            engine_free_workers(e);
            return false;
        }
        e->workers.push_back(std::move(w));
    }
    return true;
}

void engine_free_workers(Engine* e) {
    for (auto& w : e->workers) {
        if (!w) continue;
        if (w->stream) cudaStreamSynchronize(w->stream);  // engine.cc:191-215 order
        if (w->graph_dense) cudaGraphExecDestroy(w->graph_dense);
        if (w->graph_fused) cudaGraphExecDestroy(w->graph_fused);
        for (auto ev : w->events) cudaEventDestroy(ev);
        for (void* p : w->dev_allocs) cudaFree(p);
        for (void* p : w->host_allocs) cudaFreeHost(p);
        if (w->stream) cudaStreamDestroy(w->stream);
    }
    e->workers.clear();
}

}  // namespace hm
