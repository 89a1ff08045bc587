// hm_net.cu — weights ingress (.hvmw), device parameter layout, per-worker layer program.
// Layer list follows the reference model (src/architectures/rise_mobile_v3.py:820-849
// get_rise_v33_model; builder_util.py:154-178 stem, :445-483 bottleneck block, :49-80 ECA,
// :246-326 value head; rise_mobile_v3.py:38-57 shared policy heads).
#include "hm_net.h"

#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <mutex>

#include "hm_cgemm.cuh"
#include "hm_cconv.cuh"
#include "hm_lgemm.cuh"
#include "hm_mlp.cuh"
#include "hm_pgemm.cuh"
#include "hm_trunk.cuh"
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
        // overflow-safe bounds (off + nbytes may wrap for a crafted entry) and data must follow the table
        const uint64_t size = bytes.size(), table_end = 64 + static_cast<uint64_t>(n) * 104;
        if (dtype != 0 || ndim > 4 || off > size || nbytes > size - off || off < table_end || (off & 3)) {
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
// W: 16-bit [rows][kcols] K-major, box = 64 K-elements x 128 rows, 128-byte swizzle.
static bool make_tm_weights(CUtensorMap* tm, const void* w, size_t rows, int kcols, int box_rows = 128) {
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
// NCHW activation [n][C][64] as a 3-D map (square, channel, position); box = 64 squares x 64
// channels x 2 positions = one MN-major B-operand block (also the trunk's tile load / store).
// A box shifted by +-8 squares reads zeros off the board, one shifted beyond the last channel
// reads zeros too (convolution padding and K padding for free).
static bool make_tm_act3(CUtensorMap* tm, const void* x, int n, int C, int pos_stride_ch = 0, int box_ch = 64, int box_pos = 2) {
    cuuint64_t dims[3] = {64, static_cast<cuuint64_t>(C), static_cast<cuuint64_t>(n)};
    cuuint64_t strides[2] = {128, static_cast<cuuint64_t>(pos_stride_ch ? pos_stride_ch : C) * 128};
    cuuint32_t box[3] = {64, static_cast<cuuint32_t>(box_ch), static_cast<cuuint32_t>(box_pos)};
    cuuint32_t es[3] = {1, 1, 1};
    CUresult r = g_encode(tm, HM_TMA_DTYPE, 3, const_cast<void*>(x), dims, strides,
                          box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                          CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled(act3) failed: " + std::to_string(r)); return false; }
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

// conv weight [cout][cin][k][k] fp32 -> 16-bit [n_rows][k*k*cpad], K index = tap*cpad + c with
// tap = ky*k + kx (cross-correlation: in[y+ky-k/2][x+kx-k/2]).  `row_offset` lets two convs
// share one padded matrix (policy projections A|B).
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
        L.bias = upload(net, *bias);
        if (!L.bias) return false;
    }
    return make_tm_weights(&L.tm_w, L.w, L.n_rows, L.ksize * L.ksize * L.cpad);
}
// 3x3, 64 < cin <= 80: the tap-packed copy for cconv3_kernel<PACKED> (layout in hm_net.h / hm_cconv.cuh)
static bool pack_conv_tap_packed(Net& net, DenseLayer& L, const float* w) {
    std::vector<act_t> packed(static_cast<size_t>(L.n_rows) * 768, float_to_act(0.0f));
    for (int o = 0; o < L.cout; ++o)
        for (int c = 0; c < L.cin; ++c)
            for (int ky = 0; ky < 3; ++ky)
                for (int kx = 0; kx < 3; ++kx) {
                    const int col = kx * 256 + (c < 64 ? ky * 64 + c : 192 + ky * 16 + (c - 64));
                    packed[static_cast<size_t>(o) * 768 + col] = float_to_act(w[(static_cast<size_t>(o) * L.cin + c) * 9 + ky * 3 + kx]);
                }
    L.w_tap = upload(net, packed);
    return L.w_tap != nullptr && make_tm_weights(&L.tm_w_tap, L.w_tap, L.n_rows, 768);
}
// w: fp32 [cout][cin][k][k]; K per tap is padded to a multiple of 128 (cgemm stages hold two
// 64-channel blocks).
static bool make_dense_raw(Net& net, DenseLayer& L, const std::string& name, const float* w, int cout,
                           int cin, int ksize, const float* b) {
    L.name = name;
    L.cout = cout;
    L.cin = cin;
    L.ksize = ksize;
    L.cpad = round_up(L.cin, 128);
    L.n_rows = round_up(L.cout, 128);
    const int ktot = L.ksize * L.ksize * L.cpad;
    std::vector<act_t> packed(static_cast<size_t>(L.n_rows) * ktot, float_to_act(0.0f));
    pack_conv(packed, ktot, L.cpad, w, L.cout, L.cin, L.ksize, 0);
    std::vector<float> bias;
    if (b) bias.assign(b, b + cout);
    net.param_count += static_cast<size_t>(cout) * cin * ksize * ksize + (b ? cout : 0);
    if (ksize == 3 && cin > 64 && cin <= 80 && !pack_conv_tap_packed(net, L, w)) return false;
    return finish_dense(net, L, packed, b ? &bias : nullptr);
}
static bool make_dense(Net& net, DenseLayer& L, const std::string& name, const HvmwTensor* w,
                       const HvmwTensor* b) {
    if (!w || w->dims.size() < 2) { set_error("missing tensor " + name + ".w"); return false; }
    if (b && b->numel() != w->dims[0]) { set_error("bias size mismatch for " + name); return false; }
    return make_dense_raw(net, L, name, w->data, w->dims[0], w->dims[1], w->dims.size() == 4 ? w->dims[2] : 1,
                          b ? b->data : nullptr);
}
static float* upload_tensor(Net& net, const HvmwTensor* t, size_t expect, const std::string& name) {
    if (!t || t->numel() != expect) { set_error("missing / mis-sized tensor " + name); return nullptr; }
    net.param_count += expect;
    return upload(net, std::vector<float>(t->data, t->data + expect));
}

// Everything of CrossBoardRiseV3 between the trunk and the value head (rise_mobile_v3.py:299-351).
static bool build_crossboard(Net& net, const HvmwBlob& blob) {
    auto T = [&](const std::string& n) { return blob.find(n); };
    constexpr int D = 192, C = 384, BC = 37;
    const HvmwTensor *t0 = T("tok0.wt"), *t1 = T("tok1.wt");
    if (!t0 || !t1 || t0->numel() != static_cast<size_t>(D) * C || t1->numel() != static_cast<size_t>(D) * C) {
        set_error("missing token projection tensors");
        return false;
    }
    std::vector<float> stacked(static_cast<size_t>(2) * D * C);
    std::copy(t0->data, t0->data + D * C, stacked.begin());
    std::copy(t1->data, t1->data + D * C, stacked.begin() + D * C);
    if (!make_dense_raw(net, net.tok_proj, "tokens.proj", stacked.data(), 2 * D, C, 1, nullptr)) return false;
    {
        const HvmwTensor *p0 = T("tok0.wp"), *p1 = T("tok1.wp"), *b0 = T("tok0.b"), *b1 = T("tok1.b");
        if (!p0 || !p1 || !b0 || !b1 || p0->numel() != static_cast<size_t>(D) * BC || b0->numel() != D) {
            set_error("missing token projection plane weights");
            return false;
        }
        std::vector<float> wp(2 * D * BC), bb(2 * D);
        std::copy(p0->data, p0->data + D * BC, wp.begin());
        std::copy(p1->data, p1->data + D * BC, wp.begin() + D * BC);
        std::copy(b0->data, b0->data + D, bb.begin());
        std::copy(b1->data, b1->data + D, bb.begin() + D);
        net.tok_wp = upload(net, wp);
        net.tok_bias = upload(net, bb);
        net.param_count += wp.size() + bb.size();
        if (!net.tok_wp || !net.tok_bias) return false;
    }
    {   // position embedding [64 squares][192] -> channel-major [192][64]
        const HvmwTensor* pe = T("pos_emb");
        if (!pe || pe->numel() != static_cast<size_t>(64) * D) { set_error("missing pos_emb"); return false; }
        std::vector<float> t(64 * D);
        for (int sq = 0; sq < 64; ++sq)
            for (int o = 0; o < D; ++o) t[o * 64 + sq] = pe->data[sq * D + o];
        net.pos_emb = upload(net, t);
        net.param_count += t.size();
        if (!net.pos_emb) return false;
    }
    if (!(net.board_emb = upload_tensor(net, T("board_emb"), 2 * D, "board_emb"))) return false;
    {   // folded token projection: [384][384 trunk | 74 planes] + the post-activation embedding table
        const HvmwTensor *p0 = T("tok0.wp"), *p1 = T("tok1.wp"), *b0 = T("tok0.b"), *b1 = T("tok1.b");
        const HvmwTensor *pe = T("pos_emb"), *be = T("board_emb");
        constexpr int KF = C + 2 * BC;
        std::vector<float> wf(static_cast<size_t>(2) * D * KF, 0.0f), bb(2 * D), post(static_cast<size_t>(2) * D * 64);
        for (int b = 0; b < 2; ++b)
            for (int o = 0; o < D; ++o) {
                float* row = wf.data() + static_cast<size_t>(b * D + o) * KF;
                std::copy(stacked.begin() + static_cast<size_t>(b * D + o) * C, stacked.begin() + static_cast<size_t>(b * D + o + 1) * C, row);
                const float* wp = (b ? p1 : p0)->data + static_cast<size_t>(o) * BC;
                std::copy(wp, wp + BC, row + C + b * BC);
                bb[b * D + o] = (b ? b1 : b0)->data[o];
                for (int sq = 0; sq < 64; ++sq) post[static_cast<size_t>(b * D + o) * 64 + sq] = pe->data[sq * D + o] + be->data[b * D + o];
            }
        const size_t before = net.param_count;
        if (!make_dense_raw(net, net.tok_fold, "tokens.fold", wf.data(), 2 * D, KF, 1, bb.data())) return false;
        net.param_count = before;   // the same parameters were counted above
        if (!(net.tok_post = upload(net, post))) return false;
    }
    if (!(net.ctx_role = upload_tensor(net, T("ctx_role"), 7 * D, "ctx_role"))) return false;
    if (!(net.pocket_w = upload_tensor(net, T("pocket.w"), D * 5, "pocket.w"))) return false;
    if (!(net.pocket_b = upload_tensor(net, T("pocket.b"), D, "pocket.b"))) return false;
    if (!(net.scalar_w = upload_tensor(net, T("scalar.w"), D, "scalar.w"))) return false;
    if (!(net.scalar_b = upload_tensor(net, T("scalar.b"), D, "scalar.b"))) return false;
    for (int l = 0;; ++l) {
        const std::string q = "xl" + std::to_string(l) + ".";
        const HvmwTensor *inw = T(q + "in.w"), *inb = T(q + "in.b");
        if (!inw) break;
        if (!inb || inw->numel() != static_cast<size_t>(3) * D * D || inb->numel() != 3 * D) {
            set_error("malformed attention in_proj of layer " + q);
            return false;
        }
        net.xlayers.emplace_back();
        XLayer& L = net.xlayers.back();
        // nn.MultiheadAttention packs q | k | v rows in in_proj_weight
        if (!make_dense_raw(net, L.wq, q + "q", inw->data, D, D, 1, inb->data)) return false;
        if (!make_dense_raw(net, L.wkv, q + "kv", inw->data + D * D, 2 * D, D, 1, inb->data + D)) return false;
        std::vector<float> wt(static_cast<size_t>(D) * 2 * D);
        for (int o = 0; o < 2 * D; ++o)
            for (int i = 0; i < D; ++i) wt[static_cast<size_t>(i) * 2 * D + o] = inw->data[static_cast<size_t>(D + o) * D + i];
        L.wkv_t = upload(net, wt);
        L.bkv = upload(net, std::vector<float>(inb->data + D, inb->data + 3 * D));
        if (!L.wkv_t || !L.bkv) return false;
        if (!make_dense(net, L.wo, q + "out", T(q + "out.w"), T(q + "out.b"))) return false;
        if (!make_dense(net, L.fc1, q + "fc1", T(q + "fc1.w"), T(q + "fc1.b"))) return false;
        if (!make_dense(net, L.fc2, q + "fc2", T(q + "fc2.w"), T(q + "fc2.b"))) return false;
        if (!(L.qn_g = upload_tensor(net, T(q + "qn.w"), D, q + "qn.w")) || !(L.qn_b = upload_tensor(net, T(q + "qn.b"), D, q + "qn.b")) ||
            !(L.mn_g = upload_tensor(net, T(q + "mn.w"), D, q + "mn.w")) || !(L.mn_b = upload_tensor(net, T(q + "mn.b"), D, q + "mn.b")) ||
            !(L.ln_g = upload_tensor(net, T(q + "ln.w"), D, q + "ln.w")) || !(L.ln_b = upload_tensor(net, T(q + "ln.b"), D, q + "ln.b")))
            return false;
    }
    if (net.xlayers.empty()) { set_error("crossboard network without attention layers"); return false; }
    if (!make_dense(net, net.vfuse, "value.fusion", T("vfuse.w"), T("vfuse.b"))) return false;
    for (int b = 0; b < 2; ++b) {
        const std::string q = "pol" + std::to_string(b) + ".";
        if (!make_dense(net, net.pol_c1[b], "policy" + std::to_string(b) + ".conv1", T(q + "c1.w"), T(q + "c1.b"))) return false;
        if (!make_dense(net, net.pol_c2[b], "policy" + std::to_string(b) + ".conv2", T(q + "c2.w"), nullptr)) return false;
        if (net.pol_c1[b].cin != D || net.pol_c2[b].cout != 73) { set_error("unexpected policy head shape"); return false; }
    }
    return true;
}

std::unique_ptr<Net> build_net(const HvmwBlob& blob) {
    if (!tensormap_init()) return nullptr;
    auto net = std::make_unique<Net>();
    net->arch = blob.arch;
    if (blob.arch != HM_ARCH_RISEV33 && blob.arch != HM_ARCH_CROSSBOARD_RISEV33) {
        set_error("architecture " + std::to_string(blob.arch) + " not supported by this build");
        return nullptr;
    }
    auto T = [&](const std::string& n) { return blob.find(n); };
    if (!make_dense(*net, net->stem, "stem", T("stem.w"), T("stem.b"))) return nullptr;
    net->channels = net->stem.cout;
    if (net->channels != 384) { set_error("trunk width must be 384"); return nullptr; }

    // ---- residual trunk: all blocks concatenated by 128-channel chunk (hm_trunk.cuh) ----
    std::vector<act_t> w1cat, w2cat;
    std::vector<float> par, bias2;
    for (int i = 0;; ++i) {
        const std::string p = "b" + std::to_string(i) + ".";
        const HvmwTensor *w1 = T(p + "pw1.w"), *b1 = T(p + "pw1.b"), *w2 = T(p + "pw2.w"),
                         *b2 = T(p + "pw2.b"), *dw = T(p + "dw.w"), *db = T(p + "dw.b");
        if (!w1) break;
        if (!b1 || !w2 || !b2 || !dw || !db || dw->dims.size() != 3 || w1->dims.size() != 2 ||
            w1->dims[1] != 384 || w2->dims[0] != 384) {
            set_error("missing / malformed tensors of block " + p);
            return nullptr;
        }
        Block blk;
        blk.cop = dw->dims[0];
        blk.ksize = dw->dims[1];
        if (blk.cop != static_cast<int>(w1->dims[0]) || blk.cop != static_cast<int>(w2->dims[1]) ||
            (blk.ksize != 3 && blk.ksize != 5)) {
            set_error("unsupported depthwise shape in " + p);
            return nullptr;
        }
        const int kk = blk.ksize * blk.ksize;
        blk.n_chunks = (blk.cop + 127) / 128;
        blk.chunk0 = net->total_chunks;
        blk.bias2_off = static_cast<int>(bias2.size());
        net->total_chunks += blk.n_chunks;
        w1cat.resize(static_cast<size_t>(net->total_chunks) * 128 * 384, float_to_act(0.0f));
        w2cat.resize(static_cast<size_t>(net->total_chunks) * 384 * 128, float_to_act(0.0f));
        par.resize(static_cast<size_t>(net->total_chunks) * 27 * 128, 0.0f);
        for (int c = 0; c < blk.cop; ++c) {
            const int g = blk.chunk0 + c / 128, r = c % 128;
            for (int k = 0; k < 384; ++k)
                w1cat[(static_cast<size_t>(g) * 128 + r) * 384 + k] = float_to_act(w1->data[static_cast<size_t>(c) * 384 + k]);
            for (int o = 0; o < 384; ++o)
                w2cat[(static_cast<size_t>(g) * 384 + o) * 128 + r] = float_to_act(w2->data[static_cast<size_t>(o) * blk.cop + c]);
            float* pp = par.data() + static_cast<size_t>(g) * 27 * 128;
            pp[r] = b1->data[c];
            pp[128 + r] = db->data[c];
            for (int t = 0; t < kk; ++t) pp[256 + t * 128 + r] = dw->data[c * kk + t];
        }
        // The trunk kernel keeps the residual stream as an fp32 TMEM accumulator that is never cleared, so the pw2
        // biases are carried as a CUMULATIVE vector added when the accumulator is read: the sum of the biases of all
        // blocks since the stream was last written back with its bias included (the kernel start, or the ECA gate of
        // a block, which stores g * (accumulator + cumulative bias) and so re-bases the sum to zero).
        {
            const bool rebased = net->blocks.empty() || T(p + "eca.w") != nullptr;
            const size_t prev = net->blocks.empty() ? 0 : static_cast<size_t>(net->blocks.back().bias2_off);
            for (int o = 0; o < 384; ++o) bias2.push_back(b2->data[o] + (rebased ? 0.0f : bias2[prev + o]));
        }
        net->param_count += w1->numel() + b1->numel() + w2->numel() + b2->numel() + dw->numel() + db->numel();
        if (const HvmwTensor* ew = T(p + "eca.w")) {
            const HvmwTensor* eb = T(p + "eca.b");
            if (!eb || ew->dims.size() != 2 || ew->dims[0] != 384 || ew->dims[1] != 384) {
                set_error("bad ECA tensors in " + p);
                return nullptr;
            }
            blk.eca = true;
            net->param_count += ew->numel() + eb->numel();
            // in-kernel ECA (hm_trunk.cuh): Wc [384 out][384 in] rides in W1cat as 3 extra row
            // chunks (same [rows][384] K-major shape as W1), its bias in the bias2 array
            blk.eca_chunk0 = net->total_chunks;
            net->total_chunks += 3;
            w1cat.resize(static_cast<size_t>(net->total_chunks) * 128 * 384, float_to_act(0.0f));
            w2cat.resize(static_cast<size_t>(net->total_chunks) * 384 * 128, float_to_act(0.0f));
            par.resize(static_cast<size_t>(net->total_chunks) * 27 * 128, 0.0f);
            for (int o = 0; o < 384; ++o)
                for (int k = 0; k < 384; ++k)
                    w1cat[(static_cast<size_t>(blk.eca_chunk0) * 128 + o) * 384 + k] = float_to_act(ew->data[o * 384 + k]);
            blk.eca_bias_off = static_cast<int>(bias2.size());
            bias2.insert(bias2.end(), eb->data, eb->data + 384);
        }
        net->blocks.push_back(blk);
    }
    if (net->blocks.empty()) { set_error("no residual blocks in blob"); return nullptr; }
    net->w1cat = upload(*net, w1cat);
    net->w2cat = upload(*net, w2cat);
    net->trunk_par = upload(*net, par);
    net->trunk_bias2 = upload(*net, bias2);
    if (!net->w1cat || !net->w2cat || !net->trunk_par || !net->trunk_bias2) return nullptr;
    if (!make_tm_weights(&net->tm_w1, net->w1cat, static_cast<size_t>(net->total_chunks) * 128, 384)) return nullptr;
    if (!make_tm_weights(&net->tm_w2, net->w2cat, static_cast<size_t>(net->total_chunks) * 384, 128)) return nullptr;

    // value head
    if (!make_dense(*net, net->value_conv, "value.conv", T("value.conv.w"), T("value.conv.b"))) return nullptr;
    if (net->value_conv.cout != 16) { set_error("value head must have 16 channels"); return nullptr; }
    if (T("value.wdl.w")) {
        const HvmwTensor *ww = T("value.wdl.w"), *wb = T("value.wdl.b"), *pw = T("value.plys.w"),
                         *pb = T("value.plys.b");
        if (!ww || !wb || !pw || !pb || ww->numel() != 3 * 1024 || wb->numel() != 3 || pw->numel() != 1024 || pb->numel() != 1) {
            set_error("malformed WDL / plys value head tensors");
            return nullptr;
        }
        std::vector<float> w4(1024 * 4), b4(4);
        for (int i = 0; i < 1024; ++i) {  // feature index i = c*64 + sq (builder_util.py:311 flatten)
            for (int o = 0; o < 3; ++o) w4[i * 4 + o] = ww->data[o * 1024 + i];
            w4[i * 4 + 3] = pw->data[i];
        }
        for (int o = 0; o < 3; ++o) b4[o] = wb->data[o];
        b4[3] = pb->data[0];
        net->value_w4 = upload(*net, w4);
        net->value_b4 = upload(*net, b4);
        if (!net->value_w4 || !net->value_b4) return nullptr;
        net->param_count += ww->numel() + wb->numel() + pw->numel() + pb->numel();
        if (const HvmwTensor* m0 = T("value.mlp0.w")) {   // use_mlp_wdl_ply (builder_util.py:292-296)
            const HvmwTensor *m0b = T("value.mlp0.b"), *m2 = T("value.mlp2.w"), *m2b = T("value.mlp2.b");
            if (!m0b || !m2 || !m2b || m0->numel() != 32 || m0b->numel() != 8 || m2->numel() != 8 || m2b->numel() != 1) {
                set_error("malformed value MLP tensors");
                return nullptr;
            }
            std::vector<float> mlp(49);
            std::copy(m0->data, m0->data + 32, mlp.begin());
            std::copy(m0b->data, m0b->data + 8, mlp.begin() + 32);
            std::copy(m2->data, m2->data + 8, mlp.begin() + 40);
            mlp[48] = m2b->data[0];
            if (!(net->value_mlp = upload(*net, mlp))) return nullptr;
            net->param_count += 49;
        }
    } else {   // tanh value head (builder_util.py:297-301, :326)
        const HvmwTensor *f0 = T("value.final0.w"), *f0b = T("value.final0.b"), *f2 = T("value.final2.w"), *f2b = T("value.final2.b");
        if (!f0 || !f0b || !f2 || !f2b || f0->numel() != 512 * 1024 || f0b->numel() != 512 || f2->numel() != 512 || f2b->numel() != 1) {
            set_error("value head tensors missing (WDL + plys heads or the 1024-512-1 tanh head)");
            return nullptr;
        }
        std::vector<float> w1t(static_cast<size_t>(1024) * 512);
        for (int j = 0; j < 512; ++j)
            for (int i = 0; i < 1024; ++i) w1t[static_cast<size_t>(i) * 512 + j] = f0->data[static_cast<size_t>(j) * 1024 + i];
        net->value_f0t = upload(*net, w1t);
        if (!net->value_f0t || !(net->value_f0b = upload_tensor(*net, f0b, 512, "value.final0.b")) ||
            !(net->value_f2w = upload_tensor(*net, f2, 512, "value.final2.w")) ||
            !(net->value_f2b = upload_tensor(*net, f2b, 1, "value.final2.b"))) return nullptr;
        net->param_count += f0->numel();
    }
    if (net->arch == HM_ARCH_CROSSBOARD_RISEV33) {
        if (!build_crossboard(*net, blob)) return nullptr;
        return net;
    }
    if (T("pol0.c1.w")) {
        // two independent _PolicyHead(384) (shared_policy_trunk=False, rise_mobile_v3.py:198-204, builder_util.py:206-243)
        net->policy_per_board = true;
        const HvmwTensor *a = T("pol0.c1.w"), *b = T("pol1.c1.w"), *ab = T("pol0.c1.b"), *bb = T("pol1.c1.b");
        if (!a || !b || !ab || !bb || a->dims.size() != 4 || a->dims[0] != 384 || a->dims[1] != 384 || a->dims[2] != 3 ||
            b->numel() != a->numel() || ab->numel() != 384 || bb->numel() != 384) {
            set_error("per-board policy head tensors missing / malformed");
            return nullptr;
        }
        std::vector<float> w(2 * a->numel()), bias(768);
        std::copy(a->data, a->data + a->numel(), w.begin());
        std::copy(b->data, b->data + b->numel(), w.begin() + a->numel());
        std::copy(ab->data, ab->data + 384, bias.begin());
        std::copy(bb->data, bb->data + 384, bias.begin() + 384);
        if (!make_dense_raw(*net, net->pol_c1x, "policy.conv1", w.data(), 768, 384, 3, bias.data())) return nullptr;
        for (int h = 0; h < 2; ++h) {
            const std::string q = "pol" + std::to_string(h) + ".";
            if (!make_dense(*net, net->pol_c2[h], "policy" + std::to_string(h) + ".conv2", T(q + "c2.w"), nullptr)) return nullptr;
            if (net->pol_c2[h].cin != 384 || net->pol_c2[h].cout != 73 || net->pol_c2[h].ksize != 3) { set_error("unexpected policy head shape"); return nullptr; }
        }
        return net;
    }
    // policy heads (shared trunk, two 73-channel projections stacked)
    if (!make_dense(*net, net->policy_shared, "policy.shared", T("policy.shared.w"), T("policy.shared.b"))) return nullptr;
    {
        const HvmwTensor *pa = T("policy.a.w"), *pb = T("policy.b.w");
        if (!pa || !pb || pa->dims.size() != 4 || pa->dims[0] != 73 || pb->dims[0] != 73) {
            set_error("policy projection tensors missing");
            return nullptr;
        }
        DenseLayer& L = net->policy_proj;
        L.name = "policy.proj";
        L.cout = POLICY_CH;
        L.cin = pa->dims[1];
        L.ksize = pa->dims[2];
        L.cpad = round_up(L.cin, 64);
        L.n_rows = round_up(POLICY_CH, 128);
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

// Opt-in to > 48 KB of dynamic shared memory is a per-DEVICE function attribute: set once per device a worker is built on
// (a process may hold engines on several GPUs, as the reference's multi-GPU self-play does with one Engine per device).
static int current_sm_count() {
    int dev = 0, n_sm = 148;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev);
    return n_sm;
}
static bool kernels_prepare() {
    static std::mutex mutex;
    static bool done[64] = {};
    int dev = 0;
    if (!cuda_ok(cudaGetDevice(&dev), "cudaGetDevice")) return false;
    std::lock_guard<std::mutex> lock(mutex);
    if (dev >= 0 && dev < 64 && done[dev]) return true;
    auto smem = [](auto* fn, int bytes, const char* what) {
        return cuda_ok(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes), what);
    };
    const bool ok = smem(cgemm_kernel, CG_SMEM, "cudaFuncSetAttribute(cgemm)") &&
                    smem(cconv3_kernel<false>, CC_SMEM, "cudaFuncSetAttribute(cconv3)") &&
                    smem(cconv3_kernel<true>, CC_SMEM, "cudaFuncSetAttribute(cconv3 packed)") &&
                    smem(trunk_kernel, TR_SMEM, "cudaFuncSetAttribute(trunk)") &&
                    smem(lgemm_kernel, lg_smem_bytes(LG_MAX_KPAIRS), "cudaFuncSetAttribute(lgemm)") &&
                    smem(pgemm_kernel, pg_smem_bytes(160), "cudaFuncSetAttribute(pgemm)") &&
                    smem(pconv3_kernel<160>, pc_smem_bytes(160), "cudaFuncSetAttribute(pconv3<160>)") &&
                    smem(pconv3_kernel<80>, pc_smem_bytes(80), "cudaFuncSetAttribute(pconv3<80>)") &&
                    smem(mlp_kernel, ML_SMEM, "cudaFuncSetAttribute(mlp)") && xboard_prepare();
    if (ok && dev >= 0 && dev < 64) done[dev] = true;
    return ok;
}

struct TrunkRun {
    TrunkParams p;
    CUtensorMap tm_x, tm_w1, tm_w2;
    int grid;
};
bool trace_trunk_run(Worker& w, int run, unsigned long long* out_host, int* cap) {
#ifndef HM_BRINGUP
    (void)w; (void)run; (void)out_host; (void)cap;
    set_error("timeline tracing is compiled out of this build (rebuild with -DHM_BRINGUP)");
    return false;
#else
    if (run < 0 || run >= static_cast<int>(w.trunk_runs.size())) { set_error("no such trunk run"); return false; }
    const size_t bytes = 3 * TR_TRACE_CAP * sizeof(unsigned long long);
    unsigned long long* d = nullptr;
    if (!cuda_ok(cudaMalloc(&d, bytes), "cudaMalloc(trace)")) return false;
    cudaMemsetAsync(d, 0, bytes, w.stream);
    TrunkRun r = *w.trunk_runs[run];
    r.p.trace = d;
    if (const char* dbg = getenv("HM_TRUNK_DBG")) r.p.dbg = atoi(dbg);   // bring-up builds, traced launches only
    trunk_kernel<<<r.grid, TR_THREADS, TR_SMEM, w.stream>>>(r.p, r.tm_x, r.tm_w1, r.tm_w2);
    bool ok = cuda_ok(cudaMemcpyAsync(out_host, d, bytes, cudaMemcpyDeviceToHost, w.stream), "D2H trace") &&
              cuda_ok(cudaStreamSynchronize(w.stream), "trace sync");
    cudaFree(d);
    if (cap) *cap = TR_TRACE_CAP;
    return ok;
#endif
}

// A run of consecutive bottleneck blocks [first, last) as one trunk_kernel launch.
static bool add_trunk_run(Worker& w, std::vector<Op>& ops, const Net& net, int first, int last, act_t* x,
                          int n_pos) {
    if (last - first > TR_MAX_BLOCKS) { set_error("trunk run too long"); return false; }
    CUtensorMap tm_x;
    if (!make_tm_act3(&tm_x, x, n_pos, 384)) return false;
    TrunkParams p{};
    p.n_blocks = last - first;
    double macs = 0;
    for (int i = first; i < last; ++i) {
        const Block& b = net.blocks[i];
        p.blk[i - first] = TrunkBlock{b.chunk0, b.n_chunks, b.ksize, b.bias2_off,
                                      b.eca ? b.eca_chunk0 : -1, b.eca_bias_off};
        macs += b.macs_per_pos();
    }
    p.par = net.trunk_par;
    p.bias2 = net.trunk_bias2;
    p.trace = nullptr;
    p.dbg = 0;
    const int grid = n_pos / 2;
    const CUtensorMap tw1 = net.tm_w1, tw2 = net.tm_w2;
    w.trunk_runs.push_back(std::make_shared<TrunkRun>(TrunkRun{p, tm_x, tw1, tw2, grid}));
    Op op;
    op.name = "trunk.b" + std::to_string(first) + "-" + std::to_string(last - 1);
    op.klass = "dense";
    op.flops = 2.0 * macs * n_pos;
    op.bytes = 2.0 * n_pos * 64 * 384 * 2;
    op.launch = [p, tm_x, tw1, tw2, grid](cudaStream_t s) {
        trunk_kernel<<<grid, TR_THREADS, TR_SMEM, s>>>(p, tm_x, tw1, tw2);
    };
    ops.push_back(std::move(op));
    return true;
}

// One convolution as an Op: in = NCHW [in_n][in_C][64] (positions `in_stride_ch` channels apart,
// 0 = dense) -> out [in_n][out_channels][64] (0 = L.cout), optional residual (same layout as out)
// and activation (0 none, 1 ReLU, 2 GELU).
struct DenseIO {
    const act_t* in = nullptr;
    int in_n = 0, in_C = 0, in_stride_ch = 0;
    void* out = nullptr;
    int out_channels = 0;
    bool out_f32 = false;
    int act = 0;
    const act_t* residual = nullptr;
    const char* suffix = "";
    // lgemm only: the channels beyond in_C come from a second tensor [in_n][in2_C][64] (its own 128-channel K pair),
    // and post_add[out channel][64] (fp32) is added after the activation
    const act_t* in2 = nullptr;
    int in2_C = 0;
    const float* post_add = nullptr;
};
static bool add_dense_io(std::vector<Op>& ops, const DenseLayer& L, const DenseIO& io) {
    CUtensorMap tm_x;
    if (!make_tm_act3(&tm_x, io.in, io.in_n, io.in_C, io.in_stride_ch)) return false;
    CgemmParams p{};
    p.m_blocks = L.n_rows / 128;
    p.n_valid = L.cout;
    p.ksize = L.ksize;
    p.kblocks = L.cpad / 64;
    {
        const int rem = L.cin - (L.cpad / 128 - 1) * 128;   // real channels in the last 128-channel K pair
        p.ksteps_last = std::max(1, std::min(8, (rem + 15) / 16));
    }
    p.bias = L.bias;
    p.out = io.out;
    p.out_channels = io.out_channels ? io.out_channels : L.cout;
    p.relu = io.act;
    p.out_f32 = io.out_f32 ? 1 : 0;
    p.residual = io.residual;
    // a 1x1 layer has only Cpad/128 <= 3 stage loads per CTA: a deep ring buys nothing, three
    // co-resident CTAs (64 KB each, 128 TMEM columns) overlap one CTA's epilogue with another's MMAs
    p.stages = L.ksize == 1 ? 1 : CG_STAGES;
    p.kp_in2 = -1;
    p.post_add = io.post_add;
    const int smem = p.stages == 1 ? cg_smem_bytes(1, 1) : CG_SMEM;
    const int grid = (io.in_n / 2) * p.m_blocks;
    const CUtensorMap tm_w = L.tm_w;
    Op op;
    op.name = L.name + io.suffix;
    op.klass = "dense";
    op.flops = 2.0 * L.macs_per_pos() * io.in_n;
    op.bytes = 0;
    // wide shallow-K 1x1 layers (the crossboard Linears / projections): persistent CTAs with resident weights
    // and two TMEM accumulators (hm_lgemm.cuh) instead of one latency-bound CTA per tile; HM_NO_LGEMM=1 opts out
    const int kpairs = p.kblocks / 2, n_tiles = io.in_n / 2;
    if (L.ksize == 1 && L.cout >= 128 && kpairs <= LG_MAX_KPAIRS && n_tiles >= 148 && getenv("HM_NO_LGEMM") == nullptr) {
        const int n_sm = current_sm_count();
        const dim3 lgrid(static_cast<unsigned>(std::max(1, std::min(n_tiles, n_sm / p.m_blocks))), static_cast<unsigned>(p.m_blocks));
        const int lsmem = lg_smem_bytes(kpairs);
#ifdef HM_BRINGUP
        if (const char* t = getenv("HM_CG_TRACE")) p.trace = (L.name + io.suffix) == t ? 1 : 0;
#endif
        CUtensorMap tm_o = tm_x, tm_x2 = tm_x;   // 16-bit outputs leave through a staging tile and TMA stores (hm_lgemm.cuh)
        if (!io.out_f32 && !make_tm_act3(&tm_o, io.out, io.in_n, p.out_channels)) return false;
        if (io.in2 != nullptr) {
            if (io.in_C % 128 != 0) { set_error("lgemm: the first input of a two-input layer must fill whole K pairs"); return false; }
            if (!make_tm_act3(&tm_x2, io.in2, io.in_n, io.in2_C)) return false;
            p.kp_in2 = io.in_C / 128;
        }
        op.launch = [p, tm_w, tm_x, tm_o, tm_x2, lgrid, lsmem, n_tiles](cudaStream_t s) {
            lgemm_kernel<<<lgrid, LG_THREADS, lsmem, s>>>(p, n_tiles, tm_w, tm_x, tm_o, tm_x2);
        };
        ops.push_back(std::move(op));
        return true;
    }
    // 3x3 layers: one persistent CTA per SM walks the (tile, output block) work items (hm_cconv.cuh); HM_NO_CCONV=1 opts out
    if (L.ksize == 3 && p.residual == nullptr && io.act <= 1 && getenv("HM_NO_CCONV") == nullptr) {
        p.n_items = grid;
#ifdef HM_BRINGUP
        if (const char* t = getenv("HM_CG_TRACE")) p.trace = (L.name + io.suffix) == t ? 1 : 0;
#endif
        const int cgrid = std::min(grid, current_sm_count());
        CUtensorMap tm_o = tm_x;   // 16-bit outputs leave through a staging tile and TMA stores
        if (!io.out_f32 && !make_tm_act3(&tm_o, io.out, io.in_n, p.out_channels)) return false;
        if (L.w_tap != nullptr && getenv("HM_NO_TAP_PACKED") == nullptr) {
            CUtensorMap tm_x16;   // 16 channels of ONE position: the rows of the gathered K tile
            if (!make_tm_act3(&tm_x16, io.in, io.in_n, io.in_C, io.in_stride_ch, 16, 1)) return false;
            const CUtensorMap tm_wt = L.tm_w_tap;
            op.launch = [p, tm_wt, tm_x, tm_x16, tm_o, cgrid](cudaStream_t s) {
                cconv3_kernel<true><<<cgrid, CC_THREADS, CC_SMEM, s>>>(p, tm_wt, tm_x, tm_x16, tm_o);
            };
        } else {
            op.launch = [p, tm_w, tm_x, tm_o, cgrid](cudaStream_t s) { cconv3_kernel<false><<<cgrid, CC_THREADS, CC_SMEM, s>>>(p, tm_w, tm_x, tm_x, tm_o); };
        }
        ops.push_back(std::move(op));
        return true;
    }
    if (io.in2 != nullptr || io.post_add != nullptr) { set_error("two-input / post-add layers need the lgemm path"); return false; }
    op.launch = [p, tm_w, tm_x, grid, smem](cudaStream_t s) {
        cgemm_kernel<<<grid, CG_THREADS, smem, s>>>(p, tm_w, tm_x);
    };
    ops.push_back(std::move(op));
    return true;
}
// Narrow-output 3x3 convolution (policy projections) on the square-major kernel (hm_pgemm.cuh): the
// output channels sit on the MMA N side, so 146 / 73 channels cost N = 160 / 80 instead of 256 / 128 rows.
static bool add_narrow_io(std::vector<Op>& ops, const DenseLayer& L, const DenseIO& io) {
    const int npad = (L.cout + 15) & ~15;
    if (L.ksize != 3 || npad > 160 || npad > L.n_rows || getenv("HM_NO_PGEMM") != nullptr) return add_dense_io(ops, L, io);
    CUtensorMap tm_x, tm_w;
    if (!make_tm_act3(&tm_x, io.in, io.in_n, io.in_C, io.in_stride_ch)) return false;
    if (!make_tm_weights(&tm_w, L.w, L.n_rows, L.ksize * L.ksize * L.cpad, npad)) return false;
    PgemmParams p{};
    p.n_valid = L.cout;
    p.npad = npad;
    p.kblocks = L.cpad / 64;
    {
        const int rem = L.cin - (L.cpad / 128 - 1) * 128;
        p.ksteps_last = std::max(1, std::min(8, (rem + 15) / 16));
    }
    p.bias = L.bias;
    p.out = io.out;
    p.out_channels = io.out_channels ? io.out_channels : L.cout;
    p.relu = io.act;
    p.out_f32 = io.out_f32 ? 1 : 0;
    if (io.residual != nullptr || io.act > 1) { set_error("pgemm: residual / GELU epilogues are not supported"); return false; }
    const int smem = pg_smem_bytes(npad);
    const int grid = io.in_n / 2;
    Op op;
    op.name = L.name + io.suffix;
    op.klass = "dense";
    op.flops = 2.0 * L.macs_per_pos() * io.in_n;
    op.bytes = 0;
    // persistent version (one CTA per SM walks the tiles, trunk-style issue loop, 16 epilogue warps); HM_NO_PCONV=1 opts out
    if ((npad == 160 || npad == 80) && getenv("HM_NO_PCONV") == nullptr) {
        const int n_tiles = grid, cgrid = std::min(grid, current_sm_count());
        if (npad == 160) op.launch = [p, tm_w, tm_x, cgrid, n_tiles](cudaStream_t s) { pconv3_kernel<160><<<cgrid, PC_THREADS, pc_smem_bytes(160), s>>>(p, n_tiles, tm_w, tm_x); };
        else op.launch = [p, tm_w, tm_x, cgrid, n_tiles](cudaStream_t s) { pconv3_kernel<80><<<cgrid, PC_THREADS, pc_smem_bytes(80), s>>>(p, n_tiles, tm_w, tm_x); };
        ops.push_back(std::move(op));
        return true;
    }
    op.launch = [p, tm_w, tm_x, grid, smem](cudaStream_t s) { pgemm_kernel<<<grid, PG_THREADS, smem, s>>>(p, tm_w, tm_x); };
    ops.push_back(std::move(op));
    return true;
}
static bool add_dense(std::vector<Op>& ops, const DenseLayer& L, const act_t* in, int cin_buf,
                      int n_pos, void* out, bool relu, bool out_f32, const char* suffix = "") {
    DenseIO io;
    io.in = in; io.in_n = n_pos; io.in_C = cin_buf; io.out = out; io.out_f32 = out_f32;
    io.act = relu ? 1 : 0; io.suffix = suffix;
    return add_dense_io(ops, L, io);
}
static void add_simple(std::vector<Op>& ops, const std::string& name, const char* klass, double bytes,
                       std::function<void(cudaStream_t)> fn) {
    Op op;
    op.name = name;
    op.klass = klass;
    op.bytes = bytes;
    op.launch = std::move(fn);
    ops.push_back(std::move(op));
}

// crossboard-risev33 between the trunk and the heads (rise_mobile_v3.py:371-412): board tokens,
// context tokens, the attention blocks, value fusion; leaves the final tokens in *tok_out and
// the fused value input in w.d_fuse.
static bool add_crossboard(Worker& w, std::vector<Op>& ops, const Net& net, int B, act_t** tok_out) {
    // The 7 context tokens and their K|V for every attention layer depend on the input planes only
    // (rise_mobile_v3.py:353-369: the layers never update them), so they are issued FIRST in the op list and
    // captured as a graph branch that runs beside the stem and the trunk (capture_graph, "ctx." ops).
    std::vector<Op> ctx_ops;
    if (net.xlayers.size() > 2) { set_error("more than 2 cross-board layers are not supported"); return false; }
    const size_t nb = static_cast<size_t>(B);
    bool ok = dev_alloc(w, &w.d_tokpre, nb * 384 * 64) && dev_alloc(w, &w.d_tok[0], nb * 384 * 64) &&
              dev_alloc(w, &w.d_tok[1], nb * 384 * 64) && dev_alloc(w, &w.d_nq, nb * 384 * 64) &&
              dev_alloc(w, &w.d_nm, nb * 384 * 64) && dev_alloc(w, &w.d_q, nb * 384 * 64) &&
              dev_alloc(w, &w.d_kv, nb * 768 * 64) && dev_alloc(w, &w.d_att, nb * 384 * 64) &&
              dev_alloc(w, &w.d_x1, nb * 384 * 64) && dev_alloc(w, &w.d_h, nb * 768 * 64) &&
              dev_alloc(w, &w.d_fuse, nb * 384 * 64) && dev_alloc(w, &w.d_ctx, nb * 7 * 192) &&
              dev_alloc(w, &w.d_ckv, 2 * nb * 7 * 384);   // one K|V block per attention layer
    if (!ok) return false;
    // X1/X2: trunk part of both token projections on the tensor cores, the 37 board planes,
    // BN bias, ReLU and the embeddings in the follow-up kernel
    {
        const act_t* planes = w.d_planes;
        if (getenv("HM_NO_TOKFOLD") == nullptr && getenv("HM_NO_LGEMM") == nullptr && B / 2 >= 148) {
            // ONE GEMM: the 74 planes are the fourth K pair (second input tensor), bias + ReLU + embedding table in the
            // epilogue; no fp32 intermediate, no token_embed_kernel
            DenseIO io;
            io.in = w.d_trunk; io.in_n = B; io.in_C = 384; io.out = w.d_tok[0]; io.act = 1;
            io.in2 = planes; io.in2_C = HM_NB_INPUT_CHANNELS; io.post_add = net.tok_post;
            if (!add_dense_io(ops, net.tok_fold, io)) return false;
        } else {
        DenseIO io;
        io.in = w.d_trunk; io.in_n = B; io.in_C = 384; io.out = w.d_tokpre; io.out_f32 = true;
        if (!add_dense_io(ops, net.tok_proj, io)) return false;
        const float *pre = w.d_tokpre, *wp = net.tok_wp, *tb = net.tok_bias, *pe = net.pos_emb, *be = net.board_emb;
        act_t* tok = w.d_tok[0];
        add_simple(ops, "tokens.embed", "tokens", nb * 384 * 64 * 6.0,
                   [=](cudaStream_t s) { launch_token_embed(pre, planes, wp, tb, pe, be, tok, B, s); });
        }
        const float *pw = net.pocket_w, *pb = net.pocket_b, *sw = net.scalar_w, *sb = net.scalar_b, *role = net.ctx_role;
        float* ctx = w.d_ctx;
        add_simple(ctx_ops, "ctx.tokens", "tokens", nb * 7 * 192 * 4.0,
                   [=](cudaStream_t s) { launch_context_tokens(planes, pw, pb, sw, sb, role, ctx, B, s); });
    }
    w.debug_views["tokens0"] = {w.d_tok[0], {B, 384 * 64, 0}};
    w.debug_views["context"] = {w.d_ctx, {B, 7 * 192, 1}};
    int cur = 0;
    for (size_t l = 0; l < net.xlayers.size(); ++l) {
        const XLayer& L = net.xlayers[l];
        const std::string q = "xl" + std::to_string(l) + ".";
        act_t *tin = w.d_tok[cur], *tout = w.d_tok[cur ^ 1];
        act_t *nq = w.d_nq, *nm = w.d_nm, *att = w.d_att, *x1 = w.d_x1, *qb = w.d_q, *kv = w.d_kv;
        float *ctx = w.d_ctx, *ckv = w.d_ckv + (l & 1) * nb * 7 * 384;
        const float *qg = L.qn_g, *qbb = L.qn_b, *mg = L.mn_g, *mb = L.mn_b, *lg = L.ln_g, *lb = L.ln_b;
        add_simple(ops, q + "norm_qm", "tokens", nb * 384 * 64 * 6.0,
                   [=](cudaStream_t s) { launch_token_layernorm(tin, qg, qbb, nq, mg, mb, nm, 2 * B, s); });
        DenseIO io;
        io.in = nq; io.in_n = 2 * B; io.in_C = 192; io.out = qb;
        if (!add_dense_io(ops, L.wq, io)) return false;
        io.in = nm; io.out = kv;
        if (!add_dense_io(ops, L.wkv, io)) return false;
        const float *wkv_t = L.wkv_t, *bkv = L.bkv;
        add_simple(ctx_ops, "ctx.kv" + std::to_string(l), "tokens", nb * 7 * 384 * 4.0,
                   [=](cudaStream_t s) { launch_context_kv(ctx, mg, mb, wkv_t, bkv, ckv, B, s); });
        add_simple(ops, q + "attention", "attention", nb * 2 * (192.0 + 384 + 192) * 64 * 2,
                   [=](cudaStream_t s) { launch_cross_attention(qb, kv, ckv, att, B, s); });
        ops.back().flops = 2.0 * 2 * 64 * 71 * 192 * 2 * nb;  // QK^T and PV, both boards (CUDA cores)
        io = DenseIO();
        io.in = att; io.in_n = 2 * B; io.in_C = 192; io.out = x1; io.residual = tin;
        if (!add_dense_io(ops, L.wo, io)) return false;
        act_t* ln = w.d_nq;  // reuse
        add_simple(ops, q + "norm_mlp", "tokens", nb * 384 * 64 * 4.0,
                   [=](cudaStream_t s) { launch_token_layernorm(x1, lg, lb, ln, nullptr, nullptr, nullptr, 2 * B, s); });
        if (getenv("HM_NO_MLPFUSE") == nullptr && B >= 148 && L.fc1.cin == ML_D && L.fc1.cout == ML_H && L.fc2.cin == ML_H &&
            L.fc2.cout == ML_D && L.fc1.bias != nullptr && L.fc2.bias != nullptr) {
            // fc1 -> GELU -> fc2 -> + residual as ONE kernel: the 384-wide hidden tile stays in shared memory (hm_mlp.cuh)
            MlpParams mp{};
            mp.n_tiles = B;   // 2B token blocks of 64, two per tile
            mp.bias1 = L.fc1.bias;
            mp.bias2 = L.fc2.bias;
            mp.residual = x1;
#ifdef HM_BRINGUP
            if (const char* t = getenv("HM_CG_TRACE")) mp.trace = (q + "mlp") == t ? 1 : 0;
#endif
            CUtensorMap tm_lx, tm_lo;
            if (!make_tm_act3(&tm_lx, ln, 2 * B, ML_D) || !make_tm_act3(&tm_lo, tout, 2 * B, ML_D)) return false;
            const CUtensorMap tm_w1 = L.fc1.tm_w, tm_w2 = L.fc2.tm_w;
            const int mgrid = std::min(B, current_sm_count());
            Op op;
            op.name = q + "mlp";
            op.klass = "dense";
            op.flops = 2.0 * (L.fc1.macs_per_pos() + L.fc2.macs_per_pos()) * 2 * B;
            op.bytes = 0;
            op.launch = [mp, tm_w1, tm_w2, tm_lx, tm_lo, mgrid](cudaStream_t s) {
                mlp_kernel<<<mgrid, ML_THREADS, ML_SMEM, s>>>(mp, tm_w1, tm_w2, tm_lx, tm_lo);
            };
            ops.push_back(std::move(op));
        } else {
        io = DenseIO();
        io.in = ln; io.in_n = 2 * B; io.in_C = 192; io.out = w.d_h; io.act = 2;
        if (!add_dense_io(ops, L.fc1, io)) return false;
        io = DenseIO();
        io.in = w.d_h; io.in_n = 2 * B; io.in_C = 384; io.out = tout; io.residual = x1;
        if (!add_dense_io(ops, L.fc2, io)) return false;
        }
        cur ^= 1;
        w.debug_views["tokens" + std::to_string(l + 1)] = {tout, {B, 384 * 64, 0}};
    }
    *tok_out = w.d_tok[cur];
    DenseIO io;
    io.in = *tok_out; io.in_n = B; io.in_C = 384; io.out = w.d_fuse; io.act = 1;
    if (!add_dense_io(ops, net.vfuse, io)) return false;
    ops.insert(ops.begin(), std::make_move_iterator(ctx_ops.begin()), std::make_move_iterator(ctx_ops.end()));
    w.debug_views["value_fusion"] = {w.d_fuse, {B, 384 * 64, 0}};
    return true;
}

// The value head (value.* ops: fusion/conv/tail/shape) only reads the trunk (or token) output, so
// it is captured as a parallel graph branch beside the policy head: the fork is placed before the
// first value op, the join after the last op of the lists.
static bool capture_graph(Worker& w, const std::vector<const std::vector<Op>*>& lists,
                          cudaGraphExec_t* exec) {
    cudaGraph_t graph = nullptr;
    const bool branch = getenv("HM_NO_GRAPH_BRANCH") == nullptr;
    if (!cuda_ok(cudaStreamBeginCapture(w.stream, cudaStreamCaptureModeThreadLocal), "BeginCapture")) return false;
    bool forked = false, forked2 = false, joined2 = false;
    auto join2 = [&] {
        if (forked2 && !joined2) {
            cudaEventRecord(w.ev_join2, w.ctx_stream);
            cudaStreamWaitEvent(w.stream, w.ev_join2, 0);
            joined2 = true;
        }
    };
    for (auto* l : lists)
        for (const Op& op : *l) {
            const bool side = branch && op.name.rfind("value.", 0) == 0;
            // context-token branch: forked where the ctx.* ops stand (right after the planes are there),
            // joined before the first attention core, its only consumer
            const bool ctx = branch && op.name.rfind("ctx.", 0) == 0;
            if (ctx && !forked2) {
                cudaEventRecord(w.ev_fork2, w.stream);
                cudaStreamWaitEvent(w.ctx_stream, w.ev_fork2, 0);
                forked2 = true;
            }
            if (op.name.size() > 10 && op.name.compare(op.name.size() - 10, 10, ".attention") == 0) join2();
            if (side && !forked) {
                cudaEventRecord(w.ev_fork, w.stream);
                cudaStreamWaitEvent(w.side_stream, w.ev_fork, 0);
                forked = true;
            }
            op.launch(ctx ? w.ctx_stream : side ? w.side_stream : w.stream);
        }
    join2();
    if (forked) {
        cudaEventRecord(w.ev_join, w.side_stream);
        cudaStreamWaitEvent(w.stream, w.ev_join, 0);
    }
    if (!cuda_ok(cudaStreamEndCapture(w.stream, &graph), "EndCapture")) return false;
    const bool ok = cuda_ok(cudaGraphInstantiate(exec, graph, 0), "GraphInstantiate");
    cudaGraphDestroy(graph);
    return ok;
}

static bool setup_worker(Engine* e, Worker& w) {
    const Net& net = *e->net;
    const int B = w.batch;
    const size_t nb = static_cast<size_t>(B);
    w.move_cap = nb * 2 * 256;
    if (!cuda_ok(cudaStreamCreateWithFlags(&w.stream, cudaStreamNonBlocking), "cudaStreamCreate")) return false;
    if (!cuda_ok(cudaStreamCreateWithFlags(&w.side_stream, cudaStreamNonBlocking), "cudaStreamCreate")) return false;
    if (!cuda_ok(cudaEventCreateWithFlags(&w.ev_fork, cudaEventDisableTiming), "cudaEventCreate") ||
        !cuda_ok(cudaEventCreateWithFlags(&w.ev_join, cudaEventDisableTiming), "cudaEventCreate")) return false;
    if (!cuda_ok(cudaStreamCreateWithFlags(&w.ctx_stream, cudaStreamNonBlocking), "cudaStreamCreate") ||
        !cuda_ok(cudaEventCreateWithFlags(&w.ev_fork2, cudaEventDisableTiming), "cudaEventCreate") ||
        !cuda_ok(cudaEventCreateWithFlags(&w.ev_join2, cudaEventDisableTiming), "cudaEventCreate")) return false;
    // The request inputs (records | offsets | moves) and the results (values | raw5 | priors) of the fused
    // path each live in ONE device block mirrored by ONE pinned block: a request is one H2D and one D2H
    // copy of (fixed part + 4 bytes per move) instead of three each.
    auto up16 = [](size_t v) { return (v + 15) & ~static_cast<size_t>(15); };
    w.in_fixed = up16(nb * sizeof(hm_packed_pos) + (2 * nb + 1) * 4);
    w.out_fixed = up16(nb * 4 + nb * 5 * 4);
    uint8_t *d_in = nullptr, *d_out = nullptr, *h_in = nullptr, *h_out = nullptr;
    bool ok = dev_alloc(w, &d_in, w.in_fixed + w.move_cap * 4) && dev_alloc(w, &d_out, w.out_fixed + w.move_cap * 4) &&
              host_alloc(w, &h_in, w.in_fixed + w.move_cap * 4) && host_alloc(w, &h_out, w.out_fixed + w.move_cap * 4);
    if (!ok) return false;
    w.d_in = d_in; w.d_out = d_out; w.h_in = h_in; w.h_out = h_out;
    w.d_pos = reinterpret_cast<hm_packed_pos*>(d_in);
    w.d_offsets = reinterpret_cast<uint32_t*>(d_in + nb * sizeof(hm_packed_pos));
    w.d_moves = reinterpret_cast<uint32_t*>(d_in + w.in_fixed);
    w.h_pos = reinterpret_cast<hm_packed_pos*>(h_in);
    w.h_offsets = reinterpret_cast<uint32_t*>(h_in + nb * sizeof(hm_packed_pos));
    w.h_moves = reinterpret_cast<uint32_t*>(h_in + w.in_fixed);
    w.d_values = reinterpret_cast<float*>(d_out);
    w.d_raw5 = reinterpret_cast<float*>(d_out + nb * 4);
    w.d_priors = reinterpret_cast<float*>(d_out + w.out_fixed);
    w.h_values = reinterpret_cast<float*>(h_out);
    w.h_raw5 = reinterpret_cast<float*>(h_out + nb * 4);
    w.h_priors = reinterpret_cast<float*>(h_out + w.out_fixed);
    ok = dev_alloc(w, &w.d_planes, nb * HM_NB_INPUT_VALUES) &&
              dev_alloc(w, &w.d_trunk, nb * 384 * 64) && dev_alloc(w, &w.d_pol_shared, nb * (net.policy_per_board ? 768 : 384) * 64) &&
              dev_alloc(w, &w.d_logits, nb * POLICY_CH * 64) && dev_alloc(w, &w.d_pi, nb * POLICY_CH * 64) &&
              dev_alloc(w, &w.d_vfeat, nb * 16 * 64) &&
              dev_alloc(w, &w.d_value_h, nb) && dev_alloc(w, &w.d_wdl_h, nb * 3) && dev_alloc(w, &w.d_ml_h, nb) &&
              dev_alloc(w, &w.d_captures, w.move_cap) && dev_alloc(w, &w.d_noise, w.move_cap) &&
              dev_alloc(w, &w.d_hash, nb) && dev_alloc(w, &w.d_uniforms, nb * 3) &&
              dev_alloc(w, &w.d_picks, nb * 2) && dev_alloc(w, &w.d_status, nb);
    ok = ok && host_alloc(w, &w.h_obs, nb * HM_NB_INPUT_VALUES) && host_alloc(w, &w.h_value, nb) &&
         host_alloc(w, &w.h_wdl, nb * 3) && host_alloc(w, &w.h_ml, nb) &&
         host_alloc(w, &w.h_pi_a, nb * HM_NB_POLICY_VALUES) && host_alloc(w, &w.h_pi_b, nb * HM_NB_POLICY_VALUES) &&
         host_alloc(w, &w.h_captures, w.move_cap) && host_alloc(w, &w.h_hash, nb) &&
         host_alloc(w, &w.h_uniforms, nb * 3) && host_alloc(w, &w.h_picks, nb * 2) && host_alloc(w, &w.h_status, nb);
    if (!ok || !kernels_prepare()) return false;

    // ---- stem + residual trunk + value head + shared policy conv -----------------------------
    auto& ops = w.ops_trunk;
    if (!add_dense(ops, net.stem, w.d_planes, HM_NB_INPUT_CHANNELS, B, w.d_trunk, true, false)) return false;
    int nblk = static_cast<int>(net.blocks.size());
#ifdef HM_BRINGUP
    if (const char* stop = getenv("HM_TRUNK_STOP"))  // bring-up builds only (tests/diag/diagnose.py): truncate the trunk
        nblk = std::max(0, std::min(nblk, atoi(stop)));
#endif
    // the whole trunk is ONE launch: ECA gates are computed inside the kernel (a gate depends on the
    // channel means of the SAME position only, builder_util.py:77-80)
    if (nblk > TR_MAX_BLOCKS) { set_error("trunks longer than TR_MAX_BLOCKS blocks are not supported"); return false; }
    if (nblk > 0 && net.blocks[0].eca) { set_error("ECA on the first block is not supported"); return false; }
    if (nblk > 0 && !add_trunk_run(w, ops, net, 0, nblk, w.d_trunk, B)) return false;
    w.debug_views["planes"] = {w.d_planes, {B, HM_NB_INPUT_VALUES, 0}};
    w.debug_views["trunk"] = {w.d_trunk, {B, 384 * 64, 0}};
    const bool cross = net.arch == HM_ARCH_CROSSBOARD_RISEV33;
    act_t* tokens = nullptr;
    if (cross && !add_crossboard(w, ops, net, B, &tokens)) return false;
    const act_t* value_in = cross ? w.d_fuse : w.d_trunk;
    if (!add_dense(ops, net.value_conv, value_in, 384, B, w.d_vfeat, true, false)) return false;
    w.debug_views["value_feat"] = {w.d_vfeat, {B, 16 * 64, 0}};
    {
        Op op;
        op.name = "value.tail";
        op.klass = "head";
        op.bytes = static_cast<double>(B) * 1024 * 2;
        const act_t* f = w.d_vfeat;
        const float *w4 = net.value_w4, *b4 = net.value_b4, *mlp = net.value_mlp;
        const float *f0t = net.value_f0t, *f0b = net.value_f0b, *f2w = net.value_f2w, *f2b = net.value_f2b;
        float* raw5 = w.d_raw5;
        __half *vh = w.d_value_h, *wh = w.d_wdl_h, *mh = w.d_ml_h;
        if (w4 != nullptr)
            op.launch = [f, w4, b4, mlp, B, raw5, vh, wh, mh](cudaStream_t s) { launch_value_tail(f, w4, b4, mlp, B, raw5, vh, wh, mh, s); };
        else   // tanh value head: 1024 -> 512 -> 1
            op.launch = [f, f0t, f0b, f2w, f2b, B, raw5, vh, wh, mh](cudaStream_t s) { launch_value_tanh(f, f0t, f0b, f2w, f2b, B, raw5, vh, wh, mh, s); };
        ops.push_back(std::move(op));
    }
    w.debug_views["policy_shared"] = {w.d_pol_shared, {B, (net.policy_per_board ? 768 : 384) * 64, 0}};
    w.debug_views["logits"] = {w.d_logits, {B, POLICY_CH * 64, 1}};
    if (cross) {
        // per-board _PolicyHead(192) on that board's token map (rise_mobile_v3.py:336-347): board b's
        // map is channels [192b, 192b+192) of the token tensor; conv1 outputs go side by side into
        // d_pol_shared, conv2 logits into rows [73b, 73b+73) of the stacked policy tensor
        for (int b = 0; b < 2; ++b) {
            DenseIO io;
            io.in = tokens + static_cast<size_t>(b) * 192 * 64; io.in_n = B; io.in_C = 192; io.in_stride_ch = 384;
            io.out = w.d_pol_shared + static_cast<size_t>(b) * 192 * 64; io.out_channels = 384; io.act = 1;
            if (!add_dense_io(ops, net.pol_c1[b], io)) return false;
        }
        for (int b = 0; b < 2; ++b) {
            DenseIO io;
            io.in = w.d_pol_shared + static_cast<size_t>(b) * 192 * 64; io.in_n = B; io.in_C = 192; io.in_stride_ch = 384;
            io.out_channels = POLICY_CH;
            io.out = w.d_pi + static_cast<size_t>(b) * 73 * 64; io.suffix = ".f16";
            if (!add_narrow_io(w.ops_dense_out, net.pol_c2[b], io)) return false;
            io.out = w.d_logits + static_cast<size_t>(b) * 73 * 64; io.out_f32 = true; io.suffix = "";
            if (!add_narrow_io(w.ops_fused_out, net.pol_c2[b], io)) return false;
        }
    } else if (net.policy_per_board) {
        // risev33 with two independent _PolicyHead(384): one stacked 384 -> 768 first conv over the trunk (board b's
        // map = channels [384b, 384b+384) of d_pol_shared), then each board's 384 -> 73 conv into rows [73b, 73b+73)
        DenseIO io;
        io.in = w.d_trunk; io.in_n = B; io.in_C = 384; io.out = w.d_pol_shared; io.act = 1;
        if (!add_dense_io(ops, net.pol_c1x, io)) return false;
        for (int b = 0; b < 2; ++b) {
            io = DenseIO();
            io.in = w.d_pol_shared + static_cast<size_t>(b) * 384 * 64; io.in_n = B; io.in_C = 384; io.in_stride_ch = 768;
            io.out_channels = POLICY_CH;
            io.out = w.d_pi + static_cast<size_t>(b) * 73 * 64; io.suffix = ".f16";
            if (!add_narrow_io(w.ops_dense_out, net.pol_c2[b], io)) return false;
            io.out = w.d_logits + static_cast<size_t>(b) * 73 * 64; io.out_f32 = true; io.suffix = "";
            if (!add_narrow_io(w.ops_fused_out, net.pol_c2[b], io)) return false;
        }
    } else {
        if (!add_dense(ops, net.policy_shared, w.d_trunk, 384, B, w.d_pol_shared, true, false)) return false;
        // reference-shaped post: dense fp16 policy maps, index ch*64+sq (rise_mobile_v3.py:55)
        DenseIO io;
        io.in = w.d_pol_shared; io.in_n = B; io.in_C = 384;
        io.out = w.d_pi; io.suffix = ".f16";
        if (!add_narrow_io(w.ops_dense_out, net.policy_proj, io)) return false;
        io.out = w.d_logits; io.out_f32 = true; io.suffix = "";
        if (!add_narrow_io(w.ops_fused_out, net.policy_proj, io)) return false;
    }
    // ---- fused pre/post ---------------------------------------------------------------------
    {
        Op op;
        op.name = "planes.encode";
        op.klass = "planes";
        op.bytes = static_cast<double>(B) * (256 + HM_NB_INPUT_VALUES * 2);
        const hm_packed_pos* pp = w.d_pos;
        act_t* out = w.d_planes;
#ifdef HM_ACT_BF16
        const int dt = HM_DTYPE_BF16;
#else
        const int dt = HM_DTYPE_F16;
#endif
        op.launch = [pp, out, B, dt](cudaStream_t s) { launch_planes_nchw(pp, B, out, dt, s); };
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
        op.launch = [lg, pp, B, mv, off, pri](cudaStream_t s) { launch_policy_priors_packed(lg, pp, B, mv, off, pri, s); };
        w.ops_fused_out.push_back(std::move(op));
        Op ov;
        ov.name = "value.shape";
        ov.klass = "head";
        ov.bytes = static_cast<double>(B) * 24;
        const float* raw5 = w.d_raw5;
        float* vals = w.d_values;
        const hm_value_config* cfg = e->d_vcfg;  // device struct: a configuration change is a copy, not a re-capture
        ov.launch = [raw5, B, cfg, vals](cudaStream_t s) { launch_shape_values_raw5(raw5, B, cfg, vals, s); };
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
        if (w->side_stream) cudaStreamDestroy(w->side_stream);
        if (w->ev_fork) cudaEventDestroy(w->ev_fork);
        if (w->ev_join) cudaEventDestroy(w->ev_join);
        if (w->ctx_stream) cudaStreamDestroy(w->ctx_stream);
        if (w->ev_fork2) cudaEventDestroy(w->ev_fork2);
        if (w->ev_join2) cudaEventDestroy(w->ev_join2);
    }
    e->workers.clear();
}

}  // namespace hm

#ifdef HM_BRINGUP
// bring-up only: the timeline CTA 0 of the traced cgemm launch (HM_CG_TRACE=<op name>) wrote; 3 roles x CG_TRACE_CAP
extern "C" int hm_debug_cg_trace(unsigned long long* out, int cap) {
    if (cap < 3 * hm::CC_TRACE_CAP) return 0;
    if (cudaDeviceSynchronize() != cudaSuccess) return 0;
    return cudaMemcpyFromSymbol(out, hm::g_cc_trace, sizeof(unsigned long long) * 3 * hm::CC_TRACE_CAP) == cudaSuccess ? 1 : 0;
}
#endif
