// hm_abi.cu — the C ABI declared in include/hm_engine.h (drop-in boundary for the
// reference's nn/ Engine, engine/src/nn/engine.h:43-132).  Never throws; returns 0/1.
#include <chrono>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <functional>
#include <mutex>

#include <sys/stat.h>

#include "hm_kernels.h"
#include "hm_net.h"

using namespace hm;

struct hm_engine {
    Engine e;
    std::mutex util_mutex;
};

namespace {

#define HM_TRY try {
#define HM_CATCH(ret)                                   \
    }                                                   \
    catch (const std::exception& ex) {                  \
        set_error(std::string("exception: ") + ex.what()); \
        return ret;                                     \
    }                                                   \
    catch (...) {                                       \
        set_error("unknown exception");                 \
        return ret;                                     \
    }

bool use_device(hm_engine* h) {
    // every call re-issues cudaSetDevice: worker threads hop between engines (engine.cc:579)
    return cuda_ok(cudaSetDevice(h->e.device), "cudaSetDevice");
}
Worker* get_worker(hm_engine* h, size_t idx) {
    if (!h) { set_error("null engine"); return nullptr; }
    if (!h->e.net) { set_error("no network loaded"); return nullptr; }
    if (idx >= h->e.workers.size()) { set_error("worker index out of range"); return nullptr; }
    return h->e.workers[idx].get();
}
bool load_blob(hm_engine* h, HvmwBlob& blob) {
    if (!use_device(h)) return false;
    if (!blob.parse()) return false;
    for (auto& w : h->e.workers)
        if (w && w->pending) { set_error("cannot reload weights with requests in flight"); return false; }
    engine_free_workers(&h->e);
    h->e.net.reset();
    auto net = build_net(blob);
    if (!net) return false;
    h->e.net = std::move(net);
    if (!engine_setup_workers(&h->e)) { h->e.net.reset(); return false; }
    return true;
}

template <typename T>
struct DevBuf {
    T* p = nullptr;
    bool alloc(size_t n) { return cuda_ok(cudaMalloc(reinterpret_cast<void**>(&p), std::max<size_t>(n, 1) * sizeof(T)), "cudaMalloc(util)"); }
    ~DevBuf() { if (p) cudaFree(p); }
};

}  // namespace

extern "C" {

const char* hm_last_error(void) { return get_error(); }
const char* hm_version(void) { return "hivemind_b200 0.1 (sm_100a)"; }

void* hm_alloc_pinned(size_t bytes) {
    void* p = nullptr;
    if (!cuda_ok(cudaMallocHost(&p, bytes), "cudaMallocHost")) return nullptr;
    return p;
}
void hm_free_pinned(void* p) { if (p) cudaFreeHost(p); }

hm_engine* hm_engine_create(int device, int batch, int n_workers) {
    HM_TRY
    if (batch <= 0 || n_workers <= 0) { set_error("batch and n_workers must be positive"); return nullptr; }
    int count = 0;
    if (!cuda_ok(cudaGetDeviceCount(&count), "cudaGetDeviceCount")) return nullptr;
    if (device < 0 || device >= count) { set_error("no such CUDA device"); return nullptr; }
    if (!cuda_ok(cudaSetDevice(device), "cudaSetDevice")) return nullptr;  // engine.cc:186-189
    cudaDeviceProp prop;
    if (!cuda_ok(cudaGetDeviceProperties(&prop, device), "cudaGetDeviceProperties")) return nullptr;
    if (prop.major != 10) {
        set_error("hm_engine needs an sm_100 (B200) device; found sm_" + std::to_string(prop.major) + std::to_string(prop.minor));
        return nullptr;
    }
    auto* h = new hm_engine();
    h->e.device = device;
    h->e.batch = batch;
    h->e.n_workers = n_workers;
    if (!cuda_ok(cudaStreamCreateWithFlags(&h->e.util_stream, cudaStreamNonBlocking), "cudaStreamCreate")) { delete h; return nullptr; }
    if (!cuda_ok(cudaMalloc(reinterpret_cast<void**>(&h->e.d_vcfg), sizeof(hm_value_config)), "cudaMalloc(value config)") ||
        !cuda_ok(cudaMemcpy(h->e.d_vcfg, &h->e.vcfg, sizeof(hm_value_config), cudaMemcpyHostToDevice), "upload value config")) {
        hm_engine_destroy(h);
        return nullptr;
    }
    return h;
    HM_CATCH(nullptr)
}

void hm_engine_destroy(hm_engine* h) {
    if (!h) return;
    cudaSetDevice(h->e.device);
    engine_free_workers(&h->e);
    h->e.net.reset();
    if (h->e.util_stream) cudaStreamDestroy(h->e.util_stream);
    if (h->e.d_vcfg) cudaFree(h->e.d_vcfg);
    delete h;
}

int hm_engine_load_memory(hm_engine* h, const void* blob, size_t bytes) {
    HM_TRY
    if (!h || !blob) { set_error("null argument"); return 0; }
    HvmwBlob b;
    b.bytes.assign(static_cast<const uint8_t*>(blob), static_cast<const uint8_t*>(blob) + bytes);
    return load_blob(h, b) ? 1 : 0;
    HM_CATCH(0)
}

int hm_engine_load(hm_engine* h, const char* path) {
    HM_TRY
    if (!h || !path) { set_error("null argument"); return 0; }
    std::string p(path);
    const std::string onnx = ".onnx";
    if (p.size() > onnx.size() && p.compare(p.size() - onnx.size(), onnx.size(), onnx) == 0) {
        // loadNetwork(onnxFile, ...) call sites: the converted weights sit beside the export as <name>.hvmw.  A
        // re-exported .onnx NEWER than its .hvmw means the conversion was not re-run: refuse rather than silently
        // keep playing the old network (python -m hivemind_b200.weights <checkpoint> <name>.hvmw refreshes it).
        const std::string hv = p.substr(0, p.size() - onnx.size()) + ".hvmw";
        struct stat so, sh;
        if (stat(p.c_str(), &so) == 0 && stat(hv.c_str(), &sh) == 0 &&
            (so.st_mtim.tv_sec > sh.st_mtim.tv_sec ||
             (so.st_mtim.tv_sec == sh.st_mtim.tv_sec && so.st_mtim.tv_nsec > sh.st_mtim.tv_nsec))) {
            set_error(hv + " is older than " + p + ": re-run the weight conversion");
            return 0;
        }
        p = hv;
    }
    std::ifstream f(p, std::ios::binary | std::ios::ate);
    if (!f) { set_error("cannot open " + p); return 0; }
    HvmwBlob b;
    b.bytes.resize(static_cast<size_t>(f.tellg()));
    f.seekg(0);
    f.read(reinterpret_cast<char*>(b.bytes.data()), b.bytes.size());
    if (!f) { set_error("short read on " + p); return 0; }
    return load_blob(h, b) ? 1 : 0;
    HM_CATCH(0)
}

int hm_engine_batch_size(const hm_engine* h) { return h ? h->e.batch : 0; }
int hm_engine_num_workers(const hm_engine* h) { return h ? h->e.n_workers : 0; }
int hm_engine_arch(const hm_engine* h) { return (h && h->e.net) ? h->e.net->arch : -1; }

// ---- reference-shaped path --------------------------------------------------------------
int hm_engine_enqueue_half(hm_engine* h, const uint16_t* obs, size_t worker) {
    HM_TRY
    Worker* w = get_worker(h, worker);
    if (!w || !use_device(h)) return 0;
    if (!obs) { set_error("null observation buffer"); return 0; }
    if (w->pending) { set_error("worker already has a request in flight"); return 0; }  // engine.cc:584-588
    const size_t B = h->e.batch;
    cudaStream_t s = w->stream;
    // the network's input layout IS the reference's obs layout (NCHW fp16): upload in place
    static_assert(sizeof(act_t) == 2, "16-bit activations");
#ifdef HM_ACT_BF16
    set_error("the reference-shaped fp16 entry point needs the fp16 build");
    return 0;
#endif
    const size_t prow = static_cast<size_t>(POLICY_CH) * 64 * 2, pw = static_cast<size_t>(HM_NB_POLICY_VALUES) * 2;
    bool ok = cuda_ok(cudaMemcpyAsync(w->d_planes, obs, B * HM_NB_INPUT_VALUES * 2, cudaMemcpyHostToDevice, s), "H2D obs") &&
              cuda_ok(cudaGraphLaunch(w->graph_dense, s), "graph launch") &&
              cuda_ok(cudaMemcpyAsync(w->h_value, w->d_value_h, B * 2, cudaMemcpyDeviceToHost, s), "D2H value") &&
              cuda_ok(cudaMemcpy2DAsync(w->h_pi_a, pw, w->d_pi, prow, pw, B, cudaMemcpyDeviceToHost, s), "D2H piA") &&
              cuda_ok(cudaMemcpy2DAsync(w->h_pi_b, pw, w->d_pi + HM_NB_POLICY_VALUES, prow, pw, B, cudaMemcpyDeviceToHost, s), "D2H piB") &&
              cuda_ok(cudaMemcpyAsync(w->h_wdl, w->d_wdl_h, B * 3 * 2, cudaMemcpyDeviceToHost, s), "D2H wdl") &&
              cuda_ok(cudaMemcpyAsync(w->h_ml, w->d_ml_h, B * 2, cudaMemcpyDeviceToHost, s), "D2H ml");
    if (!ok) return 0;
    w->pending = true;
    w->pending_kind = 1;
    return 1;
    HM_CATCH(0)
}

int hm_engine_sync_half(hm_engine* h, size_t worker, const uint16_t** value, const uint16_t** pi_a,
                        const uint16_t** pi_b, const uint16_t** wdl, const uint16_t** moves_left) {
    HM_TRY
    Worker* w = get_worker(h, worker);
    if (!w || !use_device(h)) return 0;
    if (!w->pending || w->pending_kind != 1) { set_error("no half inference pending on this worker"); return 0; }  // engine.cc:660-664
    w->pending = false;
    if (!cuda_ok(cudaStreamSynchronize(w->stream), "stream sync")) return 0;
    if (value) *value = reinterpret_cast<const uint16_t*>(w->h_value);
    if (pi_a) *pi_a = reinterpret_cast<const uint16_t*>(w->h_pi_a);
    if (pi_b) *pi_b = reinterpret_cast<const uint16_t*>(w->h_pi_b);
    if (wdl) *wdl = reinterpret_cast<const uint16_t*>(w->h_wdl);
    if (moves_left) *moves_left = reinterpret_cast<const uint16_t*>(w->h_ml);
    return 1;
    HM_CATCH(0)
}

int hm_engine_run_half(hm_engine* h, const uint16_t* obs, size_t worker, const uint16_t** value,
                       const uint16_t** pi_a, const uint16_t** pi_b, const uint16_t** wdl,
                       const uint16_t** moves_left) {
    if (!hm_engine_enqueue_half(h, obs, worker)) return 0;
    return hm_engine_sync_half(h, worker, value, pi_a, pi_b, wdl, moves_left);
}

int hm_engine_run_float(hm_engine* h, const float* obs, float* value, float* pi_a, float* pi_b,
                        float* wdl, float* moves_left, size_t worker) {
    HM_TRY
    Worker* w = get_worker(h, worker);
    if (!w) return 0;
    if (!obs) { set_error("null observation buffer"); return 0; }
    const size_t B = h->e.batch;
    // engine.cc:537-566: float -> half on the host, half inference, half -> float
    for (size_t i = 0; i < B * HM_NB_INPUT_VALUES; ++i) w->h_obs[i] = __float2half_rn(obs[i]);
    const uint16_t *v, *a, *b, *wd, *ml;
    if (!hm_engine_run_half(h, reinterpret_cast<const uint16_t*>(w->h_obs), worker, &v, &a, &b, &wd, &ml)) return 0;
    auto cvt = [](const uint16_t* src, float* dst, size_t n) {
        if (!dst) return;
        const __half* s = reinterpret_cast<const __half*>(src);
        for (size_t i = 0; i < n; ++i) dst[i] = __half2float(s[i]);
    };
    cvt(v, value, B);
    cvt(a, pi_a, B * HM_NB_POLICY_VALUES);
    cvt(b, pi_b, B * HM_NB_POLICY_VALUES);
    cvt(wd, wdl, B * 3);
    cvt(ml, moves_left, B);
    return 1;
    HM_CATCH(0)
}

// ---- fused path -------------------------------------------------------------------------
static bool root_policy_valid(const hm_root_policy* rp) {
    if (!rp) { set_error("null root policy"); return false; }
    if (rp->temperature < 0.0 || rp->temperature != rp->temperature) {
        set_error("Policy temperature must be positive");  // selfplay.cc:168-170
        return false;
    }
    return true;
}

static int enqueue_positions_impl(hm_engine* h, const hm_packed_pos* pos, size_t n_pos, const uint32_t* moves,
                                  const uint32_t* move_offsets, const hm_root_policy* rp, size_t worker) {
    HM_TRY
    Worker* w = get_worker(h, worker);
    if (!w || !use_device(h)) return 0;
    if (!pos || !move_offsets) { set_error("null argument"); return 0; }
    if (w->pending) { set_error("worker already has a request in flight"); return 0; }
    const size_t B = w->batch;
    if (n_pos > static_cast<size_t>(h->e.batch)) { set_error("n_pos exceeds the engine batch size"); return 0; }
    const size_t n_moves = move_offsets[2 * n_pos];
    if (n_moves > w->move_cap) { set_error("move list exceeds capacity"); return 0; }
    if (n_moves > 0 && !moves) { set_error("null move list"); return 0; }
    for (size_t i = 0; i < 2 * n_pos; ++i)
        if (move_offsets[i] > move_offsets[i + 1]) { set_error("move offsets must be non-decreasing"); return 0; }
    // stage into pinned memory; unused tail rows are zeroed so stale data is never evaluated
    // (the reference pads short batches for the same reason, searchthread.cc:431-441)
    std::memcpy(w->h_pos, pos, n_pos * sizeof(hm_packed_pos));
    if (n_pos < B) std::memset(w->h_pos + n_pos, 0, (B - n_pos) * sizeof(hm_packed_pos));
    if (n_moves) std::memcpy(w->h_moves, moves, n_moves * 4);
    std::memcpy(w->h_offsets, move_offsets, (2 * n_pos + 1) * 4);
    for (size_t i = 2 * n_pos + 1; i < 2 * B + 1; ++i) w->h_offsets[i] = static_cast<uint32_t>(n_moves);
    cudaStream_t s = w->stream;
    bool ok = cuda_ok(cudaMemcpyAsync(w->d_in, w->h_in, w->in_fixed + n_moves * 4, cudaMemcpyHostToDevice, s), "H2D request");
    if (rp) {  // inputs of the root stages ride along with the request
        RootPolicyArgs a;
        a.temperature = rp->temperature;
        a.alpha = rp->dirichlet_alpha;
        a.epsilon = rp->dirichlet_epsilon;
        a.seed = rp->noise_seed;
        if (rp->position_hash && n_pos) {
            std::memcpy(w->h_hash, rp->position_hash, n_pos * 8);
            ok = ok && cuda_ok(cudaMemcpyAsync(w->d_hash, w->h_hash, n_pos * 8, cudaMemcpyHostToDevice, s), "H2D hashes");
            a.hash = w->d_hash;
        }
        if (rp->captures && n_moves) {
            std::memcpy(w->h_captures, rp->captures, n_moves);
            ok = ok && cuda_ok(cudaMemcpyAsync(w->d_captures, w->h_captures, n_moves, cudaMemcpyHostToDevice, s), "H2D captures");
            a.captures = w->d_captures;
        }
        if (rp->uniforms && n_pos) {
            std::memcpy(w->h_uniforms, rp->uniforms, n_pos * 3 * 8);
            ok = ok && cuda_ok(cudaMemcpyAsync(w->d_uniforms, w->h_uniforms, n_pos * 3 * 8, cudaMemcpyHostToDevice, s), "H2D uniforms");
            a.uniforms = w->d_uniforms;
        }
        ok = ok && cuda_ok(cudaGraphLaunch(w->graph_fused, s), "graph launch");
        if (ok && n_pos) {
            launch_root_policy(w->d_pos, static_cast<int>(n_pos), w->d_moves, w->d_offsets, w->d_priors, a, w->d_noise,
                               w->d_picks, w->d_status, s);
            ok = cuda_ok(cudaGetLastError(), "root policy kernel") &&
                 cuda_ok(cudaMemcpyAsync(w->h_picks, w->d_picks, n_pos * 2 * 4, cudaMemcpyDeviceToHost, s), "D2H picks") &&
                 cuda_ok(cudaMemcpyAsync(w->h_status, w->d_status, n_pos * 4, cudaMemcpyDeviceToHost, s), "D2H status");
        }
    } else {
        ok = ok && cuda_ok(cudaGraphLaunch(w->graph_fused, s), "graph launch");
    }
    ok = ok && cuda_ok(cudaMemcpyAsync(w->h_out, w->d_out, w->out_fixed + n_moves * 4, cudaMemcpyDeviceToHost, s), "D2H results");
    if (!ok) return 0;
    w->pending = true;
    w->pending_kind = rp ? 3 : 2;
    w->pending_npos = n_pos;
    w->pending_nmoves = n_moves;
    return 1;
    HM_CATCH(0)
}

int hm_engine_enqueue_positions(hm_engine* h, const hm_packed_pos* pos, size_t n_pos,
                                const uint32_t* moves, const uint32_t* move_offsets, size_t worker) {
    return enqueue_positions_impl(h, pos, n_pos, moves, move_offsets, nullptr, worker);
}

int hm_engine_enqueue_positions_root(hm_engine* h, const hm_packed_pos* pos, size_t n_pos,
                                     const uint32_t* moves, const uint32_t* move_offsets,
                                     const hm_root_policy* rp, size_t worker) {
    if (!root_policy_valid(rp)) return 0;
    return enqueue_positions_impl(h, pos, n_pos, moves, move_offsets, rp, worker);
}

int hm_engine_sync_root(hm_engine* h, size_t worker, const float** priors, const float** values,
                        const float** raw, const uint32_t** picks, const int32_t** status) {
    HM_TRY
    Worker* w = get_worker(h, worker);
    if (!w || !use_device(h)) return 0;
    if (!w->pending || w->pending_kind != 3) { set_error("no root-policy request pending on this worker"); return 0; }
    w->pending = false;
    if (!cuda_ok(cudaStreamSynchronize(w->stream), "stream sync")) return 0;
    if (priors) *priors = w->h_priors;
    if (values) *values = w->h_values;
    if (raw) *raw = w->h_raw5;
    if (picks) *picks = w->h_picks;
    if (status) *status = w->h_status;
    return 1;
    HM_CATCH(0)
}

int hm_root_policy_apply(hm_engine* h, const hm_packed_pos* pos, size_t n_pos, const uint32_t* moves,
                         const uint32_t* move_offsets, float* priors, const hm_root_policy* rp,
                         float* noise_out, uint32_t* picks, int32_t* status) {
    HM_TRY
    if (!h || !pos || !move_offsets || !priors) { set_error("null argument"); return 0; }
    if (!root_policy_valid(rp)) return 0;
    if (n_pos == 0) return 1;
    if (!use_device(h)) return 0;
    std::lock_guard<std::mutex> lock(h->util_mutex);
    const size_t n_moves = move_offsets[2 * n_pos];
    if (n_moves > 0 && !moves) { set_error("null move list"); return 0; }
    for (size_t i = 0; i < 2 * n_pos; ++i)
        if (move_offsets[i] > move_offsets[i + 1]) { set_error("move offsets must be non-decreasing"); return 0; }
    DevBuf<hm_packed_pos> dpos;
    DevBuf<uint32_t> dm, doff, dpick;
    DevBuf<float> dp, dn;
    DevBuf<uint8_t> dc;
    DevBuf<uint64_t> dh;
    DevBuf<double> du;
    DevBuf<int32_t> dst;
    if (!dpos.alloc(n_pos) || !dm.alloc(n_moves) || !doff.alloc(2 * n_pos + 1) || !dp.alloc(n_moves) || !dn.alloc(n_moves) ||
        !dc.alloc(n_moves) || !dh.alloc(n_pos) || !du.alloc(3 * n_pos) || !dpick.alloc(2 * n_pos) || !dst.alloc(n_pos)) return 0;
    cudaStream_t s = h->e.util_stream;
    RootPolicyArgs a;
    a.temperature = rp->temperature;
    a.alpha = rp->dirichlet_alpha;
    a.epsilon = rp->dirichlet_epsilon;
    a.seed = rp->noise_seed;
    bool ok = cuda_ok(cudaMemcpyAsync(dpos.p, pos, n_pos * sizeof(hm_packed_pos), cudaMemcpyHostToDevice, s), "H2D") &&
              cuda_ok(cudaMemcpyAsync(doff.p, move_offsets, (2 * n_pos + 1) * 4, cudaMemcpyHostToDevice, s), "H2D") &&
              cuda_ok(cudaMemsetAsync(dn.p, 0, std::max<size_t>(n_moves, 1) * 4, s), "memset");
    if (ok && n_moves) ok = cuda_ok(cudaMemcpyAsync(dm.p, moves, n_moves * 4, cudaMemcpyHostToDevice, s), "H2D") &&
                            cuda_ok(cudaMemcpyAsync(dp.p, priors, n_moves * 4, cudaMemcpyHostToDevice, s), "H2D");
    if (ok && rp->position_hash) { ok = cuda_ok(cudaMemcpyAsync(dh.p, rp->position_hash, n_pos * 8, cudaMemcpyHostToDevice, s), "H2D"); a.hash = dh.p; }
    if (ok && rp->captures && n_moves) { ok = cuda_ok(cudaMemcpyAsync(dc.p, rp->captures, n_moves, cudaMemcpyHostToDevice, s), "H2D"); a.captures = dc.p; }
    if (ok && rp->uniforms) { ok = cuda_ok(cudaMemcpyAsync(du.p, rp->uniforms, 3 * n_pos * 8, cudaMemcpyHostToDevice, s), "H2D"); a.uniforms = du.p; }
    if (!ok) return 0;
    launch_root_policy(dpos.p, static_cast<int>(n_pos), dm.p, doff.p, dp.p, a, dn.p, dpick.p, dst.p, s);
    if (!cuda_ok(cudaGetLastError(), "root policy kernel")) return 0;
    if (n_moves) ok = cuda_ok(cudaMemcpyAsync(priors, dp.p, n_moves * 4, cudaMemcpyDeviceToHost, s), "D2H");
    if (ok && noise_out && n_moves) ok = cuda_ok(cudaMemcpyAsync(noise_out, dn.p, n_moves * 4, cudaMemcpyDeviceToHost, s), "D2H");
    if (ok && picks && rp->uniforms) ok = cuda_ok(cudaMemcpyAsync(picks, dpick.p, 2 * n_pos * 4, cudaMemcpyDeviceToHost, s), "D2H");
    if (ok && status) ok = cuda_ok(cudaMemcpyAsync(status, dst.p, n_pos * 4, cudaMemcpyDeviceToHost, s), "D2H");
    return ok && cuda_ok(cudaStreamSynchronize(s), "root policy sync") ? 1 : 0;
    HM_CATCH(0)
}

int hm_engine_sync_priors(hm_engine* h, size_t worker, const float** priors, const float** values,
                          const float** raw) {
    HM_TRY
    Worker* w = get_worker(h, worker);
    if (!w || !use_device(h)) return 0;
    if (!w->pending || w->pending_kind != 2) { set_error("no position request pending on this worker"); return 0; }
    w->pending = false;
    if (!cuda_ok(cudaStreamSynchronize(w->stream), "stream sync")) return 0;
    if (priors) *priors = w->h_priors;
    if (values) *values = w->h_values;
    if (raw) *raw = w->h_raw5;
    return 1;
    HM_CATCH(0)
}

int hm_engine_set_value_config(hm_engine* h, const hm_value_config* cfg) {
    HM_TRY
    if (!h || !cfg) { set_error("null argument"); return 0; }
    // The configuration is a device-side struct the value.shape launch of every graph reads: changing it is one
    // copy — no worker is rebuilt, no pinned output pointer handed out earlier is invalidated.  Refused while a
    // request is in flight (its shaped values would be computed under an undefined mix of old and new settings).
    for (auto& w : h->e.workers)
        if (w && w->pending) { set_error("cannot change the value configuration with requests in flight"); return 0; }
    if (!use_device(h)) return 0;
    if (!cuda_ok(cudaMemcpy(h->e.d_vcfg, cfg, sizeof(hm_value_config), cudaMemcpyHostToDevice), "upload value config")) return 0;
    h->e.vcfg = *cfg;
    return 1;
    HM_CATCH(0)
}

// ---- stand-alone stages -------------------------------------------------------------------
int hm_encode_planes(hm_engine* h, const hm_packed_pos* pos, size_t n, void* out, int dtype) {
    HM_TRY
    if (!h || !pos || !out) { set_error("null argument"); return 0; }
    if (dtype < HM_DTYPE_F32 || dtype > HM_DTYPE_U8) { set_error("bad dtype"); return 0; }
    if (n == 0) return 1;
    if (!use_device(h)) return 0;
    std::lock_guard<std::mutex> lock(h->util_mutex);
    const size_t esz = dtype == HM_DTYPE_F32 ? 4 : dtype == HM_DTYPE_U8 ? 1 : 2;
    DevBuf<hm_packed_pos> dpos;
    DevBuf<uint8_t> dout;
    if (!dpos.alloc(n) || !dout.alloc(n * HM_NB_INPUT_VALUES * esz)) return 0;
    cudaStream_t s = h->e.util_stream;
    if (!cuda_ok(cudaMemcpyAsync(dpos.p, pos, n * sizeof(hm_packed_pos), cudaMemcpyHostToDevice, s), "H2D pos")) return 0;
    launch_planes_nchw(dpos.p, static_cast<int>(n), dout.p, dtype, s);
    if (!cuda_ok(cudaGetLastError(), "planes kernel")) return 0;
    if (!cuda_ok(cudaMemcpyAsync(out, dout.p, n * HM_NB_INPUT_VALUES * esz, cudaMemcpyDeviceToHost, s), "D2H planes")) return 0;
    return cuda_ok(cudaStreamSynchronize(s), "planes sync") ? 1 : 0;
    HM_CATCH(0)
}

int hm_policy_priors_half(hm_engine* h, const uint16_t* logits_a, const uint16_t* logits_b,
                          const hm_packed_pos* pos, size_t n_pos, const uint32_t* moves,
                          const uint32_t* move_offsets, float* priors_out) {
    HM_TRY
    if (!h || !logits_a || !logits_b || !pos || !move_offsets || !priors_out) { set_error("null argument"); return 0; }
    if (n_pos == 0) return 1;
    if (!use_device(h)) return 0;
    std::lock_guard<std::mutex> lock(h->util_mutex);
    const size_t n_moves = move_offsets[2 * n_pos];
    if (n_moves == 0) return 1;
    if (!moves) { set_error("null move list"); return 0; }
    DevBuf<__half> da, db;
    DevBuf<hm_packed_pos> dpos;
    DevBuf<uint32_t> dm, doff;
    DevBuf<float> dp;
    const size_t L = n_pos * HM_NB_POLICY_VALUES;
    if (!da.alloc(L) || !db.alloc(L) || !dpos.alloc(n_pos) || !dm.alloc(n_moves) || !doff.alloc(2 * n_pos + 1) || !dp.alloc(n_moves)) return 0;
    cudaStream_t s = h->e.util_stream;
    bool ok = cuda_ok(cudaMemcpyAsync(da.p, logits_a, L * 2, cudaMemcpyHostToDevice, s), "H2D") &&
              cuda_ok(cudaMemcpyAsync(db.p, logits_b, L * 2, cudaMemcpyHostToDevice, s), "H2D") &&
              cuda_ok(cudaMemcpyAsync(dpos.p, pos, n_pos * sizeof(hm_packed_pos), cudaMemcpyHostToDevice, s), "H2D") &&
              cuda_ok(cudaMemcpyAsync(dm.p, moves, n_moves * 4, cudaMemcpyHostToDevice, s), "H2D") &&
              cuda_ok(cudaMemcpyAsync(doff.p, move_offsets, (2 * n_pos + 1) * 4, cudaMemcpyHostToDevice, s), "H2D");
    if (!ok) return 0;
    launch_policy_priors_dense(da.p, db.p, dpos.p, static_cast<int>(n_pos), dm.p, doff.p, dp.p, s);
    if (!cuda_ok(cudaGetLastError(), "policy kernel")) return 0;
    if (!cuda_ok(cudaMemcpyAsync(priors_out, dp.p, n_moves * 4, cudaMemcpyDeviceToHost, s), "D2H")) return 0;
    return cuda_ok(cudaStreamSynchronize(s), "policy sync") ? 1 : 0;
    HM_CATCH(0)
}

int hm_policy_indices(hm_engine* h, const uint32_t* moves, const uint8_t* stm_black, size_t n,
                      int32_t* idx_out) {
    HM_TRY
    if (!h || !moves || !stm_black || !idx_out) { set_error("null argument"); return 0; }
    if (n == 0) return 1;
    if (!use_device(h)) return 0;
    std::lock_guard<std::mutex> lock(h->util_mutex);
    DevBuf<uint32_t> dm;
    DevBuf<uint8_t> ds;
    DevBuf<int32_t> di;
    if (!dm.alloc(n) || !ds.alloc(n) || !di.alloc(n)) return 0;
    cudaStream_t s = h->e.util_stream;
    if (!cuda_ok(cudaMemcpyAsync(dm.p, moves, n * 4, cudaMemcpyHostToDevice, s), "H2D") ||
        !cuda_ok(cudaMemcpyAsync(ds.p, stm_black, n, cudaMemcpyHostToDevice, s), "H2D")) return 0;
    launch_policy_indices(dm.p, ds.p, static_cast<int>(n), di.p, s);
    if (!cuda_ok(cudaGetLastError(), "index kernel")) return 0;
    if (!cuda_ok(cudaMemcpyAsync(idx_out, di.p, n * 4, cudaMemcpyDeviceToHost, s), "D2H")) return 0;
    return cuda_ok(cudaStreamSynchronize(s), "index sync") ? 1 : 0;
    HM_CATCH(0)
}

int hm_shape_values_half(hm_engine* h, const uint16_t* value, const uint16_t* wdl,
                         const uint16_t* moves_left, size_t n, float* out) {
    HM_TRY
    if (!h || !value || !wdl || !moves_left || !out) { set_error("null argument"); return 0; }
    if (n == 0) return 1;
    if (!use_device(h)) return 0;
    std::lock_guard<std::mutex> lock(h->util_mutex);
    DevBuf<__half> dv, dw, dm;
    DevBuf<float> dout;
    if (!dv.alloc(n) || !dw.alloc(3 * n) || !dm.alloc(n) || !dout.alloc(n)) return 0;
    cudaStream_t s = h->e.util_stream;
    if (!cuda_ok(cudaMemcpyAsync(dv.p, value, n * 2, cudaMemcpyHostToDevice, s), "H2D") ||
        !cuda_ok(cudaMemcpyAsync(dw.p, wdl, 3 * n * 2, cudaMemcpyHostToDevice, s), "H2D") ||
        !cuda_ok(cudaMemcpyAsync(dm.p, moves_left, n * 2, cudaMemcpyHostToDevice, s), "H2D")) return 0;
    launch_shape_values_half(dv.p, dw.p, dm.p, static_cast<int>(n), h->e.vcfg, dout.p, s);
    if (!cuda_ok(cudaGetLastError(), "shape kernel")) return 0;
    if (!cuda_ok(cudaMemcpyAsync(out, dout.p, n * 4, cudaMemcpyDeviceToHost, s), "D2H")) return 0;
    return cuda_ok(cudaStreamSynchronize(s), "shape sync") ? 1 : 0;
    HM_CATCH(0)
}

// ---- measurement / debug ----------------------------------------------------------------
int hm_engine_bench_resident(hm_engine* h, size_t worker, int iters, float* ms_total,
                             int* launches_per_pass) {
    HM_TRY
    Worker* w = get_worker(h, worker);
    if (!w || !use_device(h)) return 0;
    if (w->pending) { set_error("worker busy"); return 0; }
    cudaEvent_t e0, e1;
    if (!cuda_ok(cudaEventCreate(&e0), "event") || !cuda_ok(cudaEventCreate(&e1), "event")) return 0;
    bool ok = cuda_ok(cudaEventRecord(e0, w->stream), "record");
    for (int i = 0; ok && i < iters; ++i) ok = cuda_ok(cudaGraphLaunch(w->graph_fused, w->stream), "graph launch");
    ok = ok && cuda_ok(cudaEventRecord(e1, w->stream), "record") && cuda_ok(cudaEventSynchronize(e1), "event sync");
    float ms = 0.f;
    if (ok) ok = cuda_ok(cudaEventElapsedTime(&ms, e0, e1), "elapsed");
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    if (ms_total) *ms_total = ms;
    if (launches_per_pass)
        *launches_per_pass = static_cast<int>(w->ops_fused_in.size() + w->ops_trunk.size() + w->ops_fused_out.size());
    return ok ? 1 : 0;
    HM_CATCH(0)
}

int hm_engine_bench_resident_multi(hm_engine* h, int n_workers, int iters, float* ms_total,
                                   int* launches_per_pass) {
    HM_TRY
    if (!h || n_workers < 1 || n_workers > h->e.n_workers) { set_error("bad worker count"); return 0; }
    if (!use_device(h)) return 0;
    for (int i = 0; i < n_workers; ++i)
        if (h->e.workers[i]->pending) { set_error("worker busy"); return 0; }
    Worker* w0 = h->e.workers[0].get();
    cudaEvent_t e0, e1;
    std::vector<cudaEvent_t> joins(n_workers, nullptr);
    if (!cuda_ok(cudaEventCreate(&e0), "event") || !cuda_ok(cudaEventCreate(&e1), "event")) return 0;
    bool ok = cuda_ok(cudaEventRecord(e0, w0->stream), "record");
    for (int i = 1; ok && i < n_workers; ++i) {
        ok = cuda_ok(cudaEventCreateWithFlags(&joins[i], cudaEventDisableTiming), "event") &&
             cuda_ok(cudaStreamWaitEvent(h->e.workers[i]->stream, e0, 0), "fork");
    }
    for (int it = 0; ok && it < iters; ++it) {
        Worker* w = h->e.workers[it % n_workers].get();
        ok = cuda_ok(cudaGraphLaunch(w->graph_fused, w->stream), "graph launch");
    }
    for (int i = 1; ok && i < n_workers; ++i)
        ok = cuda_ok(cudaEventRecord(joins[i], h->e.workers[i]->stream), "record") &&
             cuda_ok(cudaStreamWaitEvent(w0->stream, joins[i], 0), "join");
    ok = ok && cuda_ok(cudaEventRecord(e1, w0->stream), "record") && cuda_ok(cudaEventSynchronize(e1), "event sync");
    float ms = 0.f;
    if (ok) ok = cuda_ok(cudaEventElapsedTime(&ms, e0, e1), "elapsed");
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    for (auto ev : joins) if (ev) cudaEventDestroy(ev);
    if (ms_total) *ms_total = ms;
    if (launches_per_pass)
        *launches_per_pass = static_cast<int>(w0->ops_fused_in.size() + w0->ops_trunk.size() + w0->ops_fused_out.size());
    return ok ? 1 : 0;
    HM_CATCH(0)
}

static bool time_launches(cudaStream_t s, int iters, const std::function<void()>& launch, float* ms_total) {
    cudaEvent_t e0, e1;
    if (!cuda_ok(cudaEventCreate(&e0), "event") || !cuda_ok(cudaEventCreate(&e1), "event")) return false;
    launch();  // warm-up
    bool ok = cuda_ok(cudaStreamSynchronize(s), "warm-up") && cuda_ok(cudaEventRecord(e0, s), "record");
    for (int i = 0; ok && i < iters; ++i) launch();
    ok = ok && cuda_ok(cudaGetLastError(), "bench kernel") && cuda_ok(cudaEventRecord(e1, s), "record") &&
         cuda_ok(cudaEventSynchronize(e1), "event sync");
    float ms = 0.f;
    if (ok) ok = cuda_ok(cudaEventElapsedTime(&ms, e0, e1), "elapsed");
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    if (ms_total) *ms_total = ms;
    return ok;
}

int hm_bench_planes(hm_engine* h, const hm_packed_pos* pos, size_t n, int dtype, int iters, float* ms_total) {
    HM_TRY
    if (!h || !pos || n == 0 || iters <= 0) { set_error("bad argument"); return 0; }
    if (dtype < HM_DTYPE_F32 || dtype > HM_DTYPE_U8) { set_error("bad dtype"); return 0; }
    if (!use_device(h)) return 0;
    std::lock_guard<std::mutex> lock(h->util_mutex);
    const size_t esz = dtype == HM_DTYPE_F32 ? 4 : dtype == HM_DTYPE_U8 ? 1 : 2;
    DevBuf<hm_packed_pos> dpos;
    DevBuf<uint8_t> dout;
    if (!dpos.alloc(n) || !dout.alloc(n * HM_NB_INPUT_VALUES * esz)) return 0;
    cudaStream_t s = h->e.util_stream;
    if (!cuda_ok(cudaMemcpyAsync(dpos.p, pos, n * sizeof(hm_packed_pos), cudaMemcpyHostToDevice, s), "H2D pos")) return 0;
    return time_launches(s, iters, [&] { launch_planes_nchw(dpos.p, static_cast<int>(n), dout.p, dtype, s); }, ms_total) ? 1 : 0;
    HM_CATCH(0)
}

int hm_bench_priors(hm_engine* h, const hm_packed_pos* pos, size_t n_pos, const uint32_t* moves,
                    const uint32_t* move_offsets, int iters, float* ms_total) {
    HM_TRY
    if (!h || !pos || !moves || !move_offsets || n_pos == 0 || iters <= 0) { set_error("bad argument"); return 0; }
    if (!use_device(h)) return 0;
    std::lock_guard<std::mutex> lock(h->util_mutex);
    const size_t n_moves = move_offsets[2 * n_pos];
    const size_t L = n_pos * 2 * HM_NB_POLICY_VALUES;
    DevBuf<hm_packed_pos> dpos;
    DevBuf<uint32_t> dm, doff;
    DevBuf<float> dp, dl;
    if (!dpos.alloc(n_pos) || !dm.alloc(n_moves) || !doff.alloc(2 * n_pos + 1) || !dp.alloc(n_moves) || !dl.alloc(L)) return 0;
    cudaStream_t s = h->e.util_stream;
    // synthetic logits: a cheap deterministic pattern (the bytes gathered are what is timed, not their values)
    std::vector<float> row(2 * HM_NB_POLICY_VALUES);
    for (size_t i = 0; i < row.size(); ++i) row[i] = static_cast<float>(static_cast<int>((i * 7) % 41) - 20) * 0.25f;
    bool ok = cuda_ok(cudaMemcpyAsync(dpos.p, pos, n_pos * sizeof(hm_packed_pos), cudaMemcpyHostToDevice, s), "H2D") &&
              cuda_ok(cudaMemcpyAsync(dm.p, moves, n_moves * 4, cudaMemcpyHostToDevice, s), "H2D") &&
              cuda_ok(cudaMemcpyAsync(doff.p, move_offsets, (2 * n_pos + 1) * 4, cudaMemcpyHostToDevice, s), "H2D");
    for (size_t p = 0; ok && p < n_pos; ++p)
        ok = cuda_ok(cudaMemcpyAsync(dl.p + p * row.size(), row.data(), row.size() * 4, cudaMemcpyHostToDevice, s), "H2D logits");
    if (!ok) return 0;
    return time_launches(s, iters, [&] { launch_policy_priors_packed(dl.p, dpos.p, static_cast<int>(n_pos), dm.p, doff.p, dp.p, s); }, ms_total) ? 1 : 0;
    HM_CATCH(0)
}

int hm_engine_bench_e2e(hm_engine* h, const hm_packed_pos* pos, size_t n_pos, const uint32_t* moves,
                        const uint32_t* move_offsets, int iters, int slots, double* ms_total, double* checksum) {
    HM_TRY
    if (!h || !pos || !move_offsets || iters <= 0) { set_error("bad argument"); return 0; }
    slots = std::max(1, std::min<int>(slots, static_cast<int>(h->e.workers.size())));
    const float* values = nullptr;
    const auto t0 = std::chrono::steady_clock::now();
    int in_flight = 0, next_sync = 0;
    for (int i = 0; i < iters; ++i) {
        const int w = i % slots;
        if (in_flight == slots) {
            if (!hm_engine_sync_priors(h, next_sync, nullptr, &values, nullptr)) return 0;
            next_sync = (next_sync + 1) % slots;
            --in_flight;
        }
        if (!hm_engine_enqueue_positions(h, pos, n_pos, moves, move_offsets, w)) return 0;
        ++in_flight;
    }
    while (in_flight > 0) {
        if (!hm_engine_sync_priors(h, next_sync, nullptr, &values, nullptr)) return 0;
        next_sync = (next_sync + 1) % slots;
        --in_flight;
    }
    const double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    if (ms_total) *ms_total = ms;
    if (checksum) {
        double c = 0.0;
        for (size_t i = 0; values && i < n_pos; ++i) c += values[i];
        *checksum = c;
    }
    return 1;
    HM_CATCH(0)
}

int hm_engine_profile_pass(hm_engine* h, size_t worker, hm_kernel_time* out, int cap, int* n) {
    HM_TRY
    Worker* w = get_worker(h, worker);
    if (!w || !use_device(h) || !out || !n) return 0;
    if (w->pending) { set_error("worker busy"); return 0; }
    std::vector<const Op*> ops;
    for (auto* l : {&w->ops_fused_in, &w->ops_trunk, &w->ops_fused_out})
        for (const Op& op : *l) ops.push_back(&op);
    while (w->events.size() < ops.size() + 1) {
        cudaEvent_t ev;
        if (!cuda_ok(cudaEventCreate(&ev), "event")) return 0;
        w->events.push_back(ev);
    }
    const bool checked = getenv("HM_CHECKED") != nullptr;  // bring-up aid: name the first failing launch
    cudaEventRecord(w->events[0], w->stream);
    for (size_t i = 0; i < ops.size(); ++i) {
        ops[i]->launch(w->stream);
        cudaEventRecord(w->events[i + 1], w->stream);
        if (checked) {
            cudaError_t err = cudaStreamSynchronize(w->stream);
            if (err == cudaSuccess) err = cudaGetLastError();
            if (err != cudaSuccess) {
                set_error("launch '" + ops[i]->name + "' failed: " + cudaGetErrorString(err));
                return 0;
            }
        }
    }
    if (!cuda_ok(cudaStreamSynchronize(w->stream), "profile sync")) return 0;
    if (!cuda_ok(cudaGetLastError(), "profile pass")) return 0;
    int k = 0;
    for (size_t i = 0; i < ops.size() && k < cap; ++i, ++k) {
        std::memset(&out[k], 0, sizeof(out[k]));
        std::snprintf(out[k].name, sizeof(out[k].name), "%s", ops[i]->name.c_str());
        cudaEventElapsedTime(&out[k].ms, w->events[i], w->events[i + 1]);
        out[k].flops = ops[i]->flops;
        out[k].bytes = ops[i]->bytes;
    }
    *n = k;
    return 1;
    HM_CATCH(0)
}

int hm_engine_trace_trunk(hm_engine* h, size_t worker, int run, unsigned long long* out, int* cap) {
    HM_TRY
    Worker* w = get_worker(h, worker);
    if (!w || !use_device(h) || !out) return 0;
    if (w->pending) { set_error("worker busy"); return 0; }
    return trace_trunk_run(*w, run, out, cap) ? 1 : 0;
    HM_CATCH(0)
}

int hm_engine_debug_fetch(hm_engine* h, size_t worker, const char* name, float* out,
                          size_t cap_floats, size_t* n_floats) {
    HM_TRY
    Worker* w = get_worker(h, worker);
    if (!w || !use_device(h) || !name) return 0;
    auto it = w->debug_views.find(name);
    if (it == w->debug_views.end()) { set_error(std::string("unknown activation ") + name); return 0; }
    const size_t rows = it->second.second[0], cols = it->second.second[1];
    const bool f32 = it->second.second[2] != 0;
    const size_t count = rows * cols;
    if (n_floats) *n_floats = count;
    if (cap_floats == 0) return 1;  // size query
    if (count > cap_floats) { set_error("debug buffer too small"); return 0; }
    if (!cuda_ok(cudaStreamSynchronize(w->stream), "sync")) return 0;
    if (f32) return cuda_ok(cudaMemcpy(out, it->second.first, count * 4, cudaMemcpyDeviceToHost), "D2H") ? 1 : 0;
    std::vector<act_t> tmp(count);
    if (!cuda_ok(cudaMemcpy(tmp.data(), it->second.first, count * 2, cudaMemcpyDeviceToHost), "D2H")) return 0;
    for (size_t i = 0; i < count; ++i) out[i] = act_to_float(tmp[i]);
    return 1;
    HM_CATCH(0)
}

}  // extern "C"
