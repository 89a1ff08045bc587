// hm_root.cu — root-policy post-processing on the device:
//   * apply_temperature            engine/src/tools/selfplay.cc:167-185
//   * root Dirichlet noise         engine/src/search/node.h:288-315 (Node::try_init_and_expand)
//   * raw-policy joint sampling    engine/src/tools/selfplay.cc:301-376 (sample_raw_policy_action),
//                                  legality from engine/src/environment/joint_action.h:14-110
// One warp per position.  Everything whose rounding depends on summation order is done by ONE lane in
// index order, exactly as the reference's loops / std::accumulate / std::partial_sum do, so priors and
// picks are bit-identical to the host code; the per-move work (pow, gamma variates, the mix) is spread
// over the lanes.
#include <math.h>
#include <mutex>
#include <string>
#include "hm_kernels.h"
#include "hm_net.h"   // set_error / cuda_ok

namespace hm {

// ---- Philox4x32-10 (counter-based: every (list, move, attempt) has its own substream) ------------
struct U4 { uint32_t x, y, z, w; };
__device__ __forceinline__ U4 philox4x32(U4 c, uint32_t k0, uint32_t k1) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint32_t hi0 = __umulhi(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
        const uint32_t hi1 = __umulhi(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
        c = U4{hi1 ^ c.y ^ k0, lo1, hi0 ^ c.w ^ k1, lo0};
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    return c;
}
__device__ __forceinline__ float u01(uint32_t x) { return (static_cast<float>(x >> 8) + 0.5f) * (1.0f / 16777216.0f); }

// Gamma(alpha, 1) by Marsaglia-Tsang (alpha < 1 boosted through Gamma(alpha+1) * U^(1/alpha)), the
// same distribution std::gamma_distribution<float>(alpha, 1.0f) draws from (node.h:297).  The stream
// differs from mt19937_64's: parity for the noise is distributional (tests: moments and KS).
__device__ float gamma_variate(float alpha, uint64_t key, uint32_t item) {
    const float a = alpha < 1.0f ? alpha + 1.0f : alpha;
    const float d = a - 1.0f / 3.0f, c = rsqrtf(9.0f * d);
    const uint32_t k0 = static_cast<uint32_t>(key), k1 = static_cast<uint32_t>(key >> 32);
    for (uint32_t attempt = 0; attempt < 64; ++attempt) {
        const U4 r = philox4x32(U4{item, attempt, 0u, 0u}, k0, k1);
        const float x = sqrtf(-2.0f * logf(u01(r.x))) * cospif(2.0f * u01(r.y));  // Box-Muller
        const float t = 1.0f + c * x;
        if (t <= 0.0f) continue;
        const float v = t * t * t, u = u01(r.z);
        if (logf(u) < 0.5f * x * x + d - d * v + d * logf(v)) {
            float g = d * v;
            if (alpha < 1.0f) g *= powf(u01(r.w), 1.0f / alpha);
            return g;
        }
    }
    return d;  // unreachable in practice (acceptance > 95 % per attempt)
}

// joint_action.h:14-32,84-110: jointPrior < 0 <=> the pair is an illegal sit
__device__ __forceinline__ bool joint_legal(bool sitsA, bool sitsB, bool capA, bool capB, bool tadv, bool aOn,
                                            bool bOn, bool aCan, bool bCan) {
    if (sitsA && sitsB) return tadv && (aOn != bOn);
    if (sitsA && aCan) return tadv || !(aOn && bOn) || capB;
    if (sitsB && bCan) return tadv || !(aOn && bOn) || capA;
    return true;
}

// std::discrete_distribution<size_t>(p, p+n)(u) as libstdc++ computes it (bits/random.tcc
// _M_initialize + operator()): double sum in index order, p/sum, running partial sums, last one
// forced to 1.0, first index whose cumulative value is >= u.  n < 2 never draws (returns 0).
__device__ int discrete_pick(const float* p, int n, double u, bool* bad) {
    if (n < 2) return 0;
    double sum = 0.0;
    for (int i = 0; i < n; ++i) sum += static_cast<double>(p[i]);
    if (!(sum > 0.0) || !isfinite(sum)) { *bad = true; return 0; }
    double cp = 0.0;
    for (int i = 0; i < n - 1; ++i) {
        cp += static_cast<double>(p[i]) / sum;
        if (!(cp < u)) return i;
    }
    return n - 1;
}

__global__ void __launch_bounds__(128)
root_policy_kernel(const hm_packed_pos* __restrict__ pos, int n_pos, const uint32_t* __restrict__ moves,
                   const uint32_t* __restrict__ offsets, float* __restrict__ priors, RootPolicyArgs a,
                   float* __restrict__ noise, uint32_t* __restrict__ picks, int32_t* __restrict__ status) {
    const int p = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (p >= n_pos) return;
    int st = 0;
    for (int board = 0; board < 2; ++board) {
        const uint32_t lo = offsets[2 * p + board];
        const int n = static_cast<int>(offsets[2 * p + board + 1] - lo);
        float* pr = priors + lo;
        if (n <= 0) { st = HM_ROOT_EMPTY_LIST; continue; }
        // ---- apply_temperature (selfplay.cc:167-185) ----
        if (a.temperature > 0.0) {
            const double exponent = 1.0 / a.temperature;
            for (int i = lane; i < n; i += 32)
                pr[i] = static_cast<float>(pow(static_cast<double>(fmaxf(0.0f, pr[i])), exponent));
            __syncwarp();
            double total = 0.0;
            if (lane == 0)
                for (int i = 0; i < n; ++i) total += pr[i];  // float -> double, index order
            total = __shfl_sync(0xffffffffu, total, 0);
            if (!isfinite(total) || total <= 0.0) {
                const float u = 1.0f / static_cast<float>(n);
                for (int i = lane; i < n; i += 32) pr[i] = u;
            } else {
                for (int i = lane; i < n; i += 32) pr[i] = static_cast<float>(pr[i] / total);
            }
            __syncwarp();
        }
        // ---- root Dirichlet noise (node.h:288-315) ----
        if (a.alpha > 0.0f && a.epsilon > 0.0f && a.hash != nullptr && n > 1) {
            const uint64_t salt = board == 0 ? 0x9e3779b97f4a7c15ULL : 0xbf58476d1ce4e5b9ULL;
            const uint64_t key = a.seed ^ a.hash[p] ^ salt;
            float* nz = noise + lo;
            for (int i = lane; i < n; i += 32) nz[i] = gamma_variate(a.alpha, key, static_cast<uint32_t>(i));
            __syncwarp();
            float total = 0.0f;
            if (lane == 0)
                for (int i = 0; i < n; ++i) total = __fadd_rn(total, nz[i]);
            total = __shfl_sync(0xffffffffu, total, 0);
            if (total > 0.0f) {
                const float eps = fminf(fmaxf(a.epsilon, 0.0f), 1.0f);
                for (int i = lane; i < n; i += 32)  // no FMA contraction: restated op by op
                    pr[i] = __fadd_rn(__fmul_rn(1.0f - eps, pr[i]), __fdiv_rn(__fmul_rn(eps, nz[i]), total));
            }
            __syncwarp();
        }
    }
    if (a.uniforms == nullptr || picks == nullptr) {
        if (lane == 0 && status) status[p] = st;
        return;
    }
    // ---- sample_raw_policy_action (selfplay.cc:349-376) ----
    const uint32_t loA = offsets[2 * p], loB = offsets[2 * p + 1], hiB = offsets[2 * p + 2];
    const int nA = static_cast<int>(loB - loA), nB = static_cast<int>(hiB - loB);
    if (nA <= 0 || nB <= 0) {
        if (lane == 0) { picks[2 * p] = picks[2 * p + 1] = 0; if (status) status[p] = HM_ROOT_EMPTY_LIST; }
        return;
    }
    const double* u = a.uniforms + 3 * static_cast<size_t>(p);
    bool bad = false;
    int pick = 0;
    if (lane < 2) pick = discrete_pick(priors + (lane ? loB : loA), lane ? nB : nA, u[lane], &bad);
    int iA = __shfl_sync(0xffffffffu, pick, 0), iB = __shfl_sync(0xffffffffu, pick, 1);
    if (__any_sync(0xffffffffu, bad)) st = HM_ROOT_BAD_WEIGHTS;
    if (lane != 0) return;
    const uint8_t fA = pos[p].board[0].flags, fB = pos[p].board[1].flags;
    const bool team_black = (fA & HM_FLAG_TEAM_BLACK) != 0, tadv = (fA & HM_FLAG_TIME_ADVANTAGE) != 0;
    const bool aOn = ((fA & HM_FLAG_STM_BLACK) != 0) == team_black;   // stm(A) == team
    const bool bOn = ((fB & HM_FLAG_STM_BLACK) != 0) != team_black;   // stm(B) == ~team
    const bool aCan = aOn && nA > 1, bCan = bOn && nB > 1;
    const uint32_t* mA = moves + loA;
    const uint32_t* mB = moves + loB;
    const float *pA = priors + loA, *pB = priors + loB;
    auto cap = [&](uint32_t off, int i) { return a.captures != nullptr && a.captures[off + i] != 0; };
    auto legal = [&](int i, int j) {
        return joint_legal(mA[i] == 0u, mB[j] == 0u, mA[i] != 0u && cap(loA, i), mB[j] != 0u && cap(loB, j), tadv, aOn,
                           bOn, aCan, bCan);
    };
    if (st == 0 && !legal(iA, iB)) {
        // resample from the legal pairs only, weights double(pA)*double(pB), pairs in (iA, iB) row order
        double total = 0.0;
        int lastA = -1, lastB = -1, n_legal = 0;
        for (int i = 0; i < nA; ++i)
            for (int j = 0; j < nB; ++j)
                if (legal(i, j)) { total += static_cast<double>(pA[i]) * static_cast<double>(pB[j]); lastA = i; lastB = j; ++n_legal; }
        if (!(total > 0.0) || !isfinite(total)) {
            st = HM_ROOT_NO_LEGAL_JOINT;  // the reference throws here (selfplay.cc:369-371)
        } else {
            st = n_legal >= 2 ? HM_ROOT_RESAMPLED : HM_ROOT_RESAMPLED_FORCED;
            iA = lastA; iB = lastB;
            // a single legal pair never draws (discrete_distribution of size 1)
            double cp = 0.0;
            bool done = false;
            for (int i = 0; i < nA && !done; ++i)
                for (int j = 0; j < nB; ++j) {
                    if (!legal(i, j) || (i == lastA && j == lastB)) continue;
                    cp += static_cast<double>(pA[i]) * static_cast<double>(pB[j]) / total;
                    if (!(cp < u[2])) { iA = i; iB = j; done = true; break; }
                }
        }
    }
    picks[2 * p] = static_cast<uint32_t>(iA);
    picks[2 * p + 1] = static_cast<uint32_t>(iB);
    if (status) status[p] = st;
}

void launch_root_policy(const hm_packed_pos* pos, int n_pos, const uint32_t* moves, const uint32_t* offsets,
                        float* priors, const RootPolicyArgs& a, float* noise, uint32_t* picks, int32_t* status,
                        cudaStream_t s) {
    if (n_pos <= 0) return;
    root_policy_kernel<<<(n_pos + 3) / 4, 128, 0, s>>>(pos, n_pos, moves, offsets, priors, a, noise, picks, status);
}

}  // namespace hm

// ---- allocation-free entry point (include/hm_engine.h: hm_root_policy_apply_noalloc) ---------------------------
// No counterpart in the reference; added for callers that cannot call
// cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The only scratch
// root_policy_kernel needs — the Gamma variates of the Dirichlet-noise stage, one float per move — lives in this
// file-scope device array instead of a cudaMalloc'ed buffer.  Every GPU that loads the library holds its 4 MiB until
// the process exits.  In HOST code `g_noalloc_noise` is only a stand-in for the symbol, never a device address: the
// address a kernel may be given is resolved per device with cudaGetSymbolAddress below.
static __device__ __align__(256) float g_noalloc_noise[HM_ROOT_NOALLOC_MAX_MOVES];

namespace {
constexpr int kMaxDevices = 64;
std::mutex g_noalloc_mutex;
float* g_noalloc_ptr[kMaxDevices] = {};   // per device: the device address of g_noalloc_noise on THAT device

// Resolved once per device, with that device current.
float* noalloc_scratch(int device) {
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    if (g_noalloc_ptr[device] == nullptr) {
        void* p = nullptr;
        if (!hm::cuda_ok(cudaGetSymbolAddress(&p, g_noalloc_noise), "cudaGetSymbolAddress(noalloc scratch)")) return nullptr;
        g_noalloc_ptr[device] = static_cast<float*>(p);
    }
    return g_noalloc_ptr[device];
}
}  // namespace

extern "C" int hm_root_policy_apply_noalloc(int device, const hm_packed_pos* d_pos, size_t n_pos,
                                            const uint32_t* d_moves, const uint32_t* d_move_offsets, size_t n_moves,
                                            float* d_priors, const hm_root_policy* rp, uint32_t* d_picks,
                                            int32_t* d_status, void* stream) {
    try {
        if (!d_pos || !d_move_offsets || !d_priors) { hm::set_error("null argument"); return 0; }
        if (!rp) { hm::set_error("null root policy"); return 0; }
        if (rp->temperature < 0.0 || rp->temperature != rp->temperature) {
            hm::set_error("Policy temperature must be positive");  // selfplay.cc:168-170
            return 0;
        }
        if (device < 0 || device >= kMaxDevices) { hm::set_error("no such CUDA device"); return 0; }
        if (n_moves > HM_ROOT_NOALLOC_MAX_MOVES) {
            hm::set_error("hm_root_policy_apply_noalloc: " + std::to_string(n_moves) + " moves exceed HM_ROOT_NOALLOC_MAX_MOVES (" +
                          std::to_string(HM_ROOT_NOALLOC_MAX_MOVES) + "), the size of the static scratch array");
            return 0;
        }
        if (n_pos > 0x7fffffffu / 32u) { hm::set_error("too many positions"); return 0; }
        if (n_moves > 0 && !d_moves) { hm::set_error("null move list"); return 0; }
        if (n_pos == 0) return 1;
        int current = -1;
        if (!hm::cuda_ok(cudaGetDevice(&current), "cudaGetDevice")) return 0;
        if (current != device && !hm::cuda_ok(cudaSetDevice(device), "cudaSetDevice")) return 0;
        float* noise = noalloc_scratch(device);
        if (!noise) return 0;
        hm::RootPolicyArgs a;
        a.temperature = rp->temperature;
        a.alpha = rp->dirichlet_alpha;
        a.epsilon = rp->dirichlet_epsilon;
        a.seed = rp->noise_seed;
        a.hash = rp->position_hash;   // DEVICE pointers for this entry point
        a.captures = rp->captures;
        a.uniforms = rp->uniforms;
        // Nothing in the scratch has to start at zero: a warp writes the variates of its own lists before it reads
        // them, and lists the noise stage skips are never read.  The array keeps the previous call's contents.
        hm::launch_root_policy(d_pos, static_cast<int>(n_pos), d_moves, d_move_offsets, d_priors, a, noise, d_picks, d_status,
                               static_cast<cudaStream_t>(stream));
        return hm::cuda_ok(cudaGetLastError(), "root policy kernel (noalloc)") ? 1 : 0;
    } catch (const std::exception& ex) {
        hm::set_error(std::string("exception: ") + ex.what());
        return 0;
    } catch (...) {
        hm::set_error("unknown exception");
        return 0;
    }
}
