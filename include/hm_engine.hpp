// hm_engine.hpp — header-only C++ `Engine` with the public surface of the reference's
// nn/ Engine (reference engine/src/nn/engine.h:50-82), implemented over the C ABI in
// hm_engine.h (hivemind_b200/libhm_engine.so).  It is what a maintainer of the reference drops in place
// of engine/src/nn/engine.{h,cc}: searchthread.cc, benchmark.cc and selfplay.cc compile against
// it unchanged (same class name, method names, argument order and bool error convention), and
// TensorRT / NvOnnxParser are no longer needed.  See INTEGRATION.md.
//
// Beyond the reference surface it exposes the fused B200 entry points
// (enqueuePositions / synchronizePriors: packed positions in, priors + shaped values out) that
// replace the board_to_planes -> forward -> get_normalized_probability -> value shaping
// sequence of searchthread.cc:445-619 with one device pass.
#pragma once

#include <cstddef>
#include <cstdint>
#include <iostream>
#include <string>

#if defined(__has_include)
#if __has_include(<cuda_fp16.h>)
#include <cuda_fp16.h>
#define HM_HAVE_CUDA_FP16 1
#endif
#endif
#ifndef HM_HAVE_CUDA_FP16
// Host-only builds without the CUDA toolkit: fp16 values are opaque 16-bit patterns.
struct __half {
    uint16_t x;
};
#endif

#include "hm_engine.h"

#ifndef HM_DEFAULT_BATCH_SIZE
#define HM_DEFAULT_BATCH_SIZE 8 /* SearchParams::BATCH_SIZE, search_params.h:26 */
#endif
#ifndef HM_DEFAULT_WORKERS
#define HM_DEFAULT_WORKERS 4 /* SearchParams::NUM_SEARCH_THREADS, search_params.h:30 */
#endif

// The class is called `Engine`, as in the reference, unless a build renames it (the batching build of
// integration/nn/engine.h gives that name to hm::BatchingEngine instead).
#ifndef HM_ENGINE_CLASS_NAME
#define HM_ENGINE_CLASS_NAME Engine
#endif
class HM_ENGINE_CLASS_NAME {
public:
    // engine.h:52-58
    struct HalfInferenceOutputs {
        const __half* value = nullptr;
        const __half* policyA = nullptr;
        const __half* policyB = nullptr;
        const __half* wdl = nullptr;
        const __half* movesLeft = nullptr;
    };

    // engine.h:60 (the reference fixes the worker count at compile time; here it is a
    // constructor argument with the same default)
    explicit HM_ENGINE_CLASS_NAME(int deviceId, int batchSize = HM_DEFAULT_BATCH_SIZE,
                    int numWorkers = HM_DEFAULT_WORKERS)
        : m_deviceId(deviceId), m_batchSize(batchSize) {
        m_handle = hm_engine_create(deviceId, batchSize, numWorkers);
        if (m_handle == nullptr) report("Engine");
    }
    ~HM_ENGINE_CLASS_NAME() {
        if (m_handle != nullptr) hm_engine_destroy(m_handle);
    }
    HM_ENGINE_CLASS_NAME(const HM_ENGINE_CLASS_NAME&) = delete;
    HM_ENGINE_CLASS_NAME& operator=(const HM_ENGINE_CLASS_NAME&) = delete;

    // engine.h:66.  `onnxFile` may name the reference's .onnx: the packed weights are read
    // from the sibling .hvmw written by `python -m hivemind_b200.weights`;
    // `engineFile` (the TensorRT plan cache) has no counterpart and is ignored.
    bool loadNetwork(const std::string& onnxFile, const std::string& engineFile = std::string()) {
        (void)engineFile;
        return ok(m_handle != nullptr && hm_engine_load(m_handle, onnxFile.c_str()) != 0, "loadNetwork");
    }
    bool loadNetworkFromMemory(const void* blob, size_t bytes) {
        return ok(m_handle != nullptr && hm_engine_load_memory(m_handle, blob, bytes) != 0, "loadNetworkFromMemory");
    }

    // engine.h:72-73
    bool runInference(float* obs, float* value, float* piA, float* piB, float* wdl,
                      float* movesLeft, size_t workerIndex = 0) {
        return ok(m_handle != nullptr &&
                      hm_engine_run_float(m_handle, obs, value, piA, piB, wdl, movesLeft, workerIndex) != 0,
                  "runInference");
    }

    // engine.h:75-76
    bool runInferenceHalf(const __half* obs, HalfInferenceOutputs& outputs, size_t workerIndex = 0) {
        return enqueueInferenceHalf(obs, workerIndex) && synchronizeInferenceHalf(outputs, workerIndex);
    }

    // engine.h:80
    bool enqueueInferenceHalf(const __half* obs, size_t workerIndex = 0) {
        return ok(m_handle != nullptr &&
                      hm_engine_enqueue_half(m_handle, reinterpret_cast<const uint16_t*>(obs), workerIndex) != 0,
                  "enqueueInferenceHalf");
    }
    // engine.h:81-82
    bool synchronizeInferenceHalf(HalfInferenceOutputs& outputs, size_t workerIndex = 0) {
        const uint16_t *v = nullptr, *a = nullptr, *b = nullptr, *w = nullptr, *m = nullptr;
        if (m_handle == nullptr || hm_engine_sync_half(m_handle, workerIndex, &v, &a, &b, &w, &m) == 0)
            return ok(false, "synchronizeInferenceHalf");
        outputs.value = reinterpret_cast<const __half*>(v);
        outputs.policyA = reinterpret_cast<const __half*>(a);
        outputs.policyB = reinterpret_cast<const __half*>(b);
        outputs.wdl = reinterpret_cast<const __half*>(w);
        outputs.movesLeft = reinterpret_cast<const __half*>(m);
        return true;
    }

    // engine.h:87
    int getBatchSize() const { return m_batchSize; }

    // ---- fused B200 path (no counterpart in the reference class) ----
    struct PriorOutputs {
        const float* priors = nullptr;  // one per move, same offsets as the move lists
        const float* values = nullptr;  // shaped value per position (searchthread.cc:569-619)
        const float* raw = nullptr;     // {value, wdl loss, draw, win, moves_left} per position
    };
    bool enqueuePositions(const hm_packed_pos* pos, size_t nPos, const uint32_t* moves,
                          const uint32_t* moveOffsets, size_t workerIndex = 0) {
        return ok(m_handle != nullptr &&
                      hm_engine_enqueue_positions(m_handle, pos, nPos, moves, moveOffsets, workerIndex) != 0,
                  "enqueuePositions");
    }
    bool synchronizePriors(PriorOutputs& out, size_t workerIndex = 0) {
        return ok(m_handle != nullptr &&
                      hm_engine_sync_priors(m_handle, workerIndex, &out.priors, &out.values, &out.raw) != 0,
                  "synchronizePriors");
    }
    // ---- root policy on the device: apply_temperature (selfplay.cc:167-185), root Dirichlet noise
    // (node.h:288-315), sample_raw_policy_action's joint sampling (selfplay.cc:349-392) ----
    struct RootOutputs : PriorOutputs {
        const uint32_t* picks = nullptr;  // [nPos][2]: index into board A's / board B's move list
        const int32_t* status = nullptr;  // [nPos] HM_ROOT_*
    };
    bool enqueuePositionsRoot(const hm_packed_pos* pos, size_t nPos, const uint32_t* moves,
                              const uint32_t* moveOffsets, const hm_root_policy& root, size_t workerIndex = 0) {
        return ok(m_handle != nullptr &&
                      hm_engine_enqueue_positions_root(m_handle, pos, nPos, moves, moveOffsets, &root, workerIndex) != 0,
                  "enqueuePositionsRoot");
    }
    bool synchronizeRoot(RootOutputs& out, size_t workerIndex = 0) {
        return ok(m_handle != nullptr && hm_engine_sync_root(m_handle, workerIndex, &out.priors, &out.values, &out.raw,
                                                             &out.picks, &out.status) != 0,
                  "synchronizeRoot");
    }
    // The canonical doubles std::discrete_distribution would take from `randomEngine` in
    // sample_raw_policy_action, WITHOUT advancing it past what the reference consumes: u[0] / u[1] are
    // drawn only for lists of >= 2 entries (board B first: g++ evaluates the two arguments of
    // selfplay.cc:365 right to left), u[2] comes from a copy of the engine.  After synchronizeRoot call
    // commitRootDraw(randomEngine, status) — it takes the third draw for real iff the device used it.
    template <class URNG>
    static void drawRootUniforms(URNG& randomEngine, size_t nA, size_t nB, double u[3], bool boardBFirst = true) {
        u[0] = u[1] = u[2] = 0.0;
        for (int k = 0; k < 2; ++k) {
            const int b = boardBFirst ? 1 - k : k;
            if ((b ? nB : nA) >= 2) u[b] = canonical(randomEngine);
        }
        URNG peek = randomEngine;
        u[2] = canonical(peek);
    }
    template <class URNG>
    static void commitRootDraw(URNG& randomEngine, int32_t status) {
        if (status == HM_ROOT_RESAMPLED) (void)canonical(randomEngine);
    }
    bool setValueConfig(const hm_value_config& cfg) {   // false while a request is in flight
        return m_handle != nullptr && hm_engine_set_value_config(m_handle, &cfg) != 0;
    }
    hm_engine* handle() const { return m_handle; }
    int getDeviceId() const { return m_deviceId; }

private:
    // std::generate_canonical<double, 53> for a 64-bit engine: one draw, x / 2^64 (bits/random.tcc)
    template <class URNG>
    static double canonical(URNG& e) {
        const double r = static_cast<double>(e() - URNG::min()) / 18446744073709551616.0;
        return r < 1.0 ? r : 0.99999999999999988898;
    }
    // the reference reports failures on stderr and returns false (engine.cc:217-248, 573-679)
    static void report(const char* what) { std::cerr << "[ERROR] " << what << ": " << hm_last_error() << std::endl; }
    static bool ok(bool cond, const char* what) {
        if (!cond) report(what);
        return cond;
    }
    hm_engine* m_handle = nullptr;
    int m_deviceId;
    int m_batchSize;
};
