// oracle/scripted/nn/engine.h — TEST INFRASTRUCTURE, not product code.
//
// A scripted stand-in for the reference's `Engine` (engine/src/nn/engine.h:43-132) used ONLY to pin the value
// shaping of SearchThread::process_batch (engine/src/search/searchthread.cc:569-621): it "evaluates" a batch by
// handing back fp16 value / wdl / moves_left scalars a test chose and all-zero policy maps.  With this directory
// ahead of engine/src on the include path the UNMODIFIED reference searchthread.cc / node.cc compile against it
// (oracle/Makefile target `d5`), and oracle/d5_pin.cc reads the backed-up root value of a one-leaf search.
#pragma once
#include <cuda_fp16.h>

#include <cstddef>
#include <string>
#include <vector>

#include "environment/constants.h"
#include "search/search_params.h"

class Engine {
public:
    struct HalfInferenceOutputs {
        const __half* value = nullptr;
        const __half* policyA = nullptr;
        const __half* policyB = nullptr;
        const __half* wdl = nullptr;
        const __half* movesLeft = nullptr;
    };
    explicit Engine(int /*deviceId*/, int batchSize = SearchParams::BATCH_SIZE)
        : m_batchSize(batchSize), m_value(batchSize), m_wdl(3 * static_cast<size_t>(batchSize)), m_ml(batchSize),
          m_pi(static_cast<size_t>(batchSize) * NB_POLICY_VALUES(), __float2half(0.0f)) {}
    Engine(const Engine&) = delete;
    Engine& operator=(const Engine&) = delete;

    // the script: raw fp16 bit patterns (so NaN / inf payloads pass through untouched)
    void script(unsigned short value, const unsigned short wdl[3], unsigned short movesLeft) {
        for (int i = 0; i < m_batchSize; ++i) {
            m_value[i] = __ushort_as_half(value);
            for (int k = 0; k < 3; ++k) m_wdl[3 * i + k] = __ushort_as_half(wdl[k]);
            m_ml[i] = __ushort_as_half(movesLeft);
        }
    }
    bool loadNetwork(const std::string&, const std::string&) { return true; }
    bool enqueueInferenceHalf(const __half*, size_t = 0) {
        if (m_pending) return false;
        m_pending = true;
        return true;
    }
    bool synchronizeInferenceHalf(HalfInferenceOutputs& o, size_t = 0) {
        if (!m_pending) return false;
        m_pending = false;
        o.value = m_value.data();
        o.policyA = m_pi.data();
        o.policyB = m_pi.data();
        o.wdl = m_wdl.data();
        o.movesLeft = m_ml.data();
        return true;
    }
    int getBatchSize() const { return m_batchSize; }

private:
    int m_batchSize;
    bool m_pending = false;
    std::vector<__half> m_value, m_wdl, m_ml, m_pi;
};
