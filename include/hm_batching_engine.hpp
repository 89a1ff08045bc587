// hm_batching_engine.hpp — many small callers, one large device batch.
//
// The reference's search feeds its evaluator 8 positions at a time from each of NUM_SEARCH_THREADS
// threads of ONE game (engine/src/search/searchthread.cc:661-708, search_params.h:26,30); the B200
// evaluator wants 512.  hm::BatchingEngine keeps the reference's `Engine` surface (nn/engine.h:43-132:
// ctor, loadNetwork, runInference, runInferenceHalf, enqueueInferenceHalf, synchronizeInferenceHalf,
// getBatchSize) for a SMALL batch, and coalesces the requests of all the BatchingEngines of a device —
// one per concurrent game — into passes of one shared hm_engine with a LARGE batch.  With
// -DHM_BATCHING_ENGINE, integration/nn/engine.h names it `Engine`, so the reference's Agent /
// SearchThread / run_selfplay run many games at once against one evaluator without a changed line
// (the multi-game feeding problem, solved below the interface).
//
// Mechanics.  The large batch is cut into slots of `batchSize` rows; every (engine, workerIndex) pair
// owns one slot.  enqueue copies the caller's planes into the slot's rows of a pinned staging block (in
// the caller's thread), a dispatcher thread per device worker launches the full batch when every active
// slot of its group is waiting (or after a short linger), and synchronize copies the slot's rows of the
// engine-owned outputs into slot-owned buffers before the next pass may overwrite them.
#pragma once

#include <algorithm>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstdlib>
#include <cstring>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "hm_engine.hpp"

namespace hm {

class BatchingBackend {
public:
    // One backend per device for the whole process (created on first use).
    // The registry is a deliberately leaked singleton: the backends own CUDA resources (an hm_engine, pinned
    // staging) and dispatcher threads, and destroying those from a static destructor would run after the CUDA
    // runtime may already be gone.  shutdownAll() tears them down explicitly where a process wants that.
    static BatchingBackend& forDevice(int device, int smallBatch) {
        std::lock_guard<std::mutex> lock(registryMutex());
        auto& slot = registry()[device];
        if (!slot) slot = new BatchingBackend(device, smallBatch);
        return *slot;
    }
    static void shutdownAll() {
        std::lock_guard<std::mutex> lock(registryMutex());
        for (auto& kv : registry()) delete kv.second;
        registry().clear();
    }

    bool load(const std::string& path) {
        std::lock_guard<std::mutex> lock(m_loadMutex);
        if (m_loaded) return path == m_path;  // one network per device backend
        if (!m_engine || hm_engine_load(m_engine, path.c_str()) == 0) return false;
        m_path = path;
        m_loaded = true;
        return true;
    }
    int smallBatch() const { return m_small; }
    int largeBatch() const { return m_big; }
    int slotsTotal() const { return m_slotsPerGroup * kGroups; }
    long passes() const { return m_passes.load(); }
    long rowsServed() const { return m_rowsServed.load(); }

    // -> slot id or -1 when all slots of the device are taken.  Slots given back by releaseSlot are reused first.
    int acquireSlot() {
        std::lock_guard<std::mutex> lock(m_slotMutex);
        int id;
        if (!m_freeSlots.empty()) {
            id = m_freeSlots.back();
            m_freeSlots.pop_back();
        } else {
            if (m_nextSlot >= slotsTotal()) return -1;
            id = m_nextSlot++;
        }
        Group& g = m_groups[id % kGroups];
        std::lock_guard<std::mutex> gl(g.mutex);
        ++g.active;
        return id;
    }
    // Gives a slot back (engine destroyed / game finished): the group no longer waits for it when it decides whether
    // a pass is full, and a later engine can take it.  The slot must be idle (no request in flight).
    void releaseSlot(int id) {
        if (id < 0) return;
        std::lock_guard<std::mutex> lock(m_slotMutex);
        Group& g = m_groups[id % kGroups];
        {
            std::lock_guard<std::mutex> gl(g.mutex);
            if (g.active > 0) --g.active;
        }
        g.wake.notify_one();   // pending >= active may have become true
        m_freeSlots.push_back(id);
    }

    bool enqueue(int slot, const uint16_t* obs) {
        Group& g = m_groups[slot % kGroups];
        const int local = slot / kGroups;
        if (!m_engine || !g.staging) return false;     // no CUDA device: there is no CPU fallback
        {
            std::lock_guard<std::mutex> lock(g.mutex);
            if (g.state[local] != IDLE) return false;  // one request in flight per worker (engine.cc:579-588)
        }
        std::memcpy(g.staging + static_cast<size_t>(local) * m_small * HM_NB_INPUT_VALUES, obs,
                    static_cast<size_t>(m_small) * HM_NB_INPUT_VALUES * sizeof(uint16_t));
        {
            std::lock_guard<std::mutex> lock(g.mutex);
            g.state[local] = PENDING;
            ++g.pending;
        }
        g.wake.notify_one();
        return true;
    }

    // Blocks until the slot's request has been evaluated; copies its rows out.
    bool wait(int slot, uint16_t* value, uint16_t* piA, uint16_t* piB, uint16_t* wdl, uint16_t* ml) {
        Group& g = m_groups[slot % kGroups];
        const int local = slot / kGroups;
        std::unique_lock<std::mutex> lock(g.mutex);
        if (g.state[local] == IDLE) return false;
        g.done.wait(lock, [&] { return g.state[local] == DONE || g.state[local] == FAILED; });
        const bool ok = g.state[local] == DONE;
        const uint16_t *v = g.outValue, *a = g.outPiA, *b = g.outPiB, *w = g.outWdl, *m = g.outMl;
        lock.unlock();
        if (ok) {
            const size_t r0 = static_cast<size_t>(local) * m_small, n = m_small;
            std::memcpy(value, v + r0, n * 2);
            std::memcpy(piA, a + r0 * HM_NB_POLICY_VALUES, n * HM_NB_POLICY_VALUES * 2);
            std::memcpy(piB, b + r0 * HM_NB_POLICY_VALUES, n * HM_NB_POLICY_VALUES * 2);
            std::memcpy(wdl, w + r0 * 3, n * 3 * 2);
            std::memcpy(ml, m + r0, n * 2);
        }
        lock.lock();
        g.state[local] = IDLE;
        if (--g.copying == 0) g.wake.notify_one();
        return ok;
    }

    ~BatchingBackend() {
        for (auto& g : m_groups) {
            {
                std::lock_guard<std::mutex> lock(g.mutex);
                g.stop = true;
            }
            g.wake.notify_all();
            if (g.thread.joinable()) g.thread.join();
            if (g.staging) hm_free_pinned(g.staging);
        }
        if (m_engine) hm_engine_destroy(m_engine);
    }

private:
    enum State { IDLE, PENDING, INFLIGHT, DONE, FAILED };
    static constexpr int kGroups = 2;  // device workers: one pass computes while the other fills
    struct Group {
        std::mutex mutex;
        std::condition_variable wake, done;
        std::vector<State> state;
        int active = 0, pending = 0, copying = 0;
        bool stop = false;
        uint16_t* staging = nullptr;
        const uint16_t *outValue = nullptr, *outPiA = nullptr, *outPiB = nullptr, *outWdl = nullptr, *outMl = nullptr;
        std::thread thread;
    };

    BatchingBackend(int device, int smallBatch) : m_device(device), m_small(smallBatch) {
        const char* big = std::getenv("HM_BATCHING_LARGE_BATCH");
        m_big = big ? std::atoi(big) : 512;
        m_big = std::max(m_small, m_big - m_big % m_small);
        const char* linger = std::getenv("HM_BATCHING_LINGER_US");
        m_lingerUs = linger ? std::atoi(linger) : 200;
        m_slotsPerGroup = m_big / m_small;
        m_engine = hm_engine_create(device, m_big, kGroups);
        for (int gi = 0; gi < kGroups; ++gi) {
            Group& g = m_groups[gi];
            g.state.assign(m_slotsPerGroup, IDLE);
            g.staging = static_cast<uint16_t*>(hm_alloc_pinned(static_cast<size_t>(m_big) * HM_NB_INPUT_VALUES * sizeof(uint16_t)));
            if (g.staging) std::memset(g.staging, 0, static_cast<size_t>(m_big) * HM_NB_INPUT_VALUES * sizeof(uint16_t));
            g.thread = std::thread([this, gi] { dispatch(gi); });
        }
    }

    void dispatch(int gi) {
        Group& g = m_groups[gi];
        std::vector<int> served;
        std::unique_lock<std::mutex> lock(g.mutex);
        for (;;) {
            g.wake.wait(lock, [&] { return g.stop || (g.pending > 0 && g.copying == 0); });
            if (g.stop) return;
            if (g.pending < g.active && m_lingerUs > 0)  // give the other games a moment to join the pass
                g.wake.wait_for(lock, std::chrono::microseconds(m_lingerUs), [&] { return g.stop || g.pending >= g.active; });
            if (g.stop) return;
            served.clear();
            for (int s = 0; s < m_slotsPerGroup; ++s)
                if (g.state[s] == PENDING) { g.state[s] = INFLIGHT; served.push_back(s); }
            g.pending = 0;
            lock.unlock();
            const uint16_t *v = nullptr, *a = nullptr, *b = nullptr, *w = nullptr, *m = nullptr;
            const bool ok = m_engine && g.staging && hm_engine_enqueue_half(m_engine, g.staging, gi) != 0 &&
                            hm_engine_sync_half(m_engine, gi, &v, &a, &b, &w, &m) != 0;
            m_passes.fetch_add(1);
            m_rowsServed.fetch_add(static_cast<long>(served.size()) * m_small);
            lock.lock();
            g.outValue = v; g.outPiA = a; g.outPiB = b; g.outWdl = w; g.outMl = m;
            for (int s : served) g.state[s] = ok ? DONE : FAILED;
            g.copying = static_cast<int>(served.size());
            g.done.notify_all();
        }
    }

    int m_device, m_small, m_big = 512, m_slotsPerGroup = 0, m_lingerUs = 200;
    hm_engine* m_engine = nullptr;
    Group m_groups[kGroups];
    std::mutex m_loadMutex, m_slotMutex;
    bool m_loaded = false;
    std::string m_path;
    int m_nextSlot = 0;
    std::vector<int> m_freeSlots;
    std::atomic<long> m_passes{0}, m_rowsServed{0};

    static std::mutex& registryMutex() { static std::mutex* mu = new std::mutex; return *mu; }
    static std::map<int, BatchingBackend*>& registry() { static auto* all = new std::map<int, BatchingBackend*>; return *all; }
};

// The reference's Engine surface (nn/engine.h:43-132) for a small batch, served by the device's
// BatchingBackend.  Each workerIndex of each BatchingEngine owns one slot of the large batch.
class BatchingEngine {
public:
    struct HalfInferenceOutputs {
        const __half* value = nullptr;
        const __half* policyA = nullptr;
        const __half* policyB = nullptr;
        const __half* wdl = nullptr;
        const __half* movesLeft = nullptr;
    };

    explicit BatchingEngine(int deviceId, int batchSize = HM_DEFAULT_BATCH_SIZE)
        : m_deviceId(deviceId), m_batchSize(batchSize), m_backend(BatchingBackend::forDevice(deviceId, batchSize)),
          m_workers(HM_DEFAULT_WORKERS) {}
    BatchingEngine(const BatchingEngine&) = delete;
    BatchingEngine& operator=(const BatchingEngine&) = delete;
    ~BatchingEngine() {   // hand the large-batch slots of this engine's workers back to the device backend
        for (Worker& w : m_workers) m_backend.releaseSlot(w.slot);
    }

    bool loadNetwork(const std::string& onnxFile, const std::string& engineFile = std::string()) {
        (void)engineFile;
        if (m_backend.smallBatch() != m_batchSize) return fail("loadNetwork", "all engines of a device must use one batch size");
        if (!m_backend.load(onnxFile)) return fail("loadNetwork", hm_last_error());
        return true;
    }
    int getBatchSize() const { return m_batchSize; }

    bool enqueueInferenceHalf(const __half* obs, size_t workerIndex = 0) {
        Worker* w = worker(workerIndex);
        if (!w) return fail("enqueueInferenceHalf", "no free slot on this device (raise HM_BATCHING_LARGE_BATCH)");
        if (!m_backend.enqueue(w->slot, reinterpret_cast<const uint16_t*>(obs)))
            return fail("enqueueInferenceHalf", "worker already has a request in flight");
        return true;
    }
    bool synchronizeInferenceHalf(HalfInferenceOutputs& out, size_t workerIndex = 0) {
        Worker* w = worker(workerIndex);
        if (!w || !m_backend.wait(w->slot, w->value.data(), w->piA.data(), w->piB.data(), w->wdl.data(), w->ml.data()))
            return fail("synchronizeInferenceHalf", "no request pending or the device pass failed");
        out.value = reinterpret_cast<const __half*>(w->value.data());
        out.policyA = reinterpret_cast<const __half*>(w->piA.data());
        out.policyB = reinterpret_cast<const __half*>(w->piB.data());
        out.wdl = reinterpret_cast<const __half*>(w->wdl.data());
        out.movesLeft = reinterpret_cast<const __half*>(w->ml.data());
        return true;
    }
    bool runInferenceHalf(const __half* obs, HalfInferenceOutputs& out, size_t workerIndex = 0) {
        return enqueueInferenceHalf(obs, workerIndex) && synchronizeInferenceHalf(out, workerIndex);
    }
#ifdef HM_HAVE_CUDA_FP16
    // engine.cc:537-566: through fp16 both ways
    bool runInference(float* obs, float* value, float* piA, float* piB, float* wdl, float* movesLeft,
                      size_t workerIndex = 0) {
        const size_t B = m_batchSize;
        std::vector<__half> h(B * HM_NB_INPUT_VALUES);
        for (size_t i = 0; i < h.size(); ++i) h[i] = __float2half(obs[i]);
        HalfInferenceOutputs o;
        if (!runInferenceHalf(h.data(), o, workerIndex)) return false;
        auto cvt = [](float* dst, const __half* src, size_t n) {
            if (dst) for (size_t i = 0; i < n; ++i) dst[i] = __half2float(src[i]);
        };
        cvt(value, o.value, B);
        cvt(piA, o.policyA, B * HM_NB_POLICY_VALUES);
        cvt(piB, o.policyB, B * HM_NB_POLICY_VALUES);
        cvt(wdl, o.wdl, B * 3);
        cvt(movesLeft, o.movesLeft, B);
        return true;
    }
#endif
    int getDeviceId() const { return m_deviceId; }
    BatchingBackend& backend() { return m_backend; }

private:
    struct Worker {
        int slot = -1;
        std::vector<uint16_t> value, piA, piB, wdl, ml;
    };
    Worker* worker(size_t index) {
        if (index >= m_workers.size()) return nullptr;
        std::lock_guard<std::mutex> lock(m_workerMutex);
        Worker& w = m_workers[index];
        if (w.slot < 0) {
            w.slot = m_backend.acquireSlot();
            if (w.slot < 0) return nullptr;
            const size_t B = m_batchSize;
            w.value.resize(B); w.ml.resize(B); w.wdl.resize(3 * B);
            w.piA.resize(B * HM_NB_POLICY_VALUES); w.piB.resize(B * HM_NB_POLICY_VALUES);
        }
        return &w;
    }
    static bool fail(const char* what, const char* why) {
        std::cerr << "[ERROR] " << what << ": " << why << std::endl;  // engine.cc:105-111 convention
        return false;
    }
    int m_deviceId, m_batchSize;
    BatchingBackend& m_backend;
    std::mutex m_workerMutex;
    std::vector<Worker> m_workers;
};

}  // namespace hm
