// hm_bench — the reference's `hivemind bench` (engine/src/main.cc:84-103,
// engine/src/tools/benchmark.cc:15-57) restated against include/hm_engine.hpp: a pinned fp16
// buffer of B x 74 x 8 x 8 uniform [0,1) halves, 10 warm-up calls, then `iterations` x
// Engine::runInferenceHalf timed on the wall clock; prints the same report.  `--overlap` adds
// the double-buffered pattern of the search threads (searchthread.cc:661-708: enqueue on one
// worker while the other synchronises).
//
//   g++ -O2 -std=c++17 -Iinclude tools/hm_bench.cc -Lhivemind_b200 -lhm_engine \
//       -Wl,-rpath,'$ORIGIN/../hivemind_b200' -o tools/hm_bench
//   tools/hm_bench <weights.hvmw> [iterations] [batch] [--overlap]
#include <chrono>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <string>

#include "hm_engine.hpp"

namespace {

// round-to-nearest-even float -> fp16 bit pattern for values in [0, 1]
uint16_t half_bits(float f) {
    uint32_t u;
    std::memcpy(&u, &f, 4);
    const int32_t e = static_cast<int32_t>((u >> 23) & 0xff) - 127 + 15;
    uint32_t m = u & 0x7fffffu;
    if (e <= 0) {
        if (e < -10) return 0;
        m |= 0x800000u;
        const int shift = 14 - e;
        uint32_t h = m >> shift;
        const uint32_t rem = m & ((1u << shift) - 1), halfway = 1u << (shift - 1);
        if (rem > halfway || (rem == halfway && (h & 1))) ++h;
        return static_cast<uint16_t>(h);
    }
    uint32_t h = (static_cast<uint32_t>(e) << 10) | (m >> 13);
    const uint32_t rem = m & 0x1fffu;
    if (rem > 0x1000u || (rem == 0x1000u && (h & 1))) ++h;
    return static_cast<uint16_t>(h);
}

void report(const char* title, int iterations, double total_ms, int batch) {
    const double avg_ms = total_ms / iterations;
    const double per_sec = 1000.0 / avg_ms;
    std::cout << "=== " << title << " ===\n"
              << "Iterations: " << iterations << "\n"
              << "Total time: " << total_ms << " ms\n"
              << "Average time per inference: " << avg_ms << " ms\n"
              << "Inferences per second: " << per_sec << "\n"
              << "Positions per second: " << per_sec * batch << "\n"
              << "===========================" << std::endl;
}

}  // namespace

int main(int argc, char** argv) {
    if (argc < 2) {
        std::cerr << "usage: hm_bench <weights.hvmw|model.onnx> [iterations] [batch] [--overlap]" << std::endl;
        return 2;
    }
    int iterations = 1000, batch = 512;
    bool overlap = false;
    int pos = 0;
    for (int i = 2; i < argc; ++i) {
        if (std::strcmp(argv[i], "--overlap") == 0) overlap = true;
        else if (pos++ == 0) iterations = std::atoi(argv[i]);
        else batch = std::atoi(argv[i]);
    }
    Engine engine(0, batch, 2);
    if (!engine.loadNetwork(argv[1], "")) return 1;

    const size_t n = static_cast<size_t>(engine.getBatchSize()) * HM_NB_INPUT_VALUES;
    __half* obs[2];
    for (auto& o : obs) {
        o = static_cast<__half*>(hm_alloc_pinned(n * sizeof(__half)));
        if (o == nullptr) {
            std::cerr << "Failed to allocate pinned inference benchmark input" << std::endl;
            return 1;
        }
        for (size_t i = 0; i < n; ++i) {
            const uint16_t b = half_bits(static_cast<float>(rand()) / static_cast<float>(RAND_MAX));
            std::memcpy(&o[i], &b, 2);
        }
    }
    Engine::HalfInferenceOutputs outputs;
    for (int i = 0; i < 10; ++i)
        if (!engine.runInferenceHalf(obs[0], outputs)) return 1;

    auto t0 = std::chrono::high_resolution_clock::now();
    for (int i = 0; i < iterations; ++i)
        if (!engine.runInferenceHalf(obs[0], outputs)) return 1;
    auto t1 = std::chrono::high_resolution_clock::now();
    report("Inference Benchmark", iterations, std::chrono::duration<double, std::milli>(t1 - t0).count(), batch);
    uint16_t v0;
    std::memcpy(&v0, outputs.value, 2);
    std::cout << "value[0] bits: 0x" << std::hex << v0 << std::dec << std::endl;

    if (overlap) {
        t0 = std::chrono::high_resolution_clock::now();
        if (!engine.enqueueInferenceHalf(obs[0], 0)) return 1;
        for (int i = 1; i < iterations; ++i) {
            if (!engine.enqueueInferenceHalf(obs[i & 1], i & 1)) return 1;
            if (!engine.synchronizeInferenceHalf(outputs, (i - 1) & 1)) return 1;
        }
        if (!engine.synchronizeInferenceHalf(outputs, (iterations - 1) & 1)) return 1;
        t1 = std::chrono::high_resolution_clock::now();
        report("Inference Benchmark (2 workers in flight)", iterations,
               std::chrono::duration<double, std::milli>(t1 - t0).count(), batch);
    }
    for (auto& o : obs) hm_free_pinned(o);
    return 0;
}
