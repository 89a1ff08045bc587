// integration/multigame_selfplay.cc — G self-play games at once on ONE evaluator, with the reference's own,
// unmodified game loop: every thread calls run_selfplay (engine/src/tools/selfplay.cc:558-748) with its own
// small-batch `Engine`, which in this build (-DHM_BATCHING_ENGINE, integration/nn/engine.h) is the facade of
// include/hm_batching_engine.hpp: the 8-position requests of all games' search threads are coalesced into
// large device batches (BASELINE.json configs 3/5: many concurrent games per GPU feeding batched leaf evals).
//
//   hivemind_multigame --model m.onnx --concurrent G --games-per-thread n --nodes N --output dir [--seed s]
//
// Built by `make -C oracle multigame` against the reference sources where they lie; prints one summary line.
#include <atomic>
#include <chrono>
#include <cstring>
#include <filesystem>
#include <iostream>
#include <memory>
#include <string>
#include <thread>
#include <vector>

#include "common/globals.h"
#include "nn/engine.h"
#include "tools/selfplay.h"
#include "Fairy-Stockfish/src/bitboard.h"
#include "Fairy-Stockfish/src/piece.h"
#include "Fairy-Stockfish/src/position.h"
#include "Fairy-Stockfish/src/thread.h"
#include "Fairy-Stockfish/src/types.h"

int main(int argc, char** argv) {
    std::string model;
    std::filesystem::path output = "multigame_out";
    size_t concurrent = 8, gamesPerThread = 1, nodes = 400, maxMacroPlies = 400;
    uint64_t seed = 1;
    int device = 0;
    for (int i = 1; i + 1 < argc; i += 2) {
        const std::string k = argv[i], v = argv[i + 1];
        if (k == "--model") model = v;
        else if (k == "--concurrent") concurrent = std::stoull(v);
        else if (k == "--games-per-thread") gamesPerThread = std::stoull(v);
        else if (k == "--nodes") nodes = std::stoull(v);
        else if (k == "--output") output = v;
        else if (k == "--seed") seed = std::stoull(v);
        else if (k == "--device") device = std::stoi(v);
        else if (k == "--max-macro-plies") maxMacroPlies = std::stoull(v);
        else { std::cerr << "unknown option " << k << std::endl; return 2; }
    }
    if (model.empty()) { std::cerr << "usage: --model <onnx|hvmw> [--concurrent G] [--games-per-thread n] [--nodes N] [--output dir]" << std::endl; return 2; }
    // main.cc:75-81
    Stockfish::pieceMap.init();
    Stockfish::variants.init();
    Stockfish::Bitboards::init();
    Stockfish::Position::init();
    Stockfish::Threads.set(1);
    init_policy_index();

    std::vector<std::unique_ptr<Engine>> engines;
    for (size_t g = 0; g < concurrent; ++g) {
        engines.push_back(std::make_unique<Engine>(device));
        if (!engines.back()->loadNetwork(model, "")) { std::cerr << "loadNetwork failed" << std::endl; return 1; }
    }
    std::atomic<int> failures{0};
    const auto t0 = std::chrono::steady_clock::now();
    std::vector<std::thread> threads;
    for (size_t g = 0; g < concurrent; ++g) {
        threads.emplace_back([&, g] {
            SelfPlayConfig config;
            config.games = gamesPerThread;
            config.nodes = nodes;
            config.maxMacroPlies = maxMacroPlies;
            config.seed = seed + 1000003ULL * g;
            config.outputDirectory = output / ("thread_" + std::to_string(g));
            try {
                std::vector<Engine*> mine{engines[g].get()};
                if (run_selfplay(mine, config) != 0) ++failures;
            } catch (const std::exception& e) {
                std::cerr << "thread " << g << ": " << e.what() << std::endl;
                ++failures;
            }
        });
    }
    for (auto& t : threads) t.join();
    const double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    hm::BatchingBackend& be = engines[0]->backend();
    std::cout << "\nmultigame summary: concurrent " << concurrent << " games_per_thread " << gamesPerThread << " nodes " << nodes
              << " seconds " << secs << " device_passes " << be.passes() << " rows_served " << be.rowsServed()
              << " large_batch " << be.largeBatch() << " failures " << failures.load() << std::endl;
    return failures.load() == 0 ? 0 : 1;
}
