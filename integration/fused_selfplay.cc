// integration/fused_selfplay.cc — many concurrent self-play games per GPU feeding FUSED leaf evaluations
// (BASELINE.json configs 3 / 5; SURVEY.md §8(f) rank 1: "replace the per-thread batch with a cross-game leaf queue;
// emit packed positions + move lists directly, skip CPU planes").
//
// What is the reference's and what is ours:
//   * the TREE is the reference's own, compiled UNMODIFIED where it lies (engine/src/search/searchthread.cc, node.cc):
//     SearchThread::select_and_expand (PUCT, virtual loss, progressive widening, transposition canonicalisation),
//     SearchThread::expand_leaf_node, SearchThread::backup, classify_terminal_position, Node, TranspositionTable,
//     Board / Fairy-Stockfish move generation.  One SearchThread object per game is used as that game's tree-operations
//     helper (this TU is compiled with -fno-access-control to reach its trajectory buffer; nothing is copied or patched);
//   * what REPLACES SearchThread::collect_batch / process_batch (searchthread.cc:255-639) and Agent::worker_loop
//     (agent.cc:290-370) is below, written against hm::LeafBatch (include/hm_batch.hpp): a leaf is handed to the
//     evaluator as a 256-byte packed position plus its legal-move lists — no CPU plane encoding, no 9.5 KB/position
//     H2D, no 18.7 KB/position dense-policy D2H, no CPU softmax, no second move generation in process_batch, no board
//     replay to the leaf — and comes back as priors per move and the SHAPED value (searchthread.cc:569-619 on device);
//   * the leaves of MANY games fill one request: a host thread owns G/T games in two halves that ping-pong on two
//     execution slots of the evaluator, so the device evaluates one half's batch while the thread backs up the other
//     half's results and selects its next leaves;
//   * the game loop follows run_selfplay (tools/selfplay.cc:558-748): randomised node budget, root Dirichlet noise
//     seeded per move with mix_seed, visit-count temperature sampling, checkmate / draw / resignation / ply limit,
//     optional HVM4 training chunks through the product's ChunkWriter (include/hm_chunk.hpp).
//
//   hivemind_selfplay_fused --model net.hvmw [--games 1024] [--nodes 400] [--threads T] [--batch 512]
//                           [--seconds 10 | --max-games N] [--device 0] [--seed 1] [--chunks dir]
// Prints ONE JSON line on stdout (bench.py's "selfplay" record).  Built by `make -C integration`.
#include <algorithm>
#include <array>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstring>
#include <filesystem>
#include <iostream>
#include <memory>
#include <numeric>
#include <random>
#include <string>
#include <thread>
#include <vector>

#include "common/globals.h"
#include "common/utils.h"
#include "environment/board.h"
#include "environment/joint_action.h"
#include "environment/planes.h"
#include "search/node.h"
#include "search/searchinfo.h"
#include "search/searchthread.h"   // pulls "nn/engine.h" = integration/nn/engine.h -> include/hm_engine.hpp
#include "search/transposition_table.h"
#include "Fairy-Stockfish/src/bitboard.h"
#include "Fairy-Stockfish/src/piece.h"
#include "Fairy-Stockfish/src/position.h"
#include "Fairy-Stockfish/src/thread.h"
#include "Fairy-Stockfish/src/types.h"

#include "hm_batch.hpp"
#include "hm_chunk.hpp"

namespace fused {

struct Config {
    std::string model;
    size_t games = 1024, nodes = 400, threads = 0, batch = 512, maxMacroPlies = 400, maxGames = 0, chunkSamples = 16384;
    double seconds = 10.0;
    int device = 0;
    uint64_t seed = 1;
    std::string chunkDir;
    // exploration settings of SelfPlayConfig (tools/selfplay.h:10-31)
    double mctsTemperature = 1.0, mctsTemperatureDecay = 0.93, nodeRandomFactor = 0.05, resignDisableFraction = 0.10;
    size_t mctsTemperaturePlies = 20, resignConsecutivePlies = 3;
    float resignThreshold = -0.90f, dirichletAlpha = 0.3f, dirichletEpsilon = 0.25f;
};

// tools/selfplay.cc:160-165
inline uint64_t mix_seed(uint64_t seed, uint64_t value) {
    value += 0x9e3779b97f4a7c15ULL;
    value = (value ^ (value >> 30)) * 0xbf58476d1ce4e5b9ULL;
    value = (value ^ (value >> 27)) * 0x94d049bb133111ebULL;
    return seed ^ (value ^ (value >> 31));
}

struct Stats {
    std::atomic<uint64_t> nodes{0}, evals{0}, terminals{0}, requests{0}, collisions{0}, moves{0}, games{0}, samples{0};
    // host-time attribution (TSC ticks summed over threads; --profile 1): select_and_expand | terminal classification |
    // pack + move generation (LeafBatch::add) | expand_leaf_node | backup | waiting for the device | playing moves
    std::atomic<uint64_t> tSelect{0}, tTerminal{0}, tPack{0}, tExpand{0}, tBackup{0}, tWait{0}, tMove{0};
    bool profile = false;
};
inline uint64_t ticks() { return __builtin_ia32_rdtsc(); }
struct Lap {   // adds the ticks since construction / the last lap to a counter when profiling is on
    explicit Lap(const Stats& s) : on(s.profile), t(on ? ticks() : 0) {}
    void to(std::atomic<uint64_t>& c) { if (on) { const uint64_t n = ticks(); c.fetch_add(n - t, std::memory_order_relaxed); t = n; } }
    bool on;
    uint64_t t;
};

// One leaf on its way through the evaluator (the role of LeafContext, searchthread.h:52-78).
struct Leaf {
    std::shared_ptr<Node> node;
    std::vector<TrajectoryEntry> trajectory;
    Stockfish::Color team = Stockfish::WHITE;
    bool terminal = false, reserved = false, timeAdvantage = false, aOnTurn = false, bOnTurn = false;
    float terminalValue = 0.0f;
    uint64_t hash = 0;
    int slot = -1;        // index in the LeafBatch
    int game = -1;        // index in the worker's game list
};

class Game {
public:
    Game(const Config& cfg, uint64_t id) : m_cfg(cfg), m_id(id), m_rng(mix_seed(cfg.seed, id)) { newGame(); }

    bool searching() const { return m_searching; }
    size_t inFlight() const { return m_inFlight; }

    // Select up to `want` leaves of this game's tree into `batch` / `out` (the per-leaf body of collect_batch,
    // searchthread.cc:270-430, for a fused request).  Returns the number of leaves added (terminal ones included).
    size_t collect(hm::LeafBatch& batch, std::vector<Leaf>& out, int gameIndex, size_t want, Stats& stats) {
        size_t added = 0, attempts = 0;
        const size_t first = out.size();
        // never select more simulations than the move's budget still has room for
        const size_t room = m_target > m_done + m_inFlight ? m_target - m_done - m_inFlight : 0;
        want = std::min(want, room);
        while (added < want && attempts < 2 * want && !batch.full()) {
            ++attempts;
            std::vector<TrajectoryEntry>& path = m_tree.trajectoryBuffer;
            path.clear();
            Lap lap(stats);
            LeafSelection sel = m_tree.select_and_expand(*m_board, m_timeAdvantage);
            lap.to(stats.tSelect);
            auto rewind = [&] {
                for (auto it = path.rbegin(); it != path.rend(); ++it)
                    if (it->action.moveA != Stockfish::MOVE_NONE || it->action.moveB != Stockfish::MOVE_NONE)
                        m_board->unmake_moves(it->action.moveA, it->action.moveB);
            };
            if (!sel.leaf) {   // the node is being evaluated by a request of ours that is still in flight
                stats.collisions.fetch_add(1, std::memory_order_relaxed);
                m_tree.cancel_virtual_losses(path);
                rewind();
                if (m_inFlight + added == 0) break;   // nothing to wait for: the tree has no free leaf right now
                continue;
            }
            bool repeat = false;
            for (size_t i = first; i < out.size(); ++i) repeat = repeat || out[i].node == sel.leaf;
            if (repeat) {      // a repeated leaf is not an independent simulation (searchthread.cc:296-313)
                stats.collisions.fetch_add(1, std::memory_order_relaxed);
                if (sel.hasEvaluationReservation) sel.leaf->release_evaluation_reservation();
                m_tree.cancel_virtual_losses(path);
                rewind();
                continue;
            }
            Leaf leaf;
            leaf.node = sel.leaf;
            leaf.reserved = sel.hasEvaluationReservation;
            leaf.team = sel.leaf->get_team_to_play();
            leaf.game = gameIndex;
            const Stockfish::Color rootTeam = m_root->get_team_to_play();
            const bool sitPlaneActive = (leaf.team == rootTeam) == m_timeAdvantage;
            const int ply = static_cast<int>(path.size()) - 1;
            m_info->set_max_depth(ply);
            const float drawValue = leaf.team == rootTeam ? -m_runtime.drawContempt : m_runtime.drawContempt;
            const NodeType solved = SearchParams::ENABLE_MCTS_SOLVER ? sel.leaf->get_node_type() : NodeType::UNSOLVED;
            if (solved != NodeType::UNSOLVED) {
                leaf.terminal = true;
                leaf.terminalValue = solved == NodeType::WIN ? 1.0f : solved == NodeType::LOSS ? -1.0f : drawValue;
            } else {
                int endInPly = 0;
                const TerminalOutcome outcome = classify_terminal_position(*m_board, leaf.team, rootTeam, m_timeAdvantage, ply, &endInPly);
                if (outcome != TerminalOutcome::NONE) {
                    leaf.terminal = true;
                    if (outcome == TerminalOutcome::WIN) {
                        leaf.terminalValue = 1.0f;
                        if (SearchParams::ENABLE_MCTS_SOLVER) sel.leaf->mark_as_win(endInPly);
                    } else if (outcome == TerminalOutcome::LOSS) {
                        leaf.terminalValue = -1.0f;
                        if (SearchParams::ENABLE_MCTS_SOLVER) sel.leaf->mark_as_loss(endInPly);
                    } else {
                        leaf.terminalValue = drawValue;
                        if (SearchParams::ENABLE_MCTS_SOLVER) sel.leaf->mark_as_draw(1);
                    }
                }
            }
            lap.to(stats.tTerminal);
            if (!leaf.terminal) {
                if (!leaf.reserved) {   // someone else (an in-flight request) owns the evaluation
                    stats.collisions.fetch_add(1, std::memory_order_relaxed);
                    m_tree.cancel_virtual_losses(path);
                    rewind();
                    continue;
                }
                leaf.timeAdvantage = leaf.team == rootTeam ? m_timeAdvantage : !m_timeAdvantage;
                leaf.hash = m_board->hash_key(leaf.timeAdvantage);
                leaf.aOnTurn = m_board->side_to_move(BOARD_A) == leaf.team;
                leaf.bOnTurn = m_board->side_to_move(BOARD_B) == ~leaf.team;
                // THE fused hand-over: packed record + legal-move lists (+ capture flags), generated once, here
                leaf.slot = batch.add(*m_board, leaf.team, sitPlaneActive);
                lap.to(stats.tPack);
            }
            rewind();
            leaf.trajectory = std::move(path);   // select_and_expand's buffer is cleared before its next use
            out.push_back(std::move(leaf));
            ++added;
        }
        m_inFlight += added;
        return added;
    }

    // Results of one leaf are back: expand + back up (what process_batch does per context, searchthread.cc:444-639,
    // minus everything the device already did: legal-move softmax, value shaping; and minus the board replay and the
    // second move generation, whose outputs travelled with the request).
    void absorb(Leaf& leaf, const hm::LeafBatch& batch, Stats& stats) {
        std::vector<TrajectoryEntry>& path = leaf.trajectory;
        if (leaf.terminal) {
            if (leaf.reserved) leaf.node->release_evaluation_reservation();
            m_tree.backup(path, *m_board, leaf.terminalValue);
            stats.terminals.fetch_add(1, std::memory_order_relaxed);
        } else if (SearchParams::ENABLE_MCTS_SOLVER && leaf.node->get_node_type() != NodeType::UNSOLVED) {
            leaf.node->release_evaluation_reservation();
            m_tree.backup(path, *m_board, 0.0f);
        } else {
            for (int bn = 0; bn < 2; ++bn) {
                const size_t n = batch.numActions(leaf.slot, bn);
                const uint32_t* mv = batch.actions(leaf.slot, bn);
                const float* pr = batch.priors(leaf.slot, bn);
                const uint8_t* cp = batch.captures(leaf.slot, bn);
                m_actions[bn].resize(n);
                for (size_t i = 0; i < n; ++i) m_actions[bn][i] = static_cast<Stockfish::Move>(mv[i]);
                m_priors[bn].assign(pr, pr + n);
                m_captures[bn].assign(cp, cp + n);
            }
            Lap lap(stats);
            m_tree.expand_leaf_node(leaf.node.get(), m_actions[0], m_actions[1], m_priors[0], m_priors[1], leaf.timeAdvantage,
                                    leaf.aOnTurn, leaf.bOnTurn, m_captures[0], m_captures[1], leaf.hash);
            leaf.node->release_evaluation_reservation();
            lap.to(stats.tExpand);
            m_tree.backup(path, *m_board, batch.value(leaf.slot));   // already shaped (searchthread.cc:569-619) on the device
            lap.to(stats.tBackup);
            stats.evals.fetch_add(1, std::memory_order_relaxed);
        }
        m_info->increment_nodes(1);
        stats.nodes.fetch_add(1, std::memory_order_relaxed);
        ++m_done;
        --m_inFlight;
    }

    // Called when no request of this game is in flight: if the move's node budget is spent, play the move
    // (tools/selfplay.cc:660-720) and set the next search up (a finished game is replaced by a fresh one).
    void advance(Stats& stats, hm::ChunkWriter* writer) {
        if (!m_searching || m_inFlight != 0 || m_done < m_target) return;
        std::vector<int> visits = m_root->is_expanded() ? m_root->get_child_visits() : std::vector<int>();
        const size_t edges = std::min(visits.size(), m_root->is_expanded() ? m_root->get_num_generated() : size_t{0});
        if (edges == 0) { endGame(m_team == Stockfish::WHITE ? 1 : 0, stats, writer); return; }   // "no legal action"
        const float rootQ = m_root->Q();
        if (writer) recordSample(visits, edges, rootQ);
        const size_t side = m_team == Stockfish::WHITE ? 0 : 1;
        if (m_canResign) {   // :681-692
            if (rootQ <= m_cfg.resignThreshold) {
                if (++m_resignPlies[side] >= m_cfg.resignConsecutivePlies) { endGame(m_team == Stockfish::WHITE ? 1 : 0, stats, writer); return; }
            } else {
                m_resignPlies[side] = 0;
            }
        }
        const JointActionCandidate action = m_root->get_joint_action(static_cast<int>(pickEdge(visits, edges)));
        if (action.moveA != Stockfish::MOVE_NONE) m_board->push_move(BOARD_A, action.moveA);
        if (action.moveB != Stockfish::MOVE_NONE) m_board->push_move(BOARD_B, action.moveB);
        stats.moves.fetch_add(1, std::memory_order_relaxed);
        m_team = ~m_team;
        m_timeAdvantage = !m_timeAdvantage;
        ++m_macroPly;
        beginMove(stats, writer);
    }

private:
    // select_action, tools/selfplay.cc:429-460: arg-max of the visits at temperature 0, else weights
    // exp((log v - log vmax) / T) through std::discrete_distribution
    size_t pickEdge(const std::vector<int>& visits, size_t edges) {
        const size_t best = static_cast<size_t>(std::max_element(visits.begin(), visits.begin() + edges) - visits.begin());
        double temperature = 0.0;   // mcts_temperature, :226-232
        if (!(m_cfg.mctsTemperaturePlies > 0 && m_macroPly >= m_cfg.mctsTemperaturePlies))
            temperature = m_cfg.mctsTemperature * std::pow(m_cfg.mctsTemperatureDecay, static_cast<double>(m_macroPly / 2));
        if (temperature <= 1e-6) return best;
        std::vector<double> weights(edges, 0.0);
        const int vmax = visits[best];
        double total = 0.0;
        for (size_t i = 0; i < edges; ++i) {
            if (visits[i] > 0 && vmax > 0)
                weights[i] = std::exp((std::log(static_cast<double>(visits[i])) - std::log(static_cast<double>(vmax))) / temperature);
            total += weights[i];
        }
        if (total <= 0.0) return 0;
        std::discrete_distribution<size_t> distribution(weights.begin(), weights.end());
        return distribution(m_rng);
    }

    void newGame() {
        m_board = std::make_unique<Board>();
        m_team = (m_rng() & 1ULL) == 0 ? Stockfish::WHITE : Stockfish::BLACK;   // :607-608
        m_timeAdvantage = false;
        m_macroPly = 0;
        m_samples.clear();
        m_resignPlies = {0, 0};
        m_canResign = m_cfg.resignThreshold < 0.0f &&
                      (m_cfg.resignDisableFraction <= 0.0 ||
                       std::uniform_real_distribution<double>(0.0, 1.0)(m_rng) >= m_cfg.resignDisableFraction);
        ++m_gameSerial;
        m_searching = true;
        Stats dummy;
        beginMove(dummy, nullptr);
    }

    // top of the macro-ply loop (:620-630) + search set-up (Agent::run_search, agent.cc:503-553)
    void beginMove(Stats& stats, hm::ChunkWriter* writer) {
        if (m_board->is_checkmate(m_team, m_timeAdvantage)) { endGame(m_team == Stockfish::WHITE ? 1 : 0, stats, writer); return; }
        if (m_board->is_draw() || m_macroPly >= m_cfg.maxMacroPlies) { endGame(-1, stats, writer); return; }
        std::uniform_real_distribution<double> jitter(-m_cfg.nodeRandomFactor, m_cfg.nodeRandomFactor);   // :218-224
        m_target = std::max<size_t>(1, static_cast<size_t>(std::llround(static_cast<double>(m_cfg.nodes) * (1.0 + jitter(m_rng)))));
        m_done = 0;
        m_runtime = SearchParams::RuntimeConfig();
        m_runtime.rootDirichletAlpha = m_cfg.dirichletAlpha;
        m_runtime.rootDirichletEpsilon = m_cfg.dirichletEpsilon;
        m_runtime.rootNoiseSeed = mix_seed(m_cfg.seed ^ m_id, m_gameSerial * m_cfg.maxMacroPlies + m_macroPly);   // :658
        const uint64_t hash = m_board->hash_key(m_timeAdvantage);
        m_root = std::make_shared<Node>(m_team, hash);
        m_info = std::make_unique<SearchInfo>(std::chrono::steady_clock::now(), 0);
        m_table.clear();
        const bool graph = m_runtime.enableMCGS && m_runtime.enableTranspositions;
        if (graph) m_table.insertOrGet(hash, m_root);
        m_tree.set_root_node(m_root);
        m_tree.set_search_info(m_info.get());
        m_tree.set_runtime_config(m_runtime);
        m_tree.set_transposition_table(graph ? &m_table : nullptr);
    }

    void endGame(int winner, Stats& stats, hm::ChunkWriter* writer) {
        if (writer && !m_samples.empty()) {   // :727-735
            for (size_t i = 0; i < m_samples.size(); ++i) {
                hm::TrainingSample& s = m_samples[i];
                s.outcome = winner < 0 ? 0 : (s.team == winner ? 1 : -1);
                s.wdl = static_cast<uint8_t>(s.outcome + 1);
                s.movesLeft = static_cast<uint16_t>(std::min<size_t>(m_samples.size() - i, 65535));
            }
            stats.samples.fetch_add(m_samples.size(), std::memory_order_relaxed);
            writer->append(std::move(m_samples));
            m_samples.clear();
        }
        stats.games.fetch_add(1, std::memory_order_relaxed);
        newGame();
    }

    // marginal_policy after its move -> index step (:400-426): the product's own C entry point
    static std::vector<hm::SparsePolicyEntry> sparsePolicy(const std::vector<int32_t>& index, const std::vector<int64_t>& count) {
        std::vector<uint16_t> idx(index.size());
        std::vector<float> prob(index.size());
        const size_t n = hm_sparse_policy_from_visits(index.data(), count.data(), index.size(), idx.data(), prob.data(), idx.size());
        std::vector<hm::SparsePolicyEntry> out(n);
        for (size_t i = 0; i < n; ++i) out[i] = {idx[i], prob[i]};
        return out;
    }

    // one training sample of the position just searched (:640-678); planes through the reference's own CPU encoder
    // (one position per MOVE, not per node: off the hot path), u8 quantisation of :464-476
    void recordSample(const std::vector<int>& visits, size_t edges, float rootQ) {
        hm::TrainingSample s;
        s.gameId = m_id * 1000003ULL + m_gameSerial;
        s.macroPly = static_cast<uint16_t>(std::min<size_t>(m_macroPly, 65535));
        s.team = m_team == Stockfish::WHITE ? 0 : 1;
        s.hasTimeAdvantage = m_timeAdvantage ? 1 : 0;
        std::array<float, NB_INPUT_VALUES()> raw{};
        board_to_planes(*m_board, raw.data(), m_team, m_timeAdvantage);
        for (size_t i = 0; i < raw.size(); ++i)
            s.planes[i] = static_cast<uint8_t>(std::lround(std::clamp(raw[i], 0.0f, 1.0f) * 255.0f));
        uint64_t total = 0;
        for (size_t i = 0; i < edges; ++i) total += static_cast<uint64_t>(std::max(0, visits[i]));
        s.nodes = static_cast<uint32_t>(std::min<uint64_t>(total, 0xffffffffULL));
        s.rootQ = rootQ;
        for (int bn = 0; bn < 2; ++bn) {   // marginal_policy, :400-426
            std::vector<int32_t> index(edges);
            std::vector<int64_t> count(edges);
            for (size_t i = 0; i < edges; ++i) {
                const JointActionCandidate a = m_root->get_joint_action(static_cast<int>(i));
                index[i] = get_fast_policy_index(bn == 0 ? a.moveA : a.moveB, m_board->side_to_move(bn));
                count[i] = visits[i];
            }
            (bn == 0 ? s.policyA : s.policyB) = sparsePolicy(index, count);
        }
        m_samples.push_back(std::move(s));
    }

    const Config& m_cfg;
    uint64_t m_id;
    std::mt19937_64 m_rng;
    std::unique_ptr<Board> m_board;   // Board is not assignable: a new game gets a new object
    Stockfish::Color m_team = Stockfish::WHITE;
    bool m_timeAdvantage = false, m_searching = false, m_canResign = false;
    size_t m_macroPly = 0, m_target = 0, m_done = 0, m_inFlight = 0;
    uint64_t m_gameSerial = 0;
    std::array<size_t, 2> m_resignPlies = {0, 0};
    std::shared_ptr<Node> m_root;
    std::unique_ptr<SearchInfo> m_info;
    TranspositionTable m_table;
    SearchParams::RuntimeConfig m_runtime;
    SearchThread m_tree;   // the reference's tree operations for this game (never runs an iteration of its own)
    std::array<std::vector<Stockfish::Move>, 2> m_actions;
    std::array<std::vector<float>, 2> m_priors;
    std::array<std::vector<uint8_t>, 2> m_captures;
    std::vector<hm::TrainingSample> m_samples;
};

// A host thread: its games in two halves, each half with its own LeafBatch and execution slot of the evaluator.
// While the device evaluates half h's request, the thread absorbs the other half's results, plays the moves that
// became due and selects that half's next leaves (the double-buffering of SearchThread::run_iteration,
// searchthread.cc:661-708, across games instead of inside one tree).
class Worker {
public:
    Worker(const Config& cfg, Engine& engine, int index, size_t firstGame, size_t nGames, Stats& stats)
        : m_cfg(cfg), m_engine(engine), m_index(index), m_stats(stats), m_batches{hm::LeafBatch(cfg.batch), hm::LeafBatch(cfg.batch)} {
        for (size_t g = 0; g < nGames; ++g) m_games.emplace_back(std::make_unique<Game>(cfg, firstGame + g));
        const size_t perHalf = std::max<size_t>(1, (nGames + 1) / 2);
        m_want = std::clamp<size_t>((cfg.batch + perHalf - 1) / perHalf, 1, 32);
        if (!cfg.chunkDir.empty())
            m_writer = std::make_unique<hm::ChunkWriter>(cfg.chunkDir, cfg.chunkSamples, cfg.seed * 1000 + static_cast<uint64_t>(index));
    }

    void run(const std::atomic<bool>& stop) {
        try {
            bool pending[2] = {false, false};
            while (!stop.load(std::memory_order_relaxed)) {
                for (int h = 0; h < 2; ++h) {
                    if (pending[h]) {
                        finish(h, true);
                        pending[h] = false;
                    }
                    pending[h] = launch(h);
                }
            }
            for (int h = 0; h < 2; ++h)
                if (pending[h]) finish(h, false);
            if (m_writer) m_writer->finish();
        } catch (const std::exception& e) {
            std::cerr << "[ERROR] self-play worker " << m_index << ": " << e.what() << std::endl;
            m_failed = true;
        }
    }
    bool failed() const { return m_failed; }
    double fillSum() const { return m_fillSum; }

private:
    size_t slot(int h) const { return static_cast<size_t>(2 * m_index + h); }

    // select leaves of half h's games and send them off; leaves that need no evaluation are absorbed on the spot
    bool launch(int h) {
        hm::LeafBatch& batch = m_batches[h];
        std::vector<Leaf>& leaves = m_leaves[h];
        for (int round = 0; round < 4; ++round) {
            batch.clear();
            leaves.clear();
            for (size_t g = static_cast<size_t>(h); g < m_games.size() && !batch.full(); g += 2)
                m_games[g]->collect(batch, leaves, static_cast<int>(g), m_want, m_stats);
            if (batch.size() > 0) {
                if (!batch.submit(m_engine, slot(h))) throw std::runtime_error(std::string("submit failed: ") + hm_last_error());
                m_stats.requests.fetch_add(1, std::memory_order_relaxed);
                m_fillSum += static_cast<double>(batch.size());
                return true;
            }
            if (leaves.empty()) return false;
            // only terminal / solved leaves this round: back them up, play due moves, look again
            for (Leaf& leaf : leaves) m_games[leaf.game]->absorb(leaf, batch, m_stats);
            for (size_t g = static_cast<size_t>(h); g < m_games.size(); g += 2) m_games[g]->advance(m_stats, m_writer.get());
        }
        return false;
    }

    void finish(int h, bool playMoves) {
        hm::LeafBatch& batch = m_batches[h];
        Lap lap(m_stats);
        const bool got = batch.collect(m_engine, slot(h));
        lap.to(m_stats.tWait);
        if (!got) throw std::runtime_error(std::string("collect failed: ") + hm_last_error());
        for (Leaf& leaf : m_leaves[h]) m_games[leaf.game]->absorb(leaf, batch, m_stats);
        m_leaves[h].clear();
        lap.t = lap.on ? ticks() : 0;
        if (playMoves)
            for (size_t g = static_cast<size_t>(h); g < m_games.size(); g += 2) m_games[g]->advance(m_stats, m_writer.get());
        lap.to(m_stats.tMove);
    }

    const Config& m_cfg;
    Engine& m_engine;
    int m_index;
    Stats& m_stats;
    std::vector<std::unique_ptr<Game>> m_games;
    std::array<hm::LeafBatch, 2> m_batches;
    std::array<std::vector<Leaf>, 2> m_leaves;
    std::unique_ptr<hm::ChunkWriter> m_writer;
    size_t m_want = 8;
    double m_fillSum = 0.0;
    bool m_failed = false;
};

}  // namespace fused

int main(int argc, char** argv) {
    fused::Config cfg;
    bool profile = false;
    for (int i = 1; i + 1 < argc; i += 2) {
        const std::string k = argv[i], v = argv[i + 1];
        if (k == "--model") cfg.model = v;
        else if (k == "--games") cfg.games = std::stoull(v);
        else if (k == "--nodes") cfg.nodes = std::stoull(v);
        else if (k == "--threads") cfg.threads = std::stoull(v);
        else if (k == "--batch") cfg.batch = std::stoull(v);
        else if (k == "--seconds") cfg.seconds = std::stod(v);
        else if (k == "--max-games") cfg.maxGames = std::stoull(v);
        else if (k == "--max-macro-plies") cfg.maxMacroPlies = std::stoull(v);
        else if (k == "--device") cfg.device = std::stoi(v);
        else if (k == "--seed") cfg.seed = std::stoull(v);
        else if (k == "--chunks") cfg.chunkDir = v;
        else if (k == "--chunk-samples") cfg.chunkSamples = std::stoull(v);
        else if (k == "--profile") profile = v != "0";
        else { std::cerr << "unknown option " << k << std::endl; return 2; }
    }
    if (cfg.model.empty() || cfg.games == 0 || cfg.nodes == 0 || cfg.batch == 0) {
        std::cerr << "usage: hivemind_selfplay_fused --model <net.hvmw|net.onnx> [--games G] [--nodes N] [--threads T] [--batch B]"
                     " [--seconds S] [--max-games N] [--device d] [--seed s] [--chunks dir]" << std::endl;
        return 2;
    }
    if (cfg.threads == 0) cfg.threads = std::max(1u, std::thread::hardware_concurrency());
    cfg.threads = std::min(cfg.threads, cfg.games);
    // engine/src/main.cc:75-81
    Stockfish::pieceMap.init();
    Stockfish::variants.init();
    Stockfish::Bitboards::init();
    Stockfish::Position::init();
    Stockfish::Threads.set(1);
    init_policy_index();

    Engine engine(cfg.device, static_cast<int>(cfg.batch), static_cast<int>(2 * cfg.threads));
    if (!engine.loadNetwork(cfg.model, "")) { std::cerr << "[ERROR] loadNetwork failed" << std::endl; return 1; }
    if (!cfg.chunkDir.empty()) std::filesystem::create_directories(cfg.chunkDir);

    fused::Stats stats;
    stats.profile = profile;
    std::vector<std::unique_ptr<fused::Worker>> workers;
    size_t first = 0;
    for (size_t t = 0; t < cfg.threads; ++t) {
        const size_t n = cfg.games / cfg.threads + (t < cfg.games % cfg.threads ? 1 : 0);
        workers.emplace_back(std::make_unique<fused::Worker>(cfg, engine, static_cast<int>(t), first, n, stats));
        first += n;
    }
    std::atomic<bool> stop{false};
    const auto t0 = std::chrono::steady_clock::now();
    std::vector<std::thread> threads;
    for (auto& w : workers) threads.emplace_back([&stop, &w] { w->run(stop); });
    for (;;) {
        std::this_thread::sleep_for(std::chrono::milliseconds(20));
        const double elapsed = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        if (elapsed >= cfg.seconds || (cfg.maxGames > 0 && stats.games.load() >= cfg.maxGames)) break;
    }
    const uint64_t nodes = stats.nodes.load(), evals = stats.evals.load();
    const double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    stop.store(true);
    for (auto& t : threads) t.join();
    bool failed = false;
    double fill = 0.0;
    for (auto& w : workers) { failed = failed || w->failed(); fill += w->fillSum(); }
    const uint64_t requests = stats.requests.load();
    std::cout << "{\"selfplay\": \"fused\", \"games_concurrent\": " << cfg.games << ", \"nodes_per_move\": " << cfg.nodes
              << ", \"host_threads\": " << cfg.threads << ", \"batch\": " << cfg.batch << ", \"seconds\": " << secs
              << ", \"nodes\": " << nodes << ", \"nodes_per_s\": " << static_cast<double>(nodes) / secs
              << ", \"leaf_evals\": " << evals << ", \"leaf_evals_per_s\": " << static_cast<double>(evals) / secs
              << ", \"terminal_nodes\": " << stats.terminals.load() << ", \"requests\": " << requests
              << ", \"avg_batch_fill\": " << (requests ? fill / static_cast<double>(requests) : 0.0)
              << ", \"collisions\": " << stats.collisions.load() << ", \"moves_played\": " << stats.moves.load()
              << ", \"games_finished\": " << stats.games.load() << ", \"samples_written\": " << stats.samples.load()
              << ", \"failed\": " << (failed ? "true" : "false");
    if (profile) {
        const double all = static_cast<double>(stats.tSelect + stats.tTerminal + stats.tPack + stats.tExpand + stats.tBackup + stats.tWait + stats.tMove);
        auto pct = [&](const std::atomic<uint64_t>& c) { return all > 0 ? 100.0 * static_cast<double>(c.load()) / all : 0.0; };
        std::cout << ", \"host_time_pct\": {\"select_and_expand\": " << pct(stats.tSelect) << ", \"terminal_classification\": " << pct(stats.tTerminal)
                  << ", \"pack_and_movegen\": " << pct(stats.tPack) << ", \"expand_leaf_node\": " << pct(stats.tExpand)
                  << ", \"backup\": " << pct(stats.tBackup) << ", \"wait_for_device\": " << pct(stats.tWait)
                  << ", \"play_moves\": " << pct(stats.tMove) << "}";
    }
    std::cout << "}" << std::endl;
    return failed ? 1 : 0;
}
