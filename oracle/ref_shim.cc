// ref_shim.cc — TEST INFRASTRUCTURE, not product code.
//
// A thin C API over the UNMODIFIED reference sources, compiled where they lie under
// /root/reference (see oracle/Makefile; nothing is copied into this repo).  The resulting
// oracle/_ref/libhmref.so is the reference itself for stages P (planes.cc) and D
// (utils.h / globals.cc): it pins the CPU restatement in oracle/hm_oracle.c, generates the
// golden fixtures under tests/golden/, and is the "reference" CPU baseline in bench.py.
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
// may load it.
//
// Reference entry points used:
//   board_to_planes                      engine/src/environment/planes.cc:259-265
//   Board (accessors, push/make moves)   engine/src/environment/board.h, board.cc
//   get_fast_policy_index                engine/src/common/utils.h:184-216
//   get_normalized_probability           engine/src/common/utils.h:226-243
//   normalize_logits                     engine/src/common/utils.h:127-167
//   init_policy_index / tables           engine/src/common/globals.cc:82-136
//   init sequence                        engine/src/main.cc:75-81
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <memory>
#include <random>
#include <string>
#include <thread>
#include <vector>

#include "common/globals.h"
#include "common/utils.h"
#include "environment/board.h"
#include "environment/constants.h"
#include "environment/joint_action.h"
#include "environment/planes.h"
#include "Fairy-Stockfish/src/bitboard.h"
#include "Fairy-Stockfish/src/piece.h"
#include "Fairy-Stockfish/src/position.h"
#include "Fairy-Stockfish/src/thread.h"
#include "Fairy-Stockfish/src/types.h"

#include "../include/hm_engine.h"
#include "../include/hm_batch.hpp"  // PRODUCT header under test (templates over the reference Board)

namespace {

bool g_inited = false;

Stockfish::Move find_move(Board& board, int boardNum, const std::string& uci) {
    if (uci == "pass" || uci.empty()) return Stockfish::MOVE_NONE;
    for (Stockfish::Move move : board.legal_moves(boardNum)) {
        if (board.uci_move(boardNum, move) == uci) return move;
    }
    return Stockfish::MOVE_NULL;  // sentinel: not found
}

inline Board* B(void* h) { return static_cast<Board*>(h); }

// The action list exactly as SearchThread::process_batch builds it
// (searchthread.cc:494-545): legal moves only if the board is on turn for the team,
// R/B under-promotions erased, MOVE_NONE appended.
std::vector<Stockfish::Move> search_actions(Board& b, int boardNum, Stockfish::Color team) {
    std::vector<Stockfish::Move> actions;
    const Stockfish::Color expected = (boardNum == 0) ? team : ~team;
    if (b.side_to_move(boardNum) == expected) {
        actions = b.legal_moves(boardNum);
        std::erase_if(actions, [&](Stockfish::Move m) {
            return !is_policy_move_representable(b, boardNum, m);
        });
    }
    actions.push_back(Stockfish::MOVE_NONE);
    return actions;
}

void pack_board(Board& b, int bn, hm_packed_board* o) {
    using namespace Stockfish;
    std::memset(o, 0, sizeof(*o));
    const PieceType pts[6] = {PAWN, KNIGHT, BISHOP, ROOK, QUEEN, KING};
    for (int c = 0; c < 2; ++c) {
        for (int p = 0; p < 6; ++p)
            o->pieces[c * 6 + p] = (uint64_t)b.pieces(bn, Color(c), pts[p]);
        o->promoted[c] = (uint64_t)(b.promoted_pieces(bn) & b.pieces(bn, Color(c)));
        for (int p = 0; p < 5; ++p)
            o->pocket[c * 5 + p] = (uint8_t)std::min(255, b.count_in_hand(bn, Color(c), pts[p]));
    }
    Square ep = b.ep_square(bn);
    o->ep_square = (ep == SQ_NONE) ? HM_SQ_NONE : (uint8_t)ep;
    uint8_t f = 0;
    if (b.side_to_move(bn) == BLACK) f |= HM_FLAG_STM_BLACK;
    if (b.can_castle(bn, WHITE_OO)) f |= HM_FLAG_WHITE_OO;
    if (b.can_castle(bn, WHITE_OOO)) f |= HM_FLAG_WHITE_OOO;
    if (b.can_castle(bn, BLACK_OO)) f |= HM_FLAG_BLACK_OO;
    if (b.can_castle(bn, BLACK_OOO)) f |= HM_FLAG_BLACK_OOO;
    o->flags = f;
    Move lm = b.last_move(bn);
    if (lm == MOVE_NONE) {
        o->last_from = HM_SQ_NONE;
        o->last_to = HM_SQ_NONE;
    } else {
        o->last_from = (type_of(lm) == DROP) ? HM_LAST_FROM_DROP : (uint8_t)from_sq(lm);
        o->last_to = (uint8_t)to_sq(lm);
    }
    o->rule50 = (uint8_t)std::clamp(b.rule50_count(bn), 0, 255);
    o->repetitions = (uint8_t)std::clamp(b.repetition_count(bn), 0, 255);
}

}  // namespace

extern "C" {

int hmref_init(void) {
    if (g_inited) return 1;
    Stockfish::pieceMap.init();
    Stockfish::variants.init();
    Stockfish::Bitboards::init();
    Stockfish::Position::init();
    Stockfish::Threads.set(1);
    init_policy_index();
    g_inited = true;
    return 1;
}

void* hmref_board_new(void) { return new Board(); }
void* hmref_board_copy(void* h) { return new Board(*B(h)); }
void hmref_board_free(void* h) { delete B(h); }

int hmref_board_set(void* h, const char* fen_pair) {
    try { B(h)->set(fen_pair); } catch (...) { return 0; }
    return 1;
}
int hmref_board_set_fen(void* h, int bn, const char* fen) {
    try { B(h)->set_fen(bn, fen); } catch (...) { return 0; }
    return 1;
}
int hmref_board_fen(void* h, int bn, char* buf, int cap) {
    std::string s = B(h)->fen(bn);
    if ((int)s.size() + 1 > cap) return 0;
    std::memcpy(buf, s.c_str(), s.size() + 1);
    return 1;
}
// Board::push_move after a UCI lookup, as the reference tests do (test_move_gen.cc:47-54).
int hmref_board_push_uci(void* h, int bn, const char* uci) {
    Stockfish::Move m = find_move(*B(h), bn, uci);
    if (m == Stockfish::MOVE_NULL || m == Stockfish::MOVE_NONE) return 0;
    B(h)->push_move(bn, m);
    return 1;
}
int hmref_board_pop(void* h, int bn) { B(h)->pop_move(bn); return 1; }
// Board::make_moves / unmake_moves (board.cc:316-359) with UCI strings ("pass" = MOVE_NONE).
int hmref_board_make_moves_uci(void* h, const char* a, const char* b, uint32_t* mA, uint32_t* mB) {
    Stockfish::Move ma = find_move(*B(h), 0, a), mb = find_move(*B(h), 1, b);
    if (ma == Stockfish::MOVE_NULL || mb == Stockfish::MOVE_NULL) return 0;
    try { B(h)->make_moves(ma, mb); } catch (...) { return 0; }
    if (mA) *mA = (uint32_t)ma;
    if (mB) *mB = (uint32_t)mb;
    return 1;
}
int hmref_board_unmake_moves(void* h, uint32_t mA, uint32_t mB) {
    B(h)->unmake_moves(Stockfish::Move(mA), Stockfish::Move(mB));
    return 1;
}
int hmref_board_stm(void* h, int bn) { return B(h)->side_to_move(bn) == Stockfish::BLACK; }
int hmref_board_gives_check(void* h, int bn, uint32_t m) {
    return B(h)->gives_check(bn, Stockfish::Move(m));
}

void hmref_board_planes_f32(void* h, int team_black, int tadv, float* out) {
    board_to_planes(*B(h), out, team_black ? Stockfish::BLACK : Stockfish::WHITE, tadv != 0);
}
void hmref_board_planes_f16(void* h, int team_black, int tadv, uint16_t* out) {
    board_to_planes(*B(h), reinterpret_cast<__half*>(out),
                    team_black ? Stockfish::BLACK : Stockfish::WHITE, tadv != 0);
}
// HVM4 u8 plane quantisation exactly as tools/selfplay.cc:464-476 applies it to the
// reference float planes (that helper is file-static there, so the arithmetic is restated
// here on top of the reference board_to_planes).
void hmref_board_planes_u8(void* h, int team_black, int tadv, uint8_t* out) {
    std::vector<float> raw(NB_INPUT_VALUES());
    board_to_planes(*B(h), raw.data(), team_black ? Stockfish::BLACK : Stockfish::WHITE, tadv != 0);
    for (size_t i = 0; i < raw.size(); ++i)
        out[i] = static_cast<uint8_t>(std::lround(std::clamp(raw[i], 0.0f, 1.0f) * 255.0f));
}

void hmref_board_pack(void* h, int team_black, int tadv, hm_packed_pos* out) {
    pack_board(*B(h), 0, &out->board[0]);
    pack_board(*B(h), 1, &out->board[1]);
    if (team_black) out->board[0].flags |= HM_FLAG_TEAM_BLACK;
    if (tadv) out->board[0].flags |= HM_FLAG_TIME_ADVANTAGE;
}

int hmref_board_legal(void* h, int bn, uint32_t* out, int cap) {
    auto v = B(h)->legal_moves(bn);
    if ((int)v.size() > cap) return -1;
    for (size_t i = 0; i < v.size(); ++i) out[i] = (uint32_t)v[i];
    return (int)v.size();
}
int hmref_board_search_actions(void* h, int bn, int team_black, uint32_t* out, int cap) {
    auto v = search_actions(*B(h), bn, team_black ? Stockfish::BLACK : Stockfish::WHITE);
    if ((int)v.size() > cap) return -1;
    for (size_t i = 0; i < v.size(); ++i) out[i] = (uint32_t)v[i];
    return (int)v.size();
}
int hmref_board_move_uci(void* h, int bn, uint32_t m, char* buf, int cap) {
    std::string s = B(h)->uci_move(bn, Stockfish::Move(m));
    if ((int)s.size() + 1 > cap) return 0;
    std::memcpy(buf, s.c_str(), s.size() + 1);
    return 1;
}

void hmref_board_priors_f16(void* h, int bn, const uint16_t* policy, const uint32_t* moves,
                            int n, float* out) {
    std::vector<Stockfish::Move> a(n);
    for (int i = 0; i < n; ++i) a[i] = Stockfish::Move(moves[i]);
    auto p = get_normalized_probability(reinterpret_cast<const __half*>(policy), a, bn, *B(h));
    std::copy(p.begin(), p.end(), out);
}
void hmref_board_priors_f32(void* h, int bn, const float* policy, const uint32_t* moves, int n,
                            float* out) {
    std::vector<Stockfish::Move> a(n);
    for (int i = 0; i < n; ++i) a[i] = Stockfish::Move(moves[i]);
    auto p = get_normalized_probability(policy, a, bn, *B(h));
    std::copy(p.begin(), p.end(), out);
}
void hmref_normalize_logits(const float* logits, int n, float* out) {
    std::vector<float> l(logits, logits + n);
    auto p = normalize_logits(l);
    std::copy(p.begin(), p.end(), out);
}
int hmref_policy_index(uint32_t move, int stm_black) {
    return get_fast_policy_index(Stockfish::Move(move),
                                 stm_black ? Stockfish::BLACK : Stockfish::WHITE);
}
int hmref_policy_index_of_label(const char* label) {
    auto it = POLICY_INDEX.find(label);
    return it == POLICY_INDEX.end() ? -1 : it->second;
}
int hmref_policy_label(int idx, char* buf, int cap) {
    if (idx < 0 || idx >= (int)NB_POLICY_VALUES()) return 0;
    const std::string& s = UCI_MOVES[idx];
    if ((int)s.size() + 1 > cap) return 0;
    std::memcpy(buf, s.c_str(), s.size() + 1);
    return 1;
}
void hmref_policy_tables(int32_t* normal, int32_t* drop) {
    std::memcpy(normal, POLICY_TABLE_NORMAL, sizeof(POLICY_TABLE_NORMAL));
    std::memcpy(drop, POLICY_TABLE_DROP, sizeof(POLICY_TABLE_DROP));
}

// Config-1 generator: seeded random playouts.  boards_out receives n
// Board handles; team[g] = g&1, tadv[g] = (g>>1)&1.
int hmref_gen_positions(uint64_t seed, int n, void** boards_out) {
    std::mt19937_64 rng(seed);
    for (int g = 0; g < n; ++g) {
        Board* b = new Board();
        int plies = 8 + (int)(rng() % 60);
        for (int p = 0; p < plies; ++p) {
            int bn = (int)(rng() & 1);
            auto moves = b->legal_moves(bn);
            if (moves.empty()) break;
            b->push_move(bn, moves[rng() % moves.size()]);
        }
        boards_out[g] = b;
    }
    return n;
}

// ---- CPU baseline timers: the reference's own code on `threads` host threads ----------
// Each thread owns a disjoint slice of the boards (Board is not shared across threads).
// Returns seconds for `reps` passes over all n boards.
double hmref_time_planes_f16(void** boards, int n, int reps, int threads) {
    threads = std::max(1, threads);
    auto t0 = std::chrono::steady_clock::now();
    std::vector<std::thread> pool;
    for (int t = 0; t < threads; ++t) {
        pool.emplace_back([=]() {
            std::vector<__half> buf(NB_INPUT_VALUES());
            for (int r = 0; r < reps; ++r)
                for (int i = t; i < n; i += threads)
                    board_to_planes(*B(boards[i]), buf.data(),
                                    (i & 1) ? Stockfish::BLACK : Stockfish::WHITE, ((i >> 1) & 1) != 0);
        });
    }
    for (auto& th : pool) th.join();
    return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}

// legal_moves + filter + get_normalized_probability<__half> for both boards
// (searchthread.cc:494-545), `policy` = [n][2][4672] fp16 logits.
double hmref_time_priors_f16(void** boards, int n, const uint16_t* policy, int reps, int threads) {
    threads = std::max(1, threads);
    std::atomic<double> sink{0.0};
    auto t0 = std::chrono::steady_clock::now();
    std::vector<std::thread> pool;
    for (int t = 0; t < threads; ++t) {
        pool.emplace_back([=, &sink]() {
            double acc = 0.0;
            for (int r = 0; r < reps; ++r)
                for (int i = t; i < n; i += threads) {
                    Board& b = *B(boards[i]);
                    const Stockfish::Color team = (i & 1) ? Stockfish::BLACK : Stockfish::WHITE;
                    for (int bn = 0; bn < 2; ++bn) {
                        auto a = search_actions(b, bn, team);
                        auto p = get_normalized_probability(
                            reinterpret_cast<const __half*>(policy) + ((size_t)i * 2 + bn) * NB_POLICY_VALUES(),
                            a, bn, b);
                        acc += p[0];
                    }
                }
            sink.store(sink.load() + acc);
        });
    }
    for (auto& th : pool) th.join();
    return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}

// ---- root policy stages ------------------------------------------------
int hmref_board_is_capture(void* h, int bn, uint32_t m) {
    return m != 0u && B(h)->is_capture(bn, Stockfish::Move(m)) ? 1 : 0;
}

// std::generate_canonical<double,53> draws of std::mt19937_64(seed): what std::discrete_distribution
// consumes in sample_raw_policy_action (selfplay.cc:363-365,389-390).
void hmref_canonical(uint64_t seed, int n, double* out) {
    std::mt19937_64 e(seed);
    for (int i = 0; i < n; ++i) out[i] = std::generate_canonical<double, 53>(e);
}

// tools/selfplay.cc:167-185 apply_temperature lives in an anonymous namespace of a TU that includes
// nn/engine.h (NvInfer.h, absent): restated here statement by statement with the same std:: calls.
int hmref_apply_temperature(float* p, int n, double temperature) {
    if (n <= 0 || temperature <= 0.0) return 0;
    const double exponent = 1.0 / temperature;
    double total = 0.0;
    for (int i = 0; i < n; ++i) {
        p[i] = static_cast<float>(std::pow(std::max(0.0f, p[i]), exponent));
        total += p[i];
    }
    if (!std::isfinite(total) || total <= 0.0) {
        std::fill(p, p + n, 1.0f / static_cast<float>(n));
        return 1;
    }
    for (int i = 0; i < n; ++i) p[i] = static_cast<float>(p[i] / total);
    return 1;
}

// The sampling half of sample_raw_policy_action (selfplay.cc:349-392) run on the reference's OWN
// JointActionCandidate / JointActionRules (environment/joint_action.h) with the real
// std::discrete_distribution and a std::mt19937_64(seed).  Returns 0 ok, 1 resampled, -3 no legal
// joint action; *draws = how many canonical doubles the engine handed out.
int hmref_sample_joint(const uint32_t* mA, const float* pA, const uint8_t* cA, int nA, const uint32_t* mB,
                       const float* pB, const uint8_t* cB, int nB, int aOnTurn, int bOnTurn, int tadv,
                       uint64_t seed, uint32_t* iA_out, uint32_t* iB_out, int* draws) {
    std::mt19937_64 randomEngine(seed);
    std::vector<Stockfish::Move> actionsA(nA), actionsB(nB);
    for (int i = 0; i < nA; ++i) actionsA[i] = Stockfish::Move(mA[i]);
    for (int i = 0; i < nB; ++i) actionsB[i] = Stockfish::Move(mB[i]);
    std::vector<float> probabilitiesA(pA, pA + nA), probabilitiesB(pB, pB + nB);
    JointActionRules rules;
    rules.boardAOnTurn = aOnTurn != 0;
    rules.boardBOnTurn = bOnTurn != 0;
    rules.teamHasTimeAdvantage = tadv != 0;
    rules.boardACanMove = rules.boardAOnTurn && actionsA.size() > 1;
    rules.boardBCanMove = rules.boardBOnTurn && actionsB.size() > 1;
    auto make_candidate = [&](size_t iA, size_t iB) {
        return JointActionCandidate(actionsA[iA], probabilitiesA[iA], iA, actionsB[iB], probabilitiesB[iB], iB, rules,
                                    cA[iA] != 0, cB[iB] != 0);
    };
    int used = (nA >= 2) + (nB >= 2);
    std::discrete_distribution<size_t> sampleA(probabilitiesA.begin(), probabilitiesA.end());
    std::discrete_distribution<size_t> sampleB(probabilitiesB.begin(), probabilitiesB.end());
    // the reference's literal expression (selfplay.cc:365): the two draws are function arguments, so
    // their order is the compiler's choice (g++ evaluates right to left: B draws first)
    JointActionCandidate candidate = make_candidate(sampleA(randomEngine), sampleB(randomEngine));
    int rc = 0;
    if (candidate.jointPrior < 0.0f) {
        std::vector<std::pair<size_t, size_t>> legalPairs;
        std::vector<double> weights;
        for (size_t iA = 0; iA < actionsA.size(); ++iA)
            for (size_t iB = 0; iB < actionsB.size(); ++iB) {
                if (make_candidate(iA, iB).jointPrior < 0.0f) continue;
                legalPairs.emplace_back(iA, iB);
                weights.push_back(static_cast<double>(probabilitiesA[iA]) * static_cast<double>(probabilitiesB[iB]));
            }
        if (std::accumulate(weights.begin(), weights.end(), 0.0) <= 0.0) {
            *draws = used;
            return -3;
        }
        std::discrete_distribution<size_t> legalSample(weights.begin(), weights.end());
        used += weights.size() >= 2;
        const auto [indexA, indexB] = legalPairs[legalSample(randomEngine)];
        candidate = make_candidate(indexA, indexB);
        rc = 1;
    }
    *iA_out = static_cast<uint32_t>(candidate.idxA);
    *iB_out = static_cast<uint32_t>(candidate.idxB);
    *draws = used;
    return rc;
}

// node.h:291-312 applyNoise with the real std::gamma_distribution<float> on std::mt19937_64(key):
// the reference's noise stream (ours is counter-based: compared in distribution only).
void hmref_dirichlet_noise(uint64_t key, float alpha, int n, float* noise_out) {
    std::mt19937_64 randomEngine(key);
    std::gamma_distribution<float> gamma(alpha, 1.0f);
    for (int i = 0; i < n; ++i) noise_out[i] = gamma(randomEngine);
}
// the mix of node.h:299-311 compiled with the reference's flags
void hmref_dirichlet_mix(float* priors, const float* noise, int n, float epsilonIn) {
    if (n <= 1) return;
    float total = 0.0f;
    for (int i = 0; i < n; ++i) total += noise[i];
    if (total <= 0.0f) return;
    const float epsilon = std::clamp(epsilonIn, 0.0f, 1.0f);
    for (int index = 0; index < n; ++index)
        priors[index] = (1.0f - epsilon) * priors[index] + epsilon * noise[index] / total;
}

// ---- the product's caller-side header (include/hm_batch.hpp) instantiated on the reference Board ----
void hmref_product_pack(void* h, int team_black, int tadv, hm_packed_pos* out) {
    hm::pack_position(*B(h), team_black ? Stockfish::BLACK : Stockfish::WHITE, tadv != 0, *out);
}
int hmref_product_actions(void* h, int bn, int team_black, uint32_t* moves_out, uint8_t* caps_out, int cap) {
    std::vector<uint32_t> m;
    std::vector<uint8_t> c;
    const size_t n = hm::append_actions(*B(h), bn, team_black ? Stockfish::BLACK : Stockfish::WHITE, m, &c);
    if ((int)n > cap) return -1;
    std::copy(m.begin(), m.end(), moves_out);
    std::copy(c.begin(), c.end(), caps_out);
    return (int)n;
}
int hmref_hardware_threads(void) { return (int)std::thread::hardware_concurrency(); }

}  // extern "C"
