// hm_batch.hpp — the caller's side of the fused path: Board -> hm_packed_pos, the move lists the
// reference hands to get_normalized_probability, and a leaf batch that carries both through
// Engine::enqueuePositions / synchronizePriors (data-format half).
//
// Header-only templates over the reference's own types: include it AFTER the reference's
// "environment/board.h" and "common/utils.h" (it names Stockfish::Color / Move / PieceType and calls
// is_policy_move_representable, utils.h:169-182); nothing of the reference is copied here.  What it
// replaces, per leaf, in SearchThread (engine/src/search/searchthread.cc):
//   collect_batch   :326,:431-441   board_to_planes(board, obs + i*4736, team, sitPlaneActive)
//   process_batch   :494-545        actionsA/B building + get_normalized_probability per board
//                   :547-560        capture flags for the joint-action rules
//                   :569-619        value shaping
// and in tools/selfplay.cc:283-299 (prepare_raw_policy) the same list building.
#pragma once

#include <algorithm>
#include <cstdint>
#include <cstring>
#include <vector>

#include "hm_engine.hpp"

namespace hm {

// Every Board accessor planes.cc reads (board.h:239-332) -> 128 bytes.
template <class BoardT>
inline void pack_board(BoardT& b, int boardNum, hm_packed_board& o) {
    using namespace Stockfish;
    std::memset(&o, 0, sizeof(o));
    const PieceType pts[6] = {PAWN, KNIGHT, BISHOP, ROOK, QUEEN, KING};
    for (int c = 0; c < 2; ++c) {
        const Color color = Color(c);
        for (int p = 0; p < 6; ++p) o.pieces[c * 6 + p] = static_cast<uint64_t>(b.pieces(boardNum, color, pts[p]));
        o.promoted[c] = static_cast<uint64_t>(b.promoted_pieces(boardNum) & b.pieces(boardNum, color));
        for (int p = 0; p < 5; ++p)
            o.pocket[c * 5 + p] = static_cast<uint8_t>(std::clamp(static_cast<int>(b.count_in_hand(boardNum, color, pts[p])), 0, 255));
    }
    const Square ep = b.ep_square(boardNum);
    o.ep_square = ep == SQ_NONE ? HM_SQ_NONE : static_cast<uint8_t>(ep);
    uint8_t f = 0;
    if (b.side_to_move(boardNum) == BLACK) f |= HM_FLAG_STM_BLACK;
    if (b.can_castle(boardNum, WHITE_OO)) f |= HM_FLAG_WHITE_OO;
    if (b.can_castle(boardNum, WHITE_OOO)) f |= HM_FLAG_WHITE_OOO;
    if (b.can_castle(boardNum, BLACK_OO)) f |= HM_FLAG_BLACK_OO;
    if (b.can_castle(boardNum, BLACK_OOO)) f |= HM_FLAG_BLACK_OOO;
    o.flags = f;
    const Move last = b.last_move(boardNum);
    if (last == MOVE_NONE) {
        o.last_from = o.last_to = HM_SQ_NONE;
    } else {
        o.last_from = type_of(last) == DROP ? HM_LAST_FROM_DROP : static_cast<uint8_t>(from_sq(last));
        o.last_to = static_cast<uint8_t>(to_sq(last));
    }
    o.rule50 = static_cast<uint8_t>(std::clamp(static_cast<int>(b.rule50_count(boardNum)), 0, 255));
    o.repetitions = static_cast<uint8_t>(std::clamp(static_cast<int>(b.repetition_count(boardNum)), 0, 255));
}

// board_to_planes(board, out, teamSide, hasTimeAdvantage)'s arguments (planes.cc:259-265) as a record
template <class BoardT>
inline void pack_position(BoardT& b, Stockfish::Color teamSide, bool hasTimeAdvantage, hm_packed_pos& o) {
    pack_board(b, 0, o.board[0]);
    pack_board(b, 1, o.board[1]);
    if (teamSide == Stockfish::BLACK) o.board[0].flags |= HM_FLAG_TEAM_BLACK;
    if (hasTimeAdvantage) o.board[0].flags |= HM_FLAG_TIME_ADVANTAGE;
}

// searchthread.cc:494-545 / selfplay.cc:283-299: legal moves if the board is on turn for the team
// (board A: team, board B: ~team), R/B under-promotions erased, MOVE_NONE appended.  Appends the
// Move bit patterns to `moves` (and, if wanted, the capture flags of searchthread.cc:547-560 to
// `captures`); returns the list length.
template <class BoardT>
inline size_t append_actions(BoardT& b, int boardNum, Stockfish::Color team, std::vector<uint32_t>& moves,
                             std::vector<uint8_t>* captures = nullptr) {
    const Stockfish::Color expected = boardNum == 0 ? team : ~team;
    size_t n = 0;
    if (b.side_to_move(boardNum) == expected) {
        for (Stockfish::Move move : b.legal_moves(boardNum)) {
            if (!is_policy_move_representable(b, boardNum, move)) continue;
            moves.push_back(static_cast<uint32_t>(move));
            if (captures) captures->push_back(b.is_capture(boardNum, move) ? 1 : 0);
            ++n;
        }
    }
    moves.push_back(static_cast<uint32_t>(Stockfish::MOVE_NONE));
    if (captures) captures->push_back(0);
    return n + 1;
}

// One request worth of leaves: fill with add(), submit(), collect(), read the per-leaf views.
// Not thread-safe; one LeafBatch per worker slot, like the reference's per-thread batch vectors
// (searchthread.cc:153-174).
class LeafBatch {
public:
    explicit LeafBatch(size_t capacity) : m_capacity(capacity) {
        m_packed.reserve(capacity);
        m_offsets.reserve(2 * capacity + 1);
        m_offsets.push_back(0);
        m_moves.reserve(capacity * 80);
        m_captures.reserve(capacity * 80);
    }
    void clear() {
        m_packed.clear();
        m_moves.clear();
        m_captures.clear();
        m_offsets.assign(1, 0u);
        m_out = Engine::RootOutputs();
    }
    size_t size() const { return m_packed.size(); }
    bool full() const { return m_packed.size() >= m_capacity; }

    // Adds the leaf `board` seen by `team`; returns its index in the batch or -1 when full.
    template <class BoardT>
    int add(BoardT& board, Stockfish::Color team, bool sitPlaneActive) {
        if (full()) return -1;
        m_packed.emplace_back();
        pack_position(board, team, sitPlaneActive, m_packed.back());
        append_actions(board, 0, team, m_moves, &m_captures);
        m_offsets.push_back(static_cast<uint32_t>(m_moves.size()));
        append_actions(board, 1, team, m_moves, &m_captures);
        m_offsets.push_back(static_cast<uint32_t>(m_moves.size()));
        return static_cast<int>(m_packed.size()) - 1;
    }

    // enqueueInferenceHalf's place in run_iteration (searchthread.cc:672-675)
    bool submit(Engine& engine, size_t workerIndex) {
        m_root = false;
        return engine.enqueuePositions(m_packed.data(), m_packed.size(), m_moves.data(), m_offsets.data(), workerIndex);
    }
    // the same with the root stages (temperature / Dirichlet noise / joint sampling); captures are the batch's own
    bool submitRoot(Engine& engine, size_t workerIndex, hm_root_policy root) {
        m_root = true;
        root.captures = m_captures.data();
        return engine.enqueuePositionsRoot(m_packed.data(), m_packed.size(), m_moves.data(), m_offsets.data(), root,
                                           workerIndex);
    }
    // synchronizeInferenceHalf's place (searchthread.cc:688-693)
    bool collect(Engine& engine, size_t workerIndex) {
        return m_root ? engine.synchronizeRoot(m_out, workerIndex) : engine.synchronizePriors(m_out, workerIndex);
    }

    // views, valid until the worker's next submit
    size_t numActions(size_t leaf, int boardNum) const { return m_offsets[2 * leaf + boardNum + 1] - m_offsets[2 * leaf + boardNum]; }
    const uint32_t* actions(size_t leaf, int boardNum) const { return m_moves.data() + m_offsets[2 * leaf + boardNum]; }
    const uint8_t* captures(size_t leaf, int boardNum) const { return m_captures.data() + m_offsets[2 * leaf + boardNum]; }
    const float* priors(size_t leaf, int boardNum) const { return m_out.priors + m_offsets[2 * leaf + boardNum]; }
    float value(size_t leaf) const { return m_out.values[leaf]; }        // shaped, searchthread.cc:569-619
    const float* raw(size_t leaf) const { return m_out.raw + 5 * leaf; }  // value, wdl[3], moves_left
    uint32_t pick(size_t leaf, int boardNum) const { return m_out.picks[2 * leaf + boardNum]; }
    int32_t status(size_t leaf) const { return m_out.status[leaf]; }

    const std::vector<hm_packed_pos>& packed() const { return m_packed; }
    const std::vector<uint32_t>& moves() const { return m_moves; }
    const std::vector<uint32_t>& offsets() const { return m_offsets; }
    const std::vector<uint8_t>& captureFlags() const { return m_captures; }

private:
    size_t m_capacity;
    bool m_root = false;
    std::vector<hm_packed_pos> m_packed;
    std::vector<uint32_t> m_moves, m_offsets;
    std::vector<uint8_t> m_captures;
    Engine::RootOutputs m_out;
};

}  // namespace hm
