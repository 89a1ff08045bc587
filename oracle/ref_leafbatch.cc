// ref_leafbatch.cc — TEST INFRASTRUCTURE, not product code.
//
// The one harness that needs BOTH the reference (Board, move generation: oracle/_ref/libhmref.so) and the product
// (include/hm_batch.hpp LeafBatch over include/hm_engine.hpp -> libhm_engine.so).  Kept out of libhmref.so so that
// the pure-reference library the CPU arm of bench.py loads does not map the product library.
#include <algorithm>
#include <cstdint>
#include <vector>

#include "common/utils.h"
#include "environment/board.h"
#include "environment/joint_action.h"
#include "Fairy-Stockfish/src/types.h"

#include "../include/hm_engine.h"
#include "../include/hm_batch.hpp"

namespace {
inline Board* B(void* h) { return static_cast<Board*>(h); }
}

extern "C" {

// One LeafBatch round trip through the product Engine (include/hm_engine.hpp -> libhm_engine.so) for n
// boards with the config-1 convention team = g&1, timeAdvantage = (g>>1)&1.  Needs a CUDA device.
// root != NULL runs the root stages with the batch's own capture flags.  Outputs are sized by the caller
// (moves/priors/captures: cap_moves; offsets: 2n+1; values: n; picks: 2n; status: n).  Returns the number of
// moves, or a negative error.
int hmref_leafbatch_run(int device, const void* blob, size_t blob_bytes, void** boards, int n, const hm_root_policy* root,
                        uint32_t* moves_out, uint8_t* caps_out, uint32_t* offsets_out, float* priors_out, float* values_out,
                        uint32_t* picks_out, int32_t* status_out, int cap_moves) {
    Engine engine(device, n, 1);
    if (!engine.loadNetworkFromMemory(blob, blob_bytes)) return -1;
    hm::LeafBatch batch(n);
    for (int g = 0; g < n; ++g)
        if (batch.add(*B(boards[g]), (g & 1) ? Stockfish::BLACK : Stockfish::WHITE, ((g >> 1) & 1) != 0) != g) return -2;
    if (batch.add(*B(boards[0]), Stockfish::WHITE, false) != -1) return -3;  // full
    const int total = (int)batch.moves().size();
    if (total > cap_moves) return -4;
    if (!(root ? batch.submitRoot(engine, 0, *root) : batch.submit(engine, 0))) return -5;
    if (!batch.collect(engine, 0)) return -6;
    std::copy(batch.moves().begin(), batch.moves().end(), moves_out);
    std::copy(batch.captureFlags().begin(), batch.captureFlags().end(), caps_out);
    std::copy(batch.offsets().begin(), batch.offsets().end(), offsets_out);
    for (int g = 0; g < n; ++g) {
        values_out[g] = batch.value(g);
        for (int bn = 0; bn < 2; ++bn) {
            const size_t k = batch.numActions(g, bn), lo = batch.offsets()[2 * g + bn];
            std::copy(batch.priors(g, bn), batch.priors(g, bn) + k, priors_out + lo);
            if (root && root->uniforms) picks_out[2 * g + bn] = batch.pick(g, bn);
        }
        if (root) status_out[g] = batch.status(g);
    }
    return total;
}

}  // extern "C"
