// d5_pin.cc — TEST INFRASTRUCTURE, not product code.
//
// Pins SURVEY.md §8a row D5 (value shaping) on the reference's OWN compiled code: the unmodified
// engine/src/search/searchthread.cc and node.cc are compiled against the scripted `Engine` of
// oracle/scripted/nn/engine.h, a SearchThread runs ONE iteration with batch size 1 on a fresh root (the root is
// the only leaf), SearchThread::process_batch (searchthread.cc:444-639) shapes the scripted
// value / wdl / moves_left exactly as in a real search (:569-619) and backs it up (:621, backup :197-239:
// a one-entry trajectory -> Node::update_terminal), and the shaped value is read back as root->Q()
// (node.h:433-447: valueSum / visits with visits == 1).
//
// cudaMallocHost / cudaFreeHost (searchthread.cc:153-175 allocates its pinned plane buffer with them) are
// satisfied by plain malloc here so the fixture can be generated in the GPU-less build container; nothing else
// of the CUDA runtime is referenced.
#include <chrono>
#include <cstdint>
#include <cstdlib>
#include <memory>

#include "environment/board.h"
#include "search/node.h"
#include "search/searchinfo.h"
#include "search/searchthread.h"

extern "C" {
// host-allocation stand-ins (see header comment)
cudaError_t cudaMallocHost(void** p, size_t n) { *p = std::malloc(n); return *p ? cudaSuccess : cudaErrorMemoryAllocation; }
cudaError_t cudaFreeHost(void* p) { std::free(p); return cudaSuccess; }
const char* cudaGetErrorString(cudaError_t) { return "cuda stub"; }

int hmref_init(void);  // oracle/ref_shim.cc (engine/src/main.cc:75-81 init sequence)

// cases: n x 5 fp16 bit patterns (value, wdl loss, wdl draw, wdl win, moves_left); out: n shaped values.
// Returns the number of cases whose root did not end with exactly one visit (0 = all good).
int hmref_d5_shape(const uint16_t* cases, int n, int enable_wdl, float wdl_weight, float draw_contempt,
                   float moves_left_discount, float* out) {
    hmref_init();
    int bad = 0;
    try {
    SearchParams::RuntimeConfig cfg;
    cfg.enableWdlEval = enable_wdl != 0;
    cfg.wdlValueWeight = wdl_weight;
    cfg.drawContempt = draw_contempt;
    cfg.movesLeftDiscount = moves_left_discount;
    Engine engine(0, 1);
    for (int i = 0; i < n; ++i) {
        const uint16_t* c = cases + 5 * i;
        engine.script(c[0], c + 1, c[4]);
        Board board;
        auto root = std::make_shared<Node>(Stockfish::WHITE, board.hash_key(false));
        SearchInfo info(std::chrono::steady_clock::now(), 0);
        SearchThread st;
        st.set_root_node(root);
        st.set_search_info(&info);
        st.set_runtime_config(cfg);
        st.set_inference_worker_index(0);
        st.set_transposition_table(nullptr);
        st.run_iteration(board, &engine, false);
        st.discard_pending_iteration(board, &engine);
        if (root->get_visits() != 1) ++bad;
        out[i] = root->Q();
    }
    } catch (...) {
        return -1;   // never let a C++ exception cross the C boundary
    }
    return bad;
}
}
