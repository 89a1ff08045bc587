// integration/nn/engine.h — the ONE file a maintainer of the reference replaces:
// engine/src/nn/engine.h becomes this forwarder, engine/src/nn/engine.cc leaves the build, and
// search/, tools/, interface/ and main.cc compile unchanged against the B200 evaluator.
// `make -C oracle dropin` proves it: it compiles the UNMODIFIED reference sources where they lie
// with this directory ahead of engine/src on the include path.
#pragma once
#include "search/search_params.h"
#ifndef HM_DEFAULT_BATCH_SIZE
#define HM_DEFAULT_BATCH_SIZE SearchParams::BATCH_SIZE
#endif
#ifndef HM_DEFAULT_WORKERS
#define HM_DEFAULT_WORKERS SearchParams::NUM_SEARCH_THREADS
#endif
#ifndef HM_BATCHING_ENGINE
#include "hm_engine.hpp"  // class Engine, Engine::HalfInferenceOutputs
#else
// Many concurrent games on one evaluator: `Engine` is the small-batch facade that coalesces the requests of
// all engines of a device into large device batches (include/hm_batching_engine.hpp).  The reference's
// sources still compile unchanged; integration/multigame_selfplay.cc runs its run_selfplay from many threads.
#define HM_ENGINE_CLASS_NAME DirectEngine
#include "hm_batching_engine.hpp"
class Engine : public hm::BatchingEngine {
public:
    using hm::BatchingEngine::BatchingEngine;
};
#endif
