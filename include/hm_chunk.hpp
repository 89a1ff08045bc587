// hm_chunk.hpp — header-only C++ face of include/hm_chunk.h with the reference's own surface
// (engine/src/tools/selfplay.cc:35-53 SparsePolicyEntry / TrainingSample, :69-158 ChunkWriter), so
// `run_selfplay` (selfplay.cc:558-748) can keep its calls: construct, append(std::vector<TrainingSample>),
// finish().  Failures throw std::runtime_error like the reference.
#pragma once
#include <array>
#include <cstdint>
#include <stdexcept>
#include <string>
#include <vector>

#include "hm_chunk.h"
#include "hm_engine.h"

namespace hm {

struct SparsePolicyEntry {
    uint16_t index;
    float probability;
};

struct TrainingSample {
    uint64_t gameId = 0;
    uint32_t nodes = 0;
    uint16_t macroPly = 0;
    uint16_t movesLeft = 0;
    uint8_t team = 0;
    uint8_t hasTimeAdvantage = 0;
    int8_t outcome = 0;
    uint8_t wdl = 1;
    float rootQ = 0.0f;
    std::array<uint8_t, HM_NB_INPUT_VALUES> planes{};
    std::vector<SparsePolicyEntry> policyA;
    std::vector<SparsePolicyEntry> policyB;
};

class ChunkWriter {
public:
    ChunkWriter(const std::string& directory, size_t samplesPerChunk, uint64_t runId)
        : w_(hm_chunk_writer_create(directory.c_str(), samplesPerChunk, runId)) {
        if (!w_) throw std::runtime_error(hm_last_error());
    }
    ~ChunkWriter() { hm_chunk_writer_destroy(w_); }
    ChunkWriter(const ChunkWriter&) = delete;
    ChunkWriter& operator=(const ChunkWriter&) = delete;

    void append(const std::vector<TrainingSample>& samples) {
        std::vector<uint16_t> ia, ib;
        std::vector<float> pa, pb;
        for (const TrainingSample& s : samples) {
            split(s.policyA, ia, pa);
            split(s.policyB, ib, pb);
            const hm_training_sample c{s.gameId, s.nodes, s.macroPly, s.movesLeft, s.team, s.hasTimeAdvantage,
                                       s.outcome, s.wdl, s.rootQ};
            if (!hm_chunk_writer_append(w_, &c, s.planes.data(), ia.data(), pa.data(), ia.size(), ib.data(), pb.data(),
                                        ib.size()))
                throw std::runtime_error(hm_last_error());
        }
    }
    void finish() {
        if (!hm_chunk_writer_finish(w_)) throw std::runtime_error(hm_last_error());
    }
    size_t chunksWritten() const { return hm_chunk_writer_chunks_written(w_); }

private:
    static void split(const std::vector<SparsePolicyEntry>& p, std::vector<uint16_t>& idx, std::vector<float>& prob) {
        idx.resize(p.size());
        prob.resize(p.size());
        for (size_t i = 0; i < p.size(); ++i) { idx[i] = p[i].index; prob[i] = p[i].probability; }
    }
    hm_chunk_writer* w_;
};

}  // namespace hm
