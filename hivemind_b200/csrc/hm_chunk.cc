// hm_chunk.cc — HVM4 self-play chunk writer behind include/hm_chunk.h (host code only).
// Wire format and publication protocol follow the reference's ChunkWriter
// (engine/src/tools/selfplay.cc:32-53 sample fields, :94-103 sparse policy, :104-150 flush);
// the reader that must accept the result is src/preprocessing/convert_selfplay_data.py:69-146.
#include <cerrno>
#include <cstdio>
#include <cstring>
#include <deque>
#include <map>
#include <string>
#include <vector>

#include <sys/stat.h>
#include <unistd.h>

#include "../../include/hm_chunk.h"
#include "../../include/hm_engine.h"
#include "hm_net.h"

namespace {

using hm::set_error;

// one queued sample, already in wire order: scalars | planes | policy A | policy B
struct Record {
    std::vector<uint8_t> bytes;
};

template <typename T>
void put(std::vector<uint8_t>& out, T v) {   // little-endian host (x86-64 / aarch64), like the reference
    const uint8_t* p = reinterpret_cast<const uint8_t*>(&v);
    out.insert(out.end(), p, p + sizeof(T));
}

void put_policy(std::vector<uint8_t>& out, const uint16_t* index, const float* prob, size_t n) {
    put<uint16_t>(out, static_cast<uint16_t>(n));
    for (size_t i = 0; i < n; ++i) {
        put<uint16_t>(out, index[i]);
        put<float>(out, prob[i]);
    }
}

bool make_directories(const std::string& path) {   // mkdir -p
    if (path.empty()) return true;
    std::string cur;
    size_t pos = 0;
    while (pos <= path.size()) {
        const size_t next = path.find('/', pos);
        cur = path.substr(0, next == std::string::npos ? path.size() : next);
        if (!cur.empty() && ::mkdir(cur.c_str(), 0777) != 0 && errno != EEXIST) return false;
        if (next == std::string::npos) break;
        pos = next + 1;
    }
    struct stat st {};
    return ::stat(path.c_str(), &st) == 0 && S_ISDIR(st.st_mode);
}

}  // namespace

struct hm_chunk_writer {
    std::string directory;
    size_t samples_per_chunk = 1;
    uint64_t run_id = 0;
    size_t chunk_index = 0;
    std::deque<Record> pending;

    // publish the first `count` pending samples as one chunk (selfplay.cc:104-150)
    bool flush(size_t count) {
        char name[96];
        std::snprintf(name, sizeof(name), "chunk_%llu_%06zu.hvm", static_cast<unsigned long long>(run_id), chunk_index);
        const std::string final_path = directory.empty() ? std::string(name) : directory + "/" + name;
        const std::string tmp_path = final_path + ".tmp";

        std::vector<uint8_t> header;
        const char magic[4] = {'H', 'V', 'M', '4'};
        header.insert(header.end(), magic, magic + 4);
        put<uint32_t>(header, HM_CHUNK_VERSION);
        put<uint16_t>(header, HM_NB_INPUT_CHANNELS);
        put<uint16_t>(header, HM_NB_POLICY_VALUES);
        put<uint64_t>(header, static_cast<uint64_t>(count));

        std::FILE* f = std::fopen(tmp_path.c_str(), "wb");
        if (!f) { set_error("Unable to create " + tmp_path + ": " + std::strerror(errno)); return false; }
        bool ok = std::fwrite(header.data(), 1, header.size(), f) == header.size();
        for (size_t i = 0; ok && i < count; ++i)
            ok = std::fwrite(pending[i].bytes.data(), 1, pending[i].bytes.size(), f) == pending[i].bytes.size();
        ok = (std::fclose(f) == 0) && ok;
        if (!ok) {
            ::unlink(tmp_path.c_str());
            set_error("Failed to write self-play chunk " + tmp_path);
            return false;
        }
        if (std::rename(tmp_path.c_str(), final_path.c_str()) != 0) {
            const std::string why = std::strerror(errno);
            ::unlink(tmp_path.c_str());
            set_error("Unable to publish " + final_path + ": " + why);
            return false;
        }
        ++chunk_index;
        pending.erase(pending.begin(), pending.begin() + static_cast<std::ptrdiff_t>(count));
        return true;
    }
};

extern "C" {

hm_chunk_writer* hm_chunk_writer_create(const char* directory, size_t samples_per_chunk, uint64_t run_id) {
    try {
        if (!directory) { set_error("null directory"); return nullptr; }
        auto* w = new hm_chunk_writer();
        w->directory = directory;
        while (w->directory.size() > 1 && w->directory.back() == '/') w->directory.pop_back();
        w->samples_per_chunk = samples_per_chunk < 1 ? 1 : samples_per_chunk;
        w->run_id = run_id;
        if (!make_directories(w->directory)) {
            set_error("Unable to create directory " + w->directory + ": " + std::strerror(errno));
            delete w;
            return nullptr;
        }
        return w;
    } catch (const std::exception& ex) {
        set_error(std::string("exception: ") + ex.what());
        return nullptr;
    }
}

int hm_chunk_writer_append(hm_chunk_writer* w, const hm_training_sample* s, const uint8_t* planes,
                           const uint16_t* index_a, const float* prob_a, size_t n_a,
                           const uint16_t* index_b, const float* prob_b, size_t n_b) {
    try {
        if (!w || !s || !planes) { set_error("null argument"); return 0; }
        if (n_a > 0xffff || n_b > 0xffff) { set_error("Sparse policy is too large for HVM4"); return 0; }
        if ((n_a && (!index_a || !prob_a)) || (n_b && (!index_b || !prob_b))) { set_error("null sparse policy"); return 0; }
        Record r;
        r.bytes.reserve(24 + HM_CHUNK_PLANE_BYTES + 4 + 6 * (n_a + n_b));
        put<uint64_t>(r.bytes, s->game_id);
        put<uint32_t>(r.bytes, s->nodes);
        put<uint16_t>(r.bytes, s->macro_ply);
        put<uint16_t>(r.bytes, s->moves_left);
        put<uint8_t>(r.bytes, s->team);
        put<uint8_t>(r.bytes, s->has_time_advantage);
        put<int8_t>(r.bytes, s->outcome);
        put<uint8_t>(r.bytes, s->wdl);
        put<float>(r.bytes, s->root_q);
        r.bytes.insert(r.bytes.end(), planes, planes + HM_CHUNK_PLANE_BYTES);
        put_policy(r.bytes, index_a, prob_a, n_a);
        put_policy(r.bytes, index_b, prob_b, n_b);
        w->pending.push_back(std::move(r));
        if (w->pending.size() >= w->samples_per_chunk) return w->flush(w->samples_per_chunk) ? 1 : 0;
        return 1;
    } catch (const std::exception& ex) {
        set_error(std::string("exception: ") + ex.what());
        return 0;
    }
}

int hm_chunk_writer_finish(hm_chunk_writer* w) {
    try {
        if (!w) { set_error("null writer"); return 0; }
        if (!w->pending.empty()) return w->flush(w->pending.size()) ? 1 : 0;
        return 1;
    } catch (const std::exception& ex) {
        set_error(std::string("exception: ") + ex.what());
        return 0;
    }
}

size_t hm_chunk_writer_chunks_written(const hm_chunk_writer* w) { return w ? w->chunk_index : 0; }
size_t hm_chunk_writer_pending(const hm_chunk_writer* w) { return w ? w->pending.size() : 0; }
void hm_chunk_writer_destroy(hm_chunk_writer* w) { delete w; }

size_t hm_sparse_policy_from_visits(const int32_t* policy_index, const int64_t* visits, size_t n_edges,
                                    uint16_t* index_out, float* prob_out, size_t cap) {
    try {
        if (n_edges && (!policy_index || !visits)) { set_error("null argument"); return 0; }
        std::map<uint16_t, uint64_t> by_index;   // ascending index order, like the reference's std::map
        uint64_t total = 0;
        for (size_t i = 0; i < n_edges; ++i) {
            if (visits[i] <= 0) continue;
            // the reference throws for a visited edge without a policy label (selfplay.cc:392-398); -1 (not
            // representable) or an index beyond the 4672 labels must never reach the chunk as a sparse target
            if (policy_index[i] < 0 || policy_index[i] >= HM_NB_POLICY_VALUES) {
                set_error("Move is absent from policy map");
                return 0;
            }
            by_index[static_cast<uint16_t>(policy_index[i])] += static_cast<uint64_t>(visits[i]);
            total += static_cast<uint64_t>(visits[i]);
        }
        if (total == 0) { set_error("Search returned no visited root edges"); return 0; }
        size_t n = 0;
        for (const auto& kv : by_index) {
            if (n < cap && index_out && prob_out) {
                index_out[n] = kv.first;
                prob_out[n] = static_cast<float>(kv.second) / static_cast<float>(total);
            }
            ++n;
        }
        return n;
    } catch (const std::exception& ex) {
        set_error(std::string("exception: ") + ex.what());
        return 0;
    }
}

}  // extern "C"
