#include "hm_chunk.hpp"
#include <cstdio>
int main(int argc, char** argv) {
    hm::ChunkWriter w(argv[1], 2, 42);
    std::vector<hm::TrainingSample> v(3);
    for (int i = 0; i < 3; ++i) { v[i].gameId = i; v[i].policyA = {{5, 0.25f}, {9, 0.75f}}; v[i].policyB = {{0, 1.0f}}; v[i].planes[i] = 255; }
    w.append(v);
    w.finish();
    std::printf("%zu\n", w.chunksWritten());
    try { hm::ChunkWriter bad("/proc/nope/x", 1, 1); } catch (const std::runtime_error& e) { std::printf("threw: %s\n", e.what()); }
    return 0;
}
