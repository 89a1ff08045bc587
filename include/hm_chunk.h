/*
 * hm_chunk.h — C ABI of the HVM4 self-play chunk writer.
 *
 * Replaces the reference's `ChunkWriter` (engine/src/tools/selfplay.cc:69-158) so that leaf
 * evaluations produced through hm_engine.h end up in files the reference trainer's reader
 * (src/preprocessing/convert_selfplay_data.py:69-146, `read_binary_shard`) consumes unchanged.
 * Host code only: the one device-side piece of the format, the u8 plane quantisation
 * (selfplay.cc:464-476), is `hm_encode_planes(..., HM_DTYPE_U8)` in hm_engine.h.
 *
 * Wire format (little endian, no padding; selfplay.cc:32-53,116-140):
 *   header  <4sIHHQ   magic "HVM4", version 4, channels 74, policy values 4672, sample count
 *   sample  <QIHHBBbBf  gameId, nodes, macroPly, movesLeft, team, hasTimeAdvantage, outcome, wdl, rootQ
 *           4736 x u8   planes, index c*64 + sq
 *           <H + n x <Hf   sparse policy of board A (index = channel*64 + square, probability)
 *           <H + n x <Hf   sparse policy of board B
 * Files are named chunk_<runId>_<%06d index>.hvm, written to <name>.tmp and renamed when complete
 * (selfplay.cc:104-150).
 *
 * All functions return 1 on success and 0 on failure (the reference throws std::runtime_error,
 * selfplay.cc:63-65,96-98,112-114,143-150); the message is retrievable with hm_last_error().
 */
#ifndef HM_CHUNK_H
#define HM_CHUNK_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define HM_CHUNK_VERSION 4u              /* selfplay.cc:33 */
#define HM_CHUNK_PLANE_BYTES 4736        /* NB_INPUT_VALUES(), one u8 per plane value */

/* Scalars of one training sample, `TrainingSample` selfplay.cc:40-53 (same names, same defaults:
 * wdl = 1 means "draw"). */
typedef struct hm_training_sample {
    uint64_t game_id;
    uint32_t nodes;
    uint16_t macro_ply;
    uint16_t moves_left;
    uint8_t team;
    uint8_t has_time_advantage;
    int8_t outcome;
    uint8_t wdl;
    float root_q;
} hm_training_sample;

typedef struct hm_chunk_writer hm_chunk_writer;

/* ChunkWriter(directory, samplesPerChunk, runId), selfplay.cc:71-76: creates the directory,
 * samples_per_chunk is clamped to >= 1.  NULL on failure. */
hm_chunk_writer* hm_chunk_writer_create(const char* directory, size_t samples_per_chunk, uint64_t run_id);

/* ChunkWriter::append for ONE sample, selfplay.cc:78-85: the sample is queued and a chunk of exactly
 * samples_per_chunk samples is published as soon as that many are pending.  `planes` = 4736 bytes;
 * a sparse policy is given as parallel arrays (index, probability) of n entries, n <= 65535
 * (selfplay.cc:95-98), either pointer may be NULL when its n is 0.  Everything is copied. */
int hm_chunk_writer_append(hm_chunk_writer* w, const hm_training_sample* sample, const uint8_t* planes,
                           const uint16_t* index_a, const float* prob_a, size_t n_a,
                           const uint16_t* index_b, const float* prob_b, size_t n_b);

/* ChunkWriter::finish, selfplay.cc:87-91: publishes the pending samples (if any) as a last, short chunk. */
int hm_chunk_writer_finish(hm_chunk_writer* w);

/* Bookkeeping the reference keeps private: chunks published so far / samples still pending. */
size_t hm_chunk_writer_chunks_written(const hm_chunk_writer* w);
size_t hm_chunk_writer_pending(const hm_chunk_writer* w);

/* Destroys the writer WITHOUT publishing pending samples (the reference destructor does not flush). */
void hm_chunk_writer_destroy(hm_chunk_writer* w);

/* Root visit counts -> sparse policy of one board, `marginal_policy` selfplay.cc:400-426 after its
 * move -> policy-index step (hm_policy_indices in hm_engine.h does that one, utils.h:184-216): visits of
 * edges that map to the same index are summed, edges with visits <= 0 are skipped, entries come out in
 * ascending index order with probability = float(visits) / float(total visits).  Writes at most `cap`
 * entries and returns the number of distinct indices; returns 0 and sets the error message when no edge
 * was visited (the reference throws "Search returned no visited root edges"). */
size_t hm_sparse_policy_from_visits(const int32_t* policy_index, const int64_t* visits, size_t n_edges,
                                    uint16_t* index_out, float* prob_out, size_t cap);

#ifdef __cplusplus
}
#endif
#endif /* HM_CHUNK_H */
