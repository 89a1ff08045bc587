"""make_golden_chunk.py — pins the HVM4 chunk format against the reference's OWN reader.

Runs only where /root/reference exists (this container).  Writes the deterministic samples of
oracle.chunk_oracle.golden_samples with the PRODUCT writer (hivemind_b200.chunks.ChunkWriter over the C ABI),
checks the file byte for byte against the restated writer layout, parses it with the reference's
`read_binary_shard` (src/preprocessing/convert_selfplay_data.py:69-146; `polars`, absent here and unused by that
function, is stubbed) and with oracle.chunk_oracle.read_chunk, compares both with the inputs, and commits the
file as tests/golden/chunk_7_000000.hvm.

    python oracle/make_golden_chunk.py
"""
from __future__ import annotations

import os
import shutil
import sys
import tempfile
import types

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
REFERENCE = "/root/reference"

from hivemind_b200.chunks import ChunkWriter, TrainingSample  # noqa: E402
from oracle import chunk_oracle as CO  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")


def reference_reader():
    sys.modules.setdefault("polars", types.ModuleType("polars"))   # imported at module top, unused by the reader
    sys.path.insert(0, REFERENCE)
    import src.preprocessing.convert_selfplay_data as ref  # noqa: E402
    return ref


def to_product(s: dict) -> TrainingSample:
    return TrainingSample(gameId=s["gameId"], nodes=s["nodes"], macroPly=s["macroPly"], movesLeft=s["movesLeft"],
                          team=s["team"], hasTimeAdvantage=s["hasTimeAdvantage"], outcome=s["outcome"], wdl=s["wdl"],
                          rootQ=s["rootQ"], planes=s["planes"], policyA=s["policyA"], policyB=s["policyB"])


def main():
    fixture = np.load(os.path.join(GOLD, "positions_seed1234_n256.npz"))
    samples = CO.golden_samples(fixture)
    tmp = tempfile.mkdtemp()
    try:
        w = ChunkWriter(tmp, samplesPerChunk=len(samples), runId=7)
        w.append([to_product(s) for s in samples])
        w.finish()
        path = os.path.join(tmp, CO.chunk_name(7, 0))
        data = open(path, "rb").read()
        assert data == CO.chunk_bytes(samples), "product writer differs from the restated layout"
        ref = reference_reader()
        got_ref = ref.read_binary_shard(path, q_value_ratio=0.25)
        got_ora = CO.read_chunk(path, q_value_ratio=0.25)
        assert len(got_ref) == len(samples)
        for r, o, s in zip(got_ref, got_ora, samples):
            assert r.keys() == o.keys()
            for k in r:
                assert r[k] == o[k], k
            assert r["x"] == np.asarray(s["planes"], np.uint8).tobytes()
            assert r["game_id"] == s["gameId"] and r["search_nodes"] == s["nodes"] and r["macro_ply"] == s["macroPly"]
            assert r["y_moves_left"] == s["movesLeft"] and r["team"] == s["team"] and r["y_wdl"] == s["wdl"]
            assert r["has_time_advantage"] == bool(s["hasTimeAdvantage"]) and r["outcome"] == float(s["outcome"])
            assert r["root_q"] == float(np.float32(s["rootQ"]))
            for key, pol in (("policy_a", s["policyA"]), ("policy_b", s["policyB"])):
                dense = np.frombuffer(r[key], np.float32)
                assert dense.shape == (CO.NB_POLICY_VALUES,)
                assert np.array_equal(np.nonzero(dense)[0], pol[0].astype(np.int64))
                assert np.array_equal(dense[pol[0]], pol[1])
        shutil.copy(path, os.path.join(GOLD, CO.chunk_name(7, 0)))
        print(f"reference reader accepted {len(samples)} samples ({len(data)} bytes); wrote tests/golden/{CO.chunk_name(7, 0)}")
    finally:
        shutil.rmtree(tmp, ignore_errors=True)


if __name__ == "__main__":
    main()
