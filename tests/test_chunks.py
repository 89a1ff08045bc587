"""HVM4 self-play chunk writer (include/hm_chunk.h, hivemind_b200/chunks.py) against the restated format
(oracle/chunk_oracle.py) and the golden file the reference's own reader accepted
(tests/golden/chunk_7_000000.hvm, made by oracle/make_golden_chunk.py).  Reference: engine/src/tools/selfplay.cc:32-158,
400-426, 464-476; src/preprocessing/convert_selfplay_data.py:24-146."""
import os
import re

import numpy as np
import pytest

from hivemind_b200 import chunks
from hivemind_b200.chunks import ChunkWriter, TrainingSample
from oracle import chunk_oracle as CO

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")


def fixture():
    return np.load(os.path.join(GOLD, "positions_seed1234_n256.npz"))


def to_product(s):
    return TrainingSample(gameId=s["gameId"], nodes=s["nodes"], macroPly=s["macroPly"], movesLeft=s["movesLeft"],
                          team=s["team"], hasTimeAdvantage=s["hasTimeAdvantage"], outcome=s["outcome"], wdl=s["wdl"],
                          rootQ=s["rootQ"], planes=s["planes"], policyA=s["policyA"], policyB=s["policyB"])


def test_header_declares_exported_symbols():
    from hivemind_b200 import _lib
    L = _lib.load()
    text = open(os.path.join(ROOT, "include", "hm_chunk.h")).read()
    names = set(re.findall(r"\b(hm_[a-z0-9_]+)\s*\(", text))
    assert {"hm_chunk_writer_create", "hm_chunk_writer_append", "hm_chunk_writer_finish", "hm_chunk_writer_destroy",
            "hm_chunk_writer_chunks_written", "hm_chunk_writer_pending", "hm_sparse_policy_from_visits"} <= names
    missing = [n for n in names if n not in ("hm_last_error", "hm_encode_planes", "hm_policy_indices") and not hasattr(L, n)]
    assert not missing, missing


def test_golden_chunk_is_reproduced_byte_for_byte(tmp_path):
    samples = CO.golden_samples(fixture())
    w = ChunkWriter(tmp_path / "a" / "b", samplesPerChunk=len(samples), runId=7)   # nested directory is created
    w.append([to_product(s) for s in samples])
    assert w.chunksWritten == 1 and w.pending == 0       # published as soon as samplesPerChunk are pending
    w.finish()                                            # nothing left: no empty chunk
    assert w.chunksWritten == 1
    path = tmp_path / "a" / "b" / CO.chunk_name(7, 0)
    data = path.read_bytes()
    assert data == open(os.path.join(GOLD, CO.chunk_name(7, 0)), "rb").read()
    assert data == CO.chunk_bytes(samples)
    assert sorted(os.listdir(tmp_path / "a" / "b")) == [CO.chunk_name(7, 0)]   # no .tmp left behind


def test_golden_chunk_parses_to_the_inputs():
    samples = CO.golden_samples(fixture())
    got = CO.read_chunk(os.path.join(GOLD, CO.chunk_name(7, 0)), q_value_ratio=0.5)
    assert len(got) == len(samples)
    for r, s in zip(got, samples):
        assert r["x"] == np.asarray(s["planes"], np.uint8).tobytes()
        assert (r["game_id"], r["search_nodes"], r["macro_ply"], r["y_moves_left"]) == (s["gameId"], s["nodes"], s["macroPly"], s["movesLeft"])
        assert (r["team"], r["has_time_advantage"], r["y_wdl"]) == (s["team"], bool(s["hasTimeAdvantage"]), s["wdl"])
        assert r["outcome"] == float(s["outcome"]) and r["root_q"] == float(np.float32(s["rootQ"]))
        assert r["y_value"] == 0.5 * float(s["outcome"]) + 0.5 * float(np.float32(s["rootQ"]))
        for key, pol in (("policy_a", s["policyA"]), ("policy_b", s["policyB"])):
            dense = np.frombuffer(r[key], np.float32)
            assert np.array_equal(np.nonzero(dense)[0], pol[0].astype(np.int64))
            assert np.array_equal(dense[pol[0]], pol[1])


def test_chunking_and_finish(tmp_path):
    samples = CO.golden_samples(fixture(), n=7)
    w = ChunkWriter(tmp_path, samplesPerChunk=3, runId=123456789012345)
    w.append([to_product(s) for s in samples[:4]])
    assert (w.chunksWritten, w.pending) == (1, 1)
    w.append([to_product(s) for s in samples[4:]])
    assert (w.chunksWritten, w.pending) == (2, 1)
    w.finish()
    assert (w.chunksWritten, w.pending) == (3, 0)
    names = sorted(os.listdir(tmp_path))
    assert names == [CO.chunk_name(123456789012345, i) for i in range(3)]
    sizes = [len(CO.read_chunk(tmp_path / n)) for n in names]
    assert sizes == [3, 3, 1]
    joined = [r for n in names for r in CO.read_chunk(tmp_path / n)]
    assert [r["game_id"] for r in joined] == [s["gameId"] for s in samples]    # order preserved across chunks
    for i, n in enumerate(names):
        assert (tmp_path / n).read_bytes() == CO.chunk_bytes(samples[3 * i:3 * i + 3])


def test_samples_per_chunk_is_clamped_to_one_and_defaults(tmp_path):
    w = ChunkWriter(tmp_path, samplesPerChunk=0, runId=1)
    pol = (np.array([5], np.uint16), np.array([1.0], np.float32))
    w.append([TrainingSample(policyA=pol, policyB=pol)])
    assert w.chunksWritten == 1
    (r,) = CO.read_chunk(tmp_path / CO.chunk_name(1, 0))
    assert r["y_wdl"] == 1 and r["x"] == bytes(4736) and r["game_id"] == 0      # TrainingSample defaults, selfplay.cc:40-53


def test_empty_policy_is_written_but_rejected_by_the_reader(tmp_path):
    # the writer does not validate probabilities (selfplay.cc:94-103); the reader insists on a sum of 1
    w = ChunkWriter(tmp_path, samplesPerChunk=1, runId=2)
    w.append([TrainingSample()])
    data = (tmp_path / CO.chunk_name(2, 0)).read_bytes()
    assert len(data) == 20 + 24 + 4736 + 2 + 2
    with pytest.raises(ValueError, match="sums to"):
        CO.read_chunk(tmp_path / CO.chunk_name(2, 0))


def test_errors(tmp_path):
    blocker = tmp_path / "file"
    blocker.write_bytes(b"x")
    with pytest.raises(RuntimeError, match="Unable to create directory"):
        ChunkWriter(blocker / "sub", 4, 1)
    w = ChunkWriter(tmp_path / "ok", 4, 1)
    with pytest.raises(ValueError):
        w.append([TrainingSample(planes=np.zeros(10, np.uint8))])
    big = (np.zeros(70000, np.uint16), np.zeros(70000, np.float32))
    with pytest.raises(RuntimeError, match="too large"):
        w.append([TrainingSample(policyA=big)])
    assert w.pending == 0
    w.close()
    w.close()   # idempotent


def test_cpp_chunk_writer_header(tmp_path):
    """include/hm_chunk.hpp (the reference's ChunkWriter / TrainingSample surface) builds with plain g++ and its files
    parse: 3 samples at 2 per chunk -> chunks of 2 and 1."""
    import subprocess
    from hivemind_b200 import build as hb
    lib = hb.build()
    exe = tmp_path / "chunk_writer_main"
    cmd = ["g++", "-O1", "-std=c++17", "-I" + os.path.join(ROOT, "include"), os.path.join(ROOT, "tests", "cpp", "chunk_writer_main.cc"),
           "-L" + os.path.dirname(lib), "-lhm_engine", "-Wl,-rpath," + os.path.dirname(lib), "-o", str(exe)]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    out = tmp_path / "chunks"
    r = subprocess.run([str(exe), str(out)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert r.stdout.splitlines()[0] == "2" and "threw: Unable to create directory" in r.stdout
    a = CO.read_chunk(out / CO.chunk_name(42, 0))
    b = CO.read_chunk(out / CO.chunk_name(42, 1))
    assert [x["game_id"] for x in a + b] == [0, 1, 2]
    dense = np.frombuffer(a[0]["policy_a"], np.float32)
    assert dense[5] == 0.25 and dense[9] == 0.75 and a[1]["x"][1] == 255


def test_marginal_policy_matches_restated_reference():
    rng = np.random.default_rng(5)
    for _ in range(50):
        n = int(rng.integers(1, 120))
        idx = rng.integers(0, 4672, n).astype(np.int32)
        idx[rng.integers(0, n, n // 3)] = idx[0]          # duplicates are summed
        vis = rng.integers(-2, 40, n).astype(np.int64)    # <= 0: skipped
        if (vis > 0).sum() == 0:
            vis[0] = 3
        gi, gp = chunks.marginal_policy(idx, vis)
        ei, ep = CO.marginal_policy(idx, vis)
        assert np.array_equal(gi, ei) and np.array_equal(gp, ep)
        assert np.all(np.diff(gi.astype(np.int64)) > 0) and abs(float(gp.sum()) - 1.0) < 1e-5
    with pytest.raises(RuntimeError, match="no visited root edges"):
        chunks.marginal_policy([3, 4], [0, -1])
    # a visited edge without a policy label is an error, as in the reference (selfplay.cc:392-398): the
    # unrepresentable index -1 must never become label 65535, nor an index beyond the 4672 labels a target
    for bad in (-1, 4672, 70000):
        with pytest.raises(RuntimeError, match="absent from policy map"):
            chunks.marginal_policy([3, bad], [5, 2])
    gi, gp = chunks.marginal_policy([3, -1], [5, 0])   # ... but an UNVISITED unrepresentable edge is skipped like any other
    assert gi.tolist() == [3] and gp.tolist() == [1.0]


def test_quantisation_restated_matches_reference_fixture():
    fx = fixture()
    assert np.array_equal(CO.quantize_planes(fx["planes_f32"]), fx["planes_u8"])   # planes_u8 came from the reference


@pytest.mark.gpu
def test_device_u8_planes_feed_the_writer(tmp_path):
    """Device-quantised planes (hm_encode_planes, HM_DTYPE_U8) -> chunk -> restated reader, bit-exact against the
    reference-quantised fixture."""
    from hivemind_b200.engine import Engine
    from oracle.pyoracle import POS_DTYPE
    fx = fixture()
    eng = Engine(0, 8, 1)
    pos = fx["packed"].view(POS_DTYPE).reshape(-1)
    u8 = eng.encodePlanes(pos, "u8")
    eng.close()
    assert np.array_equal(u8, fx["planes_u8"])
    samples = CO.golden_samples(fx, n=16)
    for g, s in enumerate(samples):
        s["planes"] = u8[g].reshape(-1)
    w = ChunkWriter(tmp_path, samplesPerChunk=16, runId=9)
    w.append([to_product(s) for s in samples])
    got = CO.read_chunk(tmp_path / CO.chunk_name(9, 0))
    assert [r["x"] for r in got] == [fx["planes_u8"][g].tobytes() for g in range(16)]
