"""The drop-in claim, executed: oracle/_ref/hivemind_dropin is the reference's WHOLE engine (search/, tools/,
interface/, main.cc, environment/, Fairy-Stockfish) compiled UNMODIFIED from /root/reference with
integration/nn/engine.h (a forwarder onto include/hm_engine.hpp) in place of the TensorRT nn/engine.{h,cc}
(`make -C oracle dropin`).  On the B200 it runs the reference's own `bench` and `selfplay`
commands on the CUDA evaluator.  The binary is built where /root/reference exists and travels with oracle/_ref."""
import os
import re
import subprocess
import time

import numpy as np
import pytest

from conftest import ROOT

BIN = os.path.join(ROOT, "oracle", "_ref", "hivemind_dropin")
needs_bin = pytest.mark.skipif(not os.path.exists(BIN), reason="oracle/_ref/hivemind_dropin not built (needs /root/reference)")


@needs_bin
def test_dropin_binary_links_the_product_library():
    out = subprocess.run(["ldd", BIN], capture_output=True, text=True).stdout
    assert "libhm_engine.so" in out and "nvinfer" not in out.lower() and "nvonnxparser" not in out.lower()


@pytest.fixture(scope="module")
def model_dir(tmp_path_factory):
    """./models/model.onnx is what the reference's resolveModelPath finds (nn/onnx_utils.cc:13-33); hm_engine_load maps
    it to the sibling .hvmw written by hivemind_b200.weights."""
    from hivemind_b200 import weights as W

    d = tmp_path_factory.mktemp("dropin")
    os.makedirs(d / "models")
    (d / "models" / "model.onnx").write_bytes(b"placeholder: the packed weights are in model.hvmw\n")
    (d / "models" / "model.hvmw").write_bytes(W.random_hvmw(seed=0))
    return d


@needs_bin
@pytest.mark.gpu
def test_reference_bench_command_runs_on_the_cuda_evaluator(model_dir):
    r = subprocess.run([BIN, "bench", "300"], cwd=model_dir, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    m = re.search(r"Positions per second: ([0-9.eE+]+)", r.stdout)  # tools/benchmark.cc:53
    assert m and float(m.group(1)) > 1000, r.stdout
    print("\n[dropin] reference `bench` (batch %s): %s positions/s" % (re.search(r"Batch size: (\d+)", r.stdout).group(1)
                                                                      if "Batch size" in r.stdout else "8", m.group(1)))


@needs_bin
@pytest.mark.gpu
def test_reference_selfplay_command_plays_games_and_writes_chunks(model_dir):
    """`hivemind selfplay` (tools/selfplay.cc:558-748): the reference's Agent / SearchThread / Node drive
    enqueueInferenceHalf / synchronizeInferenceHalf of the CUDA evaluator; its ChunkWriter output parses."""
    from oracle.chunk_oracle import read_chunk

    out = model_dir / "games"
    nodes, games = int(os.environ.get("HM_DROPIN_NODES", "200")), int(os.environ.get("HM_DROPIN_GAMES", "2"))
    t0 = time.time()
    r = subprocess.run([BIN, "selfplay", "--model", "models/model.onnx", "--games", str(games), "--nodes", str(nodes),
                        "--output", str(out), "--seed", "7", "--max-macro-plies", "40", "--chunk-samples", "64"],
                       cwd=model_dir, capture_output=True, text=True, timeout=900)
    dt = time.time() - t0
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    lines = re.findall(r"selfplay game (\d+)/(\d+) raw (\d+) samples (\d+) events (\d+) termination (\S+)", r.stdout)
    assert len(lines) == games, r.stdout
    samples = sum(int(l[3]) for l in lines)
    assert samples > 0
    chunks = sorted(os.path.join(b, f) for b, _, fs in os.walk(out) for f in fs if f.endswith(".hvm"))  # out/training_data/
    assert chunks, list(os.walk(out))
    assert os.path.exists(out / "games.pgn")
    parsed = sum(len(read_chunk(c)) for c in chunks)
    assert parsed == samples
    first = read_chunk(chunks[0])[0]
    assert len(first["x"]) == 4736 and first["search_nodes"] > 0
    print(f"\n[dropin] reference selfplay: {games} games, {samples} searched positions x {nodes} nodes in {dt:.1f}s "
          f"(~{samples * nodes / dt:.0f} search nodes/s incl. start-up; batch 8, the reference's own driver)")


MG = os.path.join(ROOT, "oracle", "_ref", "hivemind_multigame")


@pytest.mark.skipif(not os.path.exists(MG), reason="oracle/_ref/hivemind_multigame not built (needs /root/reference)")
@pytest.mark.gpu
def test_many_concurrent_reference_games_share_one_evaluator(model_dir):
    """include/hm_batching_engine.hpp: G threads run the reference's unmodified run_selfplay, each with its own
    small-batch `Engine` facade; their 8-position requests are coalesced into large device batches."""
    from oracle.chunk_oracle import read_chunk

    g = int(os.environ.get("HM_MG_CONCURRENT", "8"))
    nodes = int(os.environ.get("HM_MG_NODES", "200"))
    per = int(os.environ.get("HM_MG_GAMES_PER_THREAD", "1"))
    out = model_dir / f"mg_{g}_{nodes}"
    r = subprocess.run(["timeout", "600", MG, "--model", "models/model.onnx", "--concurrent", str(g), "--games-per-thread", str(per),
                        "--nodes", str(nodes), "--output", str(out), "--seed", "11", "--max-macro-plies", "40"],
                       cwd=model_dir, capture_output=True, text=True, timeout=700)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    m = re.search(r"multigame summary: concurrent (\d+) games_per_thread (\d+) nodes (\d+) seconds ([0-9.eE+-]+) device_passes (\d+) "
                  r"rows_served (\d+) large_batch (\d+) failures (\d+)", r.stdout)
    assert m, r.stdout[-2000:]
    secs, passes, rows, big, failures = float(m.group(4)), int(m.group(5)), int(m.group(6)), int(m.group(7)), int(m.group(8))
    assert failures == 0 and passes > 0
    chunks = [os.path.join(b, f) for b, _, fs in os.walk(out) for f in fs if f.endswith(".hvm")]
    samples = [s for c in chunks for s in read_chunk(c)]
    assert len({os.path.dirname(os.path.dirname(c)) for c in chunks}) == g  # every thread finished its games
    searched = sum(s["search_nodes"] for s in samples)
    assert searched > 0
    print(f"\n[multigame] {g} concurrent reference games x {per}, {nodes} nodes/move: {len(samples)} searched positions, "
          f"{searched} search nodes in {secs:.2f}s = {searched / secs:.0f} nodes/s; {passes} device passes of {big} rows, "
          f"{rows / passes:.1f} live rows per pass ({rows / secs:.0f} evaluated positions/s)")
