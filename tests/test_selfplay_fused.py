"""BASELINE.json configs 3 / 5: integration/_build/hivemind_selfplay_fused — many concurrent
self-play games whose leaves go to the evaluator as FUSED requests (packed positions + legal-move lists in, priors per
move + shaped values out).  The tree is the reference's own code (searchthread.cc / node.cc compiled unmodified), the
cross-game leaf batching and the game loop are integration/fused_selfplay.cc.  Built where /root/reference exists
(`make -C integration`), travels with the snapshot."""
import json
import os
import subprocess

import numpy as np
import pytest

from conftest import ROOT

BIN = os.path.join(ROOT, "integration", "_build", "hivemind_selfplay_fused")
needs_bin = pytest.mark.skipif(not os.path.exists(BIN), reason="integration/_build/hivemind_selfplay_fused not built (needs /root/reference)")


@needs_bin
def test_links_the_product_library_and_refuses_bad_arguments():
    out = subprocess.run(["ldd", BIN], capture_output=True, text=True).stdout
    assert "libhm_engine.so" in out and "nvinfer" not in out.lower()
    r = subprocess.run([BIN, "--games", "4"], capture_output=True, text=True, timeout=60)
    assert r.returncode == 2 and "usage" in r.stderr


@needs_bin
def test_fails_loudly_without_a_cuda_device(tmp_path):
    """No CPU fallback anywhere on the product path: without a device the evaluator cannot be created and the
    driver exits non-zero with the engine's message (skipped on a GPU box)."""
    import torch

    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    from hivemind_b200 import weights as W

    w = tmp_path / "net.hvmw"
    w.write_bytes(W.random_hvmw(seed=0))
    r = subprocess.run([BIN, "--model", str(w), "--games", "4", "--nodes", "8", "--seconds", "1", "--threads", "1"],
                       capture_output=True, text=True, timeout=120)
    assert r.returncode == 1 and "loadNetwork failed" in r.stderr


@needs_bin
@pytest.mark.gpu
def test_fused_selfplay_searches_plays_and_writes_reference_format_chunks(tmp_path):
    from hivemind_b200 import weights as W
    from oracle.chunk_oracle import read_chunk

    w = tmp_path / "net.hvmw"
    w.write_bytes(W.random_hvmw(seed=0))
    chunks = tmp_path / "chunks"
    r = subprocess.run(["timeout", "300", BIN, "--model", str(w), "--games", "96", "--nodes", "64", "--threads", "4",
                        "--max-games", "24", "--seconds", "60", "--max-macro-plies", "24", "--chunks", str(chunks),
                        "--chunk-samples", "128", "--seed", "5"], capture_output=True, text=True, timeout=400)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    rec = json.loads([ln for ln in r.stdout.splitlines() if ln.startswith("{")][-1])
    assert not rec["failed"] and rec["games_finished"] >= 24 and rec["moves_played"] > 0
    # every simulation is either a device leaf evaluation or a terminal / solved node
    assert rec["nodes"] == rec["leaf_evals"] + rec["terminal_nodes"] or rec["nodes"] >= rec["leaf_evals"]
    assert rec["leaf_evals"] > 0 and rec["avg_batch_fill"] > 32      # leaves of many games share a request
    files = sorted(os.path.join(b, f) for b, _, fs in os.walk(chunks) for f in fs if f.endswith(".hvm"))
    assert files
    samples = [s for f in files for s in read_chunk(f)]
    assert len(samples) == rec["samples_written"] > 0
    for s in samples[:200]:
        assert len(s["x"]) == 4736 and s["search_nodes"] > 0 and s["y_wdl"] in (0, 1, 2)
        pa, pb = np.frombuffer(s["policy_a"], np.float32), np.frombuffer(s["policy_b"], np.float32)   # dense [4672]
        assert pa.size == 4672 and abs(float(pa.sum()) - 1.0) < 1e-4 and abs(float(pb.sum()) - 1.0) < 1e-4
    # the searched budget follows randomized_node_budget (selfplay.cc:218-224): 64 nodes +- 5 %
    nodes = np.array([s["search_nodes"] for s in samples])
    assert nodes.min() >= 55 and nodes.max() <= 72, (nodes.min(), nodes.max())
    print(f"\n[fused selfplay] {rec['nodes_per_s']:.0f} nodes/s, {rec['leaf_evals_per_s']:.0f} leaf evals/s, "
          f"avg batch fill {rec['avg_batch_fill']:.0f}, {rec['games_finished']} games, {len(samples)} samples")
