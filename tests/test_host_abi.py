"""CPU-side checks (no GPU): the C-ABI library loads and exports every symbol include/hm_engine.h
declares, the C++ Engine shim compiles and links, the product path fails loudly without a CUDA
device, the packed-weight container round-trips, and the N>1 sharding logic is exercised with a
world_size-2 gloo group (oracle as the per-rank evaluator stand-in: checker only)."""
import ctypes as C
import os
import re
import shutil
import struct
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "hm_engine.h")


def declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(hm_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    from hivemind_b200 import _lib

    lib = _lib.load()
    names = declared_symbols()
    assert len(names) >= 25, names
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, f"declared in include/hm_engine.h but not exported: {missing}"


def test_version_and_last_error_are_callable_without_gpu():
    from hivemind_b200 import _lib

    lib = _lib.load()
    lib.hm_version.restype = C.c_char_p
    assert b"hivemind" in lib.hm_version().lower() or len(lib.hm_version()) > 0
    assert isinstance(_lib.last_error(), str)


def test_engine_fails_loudly_without_cuda_device():
    """No CPU fallback: without a CUDA device the product path refuses to construct."""
    import torch

    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from hivemind_b200.engine import Engine

    with pytest.raises(RuntimeError):
        Engine(0, 16, 1)


def test_product_package_never_imports_the_oracle():
    """hivemind_b200/, include/ and tools/ (everything shipped or run outside tests/ and bench.py's CPU legs)
    never import, include or link anything under oracle/."""
    for top in ("hivemind_b200", "include", "tools"):
        for base, _, files in os.walk(os.path.join(ROOT, top)):
            for f in files:
                if f.endswith((".py", ".cu", ".cuh", ".h", ".hpp", ".cc", ".sh")):
                    text = open(os.path.join(base, f), errors="ignore").read()
                    assert "import oracle" not in text and "from oracle" not in text and "oracle/" not in text, f


@pytest.mark.skipif(shutil.which("g++") is None, reason="no g++")
def test_cpp_engine_shim_compiles_and_links(tmp_path):
    """include/hm_engine.hpp + tools/hm_bench.cc (the reference's `bench` command restated)
    build against the shared library with nothing but g++."""
    from hivemind_b200 import build as hb

    lib = hb.build()
    out = tmp_path / "hm_bench"
    cmd = ["g++", "-O1", "-std=c++17", "-I" + os.path.join(ROOT, "include"), os.path.join(ROOT, "tools", "hm_bench.cc"),
           "-L" + os.path.dirname(lib), "-lhm_engine", "-Wl,-rpath," + os.path.dirname(lib), "-o", str(out)]
    if os.path.isdir("/usr/local/cuda/include"):
        cmd.insert(4, "-I/usr/local/cuda/include")
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    r = subprocess.run([str(out)], capture_output=True, text=True)
    assert r.returncode == 2 and "usage" in r.stderr


def test_hvmw_container_round_trip(tmp_path):
    from hivemind_b200 import weights as W

    folded = W.fold_state_dict(W.random_state_dict(seed=3))
    blob = W.write_hvmw(None, folded)
    back = W.read_hvmw(blob)
    assert set(back) == set(folded)
    for k in folded:
        assert back[k].shape == folded[k].shape
        np.testing.assert_array_equal(back[k], folded[k].astype(back[k].dtype))
    p = tmp_path / "net.hvmw"
    W.write_hvmw(str(p), folded)
    assert p.read_bytes() == blob


def test_bn_folding_matches_unfolded_forward():
    """Folding BatchNorm into the preceding conv is exact up to fp32 rounding on a 1x1 conv."""
    from hivemind_b200 import weights as W

    sd = W.random_state_dict(seed=5)
    folded = W.fold_state_dict(sd)
    n_params = sum(int(np.prod(v.shape)) for k, v in sd.items() if not k.endswith(("running_mean", "running_var", "num_batches_tracked")))
    assert n_params == 14_122_085  # RISEv3.3 parameter count
    assert all(np.isfinite(v).all() for v in folded.values())


def test_shard_range_partitions_exactly():
    from hivemind_b200.sharding import shard_range

    for n in (0, 1, 7, 512, 1000):
        for world in (1, 2, 3, 8):
            spans = [shard_range(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard_range(4, 2, 2)


_WORKER = r"""
import os, sys
import numpy as np
sys.path.insert(0, {root!r})
import torch.distributed as dist
from hivemind_b200.sharding import shard_positions, broadcast_blob, gather_ragged
from hivemind_b200 import weights as W
from oracle.pyoracle import Oracle, POS_DTYPE

dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
blob = W.random_hvmw(seed=0) if rank == 0 else None
blob = broadcast_blob(blob, 0)
g = np.load(os.path.join({root!r}, "tests", "golden", "positions_seed1234_n256.npz"))
packed = np.ascontiguousarray(g["packed"]).view(POS_DTYPE).reshape(-1)[:37]
offsets = g["offsets"][: 2 * 37 + 1]
moves = g["moves"][: offsets[-1]]
pk, mv, off = shard_positions(packed, moves, offsets, rank, world)
oracle = Oracle()
i = np.arange(4672, dtype=np.int64)
la = np.stack([(((i * 7 + int(p["board"][0]["rule50"]) * 13) % 41) - 20).astype(np.float32) * 0.25 for p in pk]).astype(np.float16) if len(pk) else np.zeros((0, 4672), np.float16)
pri = oracle.batch_priors_f16(la.view(np.uint16), la.view(np.uint16), pk, mv, off)
allp = gather_ragged(pri.astype(np.float32))
if rank == 0:
    la_all = np.stack([(((i * 7 + int(p["board"][0]["rule50"]) * 13) % 41) - 20).astype(np.float32) * 0.25 for p in packed]).astype(np.float16)
    want = oracle.batch_priors_f16(la_all.view(np.uint16), la_all.view(np.uint16), packed, moves, offsets)
    import hashlib
    ok = np.array_equal(allp, want.astype(np.float32))
    print("RESULT", ok, len(blob), hashlib.sha1(blob).hexdigest() == hashlib.sha1(W.random_hvmw(seed=0)).hexdigest())
dist.destroy_process_group()
"""


def test_two_rank_gloo_sharded_evaluation(tmp_path):
    """world_size 2 over gloo: weight blob broadcast from rank 0, positions sharded by rank with
    no data-path collective, per-rank results concatenated in rank order equal the 1-rank run."""
    script = tmp_path / "worker.py"
    script.write_text(_WORKER.format(root=ROOT))
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", OMP_NUM_THREADS="1")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                        "--master-addr", "127.0.0.1", "--master-port", "29533", str(script)],
                       capture_output=True, text=True, env=env, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    line = [ln for ln in r.stdout.splitlines() if ln.startswith("RESULT")]
    assert line and line[0].split()[1] == "True" and line[0].split()[3] == "True", r.stdout


def test_checkpoint_conversion_cli(tmp_path):
    """Reference-format checkpoint ({'model_state_dict': ...}, trainer_agent.py:871-890) -> .hvmw
    equals folding the same tensors directly; a truncated checkpoint is refused."""
    import torch

    from hivemind_b200 import weights as W

    sd = W.random_state_dict(seed=11)
    ck = tmp_path / "model.tar"
    torch.save({"model_state_dict": {"module." + k: torch.from_numpy(v) for k, v in sd.items()}}, ck)
    out = tmp_path / "net.hvmw"
    r = subprocess.run([sys.executable, "-m", "hivemind_b200.weights", str(ck), str(out)], capture_output=True, text=True, cwd=ROOT)
    assert r.returncode == 0, r.stderr
    assert out.read_bytes() == W.write_hvmw(None, W.fold_state_dict(sd))
    sd.pop(next(iter(sd)))
    torch.save({k: torch.from_numpy(v) for k, v in sd.items()}, ck)
    with pytest.raises(KeyError):
        W.convert_checkpoint(str(ck), str(out))


def test_checkpoint_conversion_picks_the_architecture_from_the_keys(tmp_path):
    """convert_checkpoint reads architecture and head variants off the checkpoint's own keys (crossboard-risev33 has
    cross_board_blocks.*, per-board policy heads policy_heads.{0,1}.body.*, a tanh value head no body_wdl): every
    variant round-trips checkpoint -> .hvmw -> read_hvmw -> ForwardOracle and gives the outputs of folding directly."""
    import torch

    from hivemind_b200 import weights as W
    from oracle.forward_oracle import ForwardOracle

    planes = np.load(os.path.join(ROOT, "tests", "golden", "positions_seed1234_n256.npz"))["planes_f32"][:2]
    for spec in (W.crossboard_spec(), W.risev33_spec(shared_policy_trunk=False, value_head="mlp"),
                 W.risev33_spec(value_head="tanh"), W.risev33_spec()):
        sd = W.random_state_dict(spec, seed=3)
        ck = tmp_path / "model.tar"
        torch.save({"model_state_dict": {k: torch.from_numpy(v) for k, v in sd.items()}}, ck)
        out = tmp_path / "net.hvmw"
        folded = W.convert_checkpoint(str(ck), str(out))
        raw = out.read_bytes()
        assert struct.unpack_from("<I", raw, 8)[0] == spec["arch"]
        back = W.read_hvmw(raw)
        assert set(back) == set(folded) and all(np.array_equal(back[k], folded[k]) for k in back)
        a = ForwardOracle(back).forward(planes)
        b = ForwardOracle(W.fold_state_dict(sd, spec)).forward(planes)
        assert all(np.array_equal(a[k], b[k], equal_nan=True) for k in a)
        assert W.spec_from_state_dict(sd) == spec
    # a crossboard checkpoint that lost its attention output projection is named as such
    sd = W.random_state_dict(W.crossboard_spec(), seed=3)
    sd.pop("cross_board_blocks.1.cross_attention.out_proj.weight")
    torch.save({k: torch.from_numpy(v) for k, v in sd.items()}, ck)
    with pytest.raises(KeyError, match="crossboard-risev33"):
        W.convert_checkpoint(str(ck), str(out))


@pytest.mark.skipif(shutil.which("g++") is None, reason="no g++")
def test_batching_engine_header_compiles_alone_and_fails_loudly_without_gpu(tmp_path):
    """include/hm_batching_engine.hpp needs nothing of the reference; without a CUDA device loadNetwork and the
    request path return false (no CPU fallback, no hang)."""
    import torch

    from hivemind_b200 import build as hb

    lib = hb.build()
    src = tmp_path / "b.cc"
    src.write_text(r'''
#include <cstdio>
#include <vector>
#include "hm_batching_engine.hpp"
int main() {
    hm::BatchingEngine a(0, 8), b(0, 8);
    // slots are handed back when an engine dies: far more engines than slots may exist over a process lifetime
    {
        std::vector<int> taken;
        hm::BatchingBackend& be = a.backend();
        for (int round = 0; round < 3; ++round) {
            for (int i = 0; i < be.slotsTotal(); ++i) { const int s = be.acquireSlot(); if (s < 0) return 3; taken.push_back(s); }
            if (be.acquireSlot() != -1) return 4;
            for (int s : taken) be.releaseSlot(s);
            taken.clear();
        }
    }
    const bool loaded = a.loadNetwork("/nonexistent/model.hvmw", "");
    std::vector<__half> obs(8 * HM_NB_INPUT_VALUES);
    hm::BatchingEngine::HalfInferenceOutputs out;
    const bool ran = a.runInferenceHalf(obs.data(), out, 0);
    const bool synced_without_request = b.synchronizeInferenceHalf(out, 1);
    std::printf("loaded %d ran %d synced %d slots %d large %d\n", loaded, ran, synced_without_request,
                a.backend().slotsTotal(), a.backend().largeBatch());
    return 0;
}
''')
    out = tmp_path / "b"
    cmd = ["g++", "-O1", "-std=c++17", "-I" + os.path.join(ROOT, "include"), str(src), "-L" + os.path.dirname(lib), "-lhm_engine",
           "-Wl,-rpath," + os.path.dirname(lib), "-lpthread", "-o", str(out)]
    if os.path.isdir("/usr/local/cuda/include"):
        cmd.insert(4, "-I/usr/local/cuda/include")
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    r = subprocess.run(["timeout", "60", str(out)], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "loaded 0" in r.stdout and "synced 0" in r.stdout and "slots 128 large 512" in r.stdout
    if not torch.cuda.is_available():
        assert "ran 0" in r.stdout
