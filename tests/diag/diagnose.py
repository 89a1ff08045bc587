"""Stage-by-stage GPU bring-up check (verbose; the pass/fail version lives in tests/).

Compares every inspectable activation of the CUDA path with the fp32 oracle, first on a
truncated network, then on the full network.  HM_TRUNK_STOP=k in the environment stops the trunk
after block k-1 (0 = stem only) so the intermediate activation is inspectable.
Usage:  python tests/diag/diagnose.py [n_blocks ...]
"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from hivemind_b200 import weights as W  # noqa: E402
from hivemind_b200.engine import Engine  # noqa: E402
from oracle.forward_oracle import ForwardOracle  # noqa: E402
from oracle.pyoracle import Oracle, POS_DTYPE  # noqa: E402


def stats(name, got, want):
    got = np.asarray(got, np.float64)
    want = np.asarray(want, np.float64)
    d = np.abs(got - want)
    rel = d.max() / (np.abs(want).max() + 1e-12)
    print(f"  {name:16s} max|d| {d.max():.4e}  mean|d| {d.mean():.4e}  rel-to-max {rel:.3e}  "
          f"ref std {want.std():.3f}  got std {got.std():.3f}  nan {int(np.isnan(got).sum())}", flush=True)
    return d.max()


def run(n_blocks, folded_full, packed, moves, offsets, oracle):
    keep = {k: v for k, v in folded_full.items() if not k.startswith("b") or int(k[1:k.index(".")]) < n_blocks}
    blob = W.write_hvmw(None, keep)
    n = packed.shape[0]
    planes = oracle.encode_planes(packed, "f32")
    ref = ForwardOracle(keep).forward(planes, return_intermediates=True)
    eng = Engine(0, n, 1)
    t = time.time()
    assert eng.loadNetworkFromMemory(blob), "load failed"
    print(f"[{n_blocks} blocks] load+graph capture {time.time() - t:.2f}s", flush=True)
    if os.environ.get("HM_CHECKED"):
        # stage the inputs without running the graph is not possible; run launch by launch
        print("  checked pass:", eng.profilePass(), flush=True)
    res = eng.evaluatePositions(packed, moves, offsets)
    assert res is not None, "evaluate failed"
    priors, values, raw = res
    stats("planes", eng.debugFetch("planes").reshape(n, 74, 8, 8), planes)
    stop = os.environ.get("HM_TRUNK_STOP")
    last = n_blocks - 1 if stop is None else min(n_blocks, int(stop)) - 1
    want_trunk = ref["inter"]["stem"] if last < 0 else ref["inter"][f"block{last}"]
    stats("stem" if last < 0 else f"block{last}", eng.debugFetch("trunk").reshape(n, 384, 8, 8), want_trunk)
    if stop is None:
        stats("value_feat", eng.debugFetch("value_feat").reshape(n, 16, 8, 8), ref["inter"]["value_feat"])
        stats("policy_shared", eng.debugFetch("policy_shared").reshape(n, 384, 8, 8), ref["inter"]["policy_shared"])
    lg = eng.debugFetch("logits").reshape(n, 2, 4672)
    stats("pi_a", lg[:, 0], ref["pi_a"])
    stats("pi_b", lg[:, 1], ref["pi_b"])
    stats("value", raw[:, 0], ref["value"])
    stats("wdl", raw[:, 1:4], ref["wdl"])
    stats("moves_left", raw[:, 4], ref["moves_left"])
    # priors vs oracle softmax of the fp32 oracle logits
    want = oracle.batch_priors_f16  # noqa (dense-half entry point checked in tests)
    kls = []
    for g in range(n):
        for bn in range(2):
            lo, hi = offsets[2 * g + bn], offsets[2 * g + bn + 1]
            stm = int(packed[g]["board"][bn]["flags"] & 1)
            p_ref = oracle.priors_f32(ref["pi_a" if bn == 0 else "pi_b"][g], moves[lo:hi], stm)
            p = priors[lo:hi]
            m = p_ref > 0
            kls.append(float(np.sum(p_ref[m] * (np.log(p_ref[m]) - np.log(np.maximum(p[m], 1e-30))))))
    print(f"  priors KL(ref||ours): max {max(kls):.3e} mean {np.mean(kls):.3e}", flush=True)
    prof = eng.profilePass()
    tot = sum(p[1] for p in prof)
    print(f"  profile pass: {len(prof)} launches, {tot:.3f} ms total (n={n})")
    eng.close()


def main():
    blocks = [int(a) for a in sys.argv[1:]] or [1, 15]
    gold = np.load(os.path.join(ROOT, "tests", "golden", "positions_seed1234_n256.npz"))
    n = int(os.environ.get("HM_DIAG_N", "64"))
    packed = np.ascontiguousarray(gold["packed"]).view(POS_DTYPE).reshape(-1)[:n]
    offsets = gold["offsets"][: 2 * n + 1]
    moves = gold["moves"][: offsets[-1]]
    oracle = Oracle()
    folded = W.fold_state_dict(W.random_state_dict(seed=0))
    for nb in blocks:
        run(nb, folded, packed, moves, offsets, oracle)


if __name__ == "__main__":
    main()
