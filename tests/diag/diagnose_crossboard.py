"""Stage-by-stage check of the crossboard-risev33 CUDA path against the fp32 oracle (verbose).
Usage:  python tests/diag/diagnose_crossboard.py [n_positions]"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests", "diag"))

from diagnose import stats  # noqa: E402
from hivemind_b200 import weights as W  # noqa: E402
from hivemind_b200.engine import Engine  # noqa: E402
from oracle.forward_oracle import ForwardOracle  # noqa: E402
from oracle.pyoracle import Oracle, POS_DTYPE  # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 32
    g = np.load(os.path.join(ROOT, "tests", "golden", "positions_seed1234_n256.npz"))
    packed = np.ascontiguousarray(g["packed"]).view(POS_DTYPE).reshape(-1)[:n]
    offsets = g["offsets"][: 2 * n + 1]
    moves = g["moves"][: offsets[-1]]
    spec = W.crossboard_spec()
    folded = W.fold_state_dict(W.random_state_dict(spec, 0), spec)
    oracle = Oracle()
    planes = oracle.encode_planes(packed, "f32")
    ref = ForwardOracle(folded).forward(planes, return_intermediates=True)
    eng = Engine(0, n, 1)
    t = time.time()
    assert eng.loadNetworkFromMemory(W.write_hvmw(None, folded, spec["arch"])), "load failed"
    print(f"load+graph capture {time.time() - t:.2f}s", flush=True)
    if os.environ.get("HM_CHECKED"):
        print("  checked pass:", eng.profilePass(), flush=True)
    res = eng.evaluatePositions(packed, moves, offsets)
    assert res is not None, "evaluate failed"
    priors, values, raw = res
    I = ref["inter"]
    stats("trunk", eng.debugFetch("trunk").reshape(n, 384, 8, 8), I["block14"])
    stats("context", eng.debugFetch("context").reshape(n, 7, 192), I["context"])
    for l in (1, 2):  # the tokens0 buffer is reused for the last layer's output
        want = I[f"tokens{l}"]  # [n, 2, 64, 192] token-major
        got = eng.debugFetch(f"tokens{l}").reshape(n, 2, 192, 64).transpose(0, 1, 3, 2)
        stats(f"tokens{l}", got, want)
    stats("value_fusion", eng.debugFetch("value_fusion").reshape(n, 384, 8, 8), I["value_fusion"])
    stats("value_feat", eng.debugFetch("value_feat").reshape(n, 16, 8, 8), I["value_feat"])
    c1 = eng.debugFetch("policy_shared").reshape(n, 2, 192, 8, 8)
    stats("policy_c1_0", c1[:, 0], I["policy_c1_0"])
    stats("policy_c1_1", c1[:, 1], I["policy_c1_1"])
    lg = eng.debugFetch("logits").reshape(n, 2, 4672)
    stats("pi_a", lg[:, 0], ref["pi_a"])
    stats("pi_b", lg[:, 1], ref["pi_b"])
    stats("value", raw[:, 0], ref["value"])
    stats("wdl", raw[:, 1:4], ref["wdl"])
    stats("moves_left", raw[:, 4], ref["moves_left"])
    # error attribution: fp32 oracle heads on OUR trunk output
    ours_trunk = eng.debugFetch("trunk").reshape(n, 384, 8, 8)
    mixed = ForwardOracle(folded).forward(planes, trunk_override=ours_trunk)
    stats("value|our trunk", mixed["value"], ref["value"])
    stats("value vs mixed", raw[:, 0], mixed["value"])
    kls = []
    for i in range(n):
        for bn in range(2):
            lo, hi = offsets[2 * i + bn], offsets[2 * i + bn + 1]
            stm = int(packed[i]["board"][bn]["flags"] & 1)
            p_ref = oracle.priors_f32(ref["pi_a" if bn == 0 else "pi_b"][i], moves[lo:hi], stm)
            p = priors[lo:hi]
            m = p_ref > 0
            kls.append(float(np.sum(p_ref[m] * (np.log(p_ref[m]) - np.log(np.maximum(p[m], 1e-30))))))
    print(f"  priors KL(ref||ours): max {max(kls):.3e} mean {np.mean(kls):.3e}", flush=True)
    prof = eng.profilePass()
    for p in prof:
        print(f"    {p[0]:24s} {p[1]:.4f} ms")
    print(f"  profile pass: {len(prof)} launches, {sum(p[1] for p in prof):.3f} ms total (n={n})")
    eng.close()


if __name__ == "__main__":
    main()
