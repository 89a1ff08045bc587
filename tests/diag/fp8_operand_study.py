"""fp8_operand_study.py — what e4m3 GEMM operands would do to the parity bar (BASELINE.json config 4:
"crossboard-risev33 ... fp8/bf16"; SURVEY.md §8d: "fp8 = e4m3 GEMM operands with per-channel weight scales,
fp32 accumulate").  CPU simulation on the 256 golden positions with the fp32 oracle as the reference:
the trunk's pointwise convolutions (97 % of the FLOPs) get their weights and/or input activations rounded
to the candidate format, everything else stays fp32, and the outputs are compared with the unrounded
forward under north_star's bar (max |dvalue| <= 1e-2, policy KL <= 1e-3).

    python tests/diag/fp8_operand_study.py [risev33|crossboard] [n_positions]

Test-side tool (imports oracle/); nothing here is on the product path.
"""
import os
import sys

import numpy as np
import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from hivemind_b200 import weights as W  # noqa: E402
from oracle.forward_oracle import ForwardOracle  # noqa: E402
from oracle.pyoracle import POS_DTYPE, Oracle  # noqa: E402


def q_e4m3(x, scale):
    """round(x / scale) to e4m3 (saturating at 448), back to fp32."""
    y = (x / scale).clamp(-448.0, 448.0).to(torch.float8_e4m3fn).to(torch.float32)
    return y * scale


def q_fp16(x):
    return x.to(torch.float16).to(torch.float32)


class Quantised(ForwardOracle):
    """ForwardOracle whose pw1 / pw2 see rounded operands.  wmode/amode in {None, 'fp16', 'e4m3'}."""

    def __init__(self, tensors, wmode, amode):
        super().__init__(tensors)
        self.amode = amode
        for i in range(self.n_blocks):
            for name in (f"b{i}.pw1.w", f"b{i}.pw2.w"):
                w = self.t[name]
                if wmode == "fp16":
                    self.t[name] = q_fp16(w)
                elif wmode == "e4m3":  # per-output-channel scale: amax -> 448
                    s = w.abs().amax(dim=1, keepdim=True).clamp_min(1e-12) / 448.0
                    self.t[name] = q_e4m3(w, s)

    def act(self, x):
        if self.amode == "fp16":
            return q_fp16(x)
        if self.amode == "e4m3":  # dynamic per-position scale (amax of the tensor -> 448)
            s = x.abs().amax(dim=(1, 2, 3), keepdim=True).clamp_min(1e-12) / 448.0
            return q_e4m3(x, s)
        return x

    @torch.no_grad()
    def trunk(self, planes):
        t = self.t
        x = F.relu(F.conv2d(torch.from_numpy(planes), t["stem.w"], t["stem.b"], padding=1))
        for i in range(self.n_blocks):
            q = f"b{i}."
            if q + "eca.w" in t:
                g = F.hardsigmoid(F.linear(x.mean(dim=(2, 3)), t[q + "eca.w"], t[q + "eca.b"]))
                x = x * g[:, :, None, None]
            cop, k = t[q + "pw1.w"].shape[0], t[q + "dw.w"].shape[-1]
            y = F.relu(F.conv2d(self.act(x), t[q + "pw1.w"][:, :, None, None], t[q + "pw1.b"]))
            y = F.relu(F.conv2d(y, t[q + "dw.w"][:, None], t[q + "dw.b"], padding=k // 2, groups=cop))
            y = F.conv2d(self.act(y), t[q + "pw2.w"][:, :, None, None], t[q + "pw2.b"])
            x = x + y
        return x.numpy()


def main():
    arch = sys.argv[1] if len(sys.argv) > 1 else "risev33"
    n = int(sys.argv[2]) if len(sys.argv) > 2 else 256
    g = np.load(os.path.join(ROOT, "tests", "golden", "positions_seed1234_n256.npz"))
    packed = np.ascontiguousarray(g["packed"]).view(POS_DTYPE).reshape(-1)[:n]
    offsets, moves = g["offsets"][: 2 * n + 1], g["moves"]
    spec = W.crossboard_spec() if arch == "crossboard" else W.risev33_spec()
    folded = W.fold_state_dict(W.random_state_dict(spec, seed=0), spec)
    oracle = Oracle()
    planes = oracle.encode_planes(packed, "f32")
    base = ForwardOracle(folded)
    ref = base.forward(planes)

    def priors(out):
        res = []
        for gi in range(n):
            for bn in range(2):
                lo, hi = offsets[2 * gi + bn], offsets[2 * gi + bn + 1]
                stm = int(packed[gi]["board"][bn]["flags"] & 1)
                res.append(oracle.priors_f32(out["pi_a" if bn == 0 else "pi_b"][gi], moves[lo:hi], stm).astype(np.float64))
        return res

    pref = priors(ref)
    print(f"{arch}, {n} positions; bar: max|dvalue| <= 1e-2, KL <= 1e-3")
    print(f"{'weights':8s} {'acts':8s}  max|dvalue|  mean|dvalue|   max KL     verdict")
    for wmode, amode in ((None, None), ("fp16", None), ("fp16", "fp16"), ("e4m3", None), ("e4m3", "fp16"), (None, "e4m3"),
                         ("e4m3", "e4m3")):
        qo = Quantised(folded, wmode, amode)
        out = base.forward(planes, trunk_override=qo.trunk(planes))
        dv = np.abs(out["value"] - ref["value"])
        kl = 0.0
        for p, q in zip(pref, priors(out)):
            m = p > 0
            kl = max(kl, float(np.sum(p[m] * (np.log(p[m]) - np.log(np.maximum(q[m], 1e-300))))))
        ok = dv.max() <= 1e-2 and kl <= 1e-3
        print(f"{str(wmode):8s} {str(amode):8s}  {dv.max():.3e}    {dv.mean():.3e}   {kl:.3e}   {'within' if ok else 'OVER the bar'}")


if __name__ == "__main__":
    main()
