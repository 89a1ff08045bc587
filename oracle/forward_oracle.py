"""forward_oracle.py — TEST INFRASTRUCTURE.  CPU fp32 restatement of the reference RISEv3.3
forward pass on the FOLDED tensors of an .hvmw container (hivemind_b200/weights.py).

Follows, layer by layer:
  RiseV3.forward                      src/architectures/rise_mobile_v3.py:206-230
  _Stem                               src/architectures/builder_util.py:154-178
  _BottlekneckResidualBlock           src/architectures/builder_util.py:445-483
  _EfficientChannelAttentionModule    src/architectures/builder_util.py:49-80
  _ValueHead (WDL + plys branch)      src/architectures/builder_util.py:246-326
  _SharedPolicyHeads                  src/architectures/rise_mobile_v3.py:38-57

Parity status: PINNED against the reference PyTorch module itself —
tests/test_forward_oracle.py (a) imports the reference model from /root/reference where that
exists and compares outputs on the same seeded weights, and (b) everywhere checks the committed
golden outputs in tests/golden/forward_risev33.npz that oracle/make_golden.py produced from the
reference model.  (The reference's own tests hold no golden network outputs)

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import
this module.
"""
from __future__ import annotations

from typing import Dict

import numpy as np
import torch
import torch.nn.functional as F


def _t(a):
    return torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32))


class ForwardOracle:
    def __init__(self, tensors: Dict[str, np.ndarray]):
        self.t = {k: _t(v) for k, v in tensors.items()}
        self.n_blocks = 0
        while f"b{self.n_blocks}.pw1.w" in self.t:
            self.n_blocks += 1

    @torch.no_grad()
    def forward(self, planes: np.ndarray, return_intermediates: bool = False, trunk_override=None):
        """planes: [B,74,8,8] float32.  Returns dict(value[B], wdl[B,3], moves_left[B],
        pi_a[B,4672], pi_b[B,4672]) as float32 numpy arrays."""
        t = self.t
        inter = {}
        x = _t(planes)
        # stem: conv3x3 + BN + ReLU (builder_util.py:171-174), BN folded
        x = F.relu(F.conv2d(x, t["stem.w"], t["stem.b"], padding=1))
        inter["stem"] = x
        for i in range(self.n_blocks):
            q = f"b{i}."
            if q + "eca.w" in t:
                # builder_util.py:77-80: avg pool -> Conv1d(k=5,pad=2) on a length-1 sequence
                # (only the centre tap meets data) -> hard sigmoid -> scale
                m = x.mean(dim=(2, 3))
                g = F.hardsigmoid(F.linear(m, t[q + "eca.w"], t[q + "eca.b"]))
                x = x * g[:, :, None, None]
            cop = t[q + "pw1.w"].shape[0]
            k = t[q + "dw.w"].shape[-1]
            # builder_util.py:465-472
            y = F.relu(F.conv2d(x, t[q + "pw1.w"][:, :, None, None], t[q + "pw1.b"]))
            y = F.relu(F.conv2d(y, t[q + "dw.w"][:, None], t[q + "dw.b"], padding=k // 2, groups=cop))
            y = F.conv2d(y, t[q + "pw2.w"][:, :, None, None], t[q + "pw2.b"])
            x = x + y  # builder_util.py:483, no activation after the add
            inter[f"block{i}"] = x
        if trunk_override is not None:   # error attribution: heads in fp32 on a given trunk output
            x = _t(trunk_override)
        cross = "tok0.wt" in t
        if cross:
            maps = self._crossboard(_t(planes), x, inter)
            x = F.relu(F.conv2d(torch.cat(maps, dim=1), t["vfuse.w"][:, :, None, None], t["vfuse.b"]))
            inter["value_fusion"] = x
        # value head (builder_util.py:311-324)
        v = F.relu(F.conv2d(x, t["value.conv.w"][:, :, None, None], t["value.conv.b"]))
        inter["value_feat"] = v
        flat = v.reshape(v.shape[0], -1)
        if "value.wdl.w" in t:
            wdl = F.linear(flat, t["value.wdl.w"], t["value.wdl.b"])
            plys = torch.sigmoid(F.linear(flat, t["value.plys.w"], t["value.plys.b"]))
            if "value.mlp0.w" in t:   # use_mlp_wdl_ply (builder_util.py:292-296, :317-319)
                hid = F.relu(F.linear(torch.cat((wdl, plys), dim=1), t["value.mlp0.w"], t["value.mlp0.b"]))
                value = torch.tanh(F.linear(hid, t["value.mlp2.w"], t["value.mlp2.b"]))[:, 0]
            else:
                sm = torch.softmax(wdl, dim=1)
                value = sm[:, 2] - sm[:, 0]
        else:   # tanh value head (builder_util.py:297-301, :326); no auxiliary heads: the evaluator reports
            # NaN wdl logits (searchthread.cc:581-583 then keeps the scalar value) and moves_left 0
            hid = F.relu(F.linear(flat, t["value.final0.w"], t["value.final0.b"]))
            value = torch.tanh(F.linear(hid, t["value.final2.w"], t["value.final2.b"]))[:, 0]
            wdl = torch.full((x.shape[0], 3), float("nan"))
            plys = torch.zeros((x.shape[0], 1))
        if cross or "pol0.c1.w" in t:
            # per-board _PolicyHead on that board's map (crossboard: rise_mobile_v3.py:336-347) or on the shared trunk
            # (risev33 with shared_policy_trunk=False: rise_mobile_v3.py:198-204, builder_util.py:206-243, :393-396)
            pis = []
            for b in range(2):
                s = F.relu(F.conv2d(maps[b] if cross else x, t[f"pol{b}.c1.w"], t[f"pol{b}.c1.b"], padding=1))
                inter[f"policy_c1_{b}"] = s
                pis.append(F.conv2d(s, t[f"pol{b}.c2.w"], None, padding=1).reshape(x.shape[0], -1))
            pi_a, pi_b = pis
        else:
            # shared policy heads (rise_mobile_v3.py:52-57)
            s = F.relu(F.conv2d(x, t["policy.shared.w"], t["policy.shared.b"], padding=1))
            inter["policy_shared"] = s
            pi_a = F.conv2d(s, t["policy.a.w"], None, padding=1).reshape(x.shape[0], -1)
            pi_b = F.conv2d(s, t["policy.b.w"], None, padding=1).reshape(x.shape[0], -1)
        out = {
            "value": value.numpy(),
            "wdl": wdl.numpy(),
            "moves_left": plys[:, 0].numpy(),
            "pi_a": pi_a.numpy(),
            "pi_b": pi_b.numpy(),
        }
        if return_intermediates:
            out["inter"] = {k: v.numpy() for k, v in inter.items()}
        return out


    def _crossboard(self, planes, trunk, inter):
        """CrossBoardRiseV3.forward between the trunk and the heads (rise_mobile_v3.py:371-401),
        written out with explicit matrix products instead of nn.MultiheadAttention."""
        t = self.t
        B, D, heads = planes.shape[0], t["pos_emb"].shape[1], 6
        dh = D // heads
        toks = []
        for b in range(2):  # :374-385 — 1x1 conv over cat(trunk, that board's 37 planes) + BN + ReLU
            w = torch.cat((t[f"tok{b}.wt"], t[f"tok{b}.wp"]), dim=1)[:, :, None, None]
            m = F.relu(F.conv2d(torch.cat((trunk, planes[:, 37 * b: 37 * b + 37]), dim=1), w, t[f"tok{b}.b"]))
            toks.append(m.flatten(2).transpose(1, 2) + t["pos_emb"][None] + t["board_emb"][b][None, None])
        inter["tokens0"] = torch.stack(toks, 1)
        # :353-369 — 4 pocket tokens + 3 scalar tokens from spatial means of input planes
        mean = planes.mean(dim=(2, 3))
        ctx = [F.linear(mean[:, lo:lo + 5], t["pocket.w"], t["pocket.b"]) for lo in (12, 17, 49, 54)]
        ctx += [F.linear(mean[:, c:c + 1], t["scalar.w"], t["scalar.b"]) for c in (25, 62, 31)]
        ctx = torch.stack(ctx, dim=1) + t["ctx_role"][None]
        inter["context"] = ctx
        n_layers = 0
        while f"xl{n_layers}.in.w" in t:
            n_layers += 1
        for l in range(n_layers):
            q_ = f"xl{l}."
            wq, wk, wv = t[q_ + "in.w"].split(D, dim=0)
            bq, bk, bv = t[q_ + "in.b"].split(D, dim=0)

            def update(board, other):  # :262-271
                mem = torch.cat((other, ctx), dim=1)
                qn = F.layer_norm(board, (D,), t[q_ + "qn.w"], t[q_ + "qn.b"])
                mn = F.layer_norm(mem, (D,), t[q_ + "mn.w"], t[q_ + "mn.b"])
                Q = F.linear(qn, wq, bq).view(B, -1, heads, dh).transpose(1, 2)
                K = F.linear(mn, wk, bk).view(B, -1, heads, dh).transpose(1, 2)
                V = F.linear(mn, wv, bv).view(B, -1, heads, dh).transpose(1, 2)
                P = torch.softmax(Q @ K.transpose(2, 3) / np.sqrt(dh), dim=-1)
                att = (P @ V).transpose(1, 2).reshape(B, -1, D)
                x1 = board + F.linear(att, t[q_ + "out.w"], t[q_ + "out.b"])
                h = F.gelu(F.linear(F.layer_norm(x1, (D,), t[q_ + "ln.w"], t[q_ + "ln.b"]), t[q_ + "fc1.w"], t[q_ + "fc1.b"]))
                return x1 + F.linear(h, t[q_ + "fc2.w"], t[q_ + "fc2.b"])

            toks = [update(toks[0], toks[1]), update(toks[1], toks[0])]  # both from pre-update tokens (:273-278)
            inter[f"tokens{l + 1}"] = torch.stack(toks, 1)
        return [tk.transpose(1, 2).reshape(B, D, 8, 8) for tk in toks]


def deployed_precision_state_dict(sd: Dict[str, np.ndarray], spec: dict) -> Dict[str, np.ndarray]:
    """The reference state_dict with every tensor-core operand weight at the precision the evaluator (and the
    reference's own fp16 TensorRT plan, engine/src/nn/engine.cc:331-354) deploys: BatchNorm is folded into the
    preceding conv in fp32 (hivemind_b200.weights.fold_state_dict), the folded conv / Linear weight is rounded to
    fp16, and the pair is written back as conv.weight = rounded weight, BN = identity scale + the folded bias
    (mean 0, var 1 - eps, gamma 1, beta = bias), so that the UNMODIFIED reference module computes
    conv_fp16w(x) + bias in fp32 arithmetic.  Depthwise taps, biases, LayerNorms, embeddings and the value-head
    Linears stay fp32, as on the device."""
    from hivemind_b200 import weights as W

    def r16(a):
        return np.asarray(a, np.float32).astype(np.float16).astype(np.float32)

    out = {k: np.array(v, copy=True) for k, v in sd.items()}

    def requant(conv_key, bn_prefix):
        w, b = W._fold(np.asarray(sd[conv_key]), sd, bn_prefix)
        out[conv_key] = r16(w)
        c = w.shape[0]
        out[bn_prefix + ".weight"] = np.ones(c, np.float32)
        out[bn_prefix + ".bias"] = b
        out[bn_prefix + ".running_mean"] = np.zeros(c, np.float32)
        out[bn_prefix + ".running_var"] = np.full(c, 1.0 - W.BN_EPS, np.float32)

    requant("body_spatial.0.body.0.weight", "body_spatial.0.body.1")
    for i, eca in enumerate(spec["eca"]):
        p = f"body_spatial.{i + 1}."
        requant(p + "body.0.weight", p + "body.1")
        requant(p + "body.6.weight", p + "body.7")
        if eca:
            out[p + "se.body.0.weight"] = r16(sd[p + "se.body.0.weight"])
    requant("value_head.body.0.weight", "value_head.body.1")
    if "policy_heads.shared_body.0.weight" in sd:
        requant("policy_heads.shared_body.0.weight", "policy_heads.shared_body.1")
        for b in range(2):
            k = f"policy_heads.board_projections.{b}.weight"
            out[k] = r16(sd[k])
    else:
        for b in range(2):
            requant(f"policy_heads.{b}.body.0.weight", f"policy_heads.{b}.body.1")
            out[f"policy_heads.{b}.body.3.weight"] = r16(sd[f"policy_heads.{b}.body.3.weight"])
    if spec["arch"] == W.ARCH_CROSSBOARD_RISEV33:
        C = spec["channels"]
        for b in range(2):
            requant(f"board_token_projections.{b}.0.weight", f"board_token_projections.{b}.1")
            # only the 384 trunk columns ride the tensor core; the 37 plane columns stay fp32 (token_embed_kernel)
            w, _ = W._fold(np.asarray(sd[f"board_token_projections.{b}.0.weight"]), sd, f"board_token_projections.{b}.1")
            out[f"board_token_projections.{b}.0.weight"][:, C:] = w[:, C:]
        requant("value_fusion.0.weight", "value_fusion.1")
        for l in range(spec["attention_layers"]):
            p = f"cross_board_blocks.{l}."
            for k in ("cross_attention.in_proj_weight", "cross_attention.out_proj.weight", "mlp.0.weight", "mlp.2.weight"):
                out[p + k] = r16(sd[p + k])
    return out


def reference_model(state_dict: Dict[str, np.ndarray]):
    """Build the UNMODIFIED reference RiseV3 (needs /root/reference; this container only)
    and load `state_dict` strictly.  Returns an eval-mode torch module."""
    import os
    import sys

    ref = os.environ.get("HM_REFERENCE", "/root/reference")
    if not os.path.isdir(ref):
        raise FileNotFoundError(ref)
    stub = os.path.join(os.path.dirname(os.path.abspath(__file__)), "timm_stub")
    for p in (ref, stub):
        if p not in sys.path:
            sys.path.insert(0, p)
    from src.architectures.rise_mobile_v3 import get_cross_board_rise_v33_model, get_rise_v33_model  # noqa: E402

    aux = "value_head.body_wdl.0.weight" in state_dict

    class Args:  # src/training/train_loop.py:95-116; the head switches follow the keys of `state_dict`
        input_shape = (74, 8, 8)
        n_labels = 0
        channels_policy_head = 73
        select_policy_from_plane = True
        use_wdl = aux
        use_plys_to_end = aux
        use_mlp_wdl_ply = aux and tuple(np.shape(state_dict["value_head.body_final.0.weight"])) == (8, 4)
        shared_policy_trunk = "policy_heads.shared_body.0.weight" in state_dict

    if "cross_board_blocks.0.query_norm.weight" in state_dict:
        model = get_cross_board_rise_v33_model(Args()).eval()  # forces shared_policy_trunk=False itself (:286-290)
    else:
        model = get_rise_v33_model(Args()).eval()
    model.load_state_dict({k: torch.from_numpy(np.asarray(v)) for k, v in state_dict.items()}, strict=True)
    return model


@torch.no_grad()
def reference_forward(model, planes: np.ndarray):
    out = model(_t(planes))
    if len(out) == 2:   # no auxiliary heads (builder_util.py:403-406): tanh value only
        value, (pi_a, pi_b) = out
        n = value.shape[0]
        return {"value": value[:, 0].numpy(), "wdl": np.full((n, 3), np.nan, np.float32),
                "moves_left": np.zeros(n, np.float32), "pi_a": pi_a.numpy(), "pi_b": pi_b.numpy()}
    value, (pi_a, pi_b), _aux, wdl, plys = out
    return {
        "value": value[:, 0].numpy(),
        "wdl": wdl.numpy(),
        "moves_left": plys[:, 0].numpy(),
        "pi_a": pi_a.numpy(),
        "pi_b": pi_b.numpy(),
    }
