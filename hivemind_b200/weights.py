"""Weights ingress for the B200 leaf evaluator.

* ``risev33_spec()``            the layer list of the reference's deployed network
  (src/architectures/rise_mobile_v3.py:820-849 ``get_rise_v33_model`` with the arguments of
  src/training/train_loop.py:95-116: WDL + plys heads, shared policy trunk).
* ``random_state_dict()``       seeded random parameters under the reference's own
  ``state_dict`` names and shapes (so ``model.load_state_dict(strict=True)`` accepts them);
  BatchNorm statistics are randomised so that a broken kernel cannot hide behind
  near-identity normalisation (SURVEY.md §7 hard part 5).
* ``fold_state_dict()``         BatchNorm folding (eval mode, eps 1e-5) into conv weight + bias,
  ECA centre tap extraction (builder_util.py:67-80: Conv1d on a length-1 sequence).
* ``write_hvmw()/read_hvmw()``  the flat container hm_engine_load reads (csrc/hm_net.cu).

numpy only; no torch import.
"""
from __future__ import annotations

import struct
from typing import Dict, List

import numpy as np

ARCH_RISEV33 = 0
ARCH_CROSSBOARD_RISEV33 = 1
BN_EPS = 1e-5


def risev33_spec(shared_policy_trunk: bool = True, value_head: str = "wdl") -> dict:
    """``shared_policy_trunk``: one 3x3 384->384 policy trunk + two 73-channel projections
    (rise_mobile_v3.py:38-57, what train_loop.py:115 deploys) or, when False, two independent
    ``_PolicyHead``s (builder_util.py:206-243, rise_mobile_v3.py:198-204).  ``value_head``:
    "wdl" = WDL + plys heads, value = p_win - p_loss (builder_util.py:311-324, deployed);
    "mlp" = the same heads plus the small value MLP on cat(wdl, plys) (``use_mlp_wdl_ply``, :292-296, :317-319);
    "tanh" = no auxiliary heads, value = tanh(Linear(512,1)(relu(Linear(1024,512)))) (:297-301, :326)."""
    if value_head not in ("wdl", "mlp", "tanh"):
        raise ValueError(f"unknown value head {value_head!r}")
    kernels = [3] * 15
    for i in (7, 11, 12, 13):
        kernels[i] = 5
    eca = [False] * 15
    for i in (5, 8, 12, 13, 14):
        eca[i] = True
    cops: List[int] = []
    cop = 256
    for idx, k in enumerate(kernels):
        # rise_mobile_v3.py:66-72: 5x5 blocks use channels_operating - 32 * (idx // 2)
        cops.append(cop - 32 * (idx // 2) if k == 5 else cop)
        cop += 64
    return {
        "arch": ARCH_RISEV33,
        "in_channels": 74,
        "channels": 384,
        "kernels": kernels,
        "eca": eca,
        "cops": cops,
        "value_channels": 16,
        "value_fc": 512,
        "policy_channels": 73,
        "shared_policy_trunk": bool(shared_policy_trunk),
        "value_head": value_head,
    }


def crossboard_spec() -> dict:
    """``crossboard-risev33`` (rise_mobile_v3.py:852-888, class CrossBoardRiseV3 :281-412): the
    same stem + 15 blocks, then board tokens, 2 cross-board attention layers, value fusion and
    per-board 192-channel policy heads."""
    spec = risev33_spec(shared_policy_trunk=False)   # forced by the class itself, rise_mobile_v3.py:286-290
    spec.update({"arch": ARCH_CROSSBOARD_RISEV33, "attention_dim": 192, "attention_heads": 6,
                 "attention_layers": 2, "board_channels": 37, "mlp_ratio": 2})
    return spec


def spec_for_arch(arch: int) -> dict:
    return crossboard_spec() if arch == ARCH_CROSSBOARD_RISEV33 else risev33_spec()


def spec_from_state_dict(sd) -> dict:
    """The layer list a reference ``state_dict`` was built with, from its key names and shapes."""
    if "value_head.body_wdl.0.weight" in sd:
        value_head = "mlp" if tuple(np.shape(sd["value_head.body_final.0.weight"])) == (8, 4) else "wdl"
    else:
        value_head = "tanh"
    if any(k.startswith("cross_board_blocks.0.") for k in sd):
        spec = crossboard_spec()
    else:
        spec = risev33_spec(shared_policy_trunk="policy_heads.shared_body.0.weight" in sd)
    spec["value_head"] = value_head
    return spec


def _bn(rng, c, prefix, sd, gamma_lo=0.5, gamma_hi=1.5):
    sd[prefix + ".weight"] = rng.uniform(gamma_lo, gamma_hi, c).astype(np.float32)
    sd[prefix + ".bias"] = (0.1 * rng.standard_normal(c)).astype(np.float32)
    sd[prefix + ".running_mean"] = (0.1 * rng.standard_normal(c)).astype(np.float32)
    sd[prefix + ".running_var"] = rng.uniform(0.5, 1.5, c).astype(np.float32)
    sd[prefix + ".num_batches_tracked"] = np.array(0, np.int64)


def _conv(rng, shape, gain=2.0):
    fan_in = int(np.prod(shape[1:]))
    return (rng.standard_normal(shape) * np.sqrt(gain / fan_in)).astype(np.float32)


def random_state_dict(spec: dict | None = None, seed: int = 0) -> Dict[str, np.ndarray]:
    """Seeded parameters keyed exactly like the reference RiseV3 ``state_dict``."""
    spec = spec or risev33_spec()
    rng = np.random.default_rng(seed)
    sd: Dict[str, np.ndarray] = {}
    C = spec["channels"]
    sd["body_spatial.0.body.0.weight"] = _conv(rng, (C, spec["in_channels"], 3, 3))
    _bn(rng, C, "body_spatial.0.body.1", sd)
    for i, (k, cop, eca) in enumerate(zip(spec["kernels"], spec["cops"], spec["eca"])):
        p = f"body_spatial.{i + 1}."
        if eca:
            sd[p + "se.body.0.weight"] = (rng.standard_normal((C, C, 5)) * (6.0 / np.sqrt(C))).astype(np.float32)
            sd[p + "se.body.0.bias"] = (1.5 + 1.0 * rng.standard_normal(C)).astype(np.float32)
        sd[p + "body.0.weight"] = _conv(rng, (cop, C, 1, 1))
        _bn(rng, cop, p + "body.1", sd)
        sd[p + "body.3.weight"] = _conv(rng, (cop, 1, k, k))
        _bn(rng, cop, p + "body.4", sd)
        sd[p + "body.6.weight"] = _conv(rng, (C, cop, 1, 1), gain=1.0)
        # residual branch kept modest so 15 blocks do not blow the trunk up
        _bn(rng, C, p + "body.7", sd, gamma_lo=0.3, gamma_hi=0.9)
    vc = spec["value_channels"]
    sd["value_head.body.0.weight"] = _conv(rng, (vc, C, 1, 1))
    _bn(rng, vc, "value_head.body.1", sd)
    nf = vc * 64
    vh = spec.get("value_head", "wdl")
    if vh != "tanh":
        sd["value_head.body_wdl.0.weight"] = (rng.standard_normal((3, nf)) * (1.0 / np.sqrt(nf))).astype(np.float32)
        sd["value_head.body_wdl.0.bias"] = (0.1 * rng.standard_normal(3)).astype(np.float32)
        sd["value_head.body_plys.0.weight"] = (rng.standard_normal((1, nf)) * (1.5 / np.sqrt(nf))).astype(np.float32)
        sd["value_head.body_plys.0.bias"] = (0.1 * rng.standard_normal(1)).astype(np.float32)
    fc = spec["value_fc"]
    if vh == "mlp":   # builder_util.py:292-296: Linear(4,8) -> act -> Linear(8,1) -> tanh on cat(wdl, plys)
        sd["value_head.body_final.0.weight"] = (rng.standard_normal((8, 4)) * 0.6).astype(np.float32)
        sd["value_head.body_final.0.bias"] = (0.2 * rng.standard_normal(8)).astype(np.float32)
        sd["value_head.body_final.2.weight"] = (rng.standard_normal((1, 8)) * 0.3).astype(np.float32)
        sd["value_head.body_final.2.bias"] = (0.1 * rng.standard_normal(1)).astype(np.float32)
    else:
        # in the "wdl" configuration these tensors exist in the state_dict but are never executed (builder_util.py:320-324)
        sd["value_head.body_final.0.weight"] = (rng.standard_normal((fc, nf)) / np.sqrt(nf)).astype(np.float32)
        sd["value_head.body_final.0.bias"] = (0.1 * rng.standard_normal(fc)).astype(np.float32)
        sd["value_head.body_final.2.weight"] = (rng.standard_normal((1, fc)) / np.sqrt(fc)).astype(np.float32)
        if vh == "tanh":   # keeps |value| spread ~0.2 so that fp16 weight rounding alone stays well inside the 1e-2 bar
            sd["value_head.body_final.2.weight"] *= np.float32(0.5)
        sd["value_head.body_final.2.bias"] = (0.1 * rng.standard_normal(1)).astype(np.float32)
    pc = spec["policy_channels"]
    if spec["arch"] == ARCH_CROSSBOARD_RISEV33:
        _random_crossboard(rng, spec, sd)
        return sd
    if spec.get("shared_policy_trunk", True):
        sd["policy_heads.shared_body.0.weight"] = _conv(rng, (C, C, 3, 3))
        _bn(rng, C, "policy_heads.shared_body.1", sd)
        for b in range(2):
            sd[f"policy_heads.board_projections.{b}.weight"] = _conv(rng, (pc, C, 3, 3), gain=1.2)
    else:   # two _PolicyHead(384) (builder_util.py:206-243): conv3x3 + BN + act, conv3x3 -> 73 planes
        for b in range(2):
            sd[f"policy_heads.{b}.body.0.weight"] = _conv(rng, (C, C, 3, 3))
            _bn(rng, C, f"policy_heads.{b}.body.1", sd)
            sd[f"policy_heads.{b}.body.3.weight"] = _conv(rng, (pc, C, 3, 3), gain=1.2)
    return sd


def _random_crossboard(rng, spec, sd):
    """Parameters of CrossBoardRiseV3 beyond the trunk and value head, under the reference's
    state_dict names (rise_mobile_v3.py:299-351)."""
    C, D, pc, bc = spec["channels"], spec["attention_dim"], spec["policy_channels"], spec["board_channels"]
    f32 = np.float32
    for b in range(2):
        sd[f"policy_heads.{b}.body.0.weight"] = _conv(rng, (D, D, 3, 3))
        _bn(rng, D, f"policy_heads.{b}.body.1", sd)
        sd[f"policy_heads.{b}.body.3.weight"] = _conv(rng, (pc, D, 3, 3), gain=1.2)
    for b in range(2):
        sd[f"board_token_projections.{b}.0.weight"] = _conv(rng, (D, C + bc, 1, 1))
        _bn(rng, D, f"board_token_projections.{b}.1", sd)
    sd["position_embedding"] = (0.3 * rng.standard_normal((1, 64, D))).astype(f32)
    sd["board_embedding"] = (0.3 * rng.standard_normal((2, 1, D))).astype(f32)
    sd["pocket_projection.weight"] = (rng.standard_normal((D, 5)) * 2.0).astype(f32)
    sd["pocket_projection.bias"] = (0.2 * rng.standard_normal(D)).astype(f32)
    sd["scalar_projection.weight"] = (rng.standard_normal((D, 1))).astype(f32)
    sd["scalar_projection.bias"] = (0.2 * rng.standard_normal(D)).astype(f32)
    sd["context_role_embedding"] = (0.3 * rng.standard_normal((1, 7, D))).astype(f32)
    H = D * spec["mlp_ratio"]
    for l in range(spec["attention_layers"]):
        p = f"cross_board_blocks.{l}."
        for n in ("query_norm", "memory_norm", "mlp_norm"):
            sd[p + n + ".weight"] = rng.uniform(0.7, 1.3, D).astype(f32)
            sd[p + n + ".bias"] = (0.1 * rng.standard_normal(D)).astype(f32)
        sd[p + "cross_attention.in_proj_weight"] = (rng.standard_normal((3 * D, D)) * (1.5 / np.sqrt(D))).astype(f32)
        sd[p + "cross_attention.in_proj_bias"] = (0.1 * rng.standard_normal(3 * D)).astype(f32)
        sd[p + "cross_attention.out_proj.weight"] = (rng.standard_normal((D, D)) / np.sqrt(D)).astype(f32)
        sd[p + "cross_attention.out_proj.bias"] = (0.1 * rng.standard_normal(D)).astype(f32)
        sd[p + "mlp.0.weight"] = (rng.standard_normal((H, D)) * np.sqrt(2.0 / D)).astype(f32)
        sd[p + "mlp.0.bias"] = (0.1 * rng.standard_normal(H)).astype(f32)
        sd[p + "mlp.2.weight"] = (rng.standard_normal((D, H)) / np.sqrt(H)).astype(f32)
        sd[p + "mlp.2.bias"] = (0.1 * rng.standard_normal(D)).astype(f32)
    sd["value_fusion.0.weight"] = _conv(rng, (C, 2 * D, 1, 1))
    _bn(rng, C, "value_fusion.1", sd)


def _fold(w: np.ndarray, sd, bn_prefix: str):
    """conv (no bias) followed by eval-mode BatchNorm -> (w', b')."""
    g = sd[bn_prefix + ".weight"].astype(np.float64)
    b = sd[bn_prefix + ".bias"].astype(np.float64)
    m = sd[bn_prefix + ".running_mean"].astype(np.float64)
    v = sd[bn_prefix + ".running_var"].astype(np.float64)
    s = g / np.sqrt(v + BN_EPS)
    wf = w.astype(np.float64) * s.reshape(-1, *([1] * (w.ndim - 1)))
    return wf.astype(np.float32), (b - m * s).astype(np.float32)


def fold_state_dict(sd: Dict[str, np.ndarray], spec: dict | None = None) -> Dict[str, np.ndarray]:
    """Reference state_dict -> the folded fp32 tensors of the .hvmw container.  The head variants
    (shared / per-board policy heads; WDL, MLP or tanh value head) are read off the key names, so the
    container says by its tensor names what the evaluator has to run (csrc/hm_net.cu build_net)."""
    sd = {k: np.asarray(v) for k, v in sd.items()}
    spec = spec or spec_from_state_dict(sd)
    out: Dict[str, np.ndarray] = {}
    out["stem.w"], out["stem.b"] = _fold(sd["body_spatial.0.body.0.weight"], sd, "body_spatial.0.body.1")
    for i, (k, cop, eca) in enumerate(zip(spec["kernels"], spec["cops"], spec["eca"])):
        p = f"body_spatial.{i + 1}."
        q = f"b{i}."
        w, b = _fold(sd[p + "body.0.weight"], sd, p + "body.1")
        out[q + "pw1.w"], out[q + "pw1.b"] = w.reshape(cop, -1), b
        w, b = _fold(sd[p + "body.3.weight"], sd, p + "body.4")
        out[q + "dw.w"], out[q + "dw.b"] = w.reshape(cop, k, k), b
        w, b = _fold(sd[p + "body.6.weight"], sd, p + "body.7")
        out[q + "pw2.w"], out[q + "pw2.b"] = w.reshape(w.shape[0], cop), b
        if eca:
            ew = sd[p + "se.body.0.weight"]
            out[q + "eca.w"] = np.ascontiguousarray(ew[:, :, ew.shape[2] // 2]).astype(np.float32)
            out[q + "eca.b"] = sd[p + "se.body.0.bias"].astype(np.float32)
    w, b = _fold(sd["value_head.body.0.weight"], sd, "value_head.body.1")
    out["value.conv.w"], out["value.conv.b"] = w.reshape(w.shape[0], -1), b
    f32 = np.float32
    if "value_head.body_wdl.0.weight" in sd:
        out["value.wdl.w"] = sd["value_head.body_wdl.0.weight"].astype(f32)
        out["value.wdl.b"] = sd["value_head.body_wdl.0.bias"].astype(f32)
        out["value.plys.w"] = sd["value_head.body_plys.0.weight"].astype(f32)
        out["value.plys.b"] = sd["value_head.body_plys.0.bias"].astype(f32)
        if tuple(sd["value_head.body_final.0.weight"].shape) == (8, 4):   # use_mlp_wdl_ply
            out["value.mlp0.w"] = sd["value_head.body_final.0.weight"].astype(f32)
            out["value.mlp0.b"] = sd["value_head.body_final.0.bias"].astype(f32)
            out["value.mlp2.w"] = sd["value_head.body_final.2.weight"].astype(f32)
            out["value.mlp2.b"] = sd["value_head.body_final.2.bias"].astype(f32)
        # else: body_final exists in the state_dict but is never executed (builder_util.py:320-324)
    else:   # tanh value head (builder_util.py:297-301, :326)
        out["value.final0.w"] = sd["value_head.body_final.0.weight"].astype(f32)
        out["value.final0.b"] = sd["value_head.body_final.0.bias"].astype(f32)
        out["value.final2.w"] = sd["value_head.body_final.2.weight"].astype(f32)
        out["value.final2.b"] = sd["value_head.body_final.2.bias"].astype(f32)
    if spec["arch"] == ARCH_CROSSBOARD_RISEV33:
        _fold_crossboard(sd, spec, out)
        return out
    if "policy_heads.shared_body.0.weight" in sd:
        out["policy.shared.w"], out["policy.shared.b"] = _fold(
            sd["policy_heads.shared_body.0.weight"], sd, "policy_heads.shared_body.1"
        )
        out["policy.a.w"] = sd["policy_heads.board_projections.0.weight"].astype(f32)
        out["policy.b.w"] = sd["policy_heads.board_projections.1.weight"].astype(f32)
    else:   # per-board _PolicyHead(384): same tensor names as the crossboard per-board heads
        for b in range(2):
            w, bias = _fold(sd[f"policy_heads.{b}.body.0.weight"], sd, f"policy_heads.{b}.body.1")
            out[f"pol{b}.c1.w"], out[f"pol{b}.c1.b"] = w, bias
            out[f"pol{b}.c2.w"] = sd[f"policy_heads.{b}.body.3.weight"].astype(f32)
    return out


def _fold_crossboard(sd, spec, out):
    C, D = spec["channels"], spec["attention_dim"]
    f32 = np.float32
    for b in range(2):
        w, bias = _fold(sd[f"board_token_projections.{b}.0.weight"], sd, f"board_token_projections.{b}.1")
        w = w.reshape(D, -1)
        out[f"tok{b}.wt"] = np.ascontiguousarray(w[:, :C])   # trunk channels
        out[f"tok{b}.wp"] = np.ascontiguousarray(w[:, C:])   # that board's 37 input planes
        out[f"tok{b}.b"] = bias
        w, bias = _fold(sd[f"policy_heads.{b}.body.0.weight"], sd, f"policy_heads.{b}.body.1")
        out[f"pol{b}.c1.w"], out[f"pol{b}.c1.b"] = w, bias
        out[f"pol{b}.c2.w"] = sd[f"policy_heads.{b}.body.3.weight"].astype(f32)
    out["pos_emb"] = sd["position_embedding"].reshape(64, D).astype(f32)
    out["board_emb"] = sd["board_embedding"].reshape(2, D).astype(f32)
    out["ctx_role"] = sd["context_role_embedding"].reshape(7, D).astype(f32)
    out["pocket.w"] = sd["pocket_projection.weight"].astype(f32)
    out["pocket.b"] = sd["pocket_projection.bias"].astype(f32)
    out["scalar.w"] = sd["scalar_projection.weight"].astype(f32)
    out["scalar.b"] = sd["scalar_projection.bias"].astype(f32)
    for l in range(spec["attention_layers"]):
        p, q = f"cross_board_blocks.{l}.", f"xl{l}."
        for src, dst in (("query_norm", "qn"), ("memory_norm", "mn"), ("mlp_norm", "ln")):
            out[q + dst + ".w"] = sd[p + src + ".weight"].astype(f32)
            out[q + dst + ".b"] = sd[p + src + ".bias"].astype(f32)
        out[q + "in.w"] = sd[p + "cross_attention.in_proj_weight"].astype(f32)
        out[q + "in.b"] = sd[p + "cross_attention.in_proj_bias"].astype(f32)
        out[q + "out.w"] = sd[p + "cross_attention.out_proj.weight"].astype(f32)
        out[q + "out.b"] = sd[p + "cross_attention.out_proj.bias"].astype(f32)
        out[q + "fc1.w"] = sd[p + "mlp.0.weight"].astype(f32)
        out[q + "fc1.b"] = sd[p + "mlp.0.bias"].astype(f32)
        out[q + "fc2.w"] = sd[p + "mlp.2.weight"].astype(f32)
        out[q + "fc2.b"] = sd[p + "mlp.2.bias"].astype(f32)
    w, bias = _fold(sd["value_fusion.0.weight"], sd, "value_fusion.1")
    out["vfuse.w"], out["vfuse.b"] = w.reshape(C, 2 * D), bias


def write_hvmw(path_or_none, tensors: Dict[str, np.ndarray], arch: int = ARCH_RISEV33) -> bytes:
    """Serialise folded tensors.  Layout (little endian): 64-byte header
    ``"HVMW", u32 version=1, u32 arch, u32 n, u64 data_offset``; n entries of 104 bytes
    ``name[64], u32 dtype(0=f32), u32 ndim, u32 dims[4], u64 offset, u64 nbytes``; data."""
    names = list(tensors.keys())
    table_end = 64 + 104 * len(names)
    data_off = (table_end + 255) // 256 * 256
    entries, blobs, off = [], [], data_off
    for n in names:
        a = np.ascontiguousarray(tensors[n], np.float32)
        if a.ndim > 4 or len(n.encode()) > 63:
            raise ValueError(f"cannot store tensor {n}")
        dims = list(a.shape) + [0] * (4 - a.ndim)
        entries.append(struct.pack("<64sII4IQQ", n.encode(), 0, a.ndim, *dims, off, a.nbytes))
        pad = (-a.nbytes) % 256
        blobs.append(a.tobytes() + b"\0" * pad)
        off += a.nbytes + pad
    header = struct.pack("<4sIIIQ", b"HVMW", 1, arch, len(names), data_off).ljust(64, b"\0")
    raw = header + b"".join(entries)
    raw = raw.ljust(data_off, b"\0") + b"".join(blobs)
    if path_or_none is not None:
        with open(path_or_none, "wb") as f:
            f.write(raw)
    return raw


def read_hvmw(raw_or_path) -> Dict[str, np.ndarray]:
    if isinstance(raw_or_path, (bytes, bytearray, memoryview)):
        raw = bytes(raw_or_path)
    else:
        with open(raw_or_path, "rb") as f:
            raw = f.read()
    magic, version, arch, n, _ = struct.unpack_from("<4sIIIQ", raw, 0)
    if magic != b"HVMW" or version != 1:
        raise ValueError("not an HVMW v1 container")
    out: Dict[str, np.ndarray] = {}
    for i in range(n):
        name, dtype, ndim, d0, d1, d2, d3, off, nbytes = struct.unpack_from("<64sII4IQQ", raw, 64 + 104 * i)
        dims = (d0, d1, d2, d3)[:ndim]
        out[name.rstrip(b"\0").decode()] = np.frombuffer(raw, np.float32, nbytes // 4, off).reshape(dims).copy()
    return out


def random_hvmw(seed: int = 0, path=None, arch: int = ARCH_RISEV33, spec: dict | None = None) -> bytes:
    """Seeded random-init weights as an .hvmw blob (what bench.py loads)."""
    spec = spec or spec_for_arch(arch)
    return write_hvmw(path, fold_state_dict(random_state_dict(spec, seed), spec), spec["arch"])


def convert_checkpoint(src: str, dst: str) -> Dict[str, np.ndarray]:
    """Reference training checkpoint -> .hvmw.  `src` is a file written by the reference's
    ``save_torch_state`` (src/training/trainer_agent.py:871-890: a dict with
    ``model_state_dict``) or a bare ``state_dict``; a ``module.`` / ``_orig_mod.`` prefix left by
    DataParallel / torch.compile wrappers is dropped.  Fails on missing or mis-shaped tensors."""
    import torch

    ckpt = torch.load(src, map_location="cpu", weights_only=True)
    sd = ckpt.get("model_state_dict", ckpt) if isinstance(ckpt, dict) else ckpt
    clean: Dict[str, np.ndarray] = {}
    for k, v in sd.items():
        for prefix in ("module.", "_orig_mod."):
            if k.startswith(prefix):
                k = k[len(prefix):]
        clean[k] = v.detach().to(torch.float32).cpu().numpy()
    # the architecture and its head variants are read off the checkpoint's own keys: crossboard-risev33 has
    # cross_board_blocks.*, per-board policy heads have policy_heads.{0,1}.body.*, a tanh value head has no body_wdl
    spec = spec_from_state_dict(clean)
    want = random_state_dict(spec, seed=0)
    missing = sorted(set(want) - set(clean))
    if missing:
        name = "crossboard-risev33" if spec["arch"] == ARCH_CROSSBOARD_RISEV33 else "risev33"
        raise KeyError(f"checkpoint lacks {len(missing)} tensors of {name}, e.g. {missing[:3]}")
    bad = [k for k in want if tuple(clean[k].shape) != tuple(want[k].shape)]
    if bad:
        raise ValueError(f"shape mismatch for {bad[:3]}")
    folded = fold_state_dict(clean, spec)
    write_hvmw(dst, folded, spec["arch"])
    return folded


if __name__ == "__main__":
    import sys

    if len(sys.argv) == 3:
        convert_checkpoint(sys.argv[1], sys.argv[2])
        print(f"wrote {sys.argv[2]}")
    elif len(sys.argv) == 4 and sys.argv[1] == "--random":
        random_hvmw(int(sys.argv[2]), sys.argv[3])
        print(f"wrote {sys.argv[3]}")
    else:
        print("usage: python -m hivemind_b200.weights <checkpoint.tar|state_dict.pt> <out.hvmw>\n"
              "       python -m hivemind_b200.weights --random <seed> <out.hvmw>")
        sys.exit(2)
