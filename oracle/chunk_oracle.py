"""TEST INFRASTRUCTURE ONLY (imported by tests/ and oracle/make_golden_chunk.py, never by the product).

CPU restatement of the HVM4 self-play chunk format:
  * `read_chunk`   follows the reference trainer's reader, src/preprocessing/convert_selfplay_data.py:36-146
                   (`read_exact`, `read_sparse_policy`, `read_binary_shard`) including its validity checks;
  * `chunk_bytes`  follows the reference writer, engine/src/tools/selfplay.cc:94-140 (`write_policy`, `flush`);
  * `quantize_planes` follows selfplay.cc:464-476 (`encode_planes`): u8 = lround(clamp(v, 0, 1) * 255);
  * `marginal_policy` follows selfplay.cc:400-426 after the move -> index step.

Pinned: oracle/make_golden_chunk.py feeds files written by the product writer AND by `chunk_bytes` to the
reference's own `read_binary_shard` (imported from /root/reference with a stub for the absent `polars`) and
compares field by field; the checked file is committed as tests/golden/chunk_7_000000.hvm.
"""
from __future__ import annotations

import struct
from typing import Dict, List, Sequence, Tuple

import numpy as np

NB_INPUT_CHANNELS = 74
NB_INPUT_VALUES = 74 * 64
NB_POLICY_VALUES = 73 * 64
HEADER = struct.Struct("<4sIHHQ")           # convert_selfplay_data.py:24
SAMPLE_METADATA_V4 = struct.Struct("<QIHHBBbBf")   # :26
POLICY_ENTRY = struct.Struct("<Hf")         # :28


def quantize_planes(planes_f32: np.ndarray) -> np.ndarray:
    """selfplay.cc:464-476.  std::lround rounds halves away from zero; inputs are >= 0 here."""
    v = np.clip(np.asarray(planes_f32, np.float32), np.float32(0), np.float32(1)) * np.float32(255.0)
    return np.floor(v.astype(np.float64) + 0.5).astype(np.uint8)


def marginal_policy(policy_index: Sequence[int], visits: Sequence[int]) -> Tuple[np.ndarray, np.ndarray]:
    """selfplay.cc:400-426: sum visits per index (edges with visits <= 0 skipped), ascending index order,
    probability = float(visits) / float(total)."""
    by_index: Dict[int, int] = {}
    total = 0
    for i, v in zip(policy_index, visits):
        if v <= 0:
            continue
        key = int(i) & 0xFFFF
        by_index[key] = by_index.get(key, 0) + int(v)
        total += int(v)
    if total == 0:
        raise RuntimeError("Search returned no visited root edges")
    keys = sorted(by_index)
    idx = np.array(keys, np.uint16)
    prob = np.array([np.float32(by_index[k]) / np.float32(total) for k in keys], np.float32)
    return idx, prob


def _policy_bytes(policy) -> bytes:
    idx, prob = policy
    if len(idx) > 0xFFFF:
        raise RuntimeError("Sparse policy is too large for HVM3")   # selfplay.cc:95-97
    out = [struct.pack("<H", len(idx))]
    for i, p in zip(idx, prob):
        out.append(POLICY_ENTRY.pack(int(i), float(np.float32(p))))
    return b"".join(out)


def chunk_bytes(samples: List[dict]) -> bytes:
    """selfplay.cc:116-138: header, then per sample the scalars, 4736 plane bytes and the two sparse policies.
    A sample is a dict with the TrainingSample field names (selfplay.cc:40-53)."""
    out = [HEADER.pack(b"HVM4", 4, NB_INPUT_CHANNELS, NB_POLICY_VALUES, len(samples))]
    for s in samples:
        out.append(SAMPLE_METADATA_V4.pack(s["gameId"], s["nodes"], s["macroPly"], s["movesLeft"], s["team"],
                                           s["hasTimeAdvantage"], s["outcome"], s["wdl"], float(np.float32(s["rootQ"]))))
        planes = np.asarray(s["planes"], np.uint8).reshape(-1)
        assert planes.size == NB_INPUT_VALUES
        out.append(planes.tobytes())
        out.append(_policy_bytes(s["policyA"]))
        out.append(_policy_bytes(s["policyB"]))
    return b"".join(out)


def chunk_name(run_id: int, chunk_index: int) -> str:
    """selfplay.cc:105-108."""
    return f"chunk_{run_id}_{chunk_index:06d}.hvm"


def _read_exact(f, size: int) -> bytes:
    payload = f.read(size)
    if len(payload) != size:
        raise ValueError(f"Truncated HVM3 chunk: expected {size} bytes, found {len(payload)}")
    return payload


def _read_sparse_policy(f) -> np.ndarray:
    """convert_selfplay_data.py:45-66: dense float32[4672]; must sum to 1 within 1e-4."""
    (n,) = struct.unpack("<H", _read_exact(f, 2))
    policy = np.zeros(NB_POLICY_VALUES, np.float32)
    for _ in range(n):
        index, probability = POLICY_ENTRY.unpack(_read_exact(f, POLICY_ENTRY.size))
        if index >= NB_POLICY_VALUES:
            raise ValueError(f"Policy index {index} exceeds {NB_POLICY_VALUES}")
        policy[index] = probability
    if not np.isclose(policy.sum(), 1.0, atol=1e-4):
        raise ValueError(f"Sparse policy sums to {policy.sum()}, expected 1")
    return policy


def read_chunk(path, q_value_ratio: float = 0.0) -> List[dict]:
    """convert_selfplay_data.py:69-146 for version-4 chunks; same keys as the reference's sample dicts."""
    samples = []
    with open(path, "rb") as f:
        magic, version, channels, policy_values, num_samples = HEADER.unpack(_read_exact(f, HEADER.size))
        if magic != b"HVM4" or version != 4:
            raise ValueError(f"Unsupported self-play chunk {path}: {magic!r} v{version}")
        if channels != NB_INPUT_CHANNELS:
            raise ValueError(f"Expected {NB_INPUT_CHANNELS} channels, found {channels}")
        if policy_values != NB_POLICY_VALUES:
            raise ValueError(f"Expected {NB_POLICY_VALUES} policy values, found {policy_values}")
        for _ in range(num_samples):
            (game_id, nodes, macro_ply, moves_left, team, has_time_advantage, outcome, wdl,
             root_q) = SAMPLE_METADATA_V4.unpack(_read_exact(f, SAMPLE_METADATA_V4.size))
            planes = _read_exact(f, NB_INPUT_VALUES)
            policy_a = _read_sparse_policy(f)
            policy_b = _read_sparse_policy(f)
            blended = (1.0 - q_value_ratio) * float(outcome) + q_value_ratio * float(root_q)
            samples.append({
                "x": planes, "policy_a": policy_a.tobytes(), "policy_b": policy_b.tobytes(),
                "y_value": float(blended), "root_q": float(root_q), "outcome": float(outcome),
                "y_wdl": int(wdl), "y_moves_left": int(moves_left), "game_id": int(game_id),
                "macro_ply": int(macro_ply), "team": int(team), "has_time_advantage": bool(has_time_advantage),
                "search_nodes": int(nodes),
            })
        if f.read(1):
            raise ValueError(f"Trailing bytes in chunk {path}")
    return samples


def golden_samples(fixture, n: int = 5) -> List[dict]:
    """Deterministic samples built from tests/golden/positions_seed1234_n256.npz (reference-quantised u8 planes,
    reference policy indices): the content of tests/golden/chunk_7_000000.hvm."""
    planes_u8 = fixture["planes_u8"]
    offsets = fixture["offsets"].astype(np.int64)
    policy_index = fixture["policy_index"]
    samples = []
    for g in range(n):
        pol = []
        for b in range(2):
            lo, hi = offsets[2 * g + b], offsets[2 * g + b + 1]
            idx = policy_index[lo:hi]
            keep = idx >= 0
            idx = idx[keep]
            visits = ((np.arange(idx.size) * 7 + g * 3 + b) % 11).astype(np.int64)   # some edges unvisited
            if visits.sum() == 0:
                visits[0] = 1
            pol.append(marginal_policy(idx, visits))
        samples.append({
            "gameId": 1000 + g, "nodes": 800 + g, "macroPly": 2 * g + 1, "movesLeft": 40 - g, "team": g & 1,
            "hasTimeAdvantage": (g >> 1) & 1, "outcome": (-1, 0, 1)[g % 3], "wdl": (0, 1, 2)[g % 3],
            "rootQ": float(np.float32(0.125 * g - 0.3)), "planes": planes_u8[g].reshape(-1),
            "policyA": pol[0], "policyB": pol[1],
        })
    return samples
