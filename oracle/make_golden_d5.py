"""make_golden_d5.py — generates tests/golden/value_shaping.npz FROM THE REFERENCE'S OWN COMPILED CODE.

SURVEY.md §8a row D5: the value shaping of SearchThread::process_batch (engine/src/search/searchthread.cc:569-621)
is inline in a function that needs a whole search around it.  `make -C oracle d5` compiles the UNMODIFIED
searchthread.cc / node.cc against a scripted Engine (oracle/scripted/nn/engine.h); oracle/d5_pin.cc runs a
one-leaf search per case and reads the value the reference backed up into the root.  This script drives it over
random and hand-picked (value, wdl, moves_left) fp16 inputs, incl. non-finite ones, under several runtime
configurations, and stores inputs + reference outputs.  Runs only where /root/reference exists.

    python oracle/make_golden_d5.py
"""
from __future__ import annotations

import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle.pyoracle import Ref, ValueConfig  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")

# (enable_wdl_eval, wdl_value_weight, moves_left_discount, draw_contempt); the first is search_params.h:122-128,148
CONFIGS = [(1, 0.25, 0.005, 0.0), (1, 1.0, 0.0, 0.1), (0, 0.25, 0.005, 0.0), (1, 0.6, 0.3, 0.05), (1, 1.7, 2.5, -0.2)]


def cases(seed: int = 7, n_random: int = 320) -> np.ndarray:
    rng = np.random.default_rng(seed)
    f = np.zeros((n_random, 5), np.float32)
    f[:, 0] = rng.uniform(-1.3, 1.3, n_random)              # value, some outside [-1, 1]
    f[:, 1:4] = rng.normal(0.0, 2.5, (n_random, 3))         # wdl logits
    f[:, 4] = rng.uniform(-0.3, 1.4, n_random)              # moves_left, some outside [0, 1]
    special = [
        [0.5, 0, 0, 0, 0.5], [1.0, -20, 0, 20, 0.0], [-1.0, 20, 0, -20, 1.0], [0.0, 0, 10, 0, 0.3],
        [np.nan, 1, 2, 3, 0.2], [np.inf, 1, 2, 3, 0.2], [-np.inf, 1, 2, 3, 0.2],
        [0.3, np.nan, 0, 0, 0.2], [0.3, 0, np.inf, 0, 0.2], [0.3, 0, 0, -np.inf, 0.2], [0.3, np.inf, np.inf, np.inf, 0.2],
        [0.3, 65504, 65504, 65504, 0.2], [0.3, -65504, -65504, -65504, 0.2], [0.3, 65504, -65504, 0, 0.2],
        [0.3, 1, 2, 3, np.nan], [0.3, 1, 2, 3, np.inf], [0.3, 1, 2, 3, -np.inf], [0.3, 1, 2, 3, 5.0], [0.3, 1, 2, 3, -5.0],
        [np.nan, np.nan, np.nan, np.nan, np.nan], [2.0, 0, 0, 5, 0.0], [-2.0, 5, 0, 0, 0.0],
        [6e-8, 0, 0, 0, 6e-8], [-0.0, -0.0, 0.0, -0.0, -0.0],
    ]
    allc = np.concatenate([f, np.array(special, np.float32)])
    with np.errstate(over="ignore", invalid="ignore"):
        return allc.astype(np.float16).view(np.uint16)


def main():
    ref = Ref()
    c = cases()
    out = {"cases_u16": c, "configs": np.array(CONFIGS, np.float32)}
    for i, (en, w, disc, contempt) in enumerate(CONFIGS):
        out[f"shaped_{i}"] = ref.d5_shape(c, ValueConfig(en, w, disc, contempt))
    np.savez_compressed(os.path.join(GOLD, "value_shaping.npz"), **out)
    print("value_shaping.npz:", c.shape[0], "cases x", len(CONFIGS), "configs;",
          "non-finite inputs:", int((~np.isfinite(c.view(np.float16).astype(np.float32))).any(axis=1).sum()))


if __name__ == "__main__":
    main()
