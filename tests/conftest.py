import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLD = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def oracle():
    from oracle.pyoracle import Oracle

    return Oracle()


@pytest.fixture(scope="session")
def ref():
    from oracle.pyoracle import Ref, ref_available

    if not ref_available():
        pytest.skip("oracle/_ref/libhmref.so not built (needs /root/reference)")
    return Ref()


@pytest.fixture(scope="session")
def gold_positions():
    from oracle.pyoracle import POS_DTYPE

    d = dict(np.load(os.path.join(GOLD, "positions_seed1234_n256.npz")))
    d["packed"] = np.ascontiguousarray(d["packed"]).view(POS_DTYPE).reshape(-1)
    return d


@pytest.fixture(scope="session")
def gold_kat():
    return dict(np.load(os.path.join(GOLD, "reference_kat.npz")))


@pytest.fixture(scope="session")
def gold_forward():
    return dict(np.load(os.path.join(GOLD, "forward_risev33.npz")))


def formula_logits(n_pos: int) -> np.ndarray:
    """Same formula as oracle/make_golden.py."""
    i = np.arange(4672, dtype=np.int64)[None, None, :]
    p = np.arange(n_pos, dtype=np.int64)[:, None, None]
    b = np.arange(2, dtype=np.int64)[None, :, None]
    v = ((i * 7 + p * 13 + b * 5) % 41 - 20).astype(np.float32) * 0.25
    return v.astype(np.float16)


def random_packed(n: int, seed: int) -> np.ndarray:
    """Synthetic packed records exercising every field (not necessarily legal chess)."""
    from oracle.pyoracle import POS_DTYPE

    rng = np.random.default_rng(seed)
    pos = np.zeros(n, POS_DTYPE)
    b = pos["board"]
    b["pieces"] = rng.integers(0, 2**63, size=(n, 2, 12), dtype=np.uint64) & rng.integers(0, 2**63, size=(n, 2, 12), dtype=np.uint64)
    b["promoted"] = rng.integers(0, 2**63, size=(n, 2, 2), dtype=np.uint64) & rng.integers(0, 2**63, size=(n, 2, 2), dtype=np.uint64)
    b["pocket"] = rng.integers(0, 17, size=(n, 2, 10), dtype=np.uint8)
    b["ep_square"] = rng.integers(0, 65, size=(n, 2), dtype=np.uint8)
    b["flags"] = rng.integers(0, 128, size=(n, 2), dtype=np.uint8)
    b["last_from"] = rng.integers(0, 66, size=(n, 2), dtype=np.uint8)
    b["last_to"] = rng.integers(0, 65, size=(n, 2), dtype=np.uint8)
    # a drop / none last_from only makes sense with a matching last_to
    none = b["last_to"] == 64
    b["last_from"][none] = 64
    b["rule50"] = rng.integers(0, 256, size=(n, 2), dtype=np.uint8)
    b["repetitions"] = rng.integers(0, 5, size=(n, 2), dtype=np.uint8)
    return pos


def kl_divergence(p, q):
    p = np.asarray(p, np.float64)
    q = np.asarray(q, np.float64)
    m = p > 0
    return float(np.sum(p[m] * (np.log(p[m]) - np.log(np.maximum(q[m], 1e-300)))))
