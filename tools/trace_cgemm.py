"""Bring-up: per-item timeline of CTA 0 of one cgemm launch (build with HM_EXTRA_NVCC_FLAGS=-DHM_BRINGUP).
usage: HM_CG_TRACE=<op name> python tools/trace_cgemm.py [arch] [batch]    (op names: stem, policy.shared, xl0.fc1 ...)"""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from hivemind_b200 import _lib, weights as W  # noqa: E402
from hivemind_b200.engine import Engine  # noqa: E402

CAP = 512
ARCH = sys.argv[1] if len(sys.argv) > 1 else "risev33"
B = int(sys.argv[2]) if len(sys.argv) > 2 else 512
pk, mv, offs = bench.workload(B)
eng = Engine(0, B, 1)
assert eng.loadNetworkFromMemory(W.random_hvmw(seed=0, arch=W.ARCH_CROSSBOARD_RISEV33 if ARCH == "crossboard" else W.ARCH_RISEV33))
L = _lib.load()
for _ in range(5):
    assert eng.evaluatePositions(pk, mv, offs, 0) is not None
# zero the buffer by reading after a fresh pass: marks are overwritten from index 0 by every launch
assert eng.evaluatePositions(pk, mv, offs, 0) is not None
buf = np.zeros(3 * CAP, np.uint64)
fn = L.hm_debug_cg_trace
fn.restype = C.c_int
fn.argtypes = [C.c_void_p, C.c_int]
assert fn(buf.ctypes.data, buf.size) == 1
roles = []
for r in range(3):
    a = buf[r * CAP:(r + 1) * CAP]
    a = a[a != 0]
    roles.append([(int(x >> np.uint64(48)), int(x & np.uint64(0xffffffffffff))) for x in a])
t0 = min(t for r in roles for _, t in r) if any(roles) else 0
names = ["producer", "mma", "epilogue"]
for r, ev in enumerate(roles):
    print(names[r], " ".join(f"{tag}@{t - t0}" for tag, t in ev[:80]))
# per item: mma issue span, drained wait, epilogue span
mma = roles[1]
starts = [t for tag, t in mma if tag == 1]
ready = [t for tag, t in mma if tag == 2]
ends = [t for tag, t in mma if tag == 4]
epi_w = [t for tag, t in roles[2] if tag == 2]
epi_d = [t for tag, t in roles[2] if tag == 3]
epi_e = [t for tag, t in roles[2] if tag == 4]
for i in range(min(len(ends), len(epi_e), 12)):
    print(f"item {i}: wait drained {ready[i] - starts[i]:6d}  issue {ends[i] - ready[i]:6d}  "
          f"| accum seen +{epi_w[i] - ends[i]:5d}  drain {epi_d[i] - epi_w[i]:5d}  stores {epi_e[i] - epi_d[i]:5d}")
if ends:
    print("total cycles", max(epi_e) - t0 if epi_e else ends[-1] - t0)
eng.close()
