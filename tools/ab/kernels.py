"""A/B timing helper (bring-up): per-kernel CUDA-event times of one pass, median of N passes, plus 2-slot throughput.
usage: python tools/ab/kernels.py [arch] [batch]"""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import bench
from hivemind_b200 import weights as W
from hivemind_b200.engine import Engine
arch = sys.argv[1] if len(sys.argv) > 1 else "risev33"
B = int(sys.argv[2]) if len(sys.argv) > 2 else 512
pk, mv, offs = bench.workload(B)
eng = Engine(0, B, 2)
assert eng.loadNetworkFromMemory(W.random_hvmw(seed=0, arch=W.ARCH_CROSSBOARD_RISEV33 if arch == "crossboard" else W.ARCH_RISEV33))
for w in range(2):
    assert eng.evaluatePositions(pk, mv, offs, w) is not None
eng.benchResidentMulti(50, 2)
per = {}
for _ in range(15):
    for name, ms, fl, nb in eng.profilePass(0):
        per.setdefault(name, []).append(ms)
ms2, _ = eng.benchResidentMulti(300, 2)
ms1, _ = eng.benchResident(300, 0)
print(f"{B * 300 / ms2:8.1f} pos/ms 2-slot  {B * 300 / ms1:8.1f} 1-slot | " + "  ".join(f"{k} {np.median(v):.4f}" for k, v in per.items()))
eng.close()
