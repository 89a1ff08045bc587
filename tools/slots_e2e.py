"""End-to-end throughput (C-driven request loop, host buffers) by number of worker slots in flight.
usage: python tools/slots_e2e.py [iters]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
from hivemind_b200 import weights as W
from hivemind_b200.engine import Engine
K = int(sys.argv[1]) if len(sys.argv) > 1 else 400
B = 512
pk, mv, offs = bench.workload(B)
eng = Engine(0, B, 8)
assert eng.loadNetworkFromMemory(W.random_hvmw(seed=0))
for s in (2, 3, 4, 5, 6, 8):
    eng.benchE2E(pk, mv, offs, 50, s)
    best = 0.0
    for _ in range(3):
        ms, _ = eng.benchE2E(pk, mv, offs, K, s)
        best = max(best, B * K / (ms * 1e-3))
    r, _ = eng.benchResidentMulti(K, s)
    print(f"slots {s}: e2e {best/1e3:7.1f} k positions/s   device-resident {B * K / (r * 1e-3) / 1e3:7.1f} k")
eng.close()
