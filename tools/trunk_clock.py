"""Effective SM clock while trunk_kernel runs inside the steady-state bench loop: the kernel's first and last CTA
record their lifetime in SM cycles and in nanoseconds.  usage: python tools/trunk_clock.py [batch]  (on a GPU box)"""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from hivemind_b200 import weights as W
from hivemind_b200.engine import Engine
import bench

B = int(sys.argv[1]) if len(sys.argv) > 1 else 512
pk, mv, offs = bench.workload(B)
eng = Engine(0, B, 1)
assert eng.loadNetworkFromMemory(W.random_hvmw(seed=0))
for rep in range(6):
    t0 = time.time()
    n = 0
    while time.time() - t0 < 0.5:
        assert eng.evaluatePositions(pk, mv, offs) is not None
        n += 1
    c = eng.debugFetch("trunk_clk")
    print(f"rep {rep}: {n} passes; CTA 0: {c[0]:.0f} cycles in {c[1]:.0f} ns = {c[0] / c[1] * 1e3:.0f} MHz; "
          f"last CTA: {c[2]:.0f} cycles in {c[3]:.0f} ns = {c[2] / c[3] * 1e3:.0f} MHz")
# isolated passes: the GPU idles 30 ms before each one (no sustained power draw)
for rep in range(4):
    time.sleep(0.03)
    assert eng.evaluatePositions(pk, mv, offs) is not None
    c = eng.debugFetch("trunk_clk")
    print(f"isolated pass {rep}: CTA 0: {c[0]:.0f} cycles in {c[1]:.0f} ns = {c[0] / c[1] * 1e3:.0f} MHz; "
          f"last CTA: {c[2]:.0f} cycles in {c[3]:.0f} ns = {c[2] / c[3] * 1e3:.0f} MHz")
print("CTA 0 phase marks (cycles since kernel entry): prologue done %.0f, X landed %.0f, issuer starts %.0f, block 0 issued %.0f, "
      "last MMA issued %.0f, last epilogue done %.0f, output stored %.0f" % (c[4], c[9], c[5], c[10], c[6], c[7], c[8]))
for rep in range(2):
    eng.traceTrunk(0)
    c = eng.debugFetch("trunk_clk")
    print("traced launch: CTA 0 %.0f cycles (%.0f MHz), last CTA %.0f; issuer %.0f -> %.0f" % (c[0], c[0] / c[1] * 1e3, c[2], c[5], c[6]))
eng.close()
