"""Timeline of one trunk_kernel launch (CTA 0): where the TMA producer, the MMA issuer and a
compute warp spend their cycles.  usage: python tools/trace_trunk.py [run ...]   (on a GPU box)"""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from hivemind_b200 import weights as W
from hivemind_b200.engine import Engine
sys.path.insert(0, ROOT)
import bench

MMA = {1: "pw1.begin", 2: "pw1.d1_empty_ok", 3: "pw1.issued", 4: "pw2.begin", 5: "pw2.z_full_ok", 6: "pw2.issued"}
CMP = {1: "chunk.begin", 2: "d1_full_ok", 3: "d1_drained", 4: "math_done", 5: "z_empty_ok", 6: "z_published",
       7: "epi.begin", 8: "epi.d2_full_ok", 9: "epi.done"}


def summarize(name, ev, names, pairs):
    if not ev:
        print(name, "no events"); return
    t0 = ev[0][1]
    print(f"{name}: {len(ev)} events, span {ev[-1][1] - t0} cycles")
    for a, b, label in pairs:
        ds = []
        last = None
        for tag, t in ev:
            if tag == a: last = t
            elif tag == b and last is not None:
                ds.append(t - last); last = None
        if ds:
            ds = np.array(ds)
            print(f"   {label:34s} n={len(ds):3d} mean {ds.mean():8.0f} med {np.median(ds):8.0f} max {ds.max():8d} sum {ds.sum():9d}")


def main():
    runs = [int(a) for a in sys.argv[1:]] or [0, 3]
    B = 512
    pk, mv, offs = bench.workload(B)
    eng = Engine(0, B, 1)
    assert eng.loadNetworkFromMemory(W.random_hvmw(seed=0))
    assert eng.evaluatePositions(pk, mv, offs) is not None
    for run in runs:
        tr = eng.traceTrunk(run)
        tr = eng.traceTrunk(run)
        print(f"==== trunk run {run}")
        warp_math = {tag - 100: t for tag, t in tr[2] if tag >= 100}
        tr[2] = [(tag, t) for tag, t in tr[2] if tag < 100]
        if warp_math:
            print("   depthwise math cycles per compute warp (warp: SP = warp % 4): " +
                  ", ".join(f"w{w}(SP{w % 4}) {c}" for w, c in sorted(warp_math.items())))
        summarize("producer", tr[0], {}, [(1, 2, "W1 group -> next W2 group issue"), (2, 1, "W2 group -> next W1 group"), (2, 2, "W2 -> W2")])
        summarize("mma", tr[1], MMA, [(1, 2, "wait D1_EMPTY"), (2, 3, "pw1 issue (6 tiles, waits ring)"), (4, 5, "wait Z_FULL"), (5, 6, "pw2 issue (6 tiles)"),
                                      (3, 4, "pw1 end -> pw2 begin"), (6, 1, "pw2 end -> pw1 begin"),
                                      (10, 11, "pw1 burst 0"), (11, 12, "pw1 burst 1"), (12, 3, "pw1 burst 2"),
                                      (20, 21, "pw2 burst 0"), (21, 22, "pw2 burst 1"), (22, 6, "pw2 burst 2")])
        summarize("compute", tr[2], CMP, [(1, 2, "wait D1_FULL"), (2, 3, "TMEM drain + bias/relu"), (3, 4, "depthwise math"), (4, 5, "wait Z_EMPTY"),
                                          (5, 6, "Z store + fence + publish"), (1, 6, "whole chunk"), (7, 8, "epi wait D2_FULL"), (8, 9, "epi part")])
        # TMA latency: producer's issue of a pair (tags 10+b / 20+b) -> issuer sees it full (same tag)
        for base, nm in ((10, "W1"), (20, "W2")):
            for b in range(3):
                pi = [t for tag, t in tr[0] if tag == base + b]
                mi = [t for tag, t in tr[1] if tag == base + b]
                n = min(len(pi), len(mi))
                if n:
                    d = np.array(mi[:n]) - np.array(pi[:n])
                    print(f"   {nm} pair {b}: load issue -> seen full: mean {d.mean():7.0f} med {np.median(d):7.0f} min {d.min():6d} max {d.max():6d}")
        if os.environ.get("HM_TRACE_DUMP"):
            lo = int(os.environ.get("HM_TRACE_DUMP"))
            t0 = tr[1][0][1] if tr[1] else 0
            merged = sorted([(t - t0, role, tag) for role in range(3) for tag, t in tr[role]])
            sel = [e for e in merged if lo <= e[0] < lo + 12000]
            for t, role, tag in sel:
                print(f"   t={t:8d} {['producer', 'mma     ', 'compute '][role]} tag {tag}")
    eng.close()


if __name__ == "__main__":
    main()
