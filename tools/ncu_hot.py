"""Print the hottest SASS instructions (warp-stall samples) of one kernel launch in an .ncu-rep.
usage: python tools/ncu_hot.py report.ncu-rep [launch_index] [top_n]"""
import csv, subprocess, sys, io
rep = sys.argv[1]; idx = int(sys.argv[2]) if len(sys.argv) > 2 else 0; top_n = int(sys.argv[3]) if len(sys.argv) > 3 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--launch-skip", str(idx), "--launch-count", "1",
                      "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
print(rows[0][:2])
hdr = rows[1]; data = [r for r in rows[2:] if len(r) == len(hdr)]
iS = hdr.index("# Samples"); iSrc = hdr.index("Source")
stall = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
tot = sum(int(r[iS]) for r in data)
print("total samples", tot, "instrs", len(data))
agg = {}
for r in data:
    for i in stall:
        agg[hdr[i][6:]] = agg.get(hdr[i][6:], 0) + int(r[i])
print("stall mix:", sorted(agg.items(), key=lambda x: -x[1])[:8])
top = sorted(range(len(data)), key=lambda k: -int(data[k][iS]))[:top_n]
for k in sorted(top):
    r = data[k]
    st = sorted(((hdr[i][6:], int(r[i])) for i in stall if int(r[i]) > 0), key=lambda x: -x[1])[:3]
    print(k, r[iS], r[iSrc].strip()[:80], st)
# cumulative profile in 20 buckets
n = len(data); b = max(1, n // 24)
print("bucketed samples:", [(i, sum(int(r[iS]) for r in data[i:i + b])) for i in range(0, n, b)])
