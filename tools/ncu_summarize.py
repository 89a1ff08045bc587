"""Turn ncu outputs into the small CSV summaries kept under profiles/.

    python tools/ncu_summarize.py launches gpurun_out/launches.csv "<header comment>" > profiles/rN_launches_summary.csv
    python tools/ncu_summarize.py full gpurun_out/trunk.ncu-rep "<header comment>" > profiles/rN_trunk_ncu_summary.csv

`launches` reads the CSV log of `ncu --metrics gpu__time_duration.sum --csv --log-file ...` and
prints per-kernel launch counts, total time and share.  `full` reads a `--set full` report with
`ncu -i ... --page raw --csv` and keeps the metrics the roofline discussion uses.
"""
from __future__ import annotations

import csv
import io
import subprocess
import sys
from collections import OrderedDict

KEEP = [
    "gpu__time_duration.sum",
    "dram__bytes_read.sum",
    "dram__bytes_write.sum",
    "lts__t_bytes.sum",
    "l1tex__m_xbar2l1tex_read_bytes.sum",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "launch__grid_size",
    "launch__block_size",
    "launch__registers_per_thread",
    "launch__shared_mem_per_block_dynamic",
    "sm__cycles_elapsed.max",
    "smsp__cycles_active.avg",
]


def launches(path: str, comment: str) -> None:
    rows = [r for r in csv.reader(open(path, newline="")) if r]
    hdr = next(i for i, r in enumerate(rows) if "Kernel Name" in r and "Metric Value" in r)
    h = rows[hdr]
    kn, mn, mv, mu = h.index("Kernel Name"), h.index("Metric Name"), h.index("Metric Value"), h.index("Metric Unit")
    agg: "OrderedDict[str, list]" = OrderedDict()
    for r in rows[hdr + 1:]:
        if len(r) <= mv or r[mn] != "gpu__time_duration.sum":
            continue
        v = float(r[mv].replace(",", ""))
        unit = r[mu]
        v *= {"ns": 1.0, "us": 1e3, "ms": 1e6, "s": 1e9}.get(unit, 1.0)
        name = r[kn].split("(")[0]
        a = agg.setdefault(name, [0, 0.0])
        a[0] += 1
        a[1] += v
    total = sum(a[1] for a in agg.values())
    print(f"# {comment}")
    print("# ncu --metrics gpu__time_duration.sum --clock-control none ; per-launch times are cold-cache and serialised: compare SHARES")
    print("kernel,launches,total_ns,avg_ns,share")
    for name, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{name},{n},{t:.0f},{t / n:.0f},{t / total:.4f}")
    print(f"TOTAL,{sum(a[0] for a in agg.values())},{total:.0f},,1.0")


def full(path: str, comment: str) -> None:
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    h, units = rows[0], rows[1]
    cols = [h.index("ID"), h.index("Kernel Name")] + [h.index(k) for k in KEEP if k in h]
    print(f"# {comment}")
    print(",".join(h[c] for c in cols))
    print("# units: " + ",".join(units[c] for c in cols))
    for r in rows[2:]:
        print(",".join(r[c].split("(")[0].replace(",", "") if c == cols[1] else r[c].replace(",", "") for c in cols))


if __name__ == "__main__":
    {"launches": launches, "full": full}[sys.argv[1]](sys.argv[2], sys.argv[3] if len(sys.argv) > 3 else "")
