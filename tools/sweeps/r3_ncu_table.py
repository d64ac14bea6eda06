#!/usr/bin/env python
"""gpurun_out/r3_ncu_*.csv (tools/sweeps/r3_ncu.sh) -> markdown rows: per kernel, the SECOND pass's launches
(the first pass of each target is cold), median duration, DRAM read + write bytes, L2 bytes, achieved DRAM GB/s."""
import collections
import csv
import json
import os
import re
import statistics
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
hbm = next((v for k, v in peak.items() if "hbm" in k.lower() and isinstance(v, (int, float))), None)
print(f"HBM peak used: {hbm} GB/s (MEASURED_PEAKS.json)" if hbm else "HBM peak: MEASURED_PEAKS.json absent")
for name in sys.argv[1:]:
    rows = collections.OrderedDict()
    for r in csv.DictReader(l for l in open(os.path.join(ROOT, "gpurun_out", f"r3_ncu_{name}.csv")) if l.startswith('"')):
        k = rows.setdefault(r["ID"], {"kernel": re.sub(r"\(.*", "", r["Kernel Name"]).split("::")[-1], "grid": r["Grid Size"]})
        v = float(r["Metric Value"].replace(",", ""))
        unit = r["Metric Unit"]
        if r["Metric Name"] == "gpu__time_duration.sum":
            v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(unit, 1.0)
        elif unit in ("Kbyte", "Mbyte", "Gbyte"):
            v *= {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[unit]
        k[r["Metric Name"]] = v
    by = collections.OrderedDict()
    for k in rows.values():
        by.setdefault((k["kernel"], k["grid"]), []).append(k)
    print(f"\n### {name}\n\n| kernel | grid | launches read | us (median) | DRAM read MB | DRAM write MB | L2 MB | DRAM GB/s | of peak |\n|---|---|---|---|---|---|---|---|---|")
    for (kern, grid), ls in by.items():
        ls = ls[len(ls) // 2:]  # second pass
        med = lambda m: statistics.median(x[m] for x in ls)  # noqa: E731
        us, rd, wr, l2 = med("gpu__time_duration.sum"), med("dram__bytes_read.sum"), med("dram__bytes_write.sum"), med("lts__t_bytes.sum")
        gbs = (rd + wr) / us / 1e3
        print(f"| `{kern}` | {grid} | {len(ls)} | {us:.2f} | {rd / 1e6:.2f} | {wr / 1e6:.2f} | {l2 / 1e6:.2f} | {gbs:.0f} | {gbs / hbm:.2f} |" if hbm else
              f"| `{kern}` | {grid} | {len(ls)} | {us:.2f} | {rd / 1e6:.2f} | {wr / 1e6:.2f} | {l2 / 1e6:.2f} | {gbs:.0f} | - |")
