#!/usr/bin/env python
"""Compact per-launch table from `ncu -i X.ncu-rep --page raw --csv` output (what profiles/*_ncu_*.md are made from)."""
import csv
import sys

M = [("us", "gpu__time_duration.sum", 1e3), ("dram%", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", 1),
     ("MBrd", "dram__bytes_read.sum", 1), ("MBwr", "dram__bytes_write.sum", 1),
     ("l1wave%", "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed", 1),
     ("lts%", "lts__throughput.avg.pct_of_peak_sustained_elapsed", 1),
     ("issue%", "smsp__issue_active.avg.pct_of_peak_sustained_active", 1),
     ("occ%", "sm__warps_active.avg.pct_of_peak_sustained_active", 1), ("regs", "launch__registers_per_thread", 1),
     ("longsb", "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", 1),
     ("lgthr", "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio", 1),
     ("mio", "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio", 1),
     ("grid", "launch__grid_size", 1)]
rows = list(csv.reader(open(sys.argv[1])))
hdr, units, data = rows[0], rows[1], rows[2:]
ix = {h: i for i, h in enumerate(hdr)}
print(f"{'kernel':58s} " + " ".join(f"{n:>8s}" for n, _, _ in M))
for r in data:
    name = r[ix["Kernel Name"]].replace("void <unnamed>::", "").split("(")[0][:58]
    out = []
    for n, m, s in M:
        v = r[ix[m]].replace(",", "") if m in ix else ""
        u = units[ix[m]] if m in ix else ""
        try:
            f = float(v) * s
            if n.startswith("MB"):
                f = float(v) * {"Gbyte": 1e3, "Mbyte": 1, "Kbyte": 1e-3, "byte": 1e-6}.get(u, 1)
            if n == "us":
                f = float(v) * {"ms": 1e3, "us": 1, "ns": 1e-3, "s": 1e6, "msecond": 1e3, "usecond": 1, "nsecond": 1e-3, "second": 1e6}.get(u, 1)
            out.append(f"{f:8.1f}")
        except ValueError:
            out.append(f"{v[:8]:>8s}")
    print(f"{name:58s} " + " ".join(out))
