#!/usr/bin/env python
"""Summarise ncu outputs into small text files for profiles/.
  tools/ncu_summary.py launches <launches.csv> <out.txt>      per-kernel totals and shares
  tools/ncu_summary.py full <prof.ncu-rep> <out.txt>          key metrics per captured launch
"""
import collections, csv, subprocess, sys

mode, src, dst = sys.argv[1:4]
out = []
if mode == "launches":
    rows = [r for r in csv.reader(open(src)) if len(r) > 5]
    hdr = [i for i, r in enumerate(rows) if r[0] == "ID"][0]
    h = rows[hdr]; data = rows[hdr + 1:]
    ki, vi, ui = h.index("Kernel Name"), h.index("Metric Value"), h.index("Metric Unit")
    agg = collections.OrderedDict()
    for r in data:
        v = float(r[vi].replace(",", ""))
        v = v / 1e3 if r[ui] in ("ns", "nsecond") else v * 1e3 if r[ui] in ("ms", "msecond") else v
        k = r[ki].split("(")[0]
        a = agg.setdefault(k, [0, 0.0]); a[0] += 1; a[1] += v
    tot = sum(v[1] for v in agg.values())
    out.append("# gpu__time_duration.sum per kernel (ncu --clock-control none; cold-cache, serialised: compare SHARES)")
    out.append("%-24s %6s %12s %8s %10s" % ("kernel", "n", "total_us", "share", "avg_us"))
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        out.append("%-24s %6d %12.1f %8.3f %10.1f" % (k, v[0], v[1], v[1] / tot, v[1] / v[0]))
else:
    txt = subprocess.run(["ncu", "-i", src, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(txt.splitlines()))
    h = rows[0]
    want = ["Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
            "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
            "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
            "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size",
            "launch__block_size", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "lts__t_sector_hit_rate.pct"]
    idx = [(w, h.index(w)) for w in want if w in h]
    units = rows[1]
    out.append("# ncu --set full --clock-control none; one block per captured launch")
    for r in rows[2:]:
        for w, i in idx:
            out.append("%-64s %s %s" % (w, r[i].split("(")[0] if w == "Kernel Name" else r[i], units[i]))
        out.append("")
open(dst, "w").write("\n".join(out) + "\n")
print("\n".join(out[:40]))
