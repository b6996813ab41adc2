#!/usr/bin/env python
"""Instruction / stall-sample shares of line ranges of one kernel from an ncu report.
  tools/ncu_regions.py <prof.ncu-rep> <kernel-regex> name:lo-hi [name:lo-hi ...]"""
import csv, subprocess, sys
rep, kern = sys.argv[1:3]
regs = []
for a in sys.argv[3:]:
    n, r = a.split(":"); lo, hi = r.split("-"); regs.append((n, int(lo), int(hi)))
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "-k", "regex:" + kern, "--print-source", "sass,cuda"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(txt.splitlines()))
hi = [i for i, r in enumerate(rows) if r and r[0] == "Line No"][0]
h = rows[hi]
isamp, iinst, isrc, iln = h.index("# Samples"), h.index("Instructions Executed"), h.index("Source"), h.index("Line No")
agg = {}; other = {}
cur = ""
for r in rows:
    if r and r[0] == "File Path": cur = r[1]; continue
    try: ln = int(r[iln]); a = int(r[isamp]); b = int(r[iinst])
    except (ValueError, IndexError): continue
    if cur.endswith(".cu"): x = agg.setdefault(ln, [0, 0])
    else: x = other.setdefault(cur.split("/")[-1] + ":" + str(ln), [0, 0])
    x[0] += a; x[1] += b
ti = sum(v[1] for v in agg.values()) + sum(v[1] for v in other.values()); ts = sum(v[0] for v in agg.values()) + sum(v[0] for v in other.values())
print("warp instr", ti, "samples", ts)
for n, lo, hi_ in regs:
    print("%-12s instr %.3f  samples %.3f" % (n, sum(v[1] for k, v in agg.items() if lo <= k <= hi_) / ti, sum(v[0] for k, v in agg.items() if lo <= k <= hi_) / ts))
for k, v in sorted(other.items(), key=lambda kv: -kv[1][1])[:8]:
    print("%-40s instr %.3f  samples %.3f" % (k, v[1] / ti, v[0] / ts))
