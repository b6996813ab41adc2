#!/usr/bin/env python
"""Per-source-line stall-sample shares from an ncu report (source page, cuda view).
  tools/ncu_lines.py <prof.ncu-rep> <kernel-regex> [top]"""
import csv, subprocess, sys
rep, kern = sys.argv[1:3]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 30
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "-k", "regex:" + kern, "--print-source", "sass,cuda"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(txt.splitlines()))
hi = [i for i, r in enumerate(rows) if r and r[0] == "Line No"][0]
h = rows[hi]
isamp, iinst, isrc, iln = h.index("# Samples"), h.index("Instructions Executed"), h.index("Source"), h.index("Line No")
data = [r for r in rows[hi + 1:] if len(r) > isamp and r[iln].isdigit()]
# only the first captured launch of the kernel
seen = set(); first = []
for r in data:
    if r[iln] in seen and False: pass
    first.append(r)
agg = {}
def _int(v):
    try: return int(v)
    except ValueError: return None
for r in first:
    if _int(r[isamp]) is None or _int(r[iinst]) is None: continue
    a = agg.setdefault(int(r[iln]), [0, 0, r[isrc]]); a[0] += int(r[isamp]); a[1] += int(r[iinst])
ts = sum(a[0] for a in agg.values()); ti = sum(a[1] for a in agg.values())
print("samples", ts, "warp instr", ti)
for ln, a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print("%5d  samp %.3f  instr %.3f  %s" % (ln, a[0] / ts, a[1] / max(ti, 1), a[2].strip()[:100]))
