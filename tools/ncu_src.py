#!/usr/bin/env python
"""Per-SASS-chunk instruction / stall-sample shares from an ncu report's source page.
  tools/ncu_src.py <prof.ncu-rep> <kernel-regex> [chunk]"""
import csv, subprocess, sys
rep, kern = sys.argv[1:3]
chunk = int(sys.argv[3]) if len(sys.argv) > 3 else 40
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "-k", "regex:" + kern], capture_output=True, text=True).stdout
rows = list(csv.reader(txt.splitlines()))
h = rows[1]; data = rows[2:]
ia, isamp, isrc = h.index("Instructions Executed"), h.index("# Samples"), h.index("Source")
tot = sum(int(r[ia]) for r in data); totS = sum(int(r[isamp]) for r in data)
print("total warp instr", tot, "samples", totS, "n sass", len(data))
for c in range(0, len(data), chunk):
    seg = data[c:c + chunk]
    print("%5d  instr %.3f  samples %.3f  %s" % (c, sum(int(r[ia]) for r in seg) / tot, sum(int(r[isamp]) for r in seg) / max(totS, 1), seg[0][isrc][:70]))
