#!/usr/bin/env python
"""Per-source-line hot spots of an .ncu-rep captured with --import-source on (kernels compiled -lineinfo).
   tools/ncu_lines.py <file.ncu-rep> [top_n]   -> lines sorted by line number, with their share of executed
   warp instructions and of stall samples (first captured launch of each kernel)."""
import csv, subprocess, sys, collections
rep = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(txt.splitlines()))
hdr = None; agg = collections.OrderedDict(); src = {}; fn = None; done = set()
for r in rows:
    if r and r[0] == "Function Name":
        if agg: break
        fn = r[1]; continue
    if r and r[0] == "Line No": hdr = r; continue
    if not hdr or len(r) < len(hdr): continue
    ln = r[0]
    if not ln.isdigit(): continue
    ie = hdr.index("Instructions Executed"); ss = hdr.index("# Samples")
    a = agg.setdefault(int(ln), [0, 0]); a[0] += int(r[ie]) if r[ie].isdigit() else 0; a[1] += int(r[ss]) if r[ss].isdigit() else 0
    src[int(ln)] = r[1]
ti = sum(v[0] for v in agg.values()) or 1; ts = sum(v[1] for v in agg.values()) or 1
print("#", fn); print("# total warp instructions %d, samples %d" % (ti, ts))
keep = sorted(agg, key=lambda k: -agg[k][1])[:top]
for k in sorted(keep):
    print("%5d inst %5.1f%% smp %5.1f%%  %s" % (k, 100.0 * agg[k][0] / ti, 100.0 * agg[k][1] / ts, src[k].strip()[:120]))
