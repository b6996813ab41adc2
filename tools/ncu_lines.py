#!/usr/bin/env python
"""Per-source-line hot spots of an .ncu-rep captured with --import-source on (kernels compiled -lineinfo).
   tools/ncu_lines.py <file.ncu-rep> [top_n] [file-substring]
   -> lines (of every source file inlined into the first captured kernel, or only the files matching the substring) with their share
   of executed warp instructions and of stall samples, hottest top_n by samples, printed in file / line order."""
import csv, os, subprocess, sys, collections
rep = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 40; only = sys.argv[3] if len(sys.argv) > 3 else ""
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
hdr = None; agg = collections.OrderedDict(); src = {}; fn = None; cur = ""
def I(s): return int(s) if s.isdigit() else 0
for r in csv.reader(txt.splitlines()):
    if r and r[0] == "File Path": cur = os.path.basename(r[1]); continue
    if r and r[0] == "Function Name":
        if fn is None: fn = r[1]
        elif r[1] != fn: break
        continue
    if r and r[0] == "Line No": hdr = r; continue
    if not hdr or len(r) < len(hdr) or not r[0].isdigit(): continue
    ie = hdr.index("Instructions Executed"); ss = hdr.index("# Samples")
    k = (cur, int(r[0]))
    a = agg.setdefault(k, [0, 0]); a[0] += I(r[ie]); a[1] += I(r[ss])
    src[k] = r[1]
ti = sum(v[0] for v in agg.values()) or 1; ts = sum(v[1] for v in agg.values()) or 1
print("#", fn); print("# total warp instructions %d, samples %d" % (ti, ts))
keys = [k for k in agg if only in k[0]]
keep = sorted(keys, key=lambda k: -agg[k][1])[:top]
for k in sorted(keep):
    print("%-22s %5d inst %5.1f%% smp %5.1f%%  %s" % (k[0][:22], k[1], 100.0 * agg[k][0] / ti, 100.0 * agg[k][1] / ts, src[k].strip()[:110]))
