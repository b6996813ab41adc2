#!/usr/bin/env python
"""Brief per-launch summary of an .ncu-rep (ncu --set full): duration, pipes, occupancy, stall reasons.
   tools/ncu_brief.py <file.ncu-rep> [out.txt]"""
import csv, subprocess, sys
txt = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(txt.splitlines()))
h, units = rows[0], rows[1]
want = ["Kernel Name", "launch__grid_size", "launch__block_size", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "smsp__inst_executed.sum", "smsp__thread_inst_executed.sum",
        "sm__inst_executed_pipe_alu.sum", "sm__inst_executed_pipe_fma.sum",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "lts__t_sector_hit_rate.pct", "launch__occupancy_limit_shared_mem",
        "launch__occupancy_limit_registers", "launch__occupancy_limit_warps", "launch__occupancy_limit_blocks"]
stall = [c for c in h if c.startswith("smsp__average_warps_issue_stalled_") and c.endswith("_per_issue_active.ratio")]
out = []
for r in rows[2:]:
    out.append("----")
    for w in want:
        if w in h:
            i = h.index(w)
            out.append("%-72s %s %s" % (w, r[i].split("(")[0] if w == "Kernel Name" else r[i], units[i]))
    st = sorted(((float(r[h.index(c)] or 0), c) for c in stall), reverse=True)[:6]
    out.append("top stalls (warps per issue-active cycle): " + ", ".join("%s %.2f" % (c[len("smsp__average_warps_issue_stalled_"):-len("_per_issue_active.ratio")], v) for v, c in st))
s = "\n".join(out) + "\n"
if len(sys.argv) > 2:
    open(sys.argv[2], "w").write("# ncu --set full --clock-control none; one block per captured launch\n" + s)
print(s)
