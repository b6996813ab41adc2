#!/usr/bin/env python
"""Per-kernel instruction and DRAM counts per unit of work, from an ncu metric capture of `bench.py --profile-run`:

  python bench.py --profile-run > plain.log 2>&1 &&
  ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,smsp__thread_inst_executed.sum,sm__inst_executed_pipe_alu.sum,\\
sm__inst_executed_pipe_fma.sum,sm__inst_executed_pipe_lsu.sum,dram__bytes_read.sum,dram__bytes_write.sum \\
      --clock-control none --csv --log-file counts.csv python bench.py --profile-run
  tools/ncu_counts.py counts.csv <pairs_per_step> <steps> "<workload>" profiles/r2_ncu_counts.json

Units: one image for the extractor kernels (2 x pairs per step), one stereo pair for the k_stereo_* kernels.
bench.py divides the ALU-pipe lane-operations (warp instructions on the ALU pipe x 32) by the live event time of the stage
and by the measured ALU-pipe peak (profiles/MEASURED_INT_PEAKS.json) to get the INT roofline fraction of each kernel."""
import collections, csv, json, sys
src, pairs, steps, workload, dst = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), sys.argv[4], sys.argv[5]
rows = [r for r in csv.reader(open(src)) if len(r) > 5]
hdr = [i for i, r in enumerate(rows) if r[0] == "ID"][0]
h = rows[hdr]
ki, mi, vi, ui = h.index("Kernel Name"), h.index("Metric Name"), h.index("Metric Value"), h.index("Metric Unit")
agg = collections.OrderedDict()
for r in rows[hdr + 1:]:
    k = r[ki].split("(")[0].replace("void ", "").split("<")[0]
    v = float(r[vi].replace(",", ""))
    u = r[ui]
    if r[mi].startswith("gpu__time_duration"):
        v = v / 1e3 if u in ("ns", "nsecond") else v * 1e3 if u in ("ms", "msecond") else v      # -> us
    if r[mi].startswith("dram__bytes"):
        v = v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)
    a = agg.setdefault(k, collections.Counter())
    a[r[mi]] += v
    if r[mi].startswith("gpu__time_duration"):
        a["launches"] += 1
out = {"workload": workload, "pairs_per_step": pairs, "steps": steps,
       "source": "ncu --metrics ... --clock-control none on `bench.py --profile-run` (batch %d), summed over the %d captured steps" % (pairs, steps),
       "kernels": {}}
for k, a in agg.items():
    units = steps * pairs * (1 if k.startswith("k_stereo") else 2)
    out["kernels"][k] = {
        "launches_per_step": a["launches"] / steps,
        "unit": "stereo pair" if k.startswith("k_stereo") else "image",
        "ncu_us_per_step": a["gpu__time_duration.sum"] / steps,
        "warp_inst_per_unit": a["smsp__inst_executed.sum"] / units,
        "thread_inst_per_unit": a["smsp__thread_inst_executed.sum"] / units,
        "alu_lane_ops_per_unit": 32.0 * a["sm__inst_executed_pipe_alu.sum"] / units,
        "fma_lane_ops_per_unit": 32.0 * a["sm__inst_executed_pipe_fma.sum"] / units,
        "lsu_warp_inst_per_unit": a["sm__inst_executed_pipe_lsu.sum"] / units,
        "dram_bytes_per_unit": (a["dram__bytes_read.sum"] + a["dram__bytes_write.sum"]) / units,
    }
json.dump(out, open(dst, "w"), indent=1)
tot = sum(v["ncu_us_per_step"] for v in out["kernels"].values())
for k, v in sorted(out["kernels"].items(), key=lambda kv: -kv[1]["ncu_us_per_step"]):
    print("%-22s %8.1f us/step (%4.1f%%)  %9.0f thread-inst/unit  alu %9.0f  dram %9.0f B/unit" % (k, v["ncu_us_per_step"], 100 * v["ncu_us_per_step"] / tot,
          v["thread_inst_per_unit"], v["alu_lane_ops_per_unit"], v["dram_bytes_per_unit"]))
