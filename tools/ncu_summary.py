"""Summarise an .ncu-rep (read here, no GPU): duration, DRAM bytes, throughputs, top stall reasons.

  python tools/ncu_summary.py gpurun_out/prof.ncu-rep > profiles/<name>.txt
"""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr, units = rows[0], rows[1]
base = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "launch__grid_size", "launch__block_size", "smsp__inst_executed.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active"]
stalls = [h for h in hdr if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio")]
for r in rows[2:]:
    print("==", r[hdr.index("Kernel Name")][:90])
    for b in base:
        if b in hdr:
            i = hdr.index(b)
            print("   %-62s %s %s" % (b, r[i], units[i]))
    top = sorted(((float(r[hdr.index(s)] or 0), s[len("smsp__average_warps_issue_stalled_"):-len("_per_issue_active.ratio")])
                  for s in stalls), reverse=True)[:6]
    print("   stalls (warps per issue-active cycle):", ", ".join("%s %.2f" % (n, v) for v, n in top))
