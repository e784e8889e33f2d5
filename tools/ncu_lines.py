"""Instructions executed per source line of one kernel from an .ncu-rep (read here, no GPU).

  python tools/ncu_lines.py <rep> <kernel regex> [N] [samples]     ("samples": sort by stall samples)
"""
import csv
import io
import subprocess
import sys
from collections import defaultdict

rep, pat = sys.argv[1], sys.argv[2]
n = int(sys.argv[3]) if len(sys.argv) > 3 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + pat, "--launch-count", "1",
                      "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr = None
per_line = defaultdict(lambda: [0, 0, ""])
cur_file = ""
for r in rows:
    if r and r[0] == "File Path":
        cur_file = r[1].split("/")[-1]
    if r and r[0] == "Line No":
        hdr = r
        continue
    if hdr is None or len(r) != len(hdr):
        continue
    try:
        ln = int(r[0])
    except ValueError:
        continue
    ie = hdr.index("Instructions Executed")
    si = hdr.index("# Samples")
    try:
        inst = int(r[ie] or 0)
        smp = int(r[si] or 0)
    except ValueError:
        continue
    key = (cur_file, ln)
    per_line[key][0] += inst
    per_line[key][1] += smp
    per_line[key][2] = r[1]
tot = sum(v[0] for v in per_line.values())
tot_s = sum(v[1] for v in per_line.values())
by = 1 if (len(sys.argv) > 4 and sys.argv[4] == "samples") else 0
print("kernel", pat, "instructions executed", tot, "stall samples", tot_s, "(sorted by %s)" % ("samples" if by else "instructions"))
print(" inst%  samples%  instructions  location")
for (f, ln), (inst, smp, src) in sorted(per_line.items(), key=lambda kv: -kv[1][by])[:n]:
    print("%5.1f%% %7.1f%% %9d  %s:%d  %s" % (100.0 * inst / max(tot, 1), 100.0 * smp / max(tot_s, 1), inst, f, ln,
                                            src.strip()[:100]))
