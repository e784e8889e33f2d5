"""Top stalled SASS lines of one kernel from an .ncu-rep source page (read here, no GPU).

  python tools/ncu_source_top.py <rep> <kernel regex> [N]
"""
import csv
import io
import subprocess
import sys

rep, pat = sys.argv[1], sys.argv[2]
n = int(sys.argv[3]) if len(sys.argv) > 3 else 30
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + pat,
                      "--launch-count", "1"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr = None
data = []
for r in rows:
    if r and r[0] == "Address":
        if hdr is not None:
            break
        hdr = r
    elif hdr is not None and len(r) == len(hdr):
        data.append(r)
si, src = hdr.index("# Samples"), hdr.index("Source")
tot = sum(int(r[si] or 0) for r in data)
print("kernel", pat, "total samples", tot, "instructions", len(data))
top = sorted(enumerate(data), key=lambda x: -int(x[1][si] or 0))[:n]
for i, r in sorted(top):
    print("%5d %6s %5.1f%%  %s" % (i, r[si], 100.0 * int(r[si] or 0) / max(tot, 1), r[src][:110]))
