"""Per-launch DRAM traffic of each stage from an `ncu --set full` report -> profiles/traffic.json.

  python tools/ncu_traffic.py gpurun_out/prof.ncu-rep [profiles/traffic.json]

bench.py reads the JSON to fill `roofline.traffic` (dram__bytes_read.sum + dram__bytes_write.sum
per launch of the dominant kernel). Runs here, without a GPU.
"""
import csv
import io
import json
import subprocess
import sys
from collections import defaultdict

STAGE_OF = {
    "preprocess_forward_kernel": "preprocess_fwd", "scan_tiles_kernel": "scan",
    "duplicate_words_kernel": "duplicate_keys", "bucket_histogram_kernel": "duplicate_keys",
    "duplicate_packed_kernel": "duplicate_keys", "msd_partition_kernel": "sort_histogram",
    "bucket_sort_kernel": "sort_passes", "packed_onesweep_kernel": "sort_passes",
    "blend_forward3_kernel": "blend_fwd", "blend_forward2_kernel": "blend_fwd",
    "blend_backward_kernel": "blend_bwd", "blend_backward2_kernel": "blend_bwd", "bucket_split_kernel": "sort_passes",
    "duplicate_big_kernel": "duplicate_keys", "depth_words_kernel": "depth_order",
    "preprocess_backward_kernel": "preprocess_bwd", "adam_multi_kernel": "adam",
}

rep = sys.argv[1]
dst = sys.argv[2] if len(sys.argv) > 2 else "profiles/traffic.json"
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr, units = rows[0], rows[1]
kn, rd, wr = hdr.index("Kernel Name"), hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
issue = hdr.index("smsp__issue_active.avg.pct_of_peak_sustained_active")
active = hdr.index("sm__warps_active.avg.pct_of_peak_sustained_active")
def col(name):
    return hdr.index(name) if name in hdr else None


inst_c = col("smsp__inst_executed.sum")
shw_c = col("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum")
shp_c = col("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed")
red_c = col("l1tex__t_requests_pipe_lsu_mem_global_op_red.sum")
dur_c = col("gpu__time_duration.sum")
acc = defaultdict(list)
util = defaultdict(list)
extra = defaultdict(list)
for r in rows[2:]:
    name = r[kn].split("(")[0].split("<")[0].replace("void ", "").replace("egs::", "").strip()
    stage = STAGE_OF.get(name)
    if stage is None:
        continue
    # the backward blend of the bench step has all three upstream gradients (<PPL, 0, 1>); the training
    # iteration's launch without depth / alpha gradients (<PPL, 0, 0>) is another kernel
    if name == "blend_backward_kernel" and ", 0>(" in r[kn].split("(")[0] + "(":
        continue
    acc[(stage, name)].append(float(r[rd]) * scale[units[rd]] + float(r[wr]) * scale[units[wr]])
    util[(stage, name)].append((float(r[issue]), float(r[active])))
    extra[(stage, name)].append(tuple(float(r[c].replace(",", "")) if c is not None and r[c] not in ("", "n/a") else None
                                      for c in (inst_c, shw_c, shp_c, red_c, dur_c)))
result = {"source": rep, "unit": "bytes per launch (dram__bytes_read.sum + dram__bytes_write.sum)", "stages": {}}
for (stage, name), v in sorted(acc.items()):
    u = util[(stage, name)]
    ex = extra[(stage, name)]

    def mean(i):
        vals = [x[i] for x in ex if x[i] is not None]
        return sum(vals) / len(vals) if vals else None
    result["stages"].setdefault(stage, {})[name] = {
        "launches_captured": len(v), "bytes_per_launch": sum(v) / len(v),
        "issue_active_pct": round(sum(x[0] for x in u) / len(u), 1),     # smsp__issue_active, % of peak
        "warps_active_pct": round(sum(x[1] for x in u) / len(u), 1),     # achieved occupancy
        "warp_instructions_per_launch": mean(0),                          # smsp__inst_executed.sum
        "shared_wavefronts_per_launch": mean(1),                          # l1tex__data_pipe_lsu_wavefronts_mem_shared.sum
        "shared_wavefronts_pct": mean(2),                                 # ... as % of the pipe's peak over the launch
        "global_red_requests_per_launch": mean(3),                        # l1tex__t_requests_pipe_lsu_mem_global_op_red.sum
        "duration_us_under_ncu": mean(4)}
json.dump(result, open(dst, "w"), indent=1)
print(json.dumps(result, indent=1))
