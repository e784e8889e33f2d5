"""Turns the CSV of `ncu --metrics gpu__time_duration.sum --clock-control none --csv` over
tools/profile_step.py --adam --trainer into the per-kernel launch list kept under profiles/.

  python tools/launch_shares.py gpurun_out/launches.csv > profiles/<round>_launch_shares.txt

The last rasterizer step (from its preprocess_forward_kernel<1, 0> to the Adam kernel) and the last whole
training iteration (from preprocess_forward_kernel<2, 1>) are listed in launch order with their share of
the section. ncu times are cold-cache and serialised: compare SHARES with bench.py's stage timers."""
import csv
import re
import sys


def short(name):
    name = re.sub(r"^void ", "", name)
    name = re.sub(r"\(.*$", "", name)
    return name[:64]


def main(path):
    rows = []
    with open(path, newline="") as f:
        lines = [l for l in f if l.startswith('"')]
    for r in csv.DictReader(lines):
        if r.get("Metric Name") == "gpu__time_duration.sum":
            v = float(r["Metric Value"].replace(",", ""))
            unit = r["Metric Unit"]
            us = v / 1e3 if unit in ("ns", "nsecond") else v if unit in ("us", "usecond") else v * 1e3
            rows.append((short(r["Kernel Name"]), us))
    print("ncu launch list (gpu__time_duration.sum, --clock-control none) of tools/profile_step.py --adam --trainer")
    print("cold-cache and serialised: compare SHARES with bench.py's stage timers, not absolutes\n")
    for title, marker in (("one rasterizer fwd+bwd step through the C-ABI (+ fused Adam)", "preprocess_forward_kernel<1, 0>"),
                          ("one whole training iteration from raw parameters (render() raw mode, image loss, stats, Adam)",
                           "preprocess_forward_kernel<2, 1>")):
        starts = [i for i, (n, _) in enumerate(rows) if marker in n]
        if not starts:
            continue
        seg = rows[starts[-1]:]
        end = next((i for i, (n, _) in enumerate(seg) if "adam_multi_kernel" in n), len(seg) - 1)
        seg = seg[:end + 1]
        total = sum(us for _, us in seg)
        print(f"-- {title} --")
        for n, us in seg:
            print(f"{n:<66s} {us:8.1f} us {100 * us / total:6.1f} %")
        print(f"{'total':<66s} {total:8.1f} us\n")


if __name__ == "__main__":
    main(sys.argv[1])
