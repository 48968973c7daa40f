"""Summarise an ncu launch list (gpu__time_duration.sum CSV) into per-kernel launches / total us / share for ONE
training step: the launches between the last two `adam_kernel` launches in the file.

    python tools/launch_summary.py gpurun_out/launches.csv [--all]
"""
import csv
import re
import sys


def short(name):
    name = re.sub(r"^void ", "", name)
    name = re.sub(r"\(.*\)$", "", name)
    return name[:110]


def main(path, all_rows=False):
    rows = []
    with open(path) as f:
        lines = [ln for ln in f if ln.startswith('"')]
    for r in csv.DictReader(lines):
        if r.get("Metric Name") == "gpu__time_duration.sum":
            rows.append((short(r["Kernel Name"]), float(r["Metric Value"]) / 1e3))
    ends = [i for i, (n, _) in enumerate(rows) if n.startswith("adam_kernel")]
    if not all_rows and len(ends) >= 2:
        rows = rows[ends[-2] + 1: ends[-1] + 1]
    agg = {}
    for n, us in rows:
        a = agg.setdefault(n, [0, 0.0])
        a[0] += 1
        a[1] += us
    tot = sum(a[1] for a in agg.values())
    print(f"# {len(rows)} launches, sum {tot / 1e3:.3f} ms")
    print("kernel,launches,total_us,share")
    for n, (c, us) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f'"{n}",{c},{us:.1f},{us / tot:.4f}')


if __name__ == "__main__":
    main(sys.argv[1], "--all" in sys.argv)
