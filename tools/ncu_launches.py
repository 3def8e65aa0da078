"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: one row per kernel with the
number of launches, mean and total duration and its share of the listed time.

Usage: python tools/ncu_launches.py launches_raw.csv summary.csv "header comment"
"""
import csv
import sys
from collections import OrderedDict


def main():
    raw, out, note = sys.argv[1], sys.argv[2], sys.argv[3] if len(sys.argv) > 3 else ""
    rows = [r for r in csv.reader(open(raw, newline="")) if r]
    start = next(i for i, r in enumerate(rows) if r[0] == "ID")
    head = rows[start]
    k_name, k_metric, k_val, k_unit = (head.index(c) for c in ("Kernel Name", "Metric Name", "Metric Value",
                                                                  "Metric Unit"))
    agg = OrderedDict()
    for r in rows[start + 1:]:
        if len(r) <= k_val or r[k_metric] != "gpu__time_duration.sum":
            continue
        v = float(r[k_val].replace(",", ""))
        scale = {"ns": 1.0, "us": 1e3, "ms": 1e6, "s": 1e9}.get(r[k_unit], 1.0)
        a = agg.setdefault(r[k_name][:110], [0, 0.0])
        a[0] += 1
        a[1] += v * scale
    total = sum(a[1] for a in agg.values()) or 1.0
    with open(out, "w", newline="") as f:
        if note:
            f.write("# " + note + "\n")
        w = csv.writer(f)
        w.writerow(["kernel", "launches", "mean_ns", "total_ns", "share"])
        for name, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            w.writerow([name, n, round(t / n, 1), round(t, 1), round(t / total, 4)])


if __name__ == "__main__":
    main()
