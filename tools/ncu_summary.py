"""Transpose `ncu -i X.ncu-rep --page raw --csv` into one row per metric, one column per launch.

Usage: python tools/ncu_summary.py raw.csv out.csv "header comment"
"""
import csv
import sys


def main():
    raw, out, note = sys.argv[1], sys.argv[2], sys.argv[3] if len(sys.argv) > 3 else ""
    rows = [r for r in csv.reader(open(raw, newline="")) if r]
    # skip ncu's "==PROF==" chatter: the header is the first row that starts with "ID"
    start = next(i for i, r in enumerate(rows) if r[0] == "ID")
    names, units, data = rows[start], rows[start + 1], rows[start + 2:]
    with open(out, "w", newline="") as f:
        if note:
            f.write("# " + note + "\n")
        w = csv.writer(f)
        w.writerow(["metric", "unit"] + ["launch_%d" % i for i in range(len(data))])
        for j, name in enumerate(names):
            if name in ("ID", "Process ID", "Process Name", "Host Name", "Context", "Stream", "Device",
                        "CC", "Section Name", "Rule Name"):
                continue
            w.writerow([name, units[j]] + [d[j] if j < len(d) else "" for d in data])


if __name__ == "__main__":
    main()
