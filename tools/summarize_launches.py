#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel (share of device time)."""
import collections
import csv
import re
import sys


def main(path, top=25):
    with open(path) as f:
        lines = [l for l in f if not l.startswith("==")]
    agg = collections.defaultdict(lambda: [0, 0.0])
    tot = 0.0
    for row in csv.DictReader(lines):
        if row.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(row["Metric Value"].replace(",", ""))
        v *= {"ns": 1.0, "us": 1e3, "ms": 1e6, "s": 1e9}.get(row["Metric Unit"], 1.0)
        k = re.sub(r"\(.*", "", row["Kernel Name"])
        agg[k][0] += 1
        agg[k][1] += v
        tot += v
    print(f"| kernel | launches | total ms | share |\n|---|---:|---:|---:|")
    for k, (n, v) in sorted(agg.items(), key=lambda x: -x[1][1])[:top]:
        print(f"| `{k[:90]}` | {n} | {v / 1e6:.3f} | {100 * v / tot:.2f}% |")
    print(f"\ntotal device time in the capture: {tot / 1e6:.1f} ms")


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 25)
