#!/usr/bin/env python
"""Turn an ncu launch-list CSV (gpu__time_duration.sum per launch) into a per-kernel share table (markdown)."""
import collections
import csv
import re
import sys


def main(path, title):
    lines = open(path).readlines()
    start = next(i for i, l in enumerate(lines) if l.startswith('"ID"'))
    agg = collections.defaultdict(lambda: [0, 0.0])
    tot = 0.0
    for r in csv.DictReader(lines[start:]):
        if r["Metric Name"] != "gpu__time_duration.sum":
            continue
        name = re.sub(r"\(.*", "", r["Kernel Name"]).replace("void ", "")
        v = float(r["Metric Value"].replace(",", ""))
        u = r["Metric Unit"]
        v = v / 1e6 if u == "ns" else (v / 1e3 if u.startswith("us") else v)
        agg[name][0] += 1
        agg[name][1] += v
        tot += v
    print(f"# {title}\n")
    print("| kernel | launches | total ms | share |\n|---|---:|---:|---:|")
    for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"| `{k}` | {n} | {t:.2f} | {100 * t / tot:.1f} % |")
    print(f"\nsum of listed launches: {tot:.2f} ms")


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2] if len(sys.argv) > 2 else sys.argv[1])
