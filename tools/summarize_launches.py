"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel name."""
import csv
import re
import sys
from collections import defaultdict


def main(path, skip_first_half=True):
    rows = []
    with open(path, newline="") as f:
        lines = [l for l in f if not l.startswith("==")]
    for r in csv.DictReader(lines):
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(r["Metric Value"].replace(",", ""))
        unit = r["Metric Unit"]
        if unit in ("nsecond", "ns"):
            v *= 1e-3
        elif unit in ("msecond", "ms"):
            v *= 1e3
        rows.append((r["Kernel Name"], v))
    if skip_first_half:
        rows = rows[len(rows) // 2:]
    agg = defaultdict(lambda: [0, 0.0])
    for name, us in rows:
        short = name.split("(")[0]
        short = re.sub(r"^void ", "", short)
        short = re.sub(r"la::\(anonymous namespace\)::", "", short)
        short = re.sub(r"la::_GLOBAL__N__\w+::", "", short)
        agg[short][0] += 1
        agg[short][1] += us
    total = sum(v[1] for v in agg.values())
    print(f"{path}: {len(rows)} launches, {total/1e3:.2f} ms GPU time (serialised, cold cache)")
    for k, (c, us) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"  {us/1e3:9.3f} ms {100*us/total:5.1f}%  {c:6d} x {us/c:8.1f} us  {k}")


if __name__ == "__main__":
    for p in sys.argv[1:]:
        main(p)
