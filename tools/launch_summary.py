"""Summarise an ncu `--metrics gpu__time_duration.sum --csv` launch list: per-kernel time, share, count.
usage: python tools/launch_summary.py launches.csv [first_id last_id]"""
import csv
import re
import sys
from collections import defaultdict


def main():
    path = sys.argv[1]
    lo = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    hi = int(sys.argv[3]) if len(sys.argv) > 3 else 1 << 60
    rows = []
    with open(path) as f:
        lines = [ln for ln in f if ln.startswith('"')]
    for r in csv.DictReader(lines):
        if r["Metric Name"] != "gpu__time_duration.sum":
            continue
        name = r["Kernel Name"]
        name = re.sub(r"^<unnamed>::", "", name)
        name = re.sub(r"\(.*$", "", name)
        rows.append((int(r["ID"]), name, float(r["Metric Value"]) / 1e3))
    sel = [x for x in rows if lo <= x[0] <= hi]
    tot = sum(x[2] for x in sel)
    agg = defaultdict(lambda: [0.0, 0])
    for _, n, us in sel:
        agg[n][0] += us
        agg[n][1] += 1
    print(f"# launches {sel[0][0]}..{sel[-1][0]}: {len(sel)} launches, {tot / 1e3:.1f} ms")
    print(f"{'us':>10}  {'share':>5} {'n':>5} {'avg us':>8}  kernel")
    for n, (us, c) in sorted(agg.items(), key=lambda kv: -kv[1][0]):
        print(f"{us:10.1f}  {100 * us / tot:4.1f}% {c:5d} {us / c:8.1f}  {n[:100]}")


if __name__ == "__main__":
    main()
