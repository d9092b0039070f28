"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: launches, total time and share per kernel.
usage: python tools/launch_summary.py profiles/r01_launches_bench.csv"""
import collections
import csv
import sys


def main(path):
    with open(path) as f:
        lines = [l for l in f if not l.startswith("==")]
    agg = collections.OrderedDict()
    for row in csv.DictReader(lines):
        a = agg.setdefault(row["Kernel Name"], [0, 0.0])
        a[0] += 1
        a[1] += float(row["Metric Value"].replace(",", ""))
    tot = sum(a[1] for a in agg.values())
    print(f"{'launches':>8} {'total ms':>10} {'avg us':>10} {'share':>7}  kernel")
    for k, a in sorted(agg.items(), key=lambda x: -x[1][1]):
        print(f"{a[0]:8d} {a[1] / 1e6:10.3f} {a[1] / a[0] / 1e3:10.1f} {100 * a[1] / tot:6.1f}%  {k[:110]}")


if __name__ == "__main__":
    main(sys.argv[1])
