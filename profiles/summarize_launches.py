"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per kernel+grid totals,
and the last full frame step in launch order.  usage: summarize_launches.py file.csv [--last N]"""
import collections
import csv
import re
import sys


def load(path):
    lines = [l for l in open(path) if l.startswith('"')]
    r = csv.reader(lines)
    hdr = next(r)
    ki, vi, ui, gi = (hdr.index(k) for k in ("Kernel Name", "Metric Value", "Metric Unit", "Grid Size"))
    out = []
    for row in r:
        v = float(row[vi].replace(",", ""))
        v = v / 1000 if row[ui] == "ns" else (v * 1000 if row[ui] == "ms" else v)
        name = re.sub(r"^void ", "", row[ki])
        name = re.sub(r"\(.*", "", name).replace("ptts::", "")
        out.append((name, row[gi], v))
    return out


def main():
    seq = load(sys.argv[1])
    last = int(sys.argv[sys.argv.index("--last") + 1]) if "--last" in sys.argv else 0
    if last:
        seq = seq[-last:]
    agg = collections.defaultdict(lambda: [0, 0.0])
    for name, grid, v in seq:
        k = f"{name} g={grid}"
        agg[k][0] += 1
        agg[k][1] += v
    tot = sum(v[1] for v in agg.values())
    print(f"{len(seq)} launches, {tot:.0f} us total")
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:60]:
        print(f"{v[1]:9.1f} us  {v[1] / tot:6.1%}  n={v[0]:4d}  avg {v[1] / v[0]:7.1f}  {k}")


if __name__ == "__main__":
    main()
