"""Aggregate an ncu `--metrics gpu__time_duration.sum` launch list by kernel: python tools/launch_summary.py file.csv
Prints count, total and mean time per kernel (names shortened) and the share of the whole capture."""
import csv, re, sys
from collections import OrderedDict

rows = list(csv.DictReader([l for l in open(sys.argv[1]) if not l.startswith("==")]))
agg, total = OrderedDict(), 0.0
for r in rows:
    t = float(r["Metric Value"].replace(",", "")); u = r["Metric Unit"]
    t = t / 1000 if u in ("ns", "nsecond") else (t * 1000 if u in ("ms", "msecond") else t)
    name = re.sub(r"\(.*", "", r["Kernel Name"]).replace("void ", "").replace("tkb::", "")
    key = (name, r["Grid Size"], r["Block Size"])
    a = agg.setdefault(key, [0, 0.0]); a[0] += 1; a[1] += t; total += t
print(f"{len(rows)} launches, {total:.1f} us in total (cold-cache, serialised ncu replays: shares, not absolutes)")
for (name, grid, block), (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{name[:58]:58s} {grid:>14} {block:>12} x{n:<4d} {t:9.1f} us  mean {t / n:7.1f}  {100 * t / total:5.1f}%")
tc = sum(t for (name, _, _), (_, t) in agg.items() if name.startswith("tc_"))
print(f"tcgen05 kernels (tc_*): {tc:.1f} us = {100 * tc / total:.1f}% of the capture")
