"""Sum an ncu `--metrics gpu__time_duration.sum --csv` launch list per kernel.
usage: python tools/launch_shares.py launches.csv > shares.txt"""
import collections
import csv
import re
import sys

lines = [l for l in open(sys.argv[1]) if not l.startswith("==")]
tot = collections.defaultdict(lambda: [0, 0.0, []])
for row in csv.DictReader(lines):
    v = float(row["Metric Value"].replace(",", ""))
    v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(row["Metric Unit"], 1.0)
    name = re.sub(r"\(.*", "", row["Kernel Name"])
    t = tot[name]
    t[0] += 1
    t[1] += v
    t[2].append(v)
total = sum(t[1] for t in tot.values())
print(f"# {sum(t[0] for t in tot.values())} launches, {total / 1e3:.3f} ms of GPU time")
for name, t in sorted(tot.items(), key=lambda kv: -kv[1][1]):
    print(f"{name[:62]:62s} n={t[0]:4d} total={t[1] / 1e3:9.3f} ms share={t[1] / total * 100:5.1f}% "
          f"avg={t[1] / t[0]:9.1f} us max={max(t[2]):9.1f} us")
