"""Aggregate an ncu `--metrics gpu__time_duration.sum --csv` launch list by kernel name: python tools/ncu_agg.py launches.csv [skip_first_n]"""
import collections
import csv
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10 and r[0].isdigit()]
skip = int(sys.argv[2]) if len(sys.argv) > 2 else 0
d = collections.defaultdict(list)
for r in rows[skip:]:
    d[r[4].split("(")[0][-70:]].append(float(r[-1]))
for k, v in sorted(d.items(), key=lambda kv: -sum(kv[1])):
    print(f"{k:72s} n={len(v):4d} avg {sum(v) / len(v) / 1000:9.2f} us  total {sum(v) / 1e6:8.3f} ms")
