#!/usr/bin/env python
"""Aggregates an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel:
    python tools/launch_summary.py launches.csv > summary.csv"""
import csv
import sys
from collections import defaultdict

rows = [r for r in csv.reader(l for l in open(sys.argv[1]) if l.startswith('"'))]
head = rows[0]
k_name, k_val = head.index("Kernel Name"), head.index("Metric Value")
tot, cnt = defaultdict(float), defaultdict(int)
for r in rows[1:]:
    tot[r[k_name]] += float(r[k_val].replace(",", "")) / 1e6
    cnt[r[k_name]] += 1
total = sum(tot.values())
print("kernel,launches,total_ms,share")
for name in sorted(tot, key=tot.get, reverse=True):
    print(f"\"{name[:110]}\",{cnt[name]},{tot[name]:.3f},{tot[name] / total:.4f}")
