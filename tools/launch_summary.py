#!/usr/bin/env python
"""Per-kernel totals of an `ncu --metrics gpu__time_duration.sum --csv` launch list: python tools/launch_summary.py file.csv"""
import collections, csv, re, sys
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 5]
hdr = rows[0]
ki, vi = hdr.index('Kernel Name'), hdr.index('Metric Value')
tot, cnt = collections.defaultdict(float), collections.Counter()
for r in rows[1:]:
    n = re.sub(r'\(.*', '', r[ki]); n = re.sub(r'^void ', '', n)
    tot[n] += float(r[vi].replace(',', '')); cnt[n] += 1
all_ = sum(tot.values())
print(f"{'kernel':60s} {'launches':>8s} {'total ms':>10s} {'share':>7s}")
for n in sorted(tot, key=tot.get, reverse=True):
    print(f"{n[:60]:60s} {cnt[n]:8d} {tot[n] / 1e6:10.3f} {tot[n] / all_:7.3f}")
