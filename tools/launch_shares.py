#!/usr/bin/env python
"""Per-kernel totals of an ncu launch list (`ncu --metrics gpu__time_duration.sum --csv --log-file X`): tools/launch_shares.py X"""
import collections
import csv
import re
import sys

rows = [r for r in csv.reader(l for l in open(sys.argv[1]) if l.startswith('"'))]
hdr = rows[0]
k, v = hdr.index("Kernel Name"), hdr.index("Metric Value")
tot = collections.defaultdict(float); cnt = collections.Counter()
for r in rows[1:]:
    n = re.sub(r"\(.*$", "", re.sub(r"^void ", "", re.sub(r"<unnamed>::", "", r[k])))[:60]
    tot[n] += float(r[v].replace(",", "")) * 1e-6; cnt[n] += 1
total = sum(tot.values())
print(f"{sum(cnt.values())} launches, {total:.1f} ms of kernel time")
for n in sorted(tot, key=tot.get, reverse=True)[:14]:
    print(f"  {n:60s} {cnt[n]:6d} launches {tot[n]:9.2f} ms {100 * tot[n] / total:5.1f}%  avg {1e3 * tot[n] / cnt[n]:8.1f} us")
