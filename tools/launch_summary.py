#!/usr/bin/env python3
"""Aggregate an ncu launch list (--metrics gpu__time_duration.sum --csv) per kernel name."""
import csv, collections, sys
rows = list(csv.reader(open(sys.argv[1])))
i = [k for k, r in enumerate(rows) if r and r[0] == 'ID'][0]
h = rows[i]; agg = collections.OrderedDict()
for r in rows[i + 1:]:
    if len(r) < len(h): continue
    n = r[h.index('Kernel Name')].split('(')[0][:44]; v = float(r[h.index('Metric Value')].replace(',', '')); u = r[h.index('Metric Unit')]
    v = v / 1e6 if u in ('ns', 'nsecond') else (v / 1e3 if u in ('us', 'usecond') else v)
    a = agg.setdefault(n, [0, 0.0]); a[0] += 1; a[1] += v
tot = sum(a[1] for a in agg.values())
for n, (c, t) in agg.items(): print('%-46s %5d launches %10.3f ms %5.1f%%' % (n, c, t, 100 * t / tot))
print('total %.3f ms' % tot)
