#!/usr/bin/env python3
"""Per-kernel duration statistics of an ncu launch list (--metrics gpu__time_duration.sum --csv).
    python tools/launch_stats.py gpurun_out/launches.csv [skip_fraction_of_tail]"""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
h = next(i for i, r in enumerate(rows) if 'Kernel Name' in r)
hdr = rows[h]; kn = hdr.index('Kernel Name'); mv = hdr.index('Metric Value')
seq = [(r[kn].split('(')[0].replace('void rtb::', ''), float(r[mv].replace(',', ''))) for r in rows[h + 1:] if len(r) > mv]
by = collections.OrderedDict()
for k, v in seq:
    by.setdefault(k, []).append(v / 1000.0)
tot = sum(sum(v) for v in by.values())
print(f"{len(seq)} launches, total {tot / 1000:.3f} ms")
for k, v in sorted(by.items(), key=lambda kv: -sum(kv[1])):
    s = sorted(v)
    print(f"{k[:60]:60s} n={len(v):5d} sum={sum(v) / 1000:8.3f} ms  share={sum(v) / tot:6.1%}  med={s[len(s) // 2]:7.1f} us  p90={s[int(len(s) * .9)]:7.1f}  max={s[-1]:7.1f}")
