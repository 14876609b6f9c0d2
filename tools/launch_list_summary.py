#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel count, total and share of the time.
(ncu times are cold-cache and serialised: compare SHARES, not absolutes.)   usage: launch_list_summary.py file.csv [skip_first_n]"""
import csv, sys, collections
rows = list(csv.reader(l for l in open(sys.argv[1]) if l.startswith('"')))
hdr = rows[0]; iname = hdr.index("Kernel Name"); ival = hdr.index("Metric Value"); iunit = hdr.index("Metric Unit")
skip = int(sys.argv[2]) if len(sys.argv) > 2 else 0
agg = collections.OrderedDict()
for r in rows[1 + skip:]:
    t = float(r[ival].replace(",", "")); u = r[iunit]
    t_us = t / 1e3 if u in ("ns", "nsecond") else (t if u in ("us", "usecond") else t * 1e3)
    name = r[iname].split("(")[0][-70:]
    c, s = agg.get(name, (0, 0.0)); agg[name] = (c + 1, s + t_us)
tot = sum(s for _, s in agg.values())
print(f"{len(rows) - 1 - skip} launches, {tot / 1e3:.3f} ms total")
for name, (c, s) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{s / tot * 100:6.2f}%  {s:10.1f} us  x{c:<4d} {name}")
