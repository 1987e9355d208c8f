#!/usr/bin/env python3
"""Per-kernel totals of an ncu launch list (--metrics gpu__time_duration.sum --csv): python tools/launch_share.py <csv> [skip_first_n]"""
import collections
import csv
import re
import sys

rows = [r for r in csv.reader(l for l in open(sys.argv[1]) if l.startswith('"'))]
hdr = rows[0]
ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
tot = collections.OrderedDict()
for r in rows[1:]:
    v = float(r[vi].replace(",", ""))
    v *= {"ns": 1e-6, "us": 1e-3, "usecond": 1e-3, "nsecond": 1e-6, "ms": 1.0, "msecond": 1.0, "s": 1e3, "second": 1e3}[r[ui]]
    name = re.sub(r"\(.*", "", r[ki]).replace("void ", "").replace("rnafold::", "").replace("unnamed>::", "")
    t = tot.setdefault(name, [0, 0.0])
    t[0] += 1
    t[1] += v
s = sum(t[1] for t in tot.values())
for name, (n, ms) in sorted(tot.items(), key=lambda kv: -kv[1][1]):
    print(f"{ms:10.3f} ms {100 * ms / s:6.2f} %  x{n:<4d} {name}")
