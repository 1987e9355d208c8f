#!/usr/bin/env python3
"""Aggregate tools/ncu_lines.py output by kernel region (marker comments in fold_wave.cu)."""
import re, sys
lines_file, src = sys.argv[1], sys.argv[2]
text = open(src).read().splitlines()
marks = []
for k, ln in enumerate(text, 1):
    m = re.search(r"// -+ ([ITMFXGE])\(?.*?:|// ---- (backtrack|emit)|for \(int d = kTurn \+ 1; d <= n \+ 1", ln)
    if m:
        marks.append((k, m.group(1) or m.group(2) or "phase ovh"))
    if re.match(r"(template|__device__ __forceinline__|__global__)", ln):
        marks.append((k, "helpers:" + (re.search(r"(\w+)\(", text[k] if 'template' in ln else ln) or [None, "?"])[1]))
marks.sort()
def region(f, l):
    if f != "fold_wave.cu":
        return f
    r = "top"
    for k, name in marks:
        if k <= l: r = name
    return r
agg = {}
for ln in open(lines_file).read().splitlines()[3:]:
    p = [x.strip() for x in ln.split("|")]
    f, l = p[0].rsplit(":", 1)
    st = p[3].split()
    a = agg.setdefault(region(f, int(l)), [0, 0, 0, 0])
    a[0] += float(p[1]); a[1] += float(p[2]); a[2] += float(st[0]); a[3] += float(st[1])
print(f"{'region':34s} inst%  samples%  barrier  no_inst")
for r, v in sorted(agg.items(), key=lambda kv: -kv[1][0]):
    print(f"{r:34s} {v[0]:5.1f}  {v[1]:5.1f}  {v[2]:8.0f} {v[3]:8.0f}")
