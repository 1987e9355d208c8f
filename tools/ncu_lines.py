#!/usr/bin/env python3
"""Attribute ncu per-instruction metrics (source page, SASS view) to CUDA source lines.

    python tools/ncu_lines.py <report.ncu-rep> <lib.so> <kernel-substring> [out.txt]

ncu's CSV source page only carries metrics in the SASS view, so the SASS is re-disassembled with
nvdisasm -g (line info from -lineinfo) and matched to ncu's rows by instruction order.
"""
import csv
import io
import os
import re
import subprocess
import sys
import tempfile
from collections import defaultdict


def line_map(so, kernel):
    tmp = tempfile.mkdtemp()
    subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(so)], cwd=tmp, capture_output=True)
    out = []
    for f in sorted(os.listdir(tmp)):
        if not f.endswith(".cubin") or "sm_100" not in f:
            continue
        txt = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, f)], capture_output=True, text=True).stdout
        cur, infn, loc = None, False, ("?", 0)
        for ln in txt.splitlines():
            m = re.match(r"\s*\.text\.(\S+):", ln)
            if m:
                infn = kernel in m.group(1)
                continue
            if not infn:
                continue
            m = re.match(r'\s*//## File "([^"]+)", line (\d+)', ln)
            if m:
                loc = (os.path.basename(m.group(1)), int(m.group(2)))
                continue
            m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
            if m:
                out.append((int(m.group(1), 16), loc, m.group(2).strip()))
        if out:
            break
    return out


def main():
    rep, so, kernel = sys.argv[1:4]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr = rows[1]
    col = {h: k for k, h in enumerate(hdr)}
    body = rows[2:]
    lm = line_map(so, kernel)
    n = min(len(lm), len(body))
    agg = defaultdict(lambda: defaultdict(float))
    keys = ["Instructions Executed", "# Samples", "stall_barrier", "stall_no_inst", "stall_wait", "stall_short_sb",
            "stall_long_sb", "stall_branch_resolving", "stall_math", "stall_mio", "stall_not_selected",
            "L1 Wavefronts Shared", "L1 Wavefronts Shared Excessive"]
    tot = defaultdict(float)
    for k in range(n):
        loc = lm[k][1]
        for key in keys:
            try:
                v = float(body[k][col[key]])
            except (ValueError, KeyError):
                v = 0.0
            agg[loc][key] += v
            tot[key] += v
    lines = [f"# {rep}: {kernel}: {len(body)} ncu SASS rows, {len(lm)} disassembled", "# totals: " +
             ", ".join(f"{k}={tot[k]:.3g}" for k in keys)]
    lines.append("file:line | inst% | samples% | barrier no_inst wait short_sb long_sb branch math mio | smem_wavefronts excessive")
    for loc, a in sorted(agg.items(), key=lambda kv: -kv[1]["# Samples"])[:400]:
        lines.append(f"{loc[0]}:{loc[1]} | {100*a['Instructions Executed']/max(tot['Instructions Executed'],1):5.1f} | "
                     f"{100*a['# Samples']/max(tot['# Samples'],1):5.1f} | "
                     + " ".join(f"{a[k]:.0f}" for k in keys[2:10]) + f" | {a['L1 Wavefronts Shared']:.3g} {a['L1 Wavefronts Shared Excessive']:.3g}")
    text = "\n".join(lines) + "\n"
    if len(sys.argv) > 4:
        open(sys.argv[4], "w").write(text)
    print(text)


if __name__ == "__main__":
    main()
