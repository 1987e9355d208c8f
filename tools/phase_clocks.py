#!/usr/bin/env python3
"""Per-phase cycle profile of the wave kernel (needs a library built with -DWAVE_PHASE_CLK).

    RNAFOLD_LIB=build_var/clk.so python tools/phase_clocks.py 100 23680
"""
import ctypes
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

from rnagent_b200 import _lib, synth  # noqa: E402

n, count = int(sys.argv[1]), int(sys.argv[2])
seq, off = synth.random_sequences(count, n, seed=5)
db = np.zeros(len(seq), np.uint8)
mfe = np.zeros(count, np.int32)
L = _lib.lib()
out = (ctypes.c_ulonglong * 640)()
for _ in range(2):
    try:
        _lib.fold_hamming_raw(seq, off, count, db=db, mfe=mfe)
    except Exception as ex:  # WAVE_SKIP builds produce inconsistent tables: timing only
        print("(fold reported:", str(ex)[:60], ")")
    L.rnafold_debug_phase_clocks(out, 640, 1)
v = np.array(list(out), dtype=np.float64) / count
tot = v.sum()
print(f"cycles per fold (CTA wall): {tot:.0f}; prologue {v[600]:.0f} backtrack {v[601]:.0f} emit {v[602]:.0f}")
for d0 in range(4, n + 2, 8):
    seg = v[d0:d0 + 8]
    print(f"d {d0:3d}..{d0 + 7:3d}: " + " ".join(f"{x:6.0f}" for x in seg) + f"   sum {seg.sum():7.0f} ({100 * seg.sum() / tot:4.1f}%)")
raw = np.array(list(out), dtype=np.float64)
names = ["E", "T", "M", "F", "X", "G-loop"]
for late in (0, 1):
    print("items, d %s 68: " % (">=" if late else "<") + "  ".join(
        f"{nm} {raw[610 + 8 * late + k] / max(raw[626 + 8 * late + k], 1):6.0f}cyc x{raw[626 + 8 * late + k] / count:5.1f}" for k, nm in enumerate(names)))
