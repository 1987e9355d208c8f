#!/usr/bin/env python3
"""Per-16-diagonal busy/total cycles of the two cluster ranks (library built with -DWAVE_PHASE_CLK).

    RNAFOLD_LIB=build_var/clk.so python tools/cluster_clocks.py 400 1024
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
    _lib.fold_hamming_raw(seq, off, count, db=db, mfe=mfe)
    L.rnafold_debug_phase_clocks(out, 640, 1)
v = np.array(list(out), dtype=np.float64) / count
print("info", _lib.last_launch_info())
print(f"cycles per fold: total {v[128:192].sum():.0f}  pair-CTA busy {v[0:64].sum():.0f}  multiloop-CTA busy {v[64:128].sum():.0f}")
for b in range((n + 17) // 16):
    print(f"d {16*b:3d}..{16*b+15:3d}: total {v[128+b]:8.0f}  pair busy {v[b]:8.0f}  ml busy {v[64+b]:8.0f}")
