#!/usr/bin/env python3
"""Per-warp time line of ONE fold in the 512-thread (latency) instance of the wave kernel (library built with -DWAVE_TRACE).

    RNAFOLD_LIB=build_var/wtrace.so [TRACE_COUNT=8] python tools/wave_trace.py 100 [d ...]

TRACE_COUNT sequences are folded (8: the latency instance, one fold per SM; >= 740: the 128-thread throughput instance,
five folds per SM -- the trace is CTA 0's first fold either way).

Stamps per phase and warp (SM cycles): 0 phase start, 1 generic-shape cells done, 2 block items done, 3 after the barrier;
4 = kind of the warp's last block item (1 E, 2 T, 3 M, 4 F, 5 X; 0 none).
"""
import ctypes
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

from rnagent_b200 import _lib, synth  # noqa: E402

n = int(sys.argv[1])
show = [int(v) for v in sys.argv[2:]]
count = int(os.environ.get("TRACE_COUNT", 8))
seq, off = synth.random_sequences(count, n, seed=5)
db = np.zeros(len(seq), np.uint8)
mfe = np.zeros(count, np.int32)
L = _lib.lib()
out = (ctypes.c_uint * (512 * 16 * 8))()
for _ in range(2):
    _lib.fold_hamming_raw(seq, off, count, db=db, mfe=mfe)
L.rnafold_debug_wave_trace(out, len(out))
t = np.array(list(out), dtype=np.float64).reshape(512, 16, 8)
nw = 16 if count <= 148 else 4
t = t[:, :nw, :]
print("info", _lib.last_launch_info())
kinds = ["-", "E", "T", "M", "F", "X"]
tot = t[n + 1, :, 3].max() - t[5, :, 0].min()
print(f"cycles for phases 5..{n + 1}: {tot:.0f}  ({tot / (n - 3):.0f} per phase)")
for b in range(0, n + 2, 16):
    ds = list(range(max(b, 5), min(b + 16, n + 1)))
    if not ds:
        continue
    ph = np.mean([t[d + 1, :, 0].min() - t[d, :, 0].min() for d in ds])
    g = np.mean([(t[d, :, 1] - t[d, :, 0].min()).max() for d in ds])
    it = np.mean([(t[d, :, 2] - t[d, :, 0].min()).max() for d in ds])
    fast = np.mean([(t[d, :, 2] - t[d, :, 0].min()).min() for d in ds])
    last = [kinds[int(t[d, int(np.argmax(t[d, :, 2])), 4])] for d in ds]
    print(f"d {ds[0]:3d}..{ds[-1]:3d}: phase {ph:6.0f}  cells done (slowest warp) {g:6.0f}  items done slowest {it:6.0f} fastest {fast:6.0f}"
          f"  last item of the slowest warp: {''.join(last)}")
for d in show:
    base = t[d, :, 0].min()
    print(f"  d={d}: cells {np.round(t[d, :, 1] - base).astype(int).tolist()}")
    print(f"        items {np.round(t[d, :, 2] - base).astype(int).tolist()}")
    print(f"        kinds {[kinds[int(v)] for v in t[d, :, 4]]}")
    print(f"        sync  {np.round(t[d, :, 3] - base).astype(int).tolist()}")
