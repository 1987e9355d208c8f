#!/usr/bin/env python3
"""Per-warp time line of one fold in the 2-CTA cluster kernel (library built with -DCLUSTER_TRACE).

    RNAFOLD_LIB=build_var/trace.so python tools/cluster_trace.py 400 [d ...]

Stamps per phase and warp (SM cycles since the fold's first phase): 0 phase start, 1 own items done, 2 after the trailing wait
for the peer's row, 3 after the CTA barrier.  Prints, per block of 16 phases, the mean phase length, the mean of the slowest
warp's item time, the mean spread between the fastest and the slowest warp, and the per-warp line of the phases asked for.
"""
import ctypes
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

from rnagent_b200 import _lib, synth  # noqa: E402

n = int(sys.argv[1])
show = [int(v) for v in sys.argv[2:]]
count = 74
seq, off = synth.random_sequences(count, n, seed=5)
db = np.zeros(len(seq), np.uint8)
mfe = np.zeros(count, np.int32)
L = _lib.lib()
out = (ctypes.c_uint * (2 * 480 * 16 * 4))()
for _ in range(2):
    _lib.fold_hamming_raw(seq, off, count, db=db, mfe=mfe)
L.rnafold_debug_cluster_trace(out, len(out))
t = np.array(list(out), dtype=np.float64).reshape(2, 480, 16, 4)
print("info", _lib.last_launch_info())
# every warp read its own time base: take the skew out with the stamps after the CTA barrier, which all warps leave together
for rank in range(2):
    skew = np.median(t[rank, 6:n, :, 3] - t[rank, 6:n, :1, 3], axis=0)
    t[rank] -= skew[None, :, None]
for rank, name in ((0, "pair"), (1, "multiloop")):
    print(f"--- {name} CTA: total {t[rank, n + 1, :, 3].max():.0f} cycles")
    for b in range(0, n + 2, 16):
        ds = [d for d in range(max(b, 5), min(b + 16, n + 1))]
        if not ds:
            continue
        start = np.array([t[rank, d, :, 0].min() for d in ds])
        nxt = np.array([t[rank, d + 1, :, 0].min() for d in ds])
        items = np.array([(t[rank, d, :, 1] - t[rank, d, :, 0].min()).max() for d in ds])
        fast = np.array([(t[rank, d, :, 1] - t[rank, d, :, 0].min()).min() for d in ds])
        wait = np.array([(t[rank, d, :, 2] - t[rank, d, :, 1]).max() for d in ds])
        sync = np.array([t[rank, d, :, 3].max() - t[rank, d, :, 2].max() for d in ds])
        print(f"d {ds[0]:3d}..{ds[-1]:3d}: phase {np.mean(nxt - start):7.0f}  slowest warp's items {items.mean():7.0f}  fastest {fast.mean():7.0f}"
              f"  longest trailing wait {wait.mean():6.0f}  barrier after the last warp {sync.mean():6.0f}")
    for d in show:
        base = t[rank, d, :, 0].min()
        print(f"  d={d}: start+ {np.round(t[rank, d, :, 0] - base).astype(int).tolist()}")
        print(f"         items {np.round(t[rank, d, :, 1] - base).astype(int).tolist()}")
        print(f"         wait  {np.round(t[rank, d, :, 2] - base).astype(int).tolist()}")
        print(f"         sync  {np.round(t[rank, d, :, 3] - base).astype(int).tolist()}")

kinds = (ctypes.c_uint * (480 * 16))()
L.rnafold_debug_cluster_kinds(kinds, len(kinds))
kv = np.array(list(kinds), dtype=np.float64).reshape(480, 16)
names = ["E", "T", "P", "F", "G loop", "prologue"]
print("--- pair CTA, per item kind (both trace passes summed): items per phase and cycles per item, mean over 16 phases")
for b in range(0, n + 2, 32):
    ds = list(range(max(b, 5), min(b + 32, n + 1)))
    line = f"d {ds[0]:3d}..{ds[-1]:3d}:"
    for k, nm in enumerate(names):
        c = kv[ds, 8 + k].sum()
        if c:
            line += f"  {nm} {c / len(ds) / 2:5.1f} x {kv[ds, k].sum() / c:6.0f}"
    print(line)
