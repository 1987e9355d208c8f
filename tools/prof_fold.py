#!/usr/bin/env python3
"""Tiny driver for ncu captures: fold `count` random sequences of length n through the C ABI `reps` times.

    python tools/prof_fold.py 100 23680 3
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

import rnagent_b200 as rb  # noqa: E402
from rnagent_b200 import _lib, synth  # noqa: E402

n, count, reps = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]) if len(sys.argv) > 3 else 2
seq, off = synth.random_sequences(count, n, seed=5)
tgt = synth.random_targets(off, seed=6)
db = np.zeros(len(seq), np.uint8)
mfe = np.zeros(count, np.int32)
hd = np.zeros(count, np.int32)
for _ in range(reps):
    _lib.fold_hamming_raw(seq, off, count, target=tgt, db=db, mfe=mfe, hd=hd)
print("ok", _lib.last_launch_info())
