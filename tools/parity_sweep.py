#!/usr/bin/env python3
"""Randomised GPU-vs-oracle sweep (GPU box): structures and energies of random sequences over a length range, for every
parameter set, in batches that mix all kernel families.  Prints one line per parameter set; exit code 1 on any difference.

    python tools/parity_sweep.py [count_per_set=40000] [nlo=5] [nhi=451] [seed=1]
"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

import oracle  # noqa: E402
import rnagent_b200  # noqa: E402
from rnagent_b200 import _lib, vienna  # noqa: E402
from rnagent_b200.vienna import pack  # noqa: E402

count = int(sys.argv[1]) if len(sys.argv) > 1 else 40000
nlo = int(sys.argv[2]) if len(sys.argv) > 2 else 5
nhi = int(sys.argv[3]) if len(sys.argv) > 3 else 451
seed = int(sys.argv[4]) if len(sys.argv) > 4 else 1
rng = np.random.default_rng(seed)
bad_total = 0
for setno, name in ((1, "turner2004"), (2, "turner1999"), (3, "andronescu2007"), (4, "langdon2018")):
    vienna.set_vienna_params(setno)
    path = os.path.join(vienna.PARAMETERS, vienna.PARAM_FILES[setno])
    P = oracle.Params(path)
    # short lengths dominate the count (they are cheap for the CPU side); every family still gets thousands of sequences
    lens = np.concatenate([rng.integers(nlo, min(nhi, 128) + 1, int(count * 0.7)), rng.integers(nlo, min(nhi, 296) + 1, int(count * 0.2)),
                           rng.integers(min(nhi, 297), nhi + 1, count - int(count * 0.7) - int(count * 0.2))])
    rng.shuffle(lens)
    alph = list("ACGU")
    t0 = time.time()
    bad = 0
    for b0 in range(0, count, 8192):
        seqs = ["".join(rng.choice(alph if (b0 // 8192) % 3 else list("GGCCAU"), int(n))) for n in lens[b0:b0 + 8192]]
        buf, off = pack(seqs)
        B = len(seqs)
        db = np.zeros_like(buf)
        mfe = np.zeros(B, dtype=np.int32)
        _lib.fold_hamming_raw(buf, off, B, db=db, mfe=mfe)
        wdb, wmfe, _ = P.fold_batch(buf, off, nthreads=0)
        diff = np.flatnonzero(mfe != wmfe)
        sdiff = [k for k in range(B) if db[off[k]:off[k + 1]].tobytes() != wdb[off[k]:off[k + 1]].tobytes()]
        bad += len(set(diff.tolist()) | set(sdiff))
        for k in (sdiff + diff.tolist())[:2]:
            print("DIFF", name, len(seqs[k]), seqs[k], int(mfe[k]), int(wmfe[k]))
    print(f"{name}: {count} sequences of {nlo}..{nhi} nt, {bad} differ, {time.time() - t0:.0f} s, last launch {_lib.last_launch_info()}", flush=True)
    bad_total += bad
vienna.set_vienna_params(1)
sys.exit(1 if bad_total else 0)
