#!/usr/bin/env python3
"""Local-search rate: 4096 near misses of n=100 under the reference's Testing.yml settings (host strings in and out)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import rnagent_b200 as rb
from rnagent_b200 import synth

rb.set_vienna_params(1)
rng = np.random.default_rng(515)
pseq, poff = synth.random_sequences(4096, 100, 5100)
base = [pseq[poff[k]:poff[k + 1]].tobytes().decode() for k in range(4096)]
targets, _ = rb.fold_batch(base)
near = []
for s in base:
    s = list(s)
    for _ in range(int(rng.integers(1, 4))):
        s[int(rng.integers(0, 100))] = "ACGU"[int(rng.integers(0, 4))]
    near.append("".join(s))
rb.permute_batch(near[:256], targets[:256], rb.permute.TESTING, seed=1)
for rep in range(4):
    t0 = time.perf_counter()
    res = rb.permute_batch(near, targets, rb.permute.TESTING, seed=rep)
    dt = time.perf_counter() - t0
    print(f"rep {rep}: {dt * 1e3:.1f} ms, {res.folds} folds, {4096 / dt / 1e3:.1f} k solutions/s, searched {(res.rounds > 0).sum()}, solved {((res.hd == 0) & (res.rounds > 0)).sum()}")
